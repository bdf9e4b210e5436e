"""HDF5 trajectory files in the reference's layout (aircraft_b200/trajectory_io.py).
The reader is pinned on two files the reference ships, written by libhdf5 through h5py
(tests/golden/simulation.h5: contiguous datasets, main/dynamics/dynamics.py:134-139;
tests/golden/mhtt_solution.h5: chunked + gzip datasets, control/base.py:89-104); the writer is checked
through the reader and structurally against the libhdf5-written messages."""
import numpy as np
import pytest

from aircraft_b200.trajectory_io import (_Reader, load_iteration, read_trajectories, save_iteration,
                                         save_rollout, write_trajectories)


def test_reader_on_reference_files(golden_dir):
    sim = read_trajectories(golden_dir / "simulation.h5")
    assert list(sim) == ["iteration_0"]
    it = sim["iteration_0"]
    assert it["state"].shape == (13, 40) and it["control"].shape == (7, 40) and it["times"].shape == (40,)
    assert it["state"].dtype == np.float64
    # the simulation starts from the script's trim state with elevator = 3 and dt = 0.1
    assert np.array_equal(it["state"][:10, 0], [0, 0, -200, 50, 0, 0, 0, 0, 0, 1])
    assert np.array_equal(it["control"][:, 0], [0, 3, 0, 0, 0, 0, 0])
    assert np.allclose(it["times"], 0.1 * np.arange(40))
    q = it["state"][6:10]
    assert np.abs((q * q).sum(0) - 1).max() < 1e-3            # a physical trajectory: |q| stays ~1
    assert np.isfinite(it["state"]).all() and 4.9 < it["state"][0, 1] < 5.1   # ~50 m/s * 0.1 s
    # chunked + gzip datasets, negative iteration numbers (SaveMixin keeps the last 10 iterates)
    mh = read_trajectories(golden_dir / "mhtt_solution.h5")
    assert len(mh) == 6 and all(k.startswith("iteration_-") for k in mh)
    for g in mh.values():
        assert g["state"].shape == (13, 31) and g["control"].shape == (6, 30)
        assert np.isfinite(g["state"]).all() and abs(g["state"][2, 0] + 200) < 1.0
    assert load_iteration(golden_dir / "simulation.h5", 0)["state"].shape == (13, 40)
    assert load_iteration(golden_dir / "simulation.h5", 3) is None


@pytest.mark.parametrize("ngroups", [1, 9, 300])       # 1 leaf / 2 leaves / two B-tree levels
def test_writer_round_trip(tmp_path, ngroups):
    rng = np.random.default_rng(ngroups)
    groups = {f"iteration_{k}": {"state": rng.normal(size=(13, 5 + k % 3)), "control": rng.normal(size=(7, 4)),
                                 "times": 0.1 * np.arange(5 + k % 3)} for k in range(ngroups)}
    groups["iteration_0"]["lam"] = rng.normal(size=(3,)).astype(np.float32)
    groups["iteration_0"]["idx"] = np.arange(6, dtype=np.int64).reshape(2, 3)
    p = tmp_path / "t.h5"
    write_trajectories(p, groups)
    back = read_trajectories(p)
    assert set(back) == set(groups)
    for g, ds in groups.items():
        assert set(back[g]) == set(ds)
        for n, a in ds.items():
            assert back[g][n].dtype == a.dtype and np.array_equal(back[g][n], a)


def test_writer_matches_libhdf5_structures(tmp_path, golden_dir):
    """Re-writing the reference's file yields the same datatype / fill-value / layout-size messages and
    superblock parameters as libhdf5 wrote."""
    ref_path = golden_dir / "simulation.h5"
    p = tmp_path / "copy.h5"
    write_trajectories(p, read_trajectories(ref_path))
    a, b = _Reader(p.read_bytes()), _Reader(ref_path.read_bytes())
    assert a.b[8:24] == b.b[8:24]                            # versions, offset/length sizes, K values, flags
    ga = a.open_group(a.open_group(a.root["ohdr"])["iteration_0"])
    gb = b.open_group(b.open_group(b.root["ohdr"])["iteration_0"])
    assert set(ga) == set(gb) == {"state", "control", "times"}
    for n in ga:
        ma, mb = dict(a.messages(ga[n])), dict(b.messages(gb[n]))
        assert ma[3] == mb[3] and ma[5] == mb[5] and ma[8][:2] == mb[8][:2] and ma[8][10:18] == mb[8][10:18]
        assert a.dataset(ga[n]).tobytes() == b.dataset(gb[n]).tobytes()
    assert a.u(40, 8) == len(a.b)                            # end-of-file address


def test_save_iteration_appends_and_save_rollout(tmp_path):
    p = tmp_path / "run.h5"
    X = np.random.default_rng(0).normal(size=(21, 13))
    U = np.random.default_rng(1).normal(size=(20, 7))
    save_rollout(p, X, U, 0.01)
    save_iteration(p, X.T * 2, U.T, times=np.arange(21) * 0.01, iteration=1, lam=np.ones((13, 20)))
    d = read_trajectories(p)
    assert set(d) == {"iteration_0", "iteration_1"}
    assert np.array_equal(d["iteration_0"]["state"], X.T) and np.array_equal(d["iteration_0"]["control"], U.T)
    assert np.allclose(d["iteration_0"]["times"], 0.01 * np.arange(21))
    assert np.array_equal(d["iteration_1"]["state"], 2 * X.T) and d["iteration_1"]["lam"].shape == (13, 20)
    with pytest.raises(RuntimeError):
        save_rollout(p, X.T, U, 0.01)
    with pytest.raises(ValueError):
        _Reader(b"not hdf5 at all")
