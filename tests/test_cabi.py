"""The C-ABI library loads and exports every symbol include/aircraft_b200.h declares; the ctypes
mirror matches the header; the host side validates arguments like the reference's Functions do.
No compute calls (no GPU here)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "aircraft_b200.h"


def _declared():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(airc_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported(built):
    from aircraft_b200 import _cabi
    lib = _cabi.load_library()
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert set(names) == set(_cabi.PROTOTYPES), set(names) ^ set(_cabi.PROTOTYPES)
    assert lib.airc_version() >= 100
    assert b"sm_100a" in lib.airc_build_info()


def test_struct_mirror(built):
    from aircraft_b200 import _cabi
    assert C.sizeof(_cabi.AircNNBlob) == 860 * 4
    assert C.sizeof(_cabi.AircParams) == 8 * (4 + 3 + 9 + 9 + 3 + 1) + 8
    assert C.sizeof(_cabi.AircCostParams) == 10 * 4
    text = HEADER.read_text()
    for name, val in (("AIRC_NX", 13), ("AIRC_NU", 7), ("AIRC_LAYOUT_SOA", 0), ("AIRC_LAYOUT_AOS", 1)):
        assert re.search(rf"#define {name} {val}\b", text)


def test_sm100_only_no_fallback(built):
    """Built for sm_100a alone: no other SASS or PTX target in the library."""
    import subprocess
    from aircraft_b200 import _cabi
    r = subprocess.run(["cuobjdump", "-lelf", "-lptx", str(_cabi.LIB_PATH)], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", r.stdout))
    assert archs == {"sm_100a"}, archs


def test_create_rejects_bad_arguments(built):
    from aircraft_b200 import _cabi
    from aircraft_b200.params import load_nn_blob, make_params
    from aircraft_b200.config import AircraftConfiguration
    lib = _cabi.load_library()
    h = C.c_void_p()
    assert lib.airc_create(None, None, 0, C.byref(h)) == _cabi.EINVAL
    p = make_params(AircraftConfiguration.from_json())
    blob = load_nn_blob()
    p.mass = 0.0
    assert lib.airc_create(C.byref(p), C.byref(blob), 0, C.byref(h)) == _cabi.EINVAL
    assert b"mass" in lib.airc_last_error(None)
    assert lib.airc_destroy(None) == 0


def test_checkpoint_packing_is_verbatim(built):
    import torch
    from aircraft_b200.params import load_nn_blob
    from aircraft_b200.config import DEFAULT_MODEL_PATH
    ck = torch.load(str(DEFAULT_MODEL_PATH), map_location="cpu", weights_only=True)
    blob = load_nn_blob()
    assert np.array_equal(np.array(blob.W2[:], np.float32).reshape(32, 16),
                          ck["model_state_dict"]["core_layers.1.weight"].numpy())
    assert np.array_equal(np.array(blob.output_std[:], np.float32), ck["output_std"].numpy().reshape(-1))


def test_facade_mirrors_reference_interface(built):
    from aircraft_b200 import Aircraft, AircraftOpts, AircraftConfiguration, COEFF_MODEL_REGISTRY
    from aircraft_b200.config import TRIM_COM
    assert set(COEFF_MODEL_REGISTRY) == {"nn", "linear", "poly", "default"}
    cfg = AircraftConfiguration.from_json()
    ac = Aircraft(AircraftOpts(coeff_model_type="nn", aircraft_config=cfg, physical_integration_substeps=1))
    assert (ac.num_states, ac.num_controls) == (13, 7)
    assert ac.mass == 4.0 and ac.physical_integration_substeps == 1
    ac.com = TRIM_COM           # scripts assign after construction (main/control/control.py:168-172)
    assert ac.com == tuple(TRIM_COM)
    assert AircraftOpts().physical_integration_substeps == 10       # base.py:12
    # the registry's other models are tables behind airc_create_model; unknown keys are refused, not defaulted
    gold = ROOT / "tests" / "golden"
    pm = Aircraft(AircraftOpts(coeff_model_type="poly", coeff_model_path=gold / "poly_model.npz", aircraft_config=cfg))
    assert pm.coefficient_model.table.shape == (211,) and pm.coefficient_model.model_id == 1
    lm = Aircraft(AircraftOpts(coeff_model_type="linear", coeff_model_path=gold / "linear_model.npz", aircraft_config=cfg))
    assert lm.coefficient_model.table.shape == (36,)
    assert Aircraft(AircraftOpts(coeff_model_type="default", aircraft_config=cfg)).coefficient_model.table is None
    with pytest.raises(RuntimeError):
        Aircraft(AircraftOpts(coeff_model_type="poly", coeff_model_path=gold / "nn_golden.npz", aircraft_config=cfg))
    with pytest.raises(NotImplementedError):
        Aircraft(AircraftOpts(coeff_model_type="neural", aircraft_config=cfg))
    # shape errors are RuntimeError, as CasADi raises
    with pytest.raises(RuntimeError):
        ac.state_update(np.zeros(12), np.zeros(7), 0.01)
    with pytest.raises(RuntimeError):
        ac.state_update(np.zeros((13, 4)), np.zeros((7, 3)), 0.01)
    with pytest.raises(RuntimeError):
        ac.rollout(np.zeros((4, 13)), np.zeros((5, 10, 7)), 0.01)
    with pytest.raises(RuntimeError):
        ac.linearize(np.zeros((4, 13)), np.zeros((4, 6)), 0.01)


def test_product_never_imports_oracle():
    """The product path must not route through the oracle or any CPU fallback."""
    pkg = ROOT / "aircraft_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.hpp")):
        src = f.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
        assert "oracle/" not in src.replace("oracle/host_emul.cpp", "") or f.suffix in (".cuh",), f


def test_rollout_result_finite_mask():
    from aircraft_b200.aircraft import RolloutResult
    xf = np.zeros((5, 13)); xf[2, 4] = np.nan
    cost = np.zeros(5); cost[4] = np.inf
    r = RolloutResult(final=xf, cost=cost)
    assert r.finite.tolist() == [True, True, False, True, False]
    X = np.zeros((7, 13, 5)); X[3, 0, 1] = np.nan                     # soa trajectory: batch last
    assert RolloutResult(trajectory=X, cost=np.zeros(5)).finite.tolist() == [True, False, True, True, True]
    import torch
    t = RolloutResult(final=torch.tensor(xf))
    assert t.finite.tolist() == [True, True, False, True, True]
