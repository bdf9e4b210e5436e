"""Generates the committed golden fixtures.  Runs ONLY in the build container, where
/root/reference is mounted; the fixtures (not this script's inputs) travel to the GPU box.

  nn_golden.npz       inputs -> outputs of the REFERENCE's own ScaledModel
                      (imported from /root/reference/src/aircraft/surrogates/models.py:101-155,
                      loaded as utils.py:30-40 does), in float32 (as shipped) and with the
                      module cast to float64; plus the first 512 rows of
                      data/processed/data_sim.csv with their wind-tunnel targets.
  todo_known_answers.json   the three (state, control) -> forces_frd triples printed by the
                      reference's scripts and pasted into /root/reference/todo.md:38,111-114,177-180
                      (default coefficient model, thrust still added, S = 0.238).

Usage:  python tests/golden/make_golden.py
"""
import json
import sys
from pathlib import Path

import numpy as np
import pandas as pd
import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent
sys.path.insert(0, str(REF / "src"))
from aircraft.surrogates.models import ScaledModel  # noqa: E402  (the reference's module)


def load_reference_model():
    ck = torch.load(REF / "data/networks/model-dynamics.pth", map_location="cpu", weights_only=True)
    scaler = (ck["input_mean"], ck["input_std"], ck["output_mean"], ck["output_std"])
    m = ScaledModel(5, 6, scaler=scaler)
    m.load_state_dict(ck["model_state_dict"])
    m.eval()
    return m


def main():
    m = load_reference_model()
    df = pd.read_csv(REF / "data/processed/data_sim.csv")
    cols_in = ["q", "alpha", "beta", "aileron", "elevator"]
    cols_out = ["CX", "CY", "CZ", "Cl", "Cm", "Cn"]
    csv_in = df[cols_in].to_numpy(np.float64)
    csv_out = df[cols_out].to_numpy(np.float64)
    g = np.random.default_rng(7)
    # random inputs, inside and beyond the training envelope (rollouts extrapolate)
    lo = np.array([100.0, -0.6, -0.6, -10.0, -10.0])
    hi = np.array([6000.0, 0.6, 0.6, 10.0, 10.0])
    rnd = lo + (hi - lo) * g.random((512, 5))
    inputs = np.concatenate([csv_in[:512], rnd], axis=0)
    with torch.no_grad():
        out32 = m(torch.tensor(inputs, dtype=torch.float32)).numpy()
        out64 = m.double()(torch.tensor(inputs, dtype=torch.float64)).numpy()
        full64 = m(torch.tensor(csv_in, dtype=torch.float64)).numpy()
    ss_res = ((csv_out - full64) ** 2).sum(0)
    ss_tot = ((csv_out - csv_out.mean(0)) ** 2).sum(0)
    r2 = 1 - ss_res / ss_tot
    np.savez_compressed(OUT / "nn_golden.npz", inputs=inputs, out_f32=out32, out_f64=out64,
                        csv_targets=csv_out[:512], r2_full_csv=r2)
    print("R2 over data_sim.csv:", r2)

    known = [
        dict(src="todo.md:38",
             x=[-4.91443960e-04, 3.00007355e+01, -1.79999243e+02, -3.40474317e+01, -3.08113900e+01,
                -1.36658629e+01, 1.07592461e-01, -6.54142188e-01, 6.65115558e-01, -3.43760932e-01,
                -3.30126758e-02, 6.28510264e-02, 8.32389179e-02],
             u6=[-0.48060452] * 6, forces_frd=[-7.48988, 4.39679, -94.534]),
        dict(src="todo.md:111-114",
             x=[2.71825097e-10, 3.00000000e+01, -1.80000000e+02, -3.73314043e+01, 6.39928972e+01,
                2.77826364e+01, -2.65774055e-01, -6.11580371e-01, 4.45413540e-01, 5.97477466e-01,
                -5.58006954e-01, 7.18165869e-01, 1.04268472e+00],
             u6=[-4.99999779, -4.99999779, -4.9999978, -5., -4.99999999, -5.00000001],
             forces_frd=[-51.5765, -156.479, -1472.69]),
        dict(src="todo.md:177-180",
             x=[3.75432597e-32, 3.00000000e+01, -1.80000000e+02, -3.76429340e+01, 6.47602149e+01,
                2.58801853e+01, -2.83827959e-01, -6.36278543e-01, 4.24668182e-01, 5.78203029e-01,
                -5.56987777e-01, 7.14029880e-01, 1.04398353e+00],
             u6=[-4.9999613, -4.99996123, -4.99996145, -4.99999856, -4.99999694, -4.99999946],
             forces_frd=[-49.6078, -157.228, -1421.45]),
    ]
    json.dump(known, open(OUT / "todo_known_answers.json", "w"), indent=1)


if __name__ == "__main__":
    main()
