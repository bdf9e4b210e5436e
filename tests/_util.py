"""Shared helpers of the parity tests."""
from __future__ import annotations

import numpy as np
import torch

from aircraft_b200.config import AircraftConfiguration, TRIM_COM
from aircraft_b200.params import load_nn_blob, make_params

# The parity bar of BASELINE.json's north_star: 1e-5 relative in FP32.  Element-wise relative
# error is meaningless for components that pass through zero, so it is judged per output column
# against the column's scale over the batch (SURVEY.md section 8c "Parity metric").
PARITY_TOL = 1e-5


def col_err(got, ref, atol=1e-3, batch_axes=(0,)) -> float:
    """max over columns of  max_batch |got-ref| / max(max_batch |ref|, atol)."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    d = np.abs(got - ref).max(axis=batch_axes)
    s = np.maximum(np.abs(ref).max(axis=batch_axes), atol)
    return float((d / s).max())


def oracle_for(com=TRIM_COM, **kw):
    from oracle.aircraft_oracle import AircraftOracle, OracleConfig
    return AircraftOracle(OracleConfig.from_problem(com=tuple(com), **kw))


def params_for(com=TRIM_COM, stall_scaling=False):
    cfg = AircraftConfiguration.from_json()
    return make_params(cfg, com=com, stall_scaling=stall_scaling), load_nn_blob()


def t64(a):
    return torch.as_tensor(np.asarray(a), dtype=torch.float64)
