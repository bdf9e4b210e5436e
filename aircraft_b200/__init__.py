"""aircraft_b200 -- B200-native (sm_100a) drop-in for the batched rollout / linearisation hot path
of wgrosche/AIrcraft: fixed-wing 6-DOF quaternion dynamics with the NN-surrogate aero model inside
RK4, plus dPhi/dx and dPhi/du.  CUDA only; there is no CPU fallback."""
from .config import (AircraftConfiguration, AircraftOpts, SixDOFOpts, DEFAULT_MODEL_PATH,
                     DEFAULT_PROBLEM_PATH, NETWORKPATH, DATAPATH, TRIM_COM)

from .dm import DM

__version__ = "0.2.0"


def __getattr__(name):  # lazy: the CUDA library is loaded on first use of the compute path
    if name in ("Aircraft", "COEFF_MODEL_REGISTRY", "NeuralModel", "CoefficientModel"):
        from . import aircraft as _a
        return getattr(_a, name)
    if name in ("ShardedRollout", "shard_range", "sweep_sampled"):
        from . import sharding as _s
        return getattr(_s, name)
    raise AttributeError(name)
