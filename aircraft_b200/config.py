"""Configuration objects with the reference's names and defaults.

Mirrors (values and spelling, not code):
  AircraftConfiguration  /root/reference/src/aircraft/utils.py:201-215
  SixDOFOpts             /root/reference/src/aircraft/dynamics/base.py:9-14
  AircraftOpts           /root/reference/src/aircraft/dynamics/aircraft.py:22-38
  NETWORKPATH/DATAPATH   /root/reference/src/aircraft/config.py:5-13
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from pathlib import Path
from typing import Sequence, Tuple

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
DATAPATH = PKG_DIR / "data"
NETWORKPATH = DATAPATH / "networks"
DEFAULT_MODEL_PATH = NETWORKPATH / "model-dynamics.pth"
DEFAULT_PROBLEM_PATH = DATAPATH / "glider" / "problem_definition.json"

# CoM offset the reference's MPC / moving-horizon scripts assign before first use
# (main/control/control.py:169-172, main/mhe/mhtt.py:51-56)
TRIM_COM = (0.0131991, -1.78875e-08, 0.00313384)


class AircraftConfiguration:
    """Same keys and defaults as the reference's AircraftConfiguration (utils.py:201-215)."""

    def __init__(self, aircraft_dict: dict | None = None):
        d = aircraft_dict or {}
        self.mass = d.get("mass", 1.0)
        self.span = d.get("span", 1.0)
        self.length = d.get("length", 1.2)
        self.chord = d.get("chord", 1.0)
        self.reference_area = d.get("reference_area", 0.238)
        self.aero_centre_offset = d.get("aero_centre_offset", [0.133, 0, 0.003])
        self.Ixx = d.get("Ixx", 0.155)
        self.Iyy = d.get("Iyy", 0.114)
        self.Izz = d.get("Izz", 0.262)
        self.Ixz = d.get("Ixz", 0.01)
        self.r_min = d.get("r_min", 10.0)
        self.glide_ratio = d.get("glide_ratio", 10.0)
        self.rudder_moment_arm = d.get("rudder_moment_arm", 0.5)

    @classmethod
    def from_json(cls, path: Path | str = DEFAULT_PROBLEM_PATH) -> "AircraftConfiguration":
        """The ``"aircraft"`` block of data/glider/problem_definition.json (:12-25), which every
        reference script loads (e.g. main/control/control.py:159-163)."""
        return cls(json.load(open(path))["aircraft"])


@dataclass
class SixDOFOpts:
    epsilon: float = 1e-6
    physical_integration_substeps: int = 10
    gravity: Sequence[float] = (0.0, 0.0, 9.81)
    mass: float = 1.0


@dataclass
class AircraftOpts(SixDOFOpts):
    coeff_model_type: str = "default"      # the reference's registry keys; this build serves "nn"
    coeff_model_path: Path | str = ""
    realtime: bool = False
    aircraft_config: AircraftConfiguration = field(default_factory=AircraftConfiguration)
    stall_angle_alpha: Tuple[float, float] = (float(np.deg2rad(-10)), float(np.deg2rad(10)))
    stall_angle_beta: Tuple[float, float] = (float(np.deg2rad(-10)), float(np.deg2rad(10)))
    stall_scaling: bool = False

    def __post_init__(self):
        self.mass = self.aircraft_config.mass


def inertia_about_com(cfg: AircraftConfiguration, com: Sequence[float]) -> tuple[np.ndarray, np.ndarray]:
    """I = I_static + m * com_term, and its inverse, in float64
    (aircraft.py:168-187 ; base.py:139-144)."""
    x, y, z = (float(c) for c in com)
    static = np.array([[cfg.Ixx, 0.0, cfg.Ixz], [0.0, cfg.Iyy, 0.0], [cfg.Ixz, 0.0, cfg.Izz]], np.float64)
    com_term = np.array([[y * y + z * z, -x * y, -x * z],
                         [-y * x, x * x + z * z, -y * z],
                         [-z * x, -z * y, x * x + y * y]], np.float64)
    inertia = static + float(cfg.mass) * com_term
    return inertia, np.linalg.inv(inertia)
