"""DM-compatible return values.

The reference's callers consume what a CasADi Function returns -- a ``casadi.DM`` -- and immediately ask it for
numpy data:

    self.dynamics(x, u, dt).toarray().flatten()          /root/reference/main/control/control.py:84-90
                                                         /root/reference/src/aircraft/control/moving_horizon.py:208-213
    aircraft.phi(state).full().flatten(); state.full()   /root/reference/main/dynamics/dynamics.py:112-118
    np.isnan(state[0])                                   /root/reference/main/dynamics/dynamics.py:108

``DM`` below is a thin ``numpy.ndarray`` subclass (host path) and ``DMTensor`` a thin ``torch.Tensor`` subclass (device
path) that answer the part of the DM interface those call sites use -- ``full()``, ``toarray()``, ``nonzeros()``,
``size1()/size2()``, ``float()`` -- so the loops run unmodified against ``aircraft_b200.Aircraft``.  They are plain
arrays / tensors otherwise: indexing, arithmetic and numpy / torch functions work as usual, and no arithmetic of the
hot path happens here (values come from the CUDA library).  ``DM(a)`` also stands in for ``casadi.DM(a)`` where a
script builds its initial state (main/control/control.py:178, main/dynamics/dynamics.py:62-70).
"""
from __future__ import annotations

import numpy as np

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


def _as_matrix(a: np.ndarray) -> np.ndarray:
    """CasADi matrices are always 2-D; a vector is a column."""
    a = np.asarray(a)
    if a.ndim == 0:
        return a.reshape(1, 1)
    if a.ndim == 1:
        return a.reshape(-1, 1)
    return a


class DM(np.ndarray):
    """numpy array with the accessors of ``casadi.DM`` that the reference's call sites use."""

    def __new__(cls, data=0.0, dtype=np.float64):
        if hasattr(data, "full") and not isinstance(data, np.ndarray):      # a real casadi.DM (or a DMTensor)
            data = data.full()
        arr = np.array(data, dtype=dtype, copy=True)
        if arr.ndim <= 1:                                                   # casadi.DM([...]) is a column
            arr = _as_matrix(arr)
        return arr.view(cls)

    # ---- casadi.DM accessors
    def full(self) -> np.ndarray:
        """dense 2-D numpy copy (vectors come back as columns), as ``DM.full()``"""
        return np.array(_as_matrix(self.view(np.ndarray)))

    def toarray(self, simplify: bool = False):
        """as ``DM.toarray()``: 2-D numpy copy; ``simplify=True`` squeezes vectors / scalars like CasADi does"""
        m = self.full()
        if simplify:
            if m.size == 1:
                return float(m[0, 0])
            if 1 in m.shape:
                return m.ravel()
        return m

    def nonzeros(self):
        return [float(v) for v in self.full().ravel(order="F")]            # CasADi is column-major

    def size1(self) -> int:
        return int(_as_matrix(self.view(np.ndarray)).shape[0])

    def size2(self) -> int:
        return int(_as_matrix(self.view(np.ndarray)).shape[1])

    def is_scalar(self) -> bool:
        return self.size == 1

    def __float__(self):
        if self.size != 1:
            raise TypeError("only a 1x1 DM converts to float")
        return float(self.view(np.ndarray).reshape(-1)[0])


def as_dm(a, like=None) -> DM:
    """View a host result as DM without copying (shape is kept: 1-D in, 1-D out)."""
    return np.asarray(a).view(DM)


if torch is not None:
    class DMTensor(torch.Tensor):
        """torch tensor (results of the device path stay on the GPU) with ``full()`` / ``toarray()``: the
        device->host copy happens only when a caller asks for numpy data, exactly where the reference's loops do."""

        @staticmethod
        def wrap(t: "torch.Tensor") -> "DMTensor":
            return t.as_subclass(DMTensor)

        def full(self) -> np.ndarray:
            return _as_matrix(self.detach().as_subclass(torch.Tensor).cpu().numpy()).astype(np.float64)

        def toarray(self, simplify: bool = False):
            m = self.full()
            if simplify:
                if m.size == 1:
                    return float(m[0, 0])
                if 1 in m.shape:
                    return m.ravel()
            return m

        def nonzeros(self):
            return [float(v) for v in self.full().ravel(order="F")]

        def size1(self) -> int:
            return int(self.shape[0]) if self.ndim else 1

        def size2(self) -> int:
            return int(self.shape[1]) if self.ndim > 1 else 1

    def as_dm_tensor(t):
        return DMTensor.wrap(t)
else:  # pragma: no cover
    DMTensor = None

    def as_dm_tensor(t):
        return t


class Function:
    """What the reference's dynamics object hands out for ``state_update``, ``state_derivative`` and the side
    functions: a named callable with the introspection of ``casadi.Function`` that the callers use
    (``ControlProblem.__init__`` reads ``dynamics.size1_in(0)`` / ``size1_in(1)``,
    /root/reference/src/aircraft/control/base.py:187-189).  Calling it evaluates on the GPU."""

    def __init__(self, name, fn, in_names, in_sizes, out_names, out_sizes):
        self._name, self._fn = name, fn
        self._in_names, self._in_sizes = tuple(in_names), tuple(in_sizes)
        self._out_names, self._out_sizes = tuple(out_names), tuple(out_sizes)
        self.__doc__ = fn.__doc__

    def __call__(self, *args, **kw):
        return self._fn(*args, **kw)

    def call(self, args, **kw):
        """``Function.call``: a list of inputs (or a dict keyed by input name) -> list (dict) of outputs"""
        if isinstance(args, dict):
            out = self._fn(*[args[n] for n in self._in_names], **kw)
            return {self._out_names[0]: out}
        return [self._fn(*args, **kw)]

    def name(self):
        return self._name

    def n_in(self):
        return len(self._in_sizes)

    def n_out(self):
        return len(self._out_sizes)

    def name_in(self, i=None):
        return list(self._in_names) if i is None else self._in_names[i]

    def name_out(self, i=None):
        return list(self._out_names) if i is None else self._out_names[i]

    def size1_in(self, i):
        return self._in_sizes[self._idx(i, self._in_names)]

    def size2_in(self, i):
        return 1

    def size_in(self, i):
        return (self.size1_in(i), 1)

    def numel_in(self, i):
        return self.size1_in(i)

    def size1_out(self, i):
        return self._out_sizes[self._idx(i, self._out_names)]

    def size2_out(self, i):
        return 1

    def size_out(self, i):
        return (self.size1_out(i), 1)

    def numel_out(self, i):
        return self.size1_out(i)

    @staticmethod
    def _idx(i, names):
        return names.index(i) if isinstance(i, str) else int(i)

    def __repr__(self):
        ins = ",".join(f"{n}[{s}]" for n, s in zip(self._in_names, self._in_sizes))
        outs = ",".join(f"{n}[{s}]" for n, s in zip(self._out_names, self._out_sizes))
        return f"Function({self._name}:({ins})->({outs}) aircraft_b200 CUDA)"
