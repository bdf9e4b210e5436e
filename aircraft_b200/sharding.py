"""Data-parallel sharding of independent rollouts / linearisation nodes over the GPUs of one box.

The reference has no parallelism at all; every rollout and every (x,u)
node is independent, so the batch is cut into contiguous slices, one per rank (one process per
GPU), with no data-path traffic.  The only exchange is the final one north_star asks for:
all ranks learn (min cost, global arg-min) and receive the winner's trajectory and control
sequence -- one all_gather of 16 bytes per rank plus one broadcast of ~16 kB over NCCL/NVLink.
All costs can optionally be gathered (4 B per trajectory).

The collective helpers take CPU tensors under gloo as well, which is how the N>1 logic is tested
without GPUs (tests/test_sharding.py).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [start, stop) of `total` items for `rank`; the remainder goes to the lowest ranks."""
    if world < 1 or not (0 <= rank < world) or total < 0:
        raise ValueError(f"bad shard request total={total} rank={rank} world={world}")
    base, rem = divmod(total, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def _orderable(cost: torch.Tensor) -> torch.Tensor:
    """float32 -> int64 key that sorts like the float (NaN last); same mapping as the argmin kernel."""
    bits = cost.detach().to(torch.float32).contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    neg = (bits & 0x80000000) != 0
    key = torch.where(neg, (~bits) & 0xFFFFFFFF, bits | 0x80000000)
    return torch.where(torch.isnan(cost.to(torch.float32)), torch.full_like(key, 0xFFFFFFFF), key)


def global_argmin(local_min: torch.Tensor, local_global_idx: torch.Tensor,
                  group=None) -> Tuple[torch.Tensor, int, int]:
    """(min cost, global index, owner-independent) over all ranks.  Ties resolve to the lower
    global index, so the result does not depend on how the batch was sharded."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    pair = torch.stack([_orderable(local_min.reshape(1))[0], local_global_idx.reshape(1)[0].to(torch.int64)])
    vals = torch.stack([local_min.reshape(1)[0].to(torch.float32)])
    if world == 1:
        return vals[0], int(pair[1].item()), 0
    pairs = [torch.empty_like(pair) for _ in range(world)]
    mins = [torch.empty_like(vals) for _ in range(world)]
    dist.all_gather(pairs, pair, group=group)
    dist.all_gather(mins, vals, group=group)
    keys = torch.stack(pairs).cpu()
    best = min(range(world), key=lambda r: (int(keys[r, 0]), int(keys[r, 1])))
    return mins[best][0], int(keys[best, 1]), best


def local_min_or_sentinel(aircraft, cost_local: torch.Tensor, start: int):
    """(min, GLOBAL arg-min) of this rank's costs for the exchange.  A rank whose shard is empty (total < world)
    contributes a sentinel that can never win -- NaN sorts last and the index is beyond any real one -- instead of
    calling the arg-min kernel with B = 0 (AIRC_EINVAL) and leaving the other ranks blocked in the all-gather."""
    if cost_local.numel() == 0:
        dev = cost_local.device
        return (torch.full((1,), float("nan"), dtype=torch.float32, device=dev),
                torch.full((1,), 2 ** 62, dtype=torch.int64, device=dev))
    lmin, lidx = aircraft.argmin(cost_local)
    return lmin, lidx + start


@dataclass
class SweepResult:
    min_cost: Any                    # 0-dim tensor (no host sync is forced)
    argmin: int                      # global trajectory index
    owner: int                       # rank that held it
    best_trajectory: Any             # [H+1,13] on every rank
    best_controls: Any               # [H,7] on every rank
    local_cost: Any                  # this rank's costs [stop-start]
    all_cost: Optional[Any] = None   # [B] on every rank when gather_costs


class ShardedRollout:
    """Sampling-MPC sweep (BASELINE config 4): B control sequences from one or many initial
    states, sharded over the ranks of the default process group."""

    def __init__(self, aircraft, group=None):
        self.aircraft = aircraft
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def local_range(self, total: int) -> Tuple[int, int]:
        return shard_range(total, self.rank, self.world)

    def sweep(self, x0_local, U_local, dt: float, total: int, *, cost=True, layout: str = "soa",
              substeps: Optional[int] = None, normalise: Optional[bool] = None,
              gather_costs: bool = False) -> SweepResult:
        """x0_local / U_local hold this rank's slice [start, stop) of the `total` rollouts in the
        given layout.  Runs the fused rollout+cost kernel, reduces (min, arg-min) on the device,
        exchanges it, and hands every rank the winner's trajectory (re-rolled by its owner: one
        extra single-trajectory launch instead of storing B trajectories)."""
        res = self.aircraft.rollout(x0_local, U_local, dt, substeps=substeps, normalise=normalise,
                                    layout=layout, return_final=False, cost=cost)
        return self.finish_sweep(res.cost, x0_local, U_local, dt, total, layout=layout,
                                 substeps=substeps, normalise=normalise, gather_costs=gather_costs)

    def finish_sweep(self, cost_local, x0_local, U_local, dt: float, total: int, *, layout: str = "soa",
                     substeps: Optional[int] = None, normalise: Optional[bool] = None,
                     gather_costs: bool = False) -> SweepResult:
        """Everything after the rollout kernel: device (min, arg-min) of this rank's costs, the
        exchange, and the winner's trajectory on every rank."""
        ac = self.aircraft
        start, stop = self.local_range(total)
        lmin, lgidx = local_min_or_sentinel(ac, cost_local, start)
        gmin, gidx, owner = global_argmin(lmin, lgidx, self.group)
        soa = layout == "soa"
        shared = U_local.ndim == 2
        H = U_local.shape[0] if (soa or shared) else U_local.shape[1]
        dev = cost_local.device
        traj = torch.empty((H + 1, 13), dtype=torch.float32, device=dev)
        ctrl = torch.empty((H, 7), dtype=torch.float32, device=dev)
        if owner == self.rank:
            j = gidx - start
            xb = (x0_local[:, j] if soa else x0_local[j]).reshape(1, 13).contiguous()
            ub = U_local if shared else (U_local[:, :, j] if soa else U_local[j])
            ub = ub.reshape(H, 7).contiguous()
            one = ac.rollout(xb, ub, dt, substeps=substeps, normalise=normalise, layout="aos",
                             return_trajectory=True, return_final=False)
            traj.copy_(one.trajectory[0])
            ctrl.copy_(ub)
        all_cost = None
        if self.world > 1:
            src = dist.get_global_rank(self.group, owner) if self.group else owner
            dist.broadcast(traj, src=src, group=self.group)
            dist.broadcast(ctrl, src=src, group=self.group)
        if gather_costs:
            all_cost = gather_varlen(cost_local, total, self.rank, self.world, self.group)
        return SweepResult(gmin, gidx, owner, traj, ctrl, cost_local, all_cost)


def select_record_torch(records: torch.Tensor) -> torch.Tensor:
    """Best of n stacked sweep records [n,R]: lowest orderable cost (word 1), ties to the lower
    global index (words 2-3), NaN last -- the rule of select_record_kernel, in torch ops, for
    records that live on the host (gloo tests, host-side callers)."""
    r = records.contiguous()
    words = r[:, 1:4].contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    keys = [(int(w[0]), int(w[1]) | (int(w[2]) << 32)) for w in words.tolist()]
    best = min(range(len(keys)), key=lambda i: keys[i])
    return r[best]


def sweep_sampled(aircraft, x0, U_nom, dt: float, total: int, *, sampler=None, cost=True,
                  substeps: Optional[int] = None, normalise: Optional[bool] = None, group=None,
                  workspace=None, return_costs: bool = False, on_local=None):
    """BASELINE config 4 over the ranks of `group`: `total` control sequences sampled in the
    kernel, contiguous index ranges per rank (the sampler is keyed on the GLOBAL trajectory
    index, so the candidate set does not depend on the sharding), one all-gather of the per-rank
    winner records (16 kB each) and a device-side selection.  No host synchronisation."""
    from .aircraft import decode_record
    from .params import make_sweep_params
    from . import _cabi
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    start, stop = shard_range(total, rank, world)
    if isinstance(sampler, _cabi.AircSweepParams):
        sp = _cabi.AircSweepParams.from_buffer_copy(sampler)
    else:
        sp = make_sweep_params(**(sampler or {}))
    base = int(sp.index_offset)
    sp.index_offset = base + start
    local = aircraft.sweep(x0, U_nom, dt, stop - start, sampler=sp, cost=cost, substeps=substeps,
                           normalise=normalise, workspace=workspace, return_costs=return_costs)
    if on_local is not None:
        on_local()          # e.g. a CUDA event that separates the local kernels from the exchange
    if world == 1:
        return local
    H = local.H
    allrec = torch.empty((world, local.raw.numel()), dtype=torch.float32, device=local.raw.device)
    dist.all_gather_into_tensor(allrec, local.raw.reshape(1, -1), group=group)
    best = aircraft.select_record(allrec) if allrec.is_cuda else select_record_torch(allrec)
    out = decode_record(best, H, local.cost)
    out.workspace = local.workspace
    return out


class SweepPlan:
    """The whole sharded sweep step -- prep + time-sliced sweep + winner re-roll (airc_sweep), the all-gather of the
    per-rank winner records and the device-side selection -- captured ONCE in a CUDA graph and replayed per step:
    one graph launch instead of five host-issued operations, so the step is not bound by Python / launch latency
    (at 8 GPUs the eager exchange adds ~35 us to a 2.3 ms sweep).  Buffers are owned by the plan; new inputs are
    copied into `x0` / `U_nom` (device tensors) before `replay()`.  GPU + NCCL only: the gloo path of
    sweep_sampled() stays eager.

        plan = SweepPlan(aircraft, H, dt, total, sampler=..., cost=...)
        plan.x0.copy_(x0); plan.U_nom.copy_(U_nom)
        rec = plan.replay()            # SweepRecord of the global winner (views of plan-owned memory)
    """

    def __init__(self, aircraft, H: int, dt: float, total: int, *, sampler=None, cost=True,
                 substeps: Optional[int] = None, normalise: Optional[bool] = None, group=None,
                 return_costs: bool = False):
        import ctypes as C
        from . import _cabi
        from .aircraft import _check
        from .params import make_cost_params, make_sweep_params
        ac = aircraft
        self.aircraft, self.group, self.H = ac, group, int(H)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        start, stop = shard_range(total, self.rank, self.world)
        self.local = stop - start
        sp = _cabi.AircSweepParams.from_buffer_copy(sampler) if isinstance(sampler, _cabi.AircSweepParams) \
            else make_sweep_params(**(sampler or {}))
        sp.index_offset = int(sp.index_offset) + start
        cp = cost if isinstance(cost, _cabi.AircCostParams) else make_cost_params(**(cost if isinstance(cost, dict) else {}))
        self._sp, self._cp = sp, cp
        lib, dev = ac._lib, torch.device("cuda", int(ac._device))
        R = int(lib.airc_sweep_record_floats(self.H))
        wsb = int(lib.airc_sweep_workspace_bytes(self.local, self.H, sp.ckpt_interval))
        self.x0 = torch.zeros(13, dtype=torch.float32, device=dev)
        self.U_nom = torch.zeros((self.H, 7), dtype=torch.float32, device=dev)
        self._ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        self._rec = torch.empty(R, dtype=torch.float32, device=dev)
        self._all = torch.empty((self.world, R), dtype=torch.float32, device=dev)
        self._best = torch.empty(R, dtype=torch.float32, device=dev)
        self.cost = torch.empty(self.local, dtype=torch.float32, device=dev) if return_costs else None
        ns, fl = ac._substeps(substeps), ac._flags(normalise)

        def enqueue():
            st = torch.cuda.current_stream(dev).cuda_stream
            _check(lib.airc_sweep(ac.handle, self.x0.data_ptr(), self.U_nom.data_ptr(), float(dt), self.local, self.H, ns, fl,
                                  C.byref(sp), C.byref(cp), self._ws.data_ptr(), self._ws.numel(),
                                  self.cost.data_ptr() if self.cost is not None else None, self._rec.data_ptr(), st),
                   lib, "airc_sweep")
            if self.world > 1:
                dist.all_gather_into_tensor(self._all, self._rec.reshape(1, -1), group=group)
                _check(lib.airc_select_record(ac.handle, self._all.data_ptr(), self.world, R, self._best.data_ptr(), st),
                       lib, "airc_select_record")

        # warm-up on a side stream (creates the NCCL communicator, loads the kernels), then capture
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(2):
                enqueue()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            enqueue()
        self.launches_per_replay = 1

    def replay(self):
        from .aircraft import decode_record
        self.graph.replay()
        out = decode_record(self._best if self.world > 1 else self._rec, self.H, self.cost)
        out.workspace = self._ws
        return out


def gather_varlen(local: torch.Tensor, total: int, rank: int, world: int, group=None) -> torch.Tensor:
    """all_gather of contiguous shards whose lengths follow shard_range (differ by at most one)."""
    if world == 1:
        return local
    longest = shard_range(total, 0, world)
    pad = longest[1] - longest[0]
    buf = torch.zeros(pad, dtype=local.dtype, device=local.device)
    buf[: local.numel()] = local
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    parts = []
    for r in range(world):
        a, b = shard_range(total, r, world)
        parts.append(out[r][: b - a])
    return torch.cat(parts)
