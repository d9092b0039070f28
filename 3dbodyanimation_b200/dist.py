"""Multi-GPU drivers for the two sharding modes of the hot path; one process per GPU, torch.distributed.

  shard_frames / shard_sequences   `3dba_single` frames (src/main_single_frame.cpp:192 loop carries no state) and independent
                                   sequences shard with NO data-path collective: contiguous frame ranges / round-robin sequences.
  fit_sequence_sharded             ONE `3dba_multi` sequence (src/main_multi_frame.cpp:109-201) over G ranks.  Stage 2 of the
                                   reference is a serial chain of windows; the sharded schedule cuts the window list into G
                                   contiguous runs, rank g solves run g from the post-stage-1 state (pass 1), receives the
                                   OVERLAP frames its left neighbour refined (poses + the already rotated R0: OVERLAP x 85
                                   doubles, one send/recv per shard edge -- the only exchange on the path; NCCL over NVLink on
                                   GPUs, gloo in the CPU tests) and re-solves its first window from them (pass 2).  With one
                                   rank it is the reference's serial chain.  The same schedule is implemented inside one GPU by
                                   b3d_fit_sequences(num_shards = G) (csrc/sequence.cu) and by the oracle (oracle/fit.cpp
                                   fit_sequence), which is what parity is checked against.

The solves go through a `solver` object.  The product solver is GpuWindowSolver (C-ABI calls on device tensors); the CPU tests
inject a deterministic stand-in to exercise the schedule and the exchange under gloo -- there is no CPU fitting path here.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist

NPF, NS, NJ, MAXKP = 76, 10, 24, 17


# ---------------------------------------------------------------------------------------------------------
# schedules (pure host logic)
# ---------------------------------------------------------------------------------------------------------
def shard_frames(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous frame range [f0, f1) of rank `rank` (single-frame mode: frames are independent)."""
    per, rem = divmod(n, world)
    f0 = rank * per + min(rank, rem)
    return f0, f0 + per + (1 if rank < rem else 0)


def shard_icp(cptr, dptr, world: int, rank: int):
    """Bodies of an ICP batch are independent (api.IcpProblem / api.icp_nearest): rank `rank` takes the contiguous body range of
    shard_frames and the matching slices of the pair and data arrays.  Returns (b0, b1, cptr_local, dptr_local, pair_slice,
    data_slice) with the local offsets rebased to 0; no collective is involved, results concatenate in rank order."""
    import numpy as np
    cptr, dptr = np.asarray(cptr, dtype=np.int64), np.asarray(dptr, dtype=np.int64)
    b0, b1 = shard_frames(len(cptr) - 1, world, rank)
    return (b0, b1, cptr[b0:b1 + 1] - cptr[b0], dptr[b0:b1 + 1] - dptr[b0], slice(int(cptr[b0]), int(cptr[b1])),
            slice(int(dptr[b0]), int(dptr[b1])))


def shard_sequences(nseq: int, world: int, rank: int) -> list[int]:
    """Independent sequences round-robin over ranks (SURVEY §8(e) last row)."""
    return list(range(rank, nseq, world))


@dataclass
class WindowPlan:
    """The window list of one sequence and its cut into shards (mirrors csrc/sequence.cu and oracle/fit.cpp)."""
    n: int
    wsize: int
    overlap: int
    skip: int
    shards: int

    def __post_init__(self):
        assert 0 <= self.overlap < self.wsize and self.skip > 0 and self.n > 0
        self.stride = self.wsize - self.overlap
        self.starts = list(range(0, self.n, self.stride))                 # main_multi_frame.cpp:165
        self.anchors = list(range(0, self.n, self.skip))                  # :109-112
        k = len(self.starts)
        self.g = max(1, min(self.shards, k))
        self.cuts = [(i * k) // self.g for i in range(self.g + 1)]        # shard i = windows [cuts[i], cuts[i+1])

    def window(self, k: int) -> tuple[int, int]:
        s = self.starts[k]
        return s, min(s + self.wsize, self.n)                             # tail windows shrink (:166)

    def windows_of(self, shard: int) -> list[tuple[int, int]]:
        return [self.window(k) for k in range(self.cuts[shard], self.cuts[shard + 1])]

    def own_range(self, shard: int) -> tuple[int, int]:
        """Frames whose final values come from this shard."""
        f0 = self.starts[self.cuts[shard]]
        f1 = self.starts[self.cuts[shard + 1]] if shard + 1 < self.g else self.n
        return f0, f1

    def halo_range(self, shard: int) -> tuple[int, int]:
        """Frames shard `shard` (> 0) receives from shard - 1: the first `overlap` frames of its first window."""
        s, e = self.window(self.cuts[shard])
        return s, min(s + self.overlap, e)

    def fix_keep_range(self, shard: int) -> tuple[int, int]:
        """Frames of the pass-2 re-solve that replace pass-1 values: up to where the shard's second window takes over."""
        s, e = self.window(self.cuts[shard])
        nwin = self.cuts[shard + 1] - self.cuts[shard]
        return s, (min(s + self.stride, e) if nwin > 1 else e)


# ---------------------------------------------------------------------------------------------------------
# solvers
# ---------------------------------------------------------------------------------------------------------
STATE_KEYS = ("kps", "nkp", "poses", "R0", "W", "r", "p")


def make_state(kps, nkp, poses, R0, W, device):
    """Per-frame state of a sequence as tensors on `device`: kps (N, 408) uint8 view of b3d_kp[17], nkp (N,) int32,
    poses (N, 76), R0 (N, 9), W (N, 10), r (N, 216), p (N, 3) float64."""
    from ._capi import KP_DTYPE
    k = np.ascontiguousarray(kps, dtype=KP_DTYPE).reshape(-1, MAXKP)
    n = k.shape[0]
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(device)
    return dict(kps=torch.from_numpy(k.view(np.uint8).reshape(n, -1).copy()).to(device), nkp=t(nkp, np.int32).reshape(n),
                poses=t(poses, np.float64).reshape(n, NPF), R0=t(R0, np.float64).reshape(n, 9), W=t(W, np.float64).reshape(n, NS),
                r=torch.zeros(n, NJ * 9, dtype=torch.float64, device=device), p=torch.zeros(n, 3, dtype=torch.float64, device=device))


class GpuWindowSolver:
    """OptimizeMultiFrame solves + write-back on device tensors through the C-ABI (b3d_fit_windows_dev, b3d_writeback_dev)."""

    def __init__(self, model, base_opts, seq_opts):
        from . import _capi as capi
        self.capi, self.lib, self.model = capi, capi.lib(), model
        mk = lambda bs, it: capi.FitOpts(1, it, seq_opts.beta_pose, bs, seq_opts.lambda_temp, base_opts.fx, base_opts.fy, base_opts.cx,
                                         base_opts.cy, base_opts.huber, base_opts.scale_lo, base_opts.scale_hi)
        # stage 1: beta_shape, it1; stage 2: the reference hard-codes beta_shape = 1e5 and 60 iterations (main_multi_frame.cpp:183-185)
        self.opts = {1: mk(seq_opts.beta_shape, seq_opts.it1), 2: mk(seq_opts.beta_shape_lock, seq_opts.it2)}
        self.solves = []          # (stage, first, count, iterations, termination, usable) per solve, read back lazily

    def solve(self, stage, first, count, st):
        dev = st["poses"].device
        assert dev.type == "cuda", "GpuWindowSolver needs device tensors (there is no CPU fitting path)"
        stat = torch.zeros(3, dtype=torch.int32, device=dev)
        us = torch.zeros(1, dtype=torch.uint8, device=dev)
        f, c = (C.c_int32 * 1)(first), (C.c_int32 * 1)(count)
        P = lambda t: C.c_void_p(t.data_ptr())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        self.capi.check(self.lib.b3d_fit_windows_dev(self.model.h, C.byref(self.opts[stage]), 1, f, c, P(st["kps"]), P(st["nkp"]), P(st["R0"]),
                                                     P(st["poses"]), P(st["W"]), P(us), None, 0, P(stat[0:1]), P(stat[1:2]), stream),
                        "b3d_fit_windows_dev")
        self.solves.append((stage, first, count, stat, us))

    def writeback(self, first, count, st):
        dev = st["poses"].device
        P = lambda t, off: C.c_void_p(t.data_ptr() + off)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        self.capi.check(self.lib.b3d_writeback_dev(count, P(st["poses"], first * NPF * 8), P(st["R0"], first * 9 * 8), P(st["r"], first * NJ * 9 * 8),
                                                   P(st["p"], first * 3 * 8), stream), "b3d_writeback_dev")

    def report(self):
        return [(s, f, c, int(stat[0]), int(stat[1]), bool(us[0])) for (s, f, c, stat, us) in self.solves]


# ---------------------------------------------------------------------------------------------------------
# one sequence over G ranks
# ---------------------------------------------------------------------------------------------------------
def _stage1(solver, st, plan):
    """Anchor stage (main_multi_frame.cpp:109-154): one solve over all anchors on copies of their parameters; what persists is
    the write-back into the avatars (R0, r, p) and the front avatar's w, which every frame then receives."""
    dev = st["poses"].device
    idx = torch.tensor(plan.anchors, dtype=torch.long, device=dev)
    a = {k: st[k].index_select(0, idx).contiguous() for k in STATE_KEYS}
    solver.solve(1, 0, len(plan.anchors), a)
    solver.writeback(0, len(plan.anchors), a)
    for k in ("R0", "r", "p"):
        st[k].index_copy_(0, idx, a[k])
    st["W"][:] = a["W"][0]


def fit_sequence_sharded(solver, st, plan: WindowPlan, rank=None, world=None, group=None, gather=True):
    """Run the sharded 3dba_multi schedule for one sequence whose full per-frame state `st` is replicated on every rank.
    Rank g solves shard g; ranks >= plan.g idle.  Returns st with, when `gather`, every rank holding the assembled result."""
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        world = dist.get_world_size(group) if dist.is_initialized() else 1
    assert plan.g <= world, "more shards than ranks"
    dev = st["poses"].device
    _stage1(solver, st, plan)          # redundantly on every rank: deterministic, and stage 2 needs its result everywhere
    orig = {k: st[k].clone() for k in ("poses", "R0", "W")}
    active = rank < plan.g
    if active:
        for (s, e) in plan.windows_of(rank):                      # pass 1: this shard's serial chain of windows
            solver.solve(2, s, e - s, st)
            solver.writeback(s, e - s, st)
    # ---- halo hand-off: OVERLAP x (76 + 9) doubles per shard edge ----
    ops, recv_buf, send_buf = [], None, None
    if active and rank + 1 < plan.g:
        h0, h1 = plan.halo_range(rank + 1)
        send_buf = torch.cat([st["poses"][h0:h1], st["R0"][h0:h1]], dim=1).contiguous()
        if world > 1 and h1 > h0:
            ops.append(dist.P2POp(dist.isend, send_buf, rank + 1 if group is None else dist.get_global_rank(group, rank + 1), group))
    if active and rank > 0:
        h0, h1 = plan.halo_range(rank)
        recv_buf = torch.empty(h1 - h0, NPF + 9, dtype=torch.float64, device=dev)
        if world > 1 and h1 > h0:
            ops.append(dist.P2POp(dist.irecv, recv_buf, rank - 1 if group is None else dist.get_global_rank(group, rank - 1), group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    # ---- pass 2: re-solve the shard's first window from the post-stage-1 state with the halo patched in ----
    if active and rank > 0:
        h0, h1 = plan.halo_range(rank)
        f = dict(kps=st["kps"], nkp=st["nkp"], poses=orig["poses"].clone(), R0=orig["R0"].clone(), W=orig["W"].clone(),
                 r=torch.zeros_like(st["r"]), p=torch.zeros_like(st["p"]))
        if h1 > h0:
            f["poses"][h0:h1] = recv_buf[:, :NPF]
            f["R0"][h0:h1] = recv_buf[:, NPF:]
        s, e = plan.window(plan.cuts[rank])
        solver.solve(2, s, e - s, f)
        solver.writeback(s, e - s, f)
        k0, k1 = plan.fix_keep_range(rank)
        for k in ("poses", "R0", "W", "r", "p"):
            st[k][k0:k1] = f[k][k0:k1]
    if gather and world > 1:
        f0, f1 = plan.own_range(rank) if active else (0, 0)
        for k in ("poses", "R0", "W", "r", "p"):       # result collection (not the data path): owned rows, zero elsewhere, summed
            t = torch.zeros_like(st[k])
            t[f0:f1] = st[k][f0:f1]
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
            st[k].copy_(t)
    return st


def emulate_sharded(solver, st, plan: WindowPlan):
    """The same schedule with the G ranks played one after the other in this process (no process group): what a single GPU
    runs to check the multi-rank driver against b3d_fit_sequences(num_shards = G) and the oracle."""
    g = plan.g
    states = [{k: v.clone() for k, v in st.items()} for _ in range(g)]
    origs = []
    for r in range(g):
        _stage1(solver, states[r], plan)
        origs.append({k: states[r][k].clone() for k in ("poses", "R0", "W")})
        for (s, e) in plan.windows_of(r):
            solver.solve(2, s, e - s, states[r])
            solver.writeback(s, e - s, states[r])
    out = {k: v.clone() for k, v in states[0].items()}
    for r in range(1, g):
        h0, h1 = plan.halo_range(r)
        f = dict(kps=st["kps"], nkp=st["nkp"], poses=origs[r]["poses"].clone(), R0=origs[r]["R0"].clone(), W=origs[r]["W"].clone(),
                 r=torch.zeros_like(st["r"]), p=torch.zeros_like(st["p"]))
        f["poses"][h0:h1] = states[r - 1]["poses"][h0:h1]
        f["R0"][h0:h1] = states[r - 1]["R0"][h0:h1]
        s, e = plan.window(plan.cuts[r])
        solver.solve(2, s, e - s, f)
        solver.writeback(s, e - s, f)
        k0, k1 = plan.fix_keep_range(r)
        f0, f1 = plan.own_range(r)
        for k in ("poses", "R0", "W", "r", "p"):
            out[k][f0:f1] = states[r][k][f0:f1]
            out[k][k0:k1] = f[k][k0:k1]
    return out
