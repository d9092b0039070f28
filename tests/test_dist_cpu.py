"""Host-side logic of the multi-GPU drivers (3dbodyanimation_b200/dist.py) on CPU: shard plans, and the sharded 3dba_multi
schedule with its halo hand-off run by 2 and 3 `gloo` ranks against the same schedule played in one process.  The LM solves
are replaced by a deterministic stand-in (the real solver needs a GPU; `-m gpu` tests cover it) whose result depends on every
input a window solve reads, so a wrong halo, keep range or ownership shows up as a mismatch."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
D = importlib.import_module("3dbodyanimation_b200.dist")


class StandInSolver:
    """Deterministic, order-sensitive replacement for the window solve / write-back (test only)."""

    def solve(self, stage, first, count, st):
        sl = slice(first, first + count)
        po, r0 = st["poses"][sl], st["R0"][sl]
        k = st["kps"][sl].to(torch.float64).mean(dim=1, keepdim=True) / 255.0
        ctx = po.mean(dim=0, keepdim=True) + 0.01 * r0.mean()            # couples the frames of the window, like the temporal terms
        st["poses"][sl] = 0.5 * po + 0.25 * ctx + 0.125 * k + 0.001 * stage + 0.01 * st["W"][first].sum()
        st["W"][first] = 0.9 * st["W"][first] + 0.01 * po.mean()

    def writeback(self, first, count, st):
        sl = slice(first, first + count)
        st["R0"][sl] = 0.99 * st["R0"][sl] + 0.01 * st["poses"][sl, 1:4].sum(dim=1, keepdim=True)   # cumulative, in place
        st["r"][sl] = st["R0"][sl].repeat(1, 24) * 0.5
        st["p"][sl] = st["poses"][sl, 4:7]


def make_state(n, seed):
    g = torch.Generator().manual_seed(seed)
    return dict(kps=torch.randint(0, 255, (n, 408), dtype=torch.uint8, generator=g), nkp=torch.full((n,), 17, dtype=torch.int32),
                poses=torch.randn(n, 76, dtype=torch.float64, generator=g), R0=torch.randn(n, 9, dtype=torch.float64, generator=g),
                W=torch.randn(n, 10, dtype=torch.float64, generator=g), r=torch.zeros(n, 216, dtype=torch.float64),
                p=torch.zeros(n, 3, dtype=torch.float64))


def test_shard_frames_and_sequences():
    for n, w in [(1000, 8), (7, 3), (3, 8), (100000, 8)]:
        ranges = [D.shard_frames(n, w, r) for r in range(w)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(w - 1))
        assert max(b - a for a, b in ranges) - min(b - a for a, b in ranges) <= 1
    assert sorted(sum((D.shard_sequences(50, 8, r) for r in range(8)), [])) == list(range(50))


def test_shard_icp_partitions_bodies_pairs_and_data():
    """ICP batches shard by body with no exchange: the ranks' slices tile the pair / data arrays, and evaluating the oracle on
    every shard reproduces the unsharded cost and gradient bit for bit."""
    import importlib
    from oracle import oracle_api as orc
    dist_mod = importlib.import_module("3dbodyanimation_b200.dist")
    synth = importlib.import_module("3dbodyanimation_b200.synth")
    model_np = synth.make_model()
    x = synth.make_icp_inputs(model_np, 5, 61, ndata=200)
    om = orc.Model(model_np)
    full = om.icp_eval(x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"])
    for world in (2, 3):
        cost, grad, pairs, pts = [], [], 0, 0
        for rank in range(world):
            b0, b1, cl, dl, ps, ds = dist_mod.shard_icp(x["cptr"], x["dptr"], world, rank)
            assert cl[0] == 0 and dl[0] == 0 and ps.start == pairs and ds.start == pts
            pairs, pts = ps.stop, ds.stop
            part = om.icp_eval(x["w"][b0:b1], x["p"][b0:b1], x["q"][b0:b1], cl, x["mpt"][ps], x["dpt"][ps], dl, x["data"][ds])
            cost.append(part["cost"]); grad.append(part["grad"])
        assert pairs == x["cptr"][-1] and pts == x["dptr"][-1]
        assert np.array_equal(np.concatenate(cost), full["cost"]) and np.array_equal(np.concatenate(grad), full["grad"])


def test_window_plan_matches_reference_schedule():
    # src/main_multi_frame.cpp:109-112, 165-166 with the README values: 2,000 frames -> 200 anchors, 134 windows; 100k -> 10,000 / 6,667
    p = D.WindowPlan(2000, 20, 5, 10, 1)
    assert len(p.anchors) == 200 and len(p.starts) == 134 and p.window(133) == (1995, 2000)
    p = D.WindowPlan(100000, 20, 5, 10, 8)
    assert len(p.anchors) == 10000 and len(p.starts) == 6667 and p.g == 8
    own = [p.own_range(g) for g in range(8)]
    assert own[0][0] == 0 and own[-1][1] == 100000 and all(own[i][1] == own[i + 1][0] for i in range(7))
    for g in range(1, 8):
        h0, h1 = p.halo_range(g)
        assert h1 - h0 == 5 and h0 == own[g][0]
        assert p.fix_keep_range(g) == (h0, h0 + 15)
    # fewer windows than shards: the surplus ranks idle
    p = D.WindowPlan(20, 20, 5, 10, 4)
    assert p.g == 2 and p.windows_of(1) == [(15, 20)] and p.fix_keep_range(1) == (15, 20)


def test_one_shard_is_the_serial_chain():
    st = make_state(83, 1)
    plan = D.WindowPlan(83, 20, 5, 10, 1)
    a = D.fit_sequence_sharded(StandInSolver(), {k: v.clone() for k, v in st.items()}, plan, rank=0, world=1)
    # by hand: stage 1, then every window in order on the same state
    b = {k: v.clone() for k, v in st.items()}
    s = StandInSolver()
    D._stage1(s, b, plan)
    for k in range(len(plan.starts)):
        f0, f1 = plan.window(k)
        s.solve(2, f0, f1 - f0, b)
        s.writeback(f0, f1 - f0, b)
    for k in ("poses", "R0", "W", "r", "p"):
        assert torch.equal(a[k], b[k]), k


def _worker(rank, world, port, n, seed, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        st = make_state(n, seed)
        plan = D.WindowPlan(n, 20, 5, 10, world)
        res = D.fit_sequence_sharded(StandInSolver(), st, plan)
        torch.save({k: res[k] for k in ("poses", "R0", "W", "r", "p")}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,n", [(2, 83), (3, 140), (2, 20)])
def test_sharded_schedule_under_gloo(tmp_path, world, n):
    mp.spawn(_worker, args=(world, _free_port(), n, 7, str(tmp_path)), nprocs=world, join=True)
    ref = D.emulate_sharded(StandInSolver(), make_state(n, 7), D.WindowPlan(n, 20, 5, 10, world))
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"rank{r}.pt"))
        for k in ("poses", "R0", "W", "r", "p"):
            assert torch.equal(got[k], ref[k]), (r, k)
    # and the sharded result differs from the serial chain only where the schedule says it may: shard 0's frames are identical
    serial = D.fit_sequence_sharded(StandInSolver(), make_state(n, 7), D.WindowPlan(n, 20, 5, 10, 1), rank=0, world=1)
    plan = D.WindowPlan(n, 20, 5, 10, world)
    f0, f1 = plan.own_range(0)
    keep = f1 if plan.g == 1 else plan.halo_range(1)[0]
    assert torch.equal(serial["poses"][f0:keep], ref["poses"][f0:keep])


def _worker_multi(rank, world, port, n, seed, per_rank, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        st = make_state(n, seed)
        plan = D.WindowPlan(n, 20, 5, 10, world * per_rank)
        res = D.fit_sequence_multishard(StandInSolver(), st, plan)
        torch.save({k: res[k] for k in ("poses", "R0", "W", "r", "p")}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,per_rank", [(2, 200, 3), (3, 260, 2), (2, 40, 4)])
def test_multishard_schedule_under_gloo(tmp_path, world, n, per_rank):
    """Several shards per rank (a rank's shards advance together, wave by wave; halos inside a rank are copies, the one per rank
    edge goes over the process group): every rank ends with the result of the same schedule played in one process."""
    mp.spawn(_worker_multi, args=(world, _free_port(), n, 11, per_rank, str(tmp_path)), nprocs=world, join=True)
    plan = D.WindowPlan(n, 20, 5, 10, world * per_rank)
    ref = D.emulate_sharded(StandInSolver(), make_state(n, 11), plan)
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"rank{r}.pt"))
        for k in ("poses", "R0", "W", "r", "p"):
            assert torch.equal(got[k], ref[k]), (r, k)
    # shard -> rank map: contiguous blocks that cover every shard once
    owned = [list(D.shards_of_rank(plan, r, world)) for r in range(world)]
    assert sum(owned, []) == list(range(plan.g))
