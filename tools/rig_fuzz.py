"""Fuzz of the skinning kernels' host tables: random and extreme rigs (<= 4 joints per vertex), every sorted float32 kernel against the
generic kernel on every (body, vertex).  usage: python tools/rig_fuzz.py [n_random]"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dbodyanimation_b200.synth")
api = importlib.import_module("3dbodyanimation_b200.api")


def rig(kind, seed, V, J):
    rng = np.random.default_rng(seed)
    wt = np.zeros((V, J), dtype=np.float32)

    def put(lo, hi, js):
        js = np.asarray(js)
        for u in range(lo, hi):
            ww = rng.dirichlet(np.ones(len(js))).astype(np.float32) + np.float32(1e-3)
            wt[u, js] = ww / ww.sum(dtype=np.float32)
    if kind == "one_joint":
        wt[:, int(rng.integers(J))] = 1.0
    elif kind == "single_weight_each":
        wt[np.arange(V), rng.integers(0, J, size=V)] = 1.0
    elif kind == "four_random_each":
        for v in range(V):
            put(v, v + 1, rng.choice(J, size=4, replace=False))
    elif kind == "alternating_two_sets":
        a, b = rng.choice(J, size=4, replace=False), rng.choice(J, size=3, replace=False)
        for v in range(V):
            put(v, v + 1, a if v % 2 == 0 else b)
    elif kind == "high_joints_only":
        v = 0
        while v < V:
            run = int(rng.integers(1, 100))
            put(v, min(V, v + run), 16 + rng.choice(8, size=int(rng.integers(1, 5)), replace=False))
            v += run
    else:   # runs
        v = 0
        hi_run = int(rng.choice([3, 12, 70, 300, 2000]))
        while v < V:
            run = int(rng.integers(1, hi_run + 1))
            js = rng.choice(J, size=int(rng.integers(1, 5)), replace=False)
            if rng.random() < 0.3 and 0 not in js:
                js[0] = 0
            put(v, min(V, v + run), js)
            v += run
    return wt


def main():
    n_random = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    base = synth.make_model()
    V, J = base["weights"].shape
    kinds = ["one_joint", "single_weight_each", "four_random_each", "alternating_two_sets", "high_joints_only"] + ["runs"] * n_random
    B = 128 + 3
    inp = synth.make_forward_inputs(B, 4242)
    worst_all, t0 = 0.0, time.time()
    for i, kind in enumerate(kinds):
        m = dict(base)
        m["weights"] = rig(kind, 9000 + i, V, J)
        gm = api.AvatarModel(m)
        ref, _, _ = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="generic")
        scale = np.max(np.abs(ref), axis=(1, 2))
        for lbs in ("tc", "sorted", "bodylane"):
            out, _, _ = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs=lbs)
            err = float((np.max(np.abs(out.astype(np.float64) - ref), axis=(1, 2)) / scale).max())
            worst_all = max(worst_all, err)
            if not err < 2e-6:
                print(f"FAIL rig {i} ({kind}) kernel {lbs}: worst rel {err:.3e}", flush=True)
        del gm
    print(f"{len(kinds)} rigs x 3 kernels in {time.time() - t0:.1f} s, worst relative difference to the generic kernel {worst_all:.3e}")


if __name__ == "__main__":
    main()
