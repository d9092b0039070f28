"""Generates tests/golden/c3_oracle_2000.npz: the CPU oracle's result for BASELINE configs[2] at its stated size — 3dba_multi on one
synthetic sequence of 2,000 frames (synth.make_sequence seed 31), anchor_skip 10, window 20, overlap 5, lambda_temp 3, beta_pose 20,
beta_shape 30, it1 = 1000, it2 = 60, beta_shape_lock 1e5 (src/main_multi_frame.cpp:29-36, 109-201).  The oracle needs minutes for
this (one core, structured solver), so the GPU parity test compares against the committed result instead of re-running it.
    python tools/make_c3_golden.py [frames]
"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dbodyanimation_b200.synth")
from oracle import oracle_api as orc  # noqa: E402

SEED, IT1, IT2 = 31, 1000, 60


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    model_np = synth.make_model()
    om = orc.Model(model_np)
    seq = synth.make_sequence(model_np, n, SEED)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    t0 = time.time()
    o = om.fit_sequence(orc.fit_options(mode=1, cam=cam, use_jets=0), orc.SeqOptions(10, 20, 5, IT1, IT2, 1, 8, 0, 20.0, 30.0, 3.0, 1e5),
                        seq["kps"], seq["nkp"], seq["poses"], seq["R0"], seq["W"])
    dt = time.time() - t0
    out = os.path.join(ROOT, "tests", "golden", f"c3_oracle_{n}.npz")
    np.savez_compressed(out, poses=o["poses"], R0=o["R0"], W=o["W"], p=o["p"], win_iters=o["win_iters"], win_term=o["win_term"],
                        nfail=np.int32(o["nfail"]), seed=np.int32(SEED), it1=np.int32(IT1), it2=np.int32(IT2), oracle_seconds=np.float64(dt))
    print(f"{out}: {n} frames in {dt:.1f} s; stage-1 iterations {o['win_iters'][0]}, window iterations mean "
          f"{o['win_iters'][1:1 + len(range(0, n, 15))].mean():.1f}, nfail {o['nfail']}")


if __name__ == "__main__":
    main()
