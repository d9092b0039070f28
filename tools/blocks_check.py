"""Diagnostic: forward parity of every float32 skinning kernel on the block-rigged synthetic model (synth.make_model(rig='blocks'))."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dbodyanimation_b200.synth")
api = importlib.import_module("3dbodyanimation_b200.api")
from oracle import oracle_api as orc


def main():
    for rig in ("cyclic", "blocks"):
        m = synth.make_model(rig=rig)
        gm, om = api.AvatarModel(m), orc.Model(m)
        B = 640 + 45
        inp = synth.make_forward_inputs(B, 77)
        w, p, r = (np.asarray(inp[k], dtype=np.float32).astype(np.float64) for k in ("w", "p", "r"))
        c0, jp0, jt0 = om.forward(w, p, r, nthreads=8)
        for lbs in ("tc", "sorted", "bodylane", "generic"):
            c, jp, jt = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs=lbs)
            d = np.abs(c.astype(np.float64) - c0)
            scale = np.max(np.abs(c0), axis=(1, 2))
            per_body = d.max(axis=(1, 2)) / scale
            bb = int(np.argmax(per_body))
            vv = int(np.argmax(d[bb].max(axis=1)))
            bad = (d.max(axis=2) / scale[:, None] > 1e-5)
            print(f"{rig:7s} {lbs:9s} worst rel {per_body.max():.3e} (body {bb}, vertex {vv}); entries over 1e-5: {int(bad.sum())} of {bad.size}"
                  f"; bad vertices (first 12): {np.unique(np.nonzero(bad)[1])[:12].tolist()}; jp rel {np.abs(jp - jp0).max() / np.abs(jp0).max():.2e}"
                  f" jt rel {np.abs(jt - jt0).max() / np.abs(jt0).max():.2e}")


if __name__ == "__main__":
    main()
