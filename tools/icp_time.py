"""Times b3d_icp_eval_dev (the fused ICP cost + gradient pass, csrc/icp.cu) on the current device.
usage: python tools/icp_time.py [B ...]    (default 4096 16384)"""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dbodyanimation_b200.synth")
api = importlib.import_module("3dbodyanimation_b200.api")


def tiled_inputs(model_np, b, seed=3, base=32):
    x = synth.make_icp_inputs(model_np, base, seed)
    reps = (b + base - 1) // base
    npair, ndata = int(x["cptr"][-1]), int(x["dptr"][-1])
    cptr = np.concatenate([x["cptr"][:-1] + r * npair for r in range(reps)] + [[reps * npair]]).astype(np.int64)
    dptr = np.concatenate([x["dptr"][:-1] + r * ndata for r in range(reps)] + [[reps * ndata]]).astype(np.int64)
    nb = reps * base
    out = dict(cptr=cptr[:b + 1], dptr=dptr[:b + 1], mpt=np.tile(x["mpt"], reps)[:cptr[b]], dpt=np.tile(x["dpt"], reps)[:cptr[b]],
               data=np.tile(x["data"].astype(np.float32), (reps, 1))[:dptr[b]])
    for k in ("w", "p", "q"):
        out[k] = np.tile(x[k], (reps,) + (1,) * (x[k].ndim - 1))[:b]
    return out


def time_eval(model, x, dtype=np.float32, iters=10):
    prob = api.IcpProblem(model, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], dtype=dtype)
    b = prob.B
    dev = torch.device("cuda:0")
    w, p, q = (torch.from_numpy(np.ascontiguousarray(x[k], dtype=np.float64)).to(dev) for k in ("w", "p", "q"))
    cost, grad = torch.empty(b, dtype=torch.float64, device=dev), torch.empty(b, 85, dtype=torch.float64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    call = lambda: api.check(api.lib().b3d_icp_eval_dev(prob.h, C.c_void_p(w.data_ptr()), C.c_void_p(p.data_ptr()), C.c_void_p(q.data_ptr()),
                                                      0, C.c_void_p(cost.data_ptr()), C.c_void_p(grad.data_ptr()), C.c_void_p(st)), "eval")
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        call()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, prob, cost.cpu().numpy()


if __name__ == "__main__":
    model_np = synth.make_model()
    model = api.AvatarModel(model_np)
    for b in [int(a) for a in sys.argv[1:]] or [4096, 16384]:
        x = tiled_inputs(model_np, b)
        ms, _, cost = time_eval(model, x)
        v = model_np["v_template"].shape[0]
        gbs = b * v * 16 / ms / 1e6
        print(f"B={b}: {ms:.3f} ms per evaluation, {b / ms * 1e3:.0f} bodies/s, {b * v / ms * 1e3:.3e} vertices/s, "
              f"{gbs:.0f} GB/s of target records (16 B per vertex and body), pairs {int(x['cptr'][-1])}, cost[0] {cost[0]:.6f}")
