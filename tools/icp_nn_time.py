"""Calls b3d_icp_nearest once per configuration on resident inputs (run under `ncu --metrics gpu__time_duration.sum` for the kernel
time; the wall clock printed here includes the call's allocations and table uploads).  usage: python tools/icp_nn_time.py [bodies]"""
import ctypes as C
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dbodyanimation_b200.synth")
api = importlib.import_module("3dbodyanimation_b200.api")
capi = importlib.import_module("3dbodyanimation_b200._capi")

if __name__ == "__main__":
    nb = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    model_np = synth.make_model()
    gm = api.AvatarModel(model_np)
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(1)
    V = synth.V
    fin = synth.make_forward_inputs(32, 3)
    cloud32 = api.smpl_forward(gm, fin["w"], fin["p"], fin["r"], dtype=np.float32)[0]
    cloud = torch.from_numpy(np.tile(cloud32, (nb // 32, 1, 1))).to(dev)
    P = lambda t: C.c_void_p(t.data_ptr())
    for nparts, coherent in ((16, False), (16, True), (1, True)):
        model_part = (np.argmax(np.asarray(model_np["weights"]), axis=1) % nparts).astype(np.int32)
        src = rng.integers(0, V, (nb, V))
        data = torch.from_numpy((np.tile(cloud32, (nb // 32, 1, 1))[np.arange(nb)[:, None], src] + rng.normal(0, 0.02, (nb, V, 3))).astype(np.float32)).to(dev)
        dpart_np = model_part[src] if coherent else rng.integers(0, nparts, (nb, V))
        if coherent:   # neighbouring data points share a part, as in a depth image
            order = np.argsort(dpart_np, axis=1, kind="stable")
            dpart_np = np.take_along_axis(dpart_np, order, axis=1)
            data = torch.from_numpy(np.take_along_axis(data.cpu().numpy(), order[:, :, None], axis=1)).to(dev)
        dpart = torch.from_numpy(np.ascontiguousarray(dpart_np.astype(np.int32))).to(dev)
        out = torch.empty(nb * V, dtype=torch.int32, device=dev)
        dptr = np.arange(nb + 1, dtype=np.int64) * V
        call = lambda: capi.check(capi.lib().b3d_icp_nearest(gm.h, nb, capi.F32, P(cloud), nparts, capi.ptr(model_part), None, capi.ptr(dptr), P(data),
                                                             P(dpart), 1, P(out)), "nearest")
        call()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        call()
        ms = (time.perf_counter() - t0) * 1e3
        print(f"bodies {nb} parts {nparts} coherent {coherent}: {ms:.3f} ms per call (wall), matched {(out >= 0).float().mean().item():.3f}")
