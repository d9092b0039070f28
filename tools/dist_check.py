"""Multi-GPU check of the sharded 3dba_multi schedule, run under torchrun with one rank per GPU (NCCL):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/dist_check.py
Every rank fits the same synthetic sequence with dist.fit_sequence_sharded (halo hand-off over NCCL send/recv) and rank 0
compares the assembled result with b3d_fit_sequences(num_shards = world) computed on its own GPU: they must be bit-identical."""
import importlib
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = lambda n: importlib.import_module("3dbodyanimation_b200." + n)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    synth, api, D = pkg("synth"), pkg("api"), pkg("dist")
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    model_np = synth.make_model()
    gm = api.AvatarModel(model_np)
    seq = synth.make_sequence(model_np, n, 31)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    base, so = api.fit_opts(cam=cam), api.seq_opts(it1=40, it2=60, num_shards=world, beta_pose=20.0, beta_shape=30.0)
    st = D.make_state(seq["kps"], seq["nkp"], seq["poses"], seq["R0"], seq["W"], dev)
    plan = D.WindowPlan(n, so.wsize, so.overlap, so.skip, world)
    solver = D.GpuWindowSolver(gm, base, so)
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = D.fit_sequence_sharded(solver, st, plan)
    torch.cuda.synchronize(); dist.barrier()
    dt = time.perf_counter() - t0
    if rank == 0:
        t1 = time.perf_counter()
        g = api.fit_sequences(gm, base, so, seq["kps"], seq["nkp"], seq["poses"], seq["R0"], seq["W"])
        dt_lib = time.perf_counter() - t1
        ok = True
        for k in ("poses", "R0", "W", "p", "r"):
            same = np.array_equal(out[k].cpu().numpy().reshape(g[k].shape), g[k])
            diff = float(np.max(np.abs(out[k].cpu().numpy().reshape(g[k].shape) - g[k])))
            print(f"{k}: bit-identical={same} max|diff|={diff:.3e}")
            ok = ok and same
        print(f"world={world} frames={n} windows={len(plan.starts)} sharded over NCCL: {dt * 1e3:.1f} ms; one-GPU library schedule: {dt_lib * 1e3:.1f} ms")
        print("DIST_CHECK", "OK" if ok else "MISMATCH")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
