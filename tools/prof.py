"""Small drivers for ncu captures and quick timings: `python tools/prof.py forward [B]`, `python tools/prof.py fit [frames]`."""
import ctypes as C
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = lambda n="": importlib.import_module("3dbodyanimation_b200" + ("." + n if n else ""))
synth, api, capi = pkg("synth"), pkg("api"), pkg("_capi")
lib = capi.lib()
dev = torch.device("cuda", 0)
P = lambda t: C.c_void_p(t.data_ptr())


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.mean(ts)), float(np.min(ts))


def forward(B, dtype="f32", reps=5, lbs=None, posedirs=False):
    gm = api.AvatarModel(synth.make_model(with_posedirs=posedirs))
    fin = synth.make_forward_inputs(B, 5)
    td = torch.float32 if dtype == "f32" else torch.float64
    code = capi.F32 if dtype == "f32" else capi.F64
    w, p, r = (torch.from_numpy(fin[k]).to(td).to(dev) for k in ("w", "p", "r"))
    cloud = torch.empty(B, gm.numPoints, 3, dtype=td, device=dev)
    jp, jt = torch.empty(B, 24, 3, dtype=td, device=dev), torch.empty(B, 24, 12, dtype=td, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    fn = lambda: capi.check(lib.b3d_smpl_forward_dev(gm.h, B, code, P(w), P(p), P(r), P(cloud), P(jp), P(jt), api.LBS_FLAGS[lbs] | (1 if posedirs else 0), C.c_void_p(st)), "fwd")
    mean, best = timed(fn, reps)
    es = 4 if dtype == "f32" else 8
    gb = B * (85036 * es / 4) / 1e9
    print(f"forward {dtype} lbs={lbs or 'auto'}{' +posedirs' if posedirs else ''} B={B}: {mean:.3f} ms (best {best:.3f}) -> {gb / mean * 1e3:.0f} GB/s algorithmic, {B * gm.numPoints / mean * 1e3:.3e} verts/s")


def fit(n, mode="C2", reps=3):
    model_np = synth.make_model()
    prior = synth.load_pose_prior(os.path.join(ROOT, "tests", "golden", "pose_prior.txt"))
    gm, gg = api.AvatarModel(model_np), api.GaussianMixture(prior=prior)
    if mode == "C2":
        fr = synth.make_frames(model_np, n, 1000, shape=True, prior=prior)
        opts, g = api.fit_opts(cam=(synth.FX, synth.FY, synth.CX, synth.CY), mode=1, max_iters=100, beta_pose=20.0, beta_shape=30.0), gg.h
    else:
        fr = synth.make_frames(model_np, n, 1000, shape=False)
        opts, g = api.fit_opts(cam=(synth.FX, synth.FY, synth.CX, synth.CY), mode=0, max_iters=100, beta_pose=20.0, beta_shape=30.0), None
    kps = torch.from_numpy(np.ascontiguousarray(fr["kps"]).view(np.uint8).reshape(n, -1)).to(dev)
    nkp, R0, s0, w0 = (torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in ("nkp", "R0", "sim3", "w"))
    s3, w = s0.clone(), w0.clone()
    jaa = torch.zeros(n, 24, 3, dtype=torch.float64, device=dev)
    us = torch.zeros(n, dtype=torch.uint8, device=dev)
    ni, te = torch.zeros(n, dtype=torch.int32, device=dev), torch.zeros(n, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def fn():
        s3.copy_(s0); w.copy_(w0)
        capi.check(lib.b3d_fit_single_dev(gm.h, g, C.byref(opts), n, P(kps), P(nkp), P(R0), P(s3), P(w), P(jaa), P(us), None, 0, P(ni), P(te),
                                          C.c_void_p(st)), "fit")
    mean, best = timed(fn, reps)
    it = ni.cpu().numpy()
    global LAST_NI
    LAST_NI = it
    print(f"fit {mode} frames={n}: {mean:.3f} ms (best {best:.3f}) -> {n / mean * 1e3:.0f} frames/s; LM iterations mean {it.mean():.1f} max {it.max()} "
          f"sum {it.sum()}; {mean * 1e3 / max(it.max(), 1):.1f} us per iteration of the slowest frame; usable {int(us.sum())}")


if __name__ == "__main__":
    what = sys.argv[1]
    if what == "forward":
        forward(int(sys.argv[2]) if len(sys.argv) > 2 else 16384, sys.argv[3] if len(sys.argv) > 3 else "f32",
                lbs=sys.argv[4] if len(sys.argv) > 4 else None)
    elif what == "pforward":   # with pose blendshapes
        forward(int(sys.argv[2]) if len(sys.argv) > 2 else 16384, "f32", lbs=sys.argv[3] if len(sys.argv) > 3 else None, posedirs=True)
    elif what == "fsweep":   # float32 forward sweep per kernel
        for lbs in (sys.argv[2:] or ["tc", "sorted"]):
            for b in (128, 512, 1024, 4096, 16384, 65536):
                forward(b, "f32", lbs=lbs)
    elif what == "fit":
        fit(int(sys.argv[2]) if len(sys.argv) > 2 else 1000, sys.argv[3] if len(sys.argv) > 3 else "C2")
    elif what == "sweep":
        for b in (1, 16, 256, 1024, 4096, 16384, 65536):
            forward(b, "f32")
        for b in (1024, 16384):
            forward(b, "f64")
        for n in (100, 1000, 8000):
            fit(n, "C2")
        fit(100, "C1"); fit(1000, "C1")


def phases(n=148, mode="C2"):
    """Per-phase cycle counts of block 0 (debug hook b3d_debug_phase_timing)."""
    names = ["loop head", "build A", "cholesky", "backsolve", "mcc", "delta+candidate", "cand evaluate", "accept copy", "eval_grad_jac total",
             "  JAC reproj", "  JAC gmm+cost", "  JAC assemble", "  cand reproj", "  cand gmm+cost", "", ""]
    lib.b3d_debug_phase_timing(1)
    lib.b3d_debug_fit_stop_at(0)          # one CTA per frame: the phase counters follow CTA 0 = frame 0
    fit(n, mode, reps=1)
    lib.b3d_debug_fit_stop_at(-1)
    out = (C.c_longlong * 32)()
    lib.b3d_debug_phase_read(out)
    tot = sum(out[i] for i in range(9))
    names += [""] * 16
    names[16:20] = ["    gmm: d = x - mu", "    gmm: h = L^T d (8 comps)", "    gmm: |h|^2, argmin", "    gmm: J^T r (JAC only)"]
    names[20:25] = ["    chol: tile load", "    chol: first factor", "    chol: solve + barrier", "    chol: update + next factor + barrier", "    chol: dump"]
    for i in list(range(14)) + [16, 17, 18, 19, 20, 21, 22, 23, 24]:
        print(f"{names[i]:22s} {out[i]:12d} cycles {out[i] / max(tot, 1) * 100:6.1f}%")
    cyc, ls = (C.c_longlong * 1024)(), (C.c_int * 1024)()
    lib.b3d_debug_block_read(cyc, ls)
    m = min(n, 1024)
    cy, lsv, ni = np.array(cyc[:m], dtype=np.float64), np.array(ls[:m]), LAST_NI[:m].astype(np.float64)
    per = cy / ni
    print("frame0 iters", int(ni[0]), "cycles", int(cy[0]))
    print("cycles per iteration: min %.0f median %.0f max %.0f" % (per.min(), np.median(per), per.max()))
    order = np.argsort(-cy)[:8]
    for b in order:
        print(f"  block {b}: {cy[b] / 1e6:.2f} Mcycles, iters {int(ni[b])}, per-iter {per[b]:.0f}, line-search fallbacks {lsv[b]}")
    print("frames with line-search fallbacks:", int((lsv > 0).sum()), "total fallbacks", int(lsv.sum()))
    lc, lt = (C.c_longlong * 1024)(), (C.c_int * 1024)()
    lib.b3d_debug_block_ls_read(lc, lt)
    lc, lt = np.array(lc[:m], dtype=np.float64), np.array(lt[:m])
    print(f"line-search fall-backs: {lc.sum() / 1e6:.1f} Mcycles of {cy.sum() / 1e6:.1f} ({100 * lc.sum() / cy.sum():.1f} %), {int(lt.sum())} trial evaluations, "
          f"{lc.sum() / max(lsv.sum(), 1):.0f} cycles per fall-back, {lt.sum() / max(lsv.sum(), 1):.2f} trials per fall-back")
    lp = (C.c_longlong * 4096)()
    lib.b3d_debug_block_lsparts_read(lp)
    lp = np.array(lp[:4 * m], dtype=np.float64).reshape(m, 4).sum(axis=0)
    print("  of which set-up %.1f, trial evaluations %.1f, polynomial minimisation %.1f, final evaluation %.1f Mcycles" % tuple(lp / 1e6))
    pc = (C.c_ulonglong * 8)()
    lib.b3d_debug_poly_read(pc)
    print("  polynomial minimisation (all calls since load): %d calls, per call interpolation %.0f, end points %.0f, roots %.0f cycles"
          % (pc[3], pc[0] / max(pc[3], 1), pc[1] / max(pc[3], 1), pc[2] / max(pc[3], 1)))
    print("  root refinements %d, iterations per refinement %.1f" % (pc[4], pc[5] / max(pc[4], 1)))
    t0, t1 = (C.c_longlong * 1024)(), (C.c_longlong * 1024)()
    lib.b3d_debug_block_times(t0, t1)
    t0, t1 = np.array(t0[:m], dtype=np.float64), np.array(t1[:m], dtype=np.float64)
    base = t0.min()
    print("kernel span %.3f ms (first frame start to last frame end)" % ((t1.max() - base) * 1e-6))
    for b in order:
        print(f"  block {b}: started at {(t0[b] - base) * 1e-6:.3f} ms, ended at {(t1[b] - base) * 1e-6:.3f} ms, busy {cy[b] / 1e6:.2f} Mcycles "
              f"-> {cy[b] / max(t1[b] - t0[b], 1.0):.3f} cycles/ns if it ran without waiting")
    late = np.argsort(-t1)[:5]
    print("last to finish:", [(int(b), round((t0[b] - base) * 1e-6, 3), round((t1[b] - base) * 1e-6, 3), int(ni[b])) for b in late])
    lib.b3d_debug_phase_timing(0)


def passes(n=1000):
    """Timing of the persistent kernel's frame queue for several checkpoint caps (0 = one CTA per frame)."""
    for cap in (0, 8, 12, 16, 20, 24, 30, 40):
        lib.b3d_debug_fit_stop_at(cap)
        print("stop_at", cap, end=": ")
        fit(n, "C2", reps=5)
    lib.b3d_debug_fit_stop_at(-1)


def winphases(F=20, iters=60):
    """Per-phase cycles of one OptimizeMultiFrame solve (block 0 of fit_window_kernel)."""
    model_np = synth.make_model()
    gm = api.AvatarModel(model_np)
    seq = synth.make_sequence(model_np, F, 31)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    o = api.fit_opts(cam=cam, max_iters=iters, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
    api.fit_window(gm, o, seq["kps"], seq["nkp"], seq["R0"], seq["poses"], seq["W"][0])
    lib.b3d_debug_win_timing(1)
    t0 = time.perf_counter()
    r = api.fit_window(gm, o, seq["kps"], seq["nkp"], seq["R0"], seq["poses"], seq["W"][0])
    dt = time.perf_counter() - t0
    out = (C.c_longlong * 16)()
    lib.b3d_debug_win_read(out)
    lib.b3d_debug_win_timing(0)
    names = ["kernel: other (diag, tolerances)", "kernel: Jacobi scaling + gradient norm after eval", "kernel: solve total", "kernel: model cost change + candidate",
             "eval: reproj (cost only)", "eval: frame load", "eval: reproj + Jacobian", "eval: 86 x 86 assembly", "eval: scatter to blocks", "solve: load rows",
             "solve: products with previous block row", "solve: Crout sweep", "solve: store + Schur", "solve: arrow block + back sweep", "kernel: candidate evaluation total", ""]
    tot = out[0] + out[1] + out[2] + out[3] + out[14]
    print(f"fit_window F={F}: {r['niters']} records, {dt * 1e3:.2f} ms wall for the call; {tot / max(r['niters'] - 1, 1) / F:.0f} cycles per frame-iteration")
    for i in range(15):
        print(f"  {names[i]:52s} {out[i]:12d} {100.0 * out[i] / max(tot, 1):6.1f}%")


def chainphases(F=200, iters=30):
    """Per-phase cycles of CTA 0 of fit_chain_kernel (grid kernel for long chains) on a chain of F anchors."""
    model_np = synth.make_model()
    gm = api.AvatarModel(model_np)
    seq = synth.make_sequence_fast(model_np, F * 10, 31)
    idx = np.arange(0, F * 10, 10)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    o = api.fit_opts(cam=cam, max_iters=iters, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
    args = (seq["kps"][idx], seq["nkp"][idx], seq["R0"][idx], seq["poses"][idx], seq["W"][0])
    lib.b3d_debug_chain_min_frames(0)
    api.fit_window(gm, o, *args)
    lib.b3d_debug_win_timing(1)
    t0 = time.perf_counter()
    r = api.fit_window(gm, o, *args)
    dt = time.perf_counter() - t0
    out = (C.c_longlong * 16)()
    lib.b3d_debug_win_read(out)
    lib.b3d_debug_win_timing(0)
    names = ["solve: working copies", "solve p1: load blocks", "solve p1: Cholesky", "solve p1: triangular solves", "eval: reproj (cost only)", "eval: frame load",
             "eval: reproj + Jacobian", "eval: 86 x 86 assembly", "eval: scatter to blocks", "solve p1: stores + new coupling", "solve: grid barrier waits",
             "solve p2: Schur updates of the survivors", "solve: root", "solve: back substitution", "solve: shape block + y", "kernel total (LM loop)"]
    tot = max(out[15], 1)
    print(f"fit_chain F={F}: {r['niters']} records, {dt * 1e3:.2f} ms wall for the call; {tot / max(r['niters'] - 1, 1):.0f} cycles per LM iteration")
    for i in range(16):
        print(f"  {names[i]:44s} {out[i]:12d} {100.0 * out[i] / tot:6.1f}%")


if __name__ == "__main__" and sys.argv[1] == "chainphases":
    chainphases(int(sys.argv[2]) if len(sys.argv) > 2 else 200)

if __name__ == "__main__" and sys.argv[1] == "winphases":
    winphases(int(sys.argv[2]) if len(sys.argv) > 2 else 20)

if __name__ == "__main__" and sys.argv[1] == "passes":
    passes(int(sys.argv[2]) if len(sys.argv) > 2 else 1000)

if __name__ == "__main__" and sys.argv[1] == "phases":
    phases(int(sys.argv[2]) if len(sys.argv) > 2 else 148, sys.argv[3] if len(sys.argv) > 3 else "C2")


def tctrace(B=16384):
    """Timeline of the tensor-core LBS kernel's CTA (0, 0): per warp and MMA tile, cycles spent in each phase."""
    lib.b3d_debug_tc_trace(1)
    forward(B, "f32", reps=1, lbs="tc")
    lib.b3d_debug_tc_trace(0)
    W, T, K = 20, 256, 8
    buf = (C.c_longlong * (W * T * K))()
    lib.b3d_debug_tc_trace_read(buf, W * T * K)
    a = np.array(buf[:], dtype=np.int64).reshape(W, T, K)
    nt = int((a[0, :, 0] > 0).sum())
    a = a[:, :nt]
    t0 = a[:, 0, 0].min()
    print("tiles traced", nt, "total cycles", int(a[:, nt - 1, 4].max() - t0))
    names = ["service+barrier wait", "D load + release", "slot loop", "copy-out (every 3rd tile)"]
    mm = a[17, 1:nt - 1]
    a = a[:16]
    print("MMA issuer (warp 17), cycles after it turns to the tile: waits over %.0f, 4 MMAs issued %.0f, 2 commits issued %.0f"
          % tuple(float((mm[:, k] - mm[:, 0]).mean()) for k in (5, 6, 7)))
    a = a[:, :, :5]
    d = np.diff(a, axis=2).astype(np.float64)       # W x nt x 4
    for k in range(4):
        print(f"{names[k]:28s} mean {d[:, :, k].mean():9.0f}  per-warp means " + " ".join(f"{x:6.0f}" for x in d[:, :, k].mean(axis=1)))
    per_tile = (a[:, 1:, 0] - a[:, :-1, 0]).astype(np.float64)
    print("tile period mean", per_tile.mean(), "; by tile % 3:", [float(per_tile[:, i::3].mean()) for i in range(3)])
    print("start skew of warps at tile 12 (cycles after the first):", (a[:, 12, 0] - a[:, 12, 0].min()).tolist())
    co = d[:, 2::3, 3]
    print("copy-out cycles (tiles 2, 5, ...): mean", co.mean(), "min", co.min(), "max", co.max())


if __name__ == "__main__" and sys.argv[1] == "tctrace":
    tctrace(int(sys.argv[2]) if len(sys.argv) > 2 else 16384)
