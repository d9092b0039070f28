"""Grid kernel (fit_chain.cuh, block cyclic reduction) against the cluster kernel on the same chains: python tools/chain_check.py [F ...]"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = lambda n="": importlib.import_module("3dbodyanimation_b200" + ("." + n if n else ""))
synth, api = pkg("synth"), pkg("api")
lib = api.lib()
model_np = synth.make_model()
gm = api.AvatarModel(model_np)
cam = (synth.FX, synth.FY, synth.CX, synth.CY)
for F in [int(a) for a in sys.argv[1:]] or [48, 61, 200]:
    seq = synth.make_sequence_fast(model_np, F * 10, 31) if F > 400 else synth.make_sequence(model_np, F * 10, 31)
    idx = np.arange(0, F * 10, 10)
    o = api.fit_opts(cam=cam, mode=1, max_iters=30, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
    args = (seq["kps"][idx], seq["nkp"][idx], seq["R0"][idx], seq["poses"][idx], seq["W"][0])
    res = {}
    for name, mf in (("cluster", 10 ** 9), ("grid", 0)):
        lib.b3d_debug_chain_min_frames(mf)
        api.fit_window(gm, o, *args)
        t0 = time.perf_counter()
        res[name] = api.fit_window(gm, o, *args)
        res[name]["dt"] = time.perf_counter() - t0
    a, b = res["cluster"], res["grid"]
    n = min(a["niters"], b["niters"])
    dc = np.max(np.abs(a["trace"]["cost"][:n] - b["trace"]["cost"][:n]) / a["trace"]["cost"][:n])
    print(f"F={F}: iterations {a['niters']} / {b['niters']}, term {a['term']} / {b['term']}, max rel cost diff {dc:.2e}, "
          f"max |d poses| {np.max(np.abs(a['poses'] - b['poses'])):.2e}, |d w| {np.max(np.abs(a['w'] - b['w'])):.2e}; "
          f"cluster {a['dt'] * 1e3:.1f} ms, grid {b['dt'] * 1e3:.1f} ms (host call incl. copies)")
lib.b3d_debug_chain_min_frames(48)
