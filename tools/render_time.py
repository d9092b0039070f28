"""Times b3d_render_overlay_dev on resident clouds (CUDA events).  python tools/render_time.py [B]"""
import ctypes as C, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
api = importlib.import_module("3dbodyanimation_b200.api"); synth = importlib.import_module("3dbodyanimation_b200.synth")
from importlib import import_module
capi = import_module("3dbodyanimation_b200._capi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
smooth = len(sys.argv) <= 2 or sys.argv[2] != "spiky"
m = synth.make_render_model() if smooth else synth.make_model()
faces = m["f"] if smooth else synth.make_mesh_faces(m); gm = api.AvatarModel(dict(m, f=faces))
fi = synth.make_forward_inputs(B, 3)
cloud, _, _ = api.smpl_forward(gm, fi["w"], fi["p"], fi["r"], dtype=np.float32)
W, H = 480, 270
dc = torch.from_numpy(cloud).cuda(); img = torch.zeros((B, H, W, 3), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def run():
    capi.check(capi.lib().b3d_render_overlay_dev(gm.h, B, 0, C.c_void_p(dc.data_ptr()), C.c_double(432.0), C.c_double(432.0), C.c_double(240.0),
                                                 C.c_double(135.0), W, H, 0, C.c_void_p(img.data_ptr()), C.c_void_p(st)), "render")
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
cov = float((img[:64] != 0).any(dim=3).float().mean())
print(f"render {'smooth' if smooth else 'spiky'} B={B}: {ms:.3f} ms per call, {B / ms * 1e3:.0f} bodies/s, {ms / B * 1e3:.2f} us/body, coverage {cov:.3f} of the frame")
