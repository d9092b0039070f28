"""C3 timing alone: python tools/c3_time.py [frames]  (same call as bench.py's c3 leg, no CPU baseline)."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import importlib
api = importlib.import_module("3dbodyanimation_b200.api")
import torch
torch.cuda.init()
a = argparse.Namespace(c3_frames=int(sys.argv[1]) if len(sys.argv) > 1 else 2000)
r = bench.c3_leg(a, api, with_cpu=False)
print(json.dumps({k: r[k] for k in ("frames", "frames_per_sec", "ms", "stage1_lm_iterations", "window_lm_iterations_mean", "failed_solves")}))
