"""C5 leg alone (one long 3dba_multi sequence, sharded windows, NCCL halo): single process or under torchrun.
   python tools/c5_time.py [frames] [shards_per_rank]"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
pkg = lambda n: importlib.import_module("3dbodyanimation_b200." + n)
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
synth, api, D = pkg("synth"), pkg("api"), pkg("dist")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
spr = int(sys.argv[2]) if len(sys.argv) > 2 else 74
model_np = synth.make_model()
seq = synth.make_sequence_fast(model_np, n, 77)
gm = api.AvatarModel(model_np)
base = api.fit_opts(cam=(synth.FX, synth.FY, synth.CX, synth.CY))
so = api.seq_opts(skip=10, wsize=20, overlap=5, it1=1000, it2=60, num_shards=world, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
r = D.bench_c5(api, gm, base, so, seq, dev, rank, world, shards_per_rank=spr)
if rank == 0:
    print(json.dumps({k: v for k, v in r.items() if k not in ("workload", "schedule_note", "stage1", "timing", "collective")}))
if world > 1:
    dist.destroy_process_group()
