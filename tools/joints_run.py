import sys, importlib, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
synth = importlib.import_module("3dbodyanimation_b200.synth"); api = importlib.import_module("3dbodyanimation_b200.api")
gm = api.AvatarModel(synth.make_model())
fin = synth.make_forward_inputs(16384, 5)
for _ in range(2):
    api.smpl_forward(gm, fin["w"], fin["p"], fin["r"], dtype=np.float32, want_cloud=False)
