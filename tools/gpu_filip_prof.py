"""Debug aid: FILIP fwd+bwd steps at B = 256, T = 512 (for an ncu launch list): python tools/gpu_filip_prof.py [ragged]"""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "protein-clip_b200"))
import torch
import bench
d = bench.Dist(1)
w = bench.FilipB256(d, ragged=len(sys.argv) > 1)
for _ in range(3):
    w.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); w.step(); e1.record(); torch.cuda.synchronize()
print("filip step ms", e0.elapsed_time(e1))
