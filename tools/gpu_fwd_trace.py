import sys, os
sys.path.insert(0, "protein-clip_b200")
import torch
from pclip import ops, synth
B = 65536
x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=4)
plan = ops.ClipPlan(B, B, 128, ops.BF16, "cuda")
plan.xl.copy_(x.cuda()); plan.yl.copy_(y.cuda())
plan.sumexp(); torch.cuda.synchronize()
print(open(os.environ["PCLIP_FWD_TRACE"]).read())
