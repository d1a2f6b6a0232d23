"""Debug aid: pipeline timeline of one CTA of the CLIP forward kernel (PCLIP_FWD_TRACE=file)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth
B = 65536
x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=4)
plan = ops.ClipPlan(B, B, 128, ops.BF16, "cuda")
plan.step([x.cuda()], [y.cuda()]); torch.cuda.synchronize()
print(open(os.environ["PCLIP_FWD_TRACE"]).read())
