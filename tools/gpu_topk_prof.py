"""Debug aid: one top-k pass at reduced Q (for ncu): python tools/gpu_topk_prof.py [Q]"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth
Q = int(sys.argv[1]) if len(sys.argv) > 1 else 37888
N = 1 << 22
bank, q, pos = synth.retrieval_bank(N, Q, cfg_id=5, device="cuda")
plan = ops.TopkPlan(Q, N, 128, 10, ops.BF16, "cuda")
for _ in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); plan.run(q, bank); e1.record(); torch.cuda.synchronize()
    print("Q", Q, "ms", e0.elapsed_time(e1), "queries/s", Q / e0.elapsed_time(e1) * 1e3)
