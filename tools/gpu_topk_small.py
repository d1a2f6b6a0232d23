"""Small-batch top-k (Q = 128 against the 2^22-row prepared bank): time per call. python tools/gpu_topk_small.py [Q]"""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "protein-clip_b200"))
import torch
from pclip import ops, synth
Q = int(sys.argv[1]) if len(sys.argv) > 1 else 128
N = 1 << 22
bank, q, pos = synth.retrieval_bank(N, Q, cfg_id=5, device="cuda")
plan = ops.TopkPlan(Q, N, 128, 10, ops.BF16, "cuda").prepare_bank(bank)
for _ in range(3):
    plan.run(q, bank)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    plan.run(q, bank)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
i32, s32, _ = ops.topk(q, bank, 10, compute=ops.F32)
print(f"Q={Q} ms/call {ms:.4f}  {N*128*2/ms/1e6:.1f} GB/s  top1 ok {float((plan.idx[:,0]==pos).float().mean()):.3f} "
      f"equal fp32 {bool(torch.equal(plan.idx, i32) and torch.equal(plan.score, s32))} unproven {int(plan.n_unsafe)}")
