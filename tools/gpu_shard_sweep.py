"""Debug aid: the rank-local problem of an 8-GPU run (Ma = B / 8 rows against all B) on ONE GPU, swept over the column
split count (PCLIP_NSPLIT) of the forward and gradient kernels."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth
B, P = 65536, int(sys.argv[1]) if len(sys.argv) > 1 else 8
Ma = B // P
x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=4)
xd, yd = x.cuda(), y.cuda()
full = ops.ClipPlan(B, B, 128, ops.BF16, "cuda")
full.pack([xd], [yd])
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for ns in [0, 1, 2, 3, 4, 5, 7, 9, 13, 18]:
    if ns: os.environ["PCLIP_NSPLIT"] = str(ns)
    p = ops.ClipPlan(Ma, B, 128, ops.BF16, "cuda", diag_offset=0)
    p.pack([xd[:Ma]], [yd[:Ma]]); p.xf.copy_(full.xl); p.yf.copy_(full.yl)
    p.sumexp(); stats = torch.stack([p.stats_l] * 1)
    p.colsum_full = None
    tf = t(p.sumexp)
    p.rinv_f.fill_(1e-5); p.cinv_f.fill_(1e-5)
    tg = t(lambda: p.grad(B))
    print(f"P={P} nsplit={'auto' if not ns else ns}: fwd {tf:.4f} ms  grad {tg:.4f} ms", flush=True)
    del p
