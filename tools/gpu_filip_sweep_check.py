"""Single-sweep FILIP forward (filip_tc1.cu) against the fp32 SIMT path in one process: sims, stored maximisers.
python tools/gpu_filip_sweep_check.py   (debug aid for the GPU box; prints where the maximisers differ)"""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "protein-clip_b200"))
import torch
from pclip import ops

dev = torch.device("cuda", 0)
torch.manual_seed(0)
cases = [(4, 4, 96, 160, False), (3, 5, 512, 512, True), (5, 3, 300, 77, True), (2, 2, 128, 128, False),
         (6, 6, 512, 300, True), (2, 7, 33, 640, True), (2, 3, 640, 200, True), (16, 16, 512, 512, False)]
bad = 0
for Ba, Bb, TA, TB, ragged in cases:
    hA = torch.nn.functional.normalize(torch.randn(Ba, TA, 128, device=dev), dim=-1)
    hB = torch.nn.functional.normalize(torch.randn(Bb, TB, 128, device=dev), dim=-1)
    # the bf16 path rounds the tokens: feed both paths the rounded values so that the maximisers are comparable
    hA, hB = hA.bfloat16().float(), hB.bfloat16().float()
    lenA = torch.full((Ba,), TA, dtype=torch.int32, device=dev)
    lenB = torch.full((Bb,), TB, dtype=torch.int32, device=dev)
    if ragged:
        lenA = torch.randint(1, TA + 1, (Ba,), dtype=torch.int32, device=dev)
        lenB = torch.randint(1, TB + 1, (Bb,), dtype=torch.int32, device=dev)
        lenA[0], lenB[0] = TA, TB
        if Ba > 2:
            lenA[1] = 0
    tau = torch.tensor([0.07], device=dev)
    out = {}
    for name, mode in (("f32", ops.F32), ("bf16", ops.BF16)):
        out[name] = ops.filip_similarity_raw(hA, hB, lenA, lenB, tau, mode)
    torch.cuda.synchronize()
    sa0, sb0, aa0, ab0 = out["f32"]
    sa1, sb1, aa1, ab1 = out["bf16"]
    fin = torch.isfinite(sa0) & torch.isfinite(sb0)
    ea = float(((sa0 - sa1).abs()[fin]).max() / sa0[fin].abs().max())
    eb = float(((sb0 - sb1).abs()[fin]).max() / sb0[fin].abs().max())
    tA = torch.arange(TA, device=dev)[None, None, :] < lenA[:, None, None]
    tA = tA & (lenB[None, :, None] > 0)
    tB = torch.arange(TB, device=dev)[None, None, :] < lenB[:, None, None]
    tB = tB & (lenA[None, :, None] > 0)
    da = ((aa0 != aa1) & tA).sum().item(), int(tA.sum())
    db = ((ab0 != ab1) & tB).sum().item(), int(tB.sum())
    nonfin = int((~torch.isfinite(sa1)).sum() - (~torch.isfinite(sa0)).sum()), int((~torch.isfinite(sb1)).sum() - (~torch.isfinite(sb0)).sum())
    print(f"Ba={Ba} Bb={Bb} TA={TA} TB={TB} ragged={ragged}: simA err {ea:.2e} simB err {eb:.2e} "
          f"argA diff {da[0]}/{da[1]} argB diff {db[0]}/{db[1]} nonfinite delta {nonfin}", flush=True)
    if ea > 2e-3 or eb > 2e-3 or da[0] > 0.002 * da[1] + 2 or db[0] > 0.002 * db[1] + 2:
        bad += 1
        for nm, x0, x1, m in (("argA", aa0, aa1, tA), ("argB", ab0, ab1, tB)):
            idx = ((x0 != x1) & m).nonzero()[:12]
            for i in idx.tolist():
                print("   ", nm, i, int(x0[tuple(i)]), int(x1[tuple(i)]))
print("BAD CASES", bad)
sys.exit(1 if bad else 0)
