"""Prints kernel-vs-torch numbers for the single-pass CLIP forward (pclip_clip_sumexp); debug aid for gpurun."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth

torch.set_printoptions(precision=5, linewidth=200)
bad = 0
QUICK = os.environ.get("QUICK") == "1"
for (Ma, Nb, off) in () if QUICK else ((16, 16, 0), (128, 128, 0), (129, 129, 0), (200, 200, 0), (256, 256, 0), (1000, 1000, 0),
                      (4096, 4096, 0), (8256, 8256, 0), (512, 2048, 512), (300, 1200, 900), (2048, 16384, 6144)):
    x, y = synth.clip_scaled_embeddings(Nb, 128, cfg_id=Nb + 7, trained_like=(Nb % 3 == 0))
    xb, yb = x.bfloat16().cuda(), y.bfloat16().cuda()
    plan = ops.ClipPlan(Ma, Nb, 128, ops.BF16, "cuda", diag_offset=off)
    assert plan.single_pass
    plan.xl.copy_(xb[off:off + Ma]); plan.yl.copy_(yb[off:off + Ma])
    if not plan.local_is_full:
        plan.xf.copy_(xb); plan.yf.copy_(yb)
    plan.sumexp()
    torch.cuda.synchronize()
    s = xb[off:off + Ma].double() @ yb.double().t()
    e = s.exp()
    R, Cc, dg = e.sum(1), e.sum(0), s[torch.arange(Ma), torch.arange(Ma) + off]
    er = float(((plan.rowsum.double() - R) / R).abs().max())
    ec = float(((plan.colsum_f.double() - Cc) / Cc).abs().max())
    ed = float((plan.diag.double() - dg).abs().max())
    flag = "" if (er < 2e-4 and ec < 2e-3 and ed < 1e-4) else "   <<<<<< BAD"
    bad += bool(flag)
    print(f"[Ma={Ma} Nb={Nb} off={off}] rowsum rel {er:.2e} colsum rel {ec:.2e} diag abs {ed:.2e}{flag}")
    if flag:
        print("  rowsum k", plan.rowsum[:6].tolist(), "\n  rowsum r", R[:6].tolist())
        print("  colsum k", plan.colsum_f[:6].tolist(), "\n  colsum r", Cc[:6].tolist())
        print("  colsum k tail", plan.colsum_f[-4:].tolist(), "\n  colsum r tail", Cc[-4:].tolist())

# timing at the bench size
B = 65536
x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=4)
plan = ops.ClipPlan(B, B, 128, ops.BF16, "cuda")
plan.xl.copy_(x.cuda()); plan.yl.copy_(y.cuda())
for name, fn in (("sumexp(single pass)", plan.sumexp), ("rowsumexp(two sided)", plan.rowsumexp)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        fn()
    b.record(); torch.cuda.synchronize()
    print(f"{name}: {a.elapsed_time(b) / 10:.3f} ms at B={B}")
    if name.startswith("sumexp"):
        r1, c1 = plan.rowsum.clone(), plan.colsum.clone()
    else:
        print("  single vs two-sided: rowsum rel", float(((r1 - plan.rowsum) / plan.rowsum).abs().max()),
              "colsum rel", float(((c1 - plan.colsum) / plan.colsum).abs().max()))
print("debug done, bad =", bad)
sys.exit(1 if bad else 0)
