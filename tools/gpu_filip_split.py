"""FILIP fwd / bwd split at B = 256, T = 512 with CUDA events: python tools/gpu_filip_split.py [ragged] [B] [T]
Prints ms of: normalise fwd, similarity fwd (C-ABI pclip_filip_fwd), loss tail, similarity bwd (pclip_filip_bwd), rest."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "protein-clip_b200"))
import torch
from pclip import ops, synth

ragged = len(sys.argv) > 1 and sys.argv[1] == "ragged"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
T = int(sys.argv[3]) if len(sys.argv) > 3 else 512
dev = torch.device("cuda", 0)
hA, hB, mA, mB, tau = synth.filip_tokens(B, T, T, cfg_id=3, ragged=ragged)
hA, hB = hA.to(dev).requires_grad_(), hB.to(dev).requires_grad_()
lenA, lenB = mA.sum(1).to(torch.int32).to(dev), mB.sum(1).to(torch.int32).to(dev)
tau = tau.to(dev).requires_grad_()
pair_tokens = float((mA.sum(1).double()[:, None] * mB.sum(1).double()[None, :]).sum())


def ev():
    return torch.cuda.Event(enable_timing=True)


def step(rec=None):
    hA.grad = hB.grad = tau.grad = None
    e = [ev() for _ in range(6)]
    e[0].record()
    nA = ops.normalize_scale(hA, None, eps=0.0)
    nB = ops.normalize_scale(hB, None, eps=0.0)
    e[1].record()
    simA, simB = ops.filip_similarity(nA, nB, lenA, lenB, tau, compute=ops.BF16)
    e[2].record()
    loss = ops.filip_loss(simA, simB)
    e[3].record()
    # backward of the loss tail only, then the similarity backward, then the normalise backward
    gA, gB = torch.autograd.grad(loss, (simA, simB), retain_graph=True)
    e[4].record()
    torch.autograd.backward((simA, simB), (gA, gB))
    e[5].record()
    if rec is not None:
        rec.append(e)
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
rec = []
for _ in range(5):
    step(rec)
torch.cuda.synchronize()
names = ["normalize_fwd", "filip_fwd", "loss_tail_fwd", "loss_tail_bwd", "filip_bwd+normalize_bwd"]
tot = 0.0
for i, n in enumerate(names):
    ms = min(r[i].elapsed_time(r[i + 1]) for r in rec)
    tot += ms
    print(f"{n:28s} {ms:8.3f} ms")
print(f"{'total':28s} {tot:8.3f} ms   ({'ragged' if ragged else 'full'}, B={B}, T={T}; "
      f"algorithmic {2 * pair_tokens * 128 / 1e12:.3f} TFLOP -> {2 * pair_tokens * 128 / (tot * 1e-3) / 1e12:.1f} TFLOP/s)")
