"""Hashes of the FILIP gradients on a few seeded batches (debug aid: two builds / env settings of a deterministic
kernel must print the same lines): python tools/gpu_filip_gradhash.py"""
import hashlib, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "protein-clip_b200"))
import torch
from pclip import ops, synth

dev = torch.device("cuda", 0)
for B, TA, TB, ragged in ((6, 512, 300, True), (4, 96, 160, False), (16, 512, 512, True), (3, 33, 500, True), (5, 301, 77, True)):
    hA, hB, mA, mB, tau = synth.filip_tokens(B, TA, TB, cfg_id=700 + TA, ragged=ragged)
    if ragged and B > 2:
        mA[1, 3:] = 0
        mB[2, 1:] = 0
    hA, hB, tau = hA.to(dev).requires_grad_(), hB.to(dev).requires_grad_(), tau.to(dev).requires_grad_()
    nA, nB = ops.normalize_scale(hA, None, eps=0.0), ops.normalize_scale(hB, None, eps=0.0)
    simA, simB = ops.filip_similarity(nA, nB, mA.to(dev), mB.to(dev), tau, compute=ops.BF16)
    loss = ops.filip_loss(simA, simB)
    loss.backward()
    h = hashlib.sha1()
    for t in (hA.grad, hB.grad, tau.grad, loss.detach()):
        h.update(t.detach().cpu().numpy().tobytes())
    print(f"B={B} TA={TA} TB={TB} ragged={ragged}: loss {float(loss):.9g} grads {h.hexdigest()[:16]}", flush=True)
