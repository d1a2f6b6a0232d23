"""Real NCCL runs (one process per GPU) of the sharded host logic in pclip/sharded.py against the single-GPU results.
Needs >= 2 GPUs on the box (skipped otherwise; run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multirank.py -m gpu`)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, ret):
    for p in (REPO, os.path.join(REPO, "protein-clip_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from oracle import restate as R
        from pclip import ops, sharded, synth
        dev = torch.device("cuda", rank)
        out = {}
        # ---- CLIP, bf16, fused one-collective exchange ----
        B = 2048
        x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=31, trained_like=True)
        Ma = B // world
        sl = slice(rank * Ma, (rank + 1) * Ma)
        job = sharded.ShardedClip(B, 128, ops.BF16, dev)
        loss, dx, dy = job.step([x[sl].to(dev)], [y[sl].to(dev)])
        lo, dxo, dyo, _, _ = R.clip_loss_and_grads(R.bf16_round(x).double(), R.bf16_round(y).double())
        out["clip"] = (job.fused and abs(float(loss) - float(lo)) <= 1e-2 * abs(float(lo))
                       and float((dx.double().cpu() - dxo[sl]).abs().max() / dxo.abs().max()) <= 4e-3
                       and float((dy.double().cpu() - dyo[sl]).abs().max() / dyo.abs().max()) <= 4e-3)
        # ---- FILIP, A rows sharded ----
        g = torch.Generator().manual_seed(9)
        Bf, T, D = 8 * world, 70, 128
        hA = R.filip_normalize(torch.randn(Bf, T, D, generator=g))
        hB = R.filip_normalize(torch.randn(Bf, T, D, generator=g))
        lenA, lenB = torch.randint(3, T + 1, (Bf,), generator=g), torch.randint(3, T + 1, (Bf,), generator=g)
        mA = (torch.arange(T)[None, :] < lenA[:, None]).long()
        mB = (torch.arange(T)[None, :] < lenB[:, None]).long()
        tau = torch.tensor(0.9)
        Bl = Bf // world
        fs = slice(rank * Bl, (rank + 1) * Bl)
        lf, dhA, dhB, dtau = sharded.ShardedFilip().step(hA[fs].to(dev), hB[fs].to(dev), mA[fs].to(dev), mB[fs].to(dev),
                                                         tau.to(dev))
        hA1, hB1, t1 = hA.to(dev).requires_grad_(), hB.to(dev).requires_grad_(), tau.to(dev).requires_grad_()
        sA, sB = ops.filip_similarity(hA1, hB1, mA.to(dev), mB.to(dev), t1)
        l1 = ops.filip_loss(sA, sB)
        l1.backward()
        out["filip"] = (abs(float(lf) - float(l1)) <= 1e-5 * abs(float(l1))
                        and float((dhA - hA1.grad[fs]).abs().max()) <= 1e-5 * float(hA1.grad.abs().max())
                        and float((dhB - hB1.grad[fs]).abs().max()) <= 1e-5 * float(hB1.grad.abs().max())
                        and abs(float(dtau) - float(t1.grad)) <= 1e-4 * abs(float(t1.grad)))
        # ---- top-k, both shardings ----
        N, Q, k = 8192 * world, 512 * world, 10
        bank, q, _ = synth.retrieval_bank(N, Q, cfg_id=6, device="cpu")
        i1, s1, _ = ops.topk(q.to(dev), bank.to(dev), k)
        Ql, Nl = Q // world, N // world
        iq, sq = sharded.ShardedTopk(k, "queries").run(q[rank * Ql:(rank + 1) * Ql].to(dev), bank.to(dev))
        ib, sb = sharded.ShardedTopk(k, "bank").run(q.to(dev), bank[rank * Nl:(rank + 1) * Nl].contiguous().to(dev))
        out["topk"] = (torch.equal(iq, i1[rank * Ql:(rank + 1) * Ql]) and torch.equal(sq, s1[rank * Ql:(rank + 1) * Ql])
                       and torch.equal(ib, i1) and torch.equal(sb, s1))
        ret[rank] = out
    finally:
        dist.destroy_process_group()


def test_nccl_ranks_match_single_gpu():
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29800 + os.getpid() % 150
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {r: {"clip": True, "filip": True, "topk": True} for r in range(world)}
