"""World-size-2 (and 3) gloo runs of the row-sharded CLIP host logic (pclip/sharded.py) on CPU ranks: offsets, gathers,
the loss all-reduce and the 1/R, 1/C exchange, checked against the unsharded float64 oracle.  The stage arithmetic is
the oracle's (tests/_oracle_stages.py); on the GPU box the same host logic drives libpclip.so (tests/test_gpu_sharded.py)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, B, single_pass, ret):
    for p in (REPO, os.path.join(REPO, "protein-clip_b200"), os.path.join(REPO, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from _oracle_stages import OracleStages
        from oracle import restate as R
        from pclip import sharded, synth
        x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=11, trained_like=True)
        Ma = B // world
        job = sharded.ShardedClip(B, 128, 0, "cpu", stages=OracleStages(Ma, B, 128, rank * Ma, single_pass=single_pass))
        sl = slice(rank * Ma, (rank + 1) * Ma)
        # two uneven chunks per rank, like GradCache leaves
        cut = Ma // 3
        loss, dx, dy = job.step([x[sl][:cut], x[sl][cut:]], [y[sl][:cut], y[sl][cut:]], gscale=torch.tensor(65536.0))
        lo, dxo, dyo, _, _ = R.clip_loss_and_grads(x.double(), y.double(), grad_output=65536.0)
        ok = (abs(float(loss) - float(lo)) < 1e-9 * abs(float(lo))
              and float((dx - dxo[sl]).abs().max()) < 1e-9 * float(dxo.abs().max())
              and float((dy - dyo[sl]).abs().max()) < 1e-9 * float(dyo.abs().max()))
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("single_pass", [False, True], ids=["two_sided", "single_pass"])
@pytest.mark.parametrize("world,B", [(2, 64), (3, 96)])
def test_row_sharded_clip_matches_unsharded(world, B, single_pass):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29600 + world + 10 * int(single_pass) + (os.getpid() % 200)
    procs = [ctx.Process(target=_worker, args=(r, world, port, B, single_pass, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {r: True for r in range(world)}


def test_indivisible_batch_is_rejected():
    class FakeGroup:
        pass
    import importlib
    for p in (REPO, os.path.join(REPO, "protein-clip_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    sharded = importlib.import_module("pclip.sharded")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29900 + os.getpid() % 90))
    dist.init_process_group("gloo", rank=0, world_size=1)
    try:
        job = sharded.ShardedClip(10, 128, 0, "cpu", stages=object())   # world 1 divides anything
        assert job.Ma == 10
    finally:
        dist.destroy_process_group()
