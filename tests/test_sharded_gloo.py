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


# ---------------------------------------------------------------------------------------------------------
# FILIP (A-rows sharded, hB all-gathered, dhB reduce-scattered) and top-k (queries or bank sharded)
# ---------------------------------------------------------------------------------------------------------
class _OracleFilipBackend:
    """pclip.ops stand-in for ShardedFilip: the oracle's differentiable float64 restatement."""

    @staticmethod
    def filip_similarity(hA, hB, maskA, maskB, tau):
        from oracle import restate as R
        return R.filip_similarity(hA, hB, maskA, maskB, tau)

    @staticmethod
    def filip_loss(simA, simB, diag_offset=0):
        rows = torch.arange(simA.shape[0])

        def one(s):     # ref: training_utils_filip.py:15-23 on a [Bl, B] row block whose positives sit at diag_offset + i
            return (-s[rows, diag_offset + rows] + s.exp().sum(dim=1).log()).mean()
        return (one(simA) + one(simB)) / 2


class _OracleTopkBackend:
    @staticmethod
    def topk(q, bank, k, index_base=0):
        from oracle import restate as R
        idx, score = R.topk_stable(q.double(), bank.double(), k)
        return idx + index_base, score, None

    @staticmethod
    def topk_merge(idx_lists, score_lists):
        n, Q, k = idx_lists.shape
        idx = idx_lists.permute(1, 0, 2).reshape(Q, n * k)
        score = score_lists.permute(1, 0, 2).reshape(Q, n * k)
        order = torch.argsort(idx, dim=1, stable=True)                       # lower global index first ...
        idx, score = idx.gather(1, order), score.gather(1, order)
        order = torch.argsort(score, dim=1, descending=True, stable=True)    # ... among equal scores
        return idx.gather(1, order)[:, :k], score.gather(1, order)[:, :k]


def _filip_worker(rank, world, port, ret):
    for p in (REPO, os.path.join(REPO, "protein-clip_b200"), os.path.join(REPO, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import restate as R
        from pclip import sharded
        g = torch.Generator().manual_seed(77)
        B, TA, TB, D = 2 * world, 7, 9, 16
        hA = R.filip_normalize(torch.randn(B, TA, D, generator=g, dtype=torch.float64))
        hB = R.filip_normalize(torch.randn(B, TB, D, generator=g, dtype=torch.float64))
        lenA = torch.randint(2, TA + 1, (B,), generator=g)
        lenB = torch.randint(2, TB + 1, (B,), generator=g)
        mA = (torch.arange(TA)[None, :] < lenA[:, None]).long()
        mB = (torch.arange(TB)[None, :] < lenB[:, None]).long()
        tau = torch.tensor(0.7, dtype=torch.float64)
        Bl = B // world
        sl = slice(rank * Bl, (rank + 1) * Bl)
        job = sharded.ShardedFilip(backend=_OracleFilipBackend)
        loss, dhA, dhB, dtau = job.step(hA[sl], hB[sl], mA[sl], mB[sl], tau)
        want = R.filip_loss_and_grads(hA, hB, mA, mB, tau, normalize=False)
        ok = (abs(float(loss) - float(want["loss"])) < 1e-10
              and float((dhA - want["dhA"][sl]).abs().max()) < 1e-10
              and float((dhB - want["dhB"][sl]).abs().max()) < 1e-10
              and abs(float(dtau) - float(want["dtau"])) < 1e-10)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def _topk_worker(rank, world, port, ret):
    for p in (REPO, os.path.join(REPO, "protein-clip_b200"), os.path.join(REPO, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import restate as R
        from pclip import sharded
        g = torch.Generator().manual_seed(5)
        N, Q, k = 60 * world, 6 * world, 5
        bank = torch.randn(N, 16, generator=g).round()          # small integers: plenty of exact ties
        bank[N // 2] = bank[3]                                  # and a duplicated row across shards
        q = torch.randn(Q, 16, generator=g).round()
        want_i, want_s = R.topk_stable(q.double(), bank.double(), k)
        Ql, Nl = Q // world, N // world
        iq, sq = sharded.ShardedTopk(k, "queries", backend=_OracleTopkBackend).run(q[rank * Ql:(rank + 1) * Ql], bank)
        ib, sb = sharded.ShardedTopk(k, "bank", backend=_OracleTopkBackend).run(q, bank[rank * Nl:(rank + 1) * Nl])
        ok = (torch.equal(iq, want_i[rank * Ql:(rank + 1) * Ql]) and torch.equal(sq, want_s[rank * Ql:(rank + 1) * Ql])
              and torch.equal(ib, want_i) and torch.equal(sb, want_s))
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("worker", [_filip_worker, _topk_worker], ids=["filip", "topk"])
@pytest.mark.parametrize("world", [2, 3])
def test_row_sharded_filip_and_topk_match_unsharded(world, worker):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29700 + world + (17 if worker is _topk_worker else 0) + (os.getpid() % 200)
    procs = [ctx.Process(target=worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {r: True for r in range(world)}
