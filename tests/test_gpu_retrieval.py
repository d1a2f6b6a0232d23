"""GPU parity of the similarity / rank-of-positive / top-k path (libpclip.so through the C ABI).

Bar (north_star): top-k indices bit-exact up to ties, ties broken by the lower index.  The oracle is the float64
stable-sort top-k of oracle/restate.py; "up to ties" is made precise by `check_topk`: every returned index must be an
oracle index, or its float64 score must lie within an fp32 rounding band of the oracle's k-th score."""
import pytest
import torch

from oracle import restate as R

pytestmark = pytest.mark.gpu
BAND = 2e-6  # |fp32 FFMA-chain score - float64 score| for unit-ish vectors, D = 128


@pytest.fixture(scope="module")
def ops():
    from pclip import ops as o
    return o


def check_topk(idx, score, q, bank, k):
    """idx/score: kernel output (cpu).  Returns the number of band (near-tie) substitutions."""
    s64 = q.double() @ bank.double().t()
    oi, ov = R.topk_stable(q.double(), bank.double(), min(k, bank.shape[0]))
    n = bank.shape[0]
    band = 0
    for r in range(q.shape[0]):
        got = idx[r].tolist()
        valid = [g for g in got if g >= 0]
        assert len(valid) == min(k, n) and len(set(valid)) == len(valid), (r, got)
        assert all(g == -1 for g in got[len(valid):])
        sc = score[r, :len(valid)]
        assert bool((sc[:-1] >= sc[1:]).all()), f"row {r}: scores not descending"
        assert float((sc.double() - s64[r, valid]).abs().max()) <= BAND
        kth = float(ov[r, -1])
        want = set(oi[r].tolist())
        for g in valid:
            if g not in want:
                band += 1
                assert abs(float(s64[r, g]) - kth) <= 2 * BAND, f"row {r}: index {g} is not a top-{k} row"
        # exact fp32 ties must be ordered by index
        for a in range(len(valid) - 1):
            if float(sc[a]) == float(sc[a + 1]):
                assert valid[a] < valid[a + 1], f"row {r}: tie not broken by lower index"
    return band


def test_cosine_matrix_and_get_logits(ops, golden):
    g = golden("cosine_16")
    cos = ops.scores(g["pep"].cuda(), g["rec"].cuda(), "dot", neg_tau=g["tau"].cuda())
    assert float((cos.cpu() - g["cos"]).abs().max()) <= 2e-6
    gc = golden("clip_gradcache_16x16")
    x, y = gc["x"].reshape(256, 128).cuda(), gc["y"].reshape(256, 128).cuda()
    e = ops.scores(x, y, "exp")
    assert float(((e.cpu() - gc["exp_logits"]) / gc["exp_logits"]).abs().max()) <= 1e-5


def test_rank_of_positive_golden(ops, golden):
    """ref: eval_gc_allrec_onepep (training_utils.py:219-222) run end to end by oracle/make_golden.py."""
    g = golden("rank_of_positive")
    n = g["pep"].shape[0]
    rank = ops.rank_of_positive(g["pep"].cuda(), g["rec"].cuda(), torch.arange(n).cuda())
    assert rank.cpu().tolist() == g["ranks"].tolist()
    acc = R.topk_accuracy_from_ranks([r for r in rank.cpu().tolist() if r < n], n)
    assert acc[-1] <= 1.0 and all(a <= b for a, b in zip(acc, acc[1:]))


@pytest.mark.parametrize("Q,N", [(5, 77), (130, 3000), (64, 20000)])
def test_rank_matches_own_scores(ops, Q, N):
    from pclip import synth
    bank, q, pos = synth.retrieval_bank(N, Q, cfg_id=50 + Q, dup_every=7)
    s = ops.scores(q.cuda(), bank.cuda())
    want = (s > s.gather(1, pos.cuda()[:, None])).sum(1) + 1
    rank = ops.rank_of_positive(q.cuda(), bank.cuda(), pos.cuda())
    assert torch.equal(rank.long(), want)


@pytest.mark.parametrize("Q,N,k", [(3, 5, 10), (70, 1000, 1), (300, 5000, 10), (257, 12345, 10), (1000, 40000, 16),
                                   (500, 30000, 6), (500, 30000, 7), (300, 20000, 14), (200, 8000, 32)])   # 16 / 24 / 32-candidate kernels
def test_topk_f32_and_bf16_vs_oracle(ops, Q, N, k):
    from pclip import synth
    bank, q, pos = synth.retrieval_bank(N, Q, cfg_id=60 + k)
    i32, s32, _ = ops.topk(q.cuda(), bank.cuda(), k, compute=ops.F32)
    check_topk(i32.cpu(), s32.cpu(), q, bank, k)
    i16, s16, unsafe = ops.topk(q.cuda(), bank.cuda(), k, compute=ops.BF16)
    # the tensor-core path re-scores with the exact path's FFMA chain: results are IDENTICAL, not merely close
    assert torch.equal(i16, i32) and torch.equal(s16, s32)
    if N >= 1000 and 2 * k + 4 <= 32:   # beyond k = 14 the candidate list has less than 2k + 4 slots
        assert int(unsafe) <= Q // 20  # random unit vectors: the candidate margin proves (almost) every query


def test_topk_ties_lower_index_wins(ops):
    from pclip import synth
    bank, q, pos = synth.retrieval_bank(6000, 200, cfg_id=71, dup_every=3)   # every third row is duplicated
    for mode in (ops.F32, ops.BF16):
        idx, sc, _ = ops.topk(q.cuda(), bank.cuda(), 10, compute=mode)
        check_topk(idx.cpu(), sc.cpu(), q, bank, 10)
        idx = idx.cpu()
        for r in range(200):
            row = idx[r].tolist()
            for j, g in enumerate(row):
                if g % 3 == 1:                  # the copy: its source (g - 1) has the identical score
                    assert j > 0 and row[j - 1] == g - 1, (r, row)


@pytest.mark.parametrize("Q", [40, 700])   # 700: several 64-row blocks of listed queries, device-planned bank split
def test_topk_unprovable_rows_are_recomputed_exactly(ops, Q):
    """Bank rows that differ by less than bf16 resolution: the margin test must fail and the exact kernel take over."""
    g = torch.Generator().manual_seed(7)
    base = torch.randn(1, 128, generator=g)
    bank = base + 1e-4 * torch.randn(3000, 128, generator=g)
    bank = bank / bank.norm(dim=-1, keepdim=True)
    q = base + 0.1 * torch.randn(Q, 128, generator=g)
    q = q / q.norm(dim=-1, keepdim=True)
    i32, s32, _ = ops.topk(q.cuda(), bank.cuda(), 10, compute=ops.F32)
    i16, s16, unsafe = ops.topk(q.cuda(), bank.cuda(), 10, compute=ops.BF16)
    assert int(unsafe) > 0
    assert torch.equal(i16, i32) and torch.equal(s16, s32)
    check_topk(i32.cpu(), s32.cpu(), q, bank, 10)


def test_topk_adversarial_bf16_midpoints(ops):
    """every entry of the query and of one bank row sits just above a bf16 rounding midpoint (1 + 2^-8 rounds
    to 1), all signs aligned, so the bf16 score of that row falls 2^-7 (relative) below its fp32 score -- the worst
    case of the exactness bound.  30 decoy rows are exact in bf16 and score higher than the victim in bf16 but lower in
    fp32: the victim is pushed out of the 24-entry candidate list although it is the true top-1.  With the round-1
    bound (2^-8: half the rigorous one) the query passed the margin test and the wrong answer was returned; with the
    rigorous bound the margin test fails and the exact kernel recomputes the row."""
    a = 0.0625
    up = 1.0 + 2.0 ** -8
    g = torch.Generator().manual_seed(11)
    N, Q, k = 4096, 8, 10
    bank = 0.01 * torch.randn(N, 128, generator=g)
    victim = 40                               # victim and decoys share one 128-row bank tile (one split)
    bank[victim] = a * up
    for i in range(30):                       # decoy i: bf16-exact, 8 entries one bf16 ulp up
        row = torch.full((128,), a)
        row[4 * i: 4 * i + 8] = a * (1.0 + 2.0 ** -7)
        bank[1 + i] = row
    q = torch.randn(Q, 128, generator=g)
    q[:4] = a * up
    assert float((q[0].double() @ bank[victim].double())) > float((q[0].double() @ bank[1].double()))
    assert float(q[0].bfloat16().double() @ bank[victim].bfloat16().double()) < \
        float(q[0].bfloat16().double() @ bank[1].bfloat16().double())
    i32, s32, _ = ops.topk(q.cuda(), bank.cuda(), k, compute=ops.F32)
    i16, s16, unsafe = ops.topk(q.cuda(), bank.cuda(), k, compute=ops.BF16)
    assert i32[:4, 0].tolist() == [victim] * 4
    assert torch.equal(i16, i32) and torch.equal(s16, s32)
    assert int(unsafe) >= 4
    check_topk(i16.cpu(), s16.cpu(), q, bank, k)


def test_topk_bank_sharded_merge(ops):
    from pclip import synth
    bank, q, pos = synth.retrieval_bank(9000, 150, cfg_id=72, dup_every=5)
    full_i, full_s, _ = ops.topk(q.cuda(), bank.cuda(), 10, compute=ops.BF16)
    parts = [ops.topk(q.cuda(), bank[a:b].cuda(), 10, compute=ops.BF16, index_base=a)
             for a, b in ((0, 3000), (3000, 7001), (7001, 9000))]
    mi, ms = ops.topk_merge(torch.stack([p[0] for p in parts]), torch.stack([p[1] for p in parts]))
    assert torch.equal(mi, full_i) and torch.equal(ms, full_s)


def test_topk_large_bank_properties(ops):
    """A 2^20-row bank (quarter of configs[4]'s) with 2048 queries: recall of the planted positive and agreement with
    the float64 oracle on a query sample."""
    from pclip import synth
    N, Q = 1 << 20, 2048
    bank, q, pos = synth.retrieval_bank(N, Q, cfg_id=5, device="cuda")
    idx, sc, unsafe = ops.topk(q, bank, 10, compute=ops.BF16)
    assert float((idx[:, 0] == pos).float().mean()) > 0.99   # noisy copies: the source row is the best match
    sel = torch.arange(0, Q, 64)
    check_topk(idx[sel].cpu(), sc[sel].cpu(), q[sel].cpu(), bank.cpu(), 10)
    assert int(unsafe) < Q // 50
