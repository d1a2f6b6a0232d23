"""CPU restatement of the reference arithmetic on the contrastive hot path (TEST INFRASTRUCTURE, see __init__).

Every function cites the reference lines it follows ("ref:" paths are relative to the reference checkout).
All functions are dtype-generic torch-CPU code: call them in float64 for the gold value and in float32 for the
"reference dtype" value.  Gradients are closed forms (no autograd), so that the oracle can be evaluated in row
chunks at sizes where the reference itself cannot run (B = 65 536 CLIP needs 3 x 17 GiB, FILIP B = 256 / T = 512
needs 4 x 64 GiB); tests/test_oracle_golden.py checks the closed forms against the reference's autograd output.
"""
from __future__ import annotations

import math

import torch

F32_MIN = torch.finfo(torch.float32).min


# ---------------------------------------------------------------------------------------------------------
# a1  Encoder tail.  ref: modules/models.py:50-51
# ---------------------------------------------------------------------------------------------------------
def normalize_scale(e: torch.Tensor, tau: torch.Tensor | float, eps: float = 1e-12) -> torch.Tensor:
    """F.normalize(e, dim=-1) * exp(tau / 2): x / max(||x||, eps), then the per-side temperature factor."""
    tau = torch.as_tensor(tau, dtype=e.dtype)
    nrm = e.pow(2).sum(-1, keepdim=True).sqrt().clamp_min(eps)
    return e / nrm * torch.exp(tau / 2)


def normalize_scale_bwd(e, tau, g, eps: float = 1e-12):
    """Gradient of normalize_scale wrt (e, tau) for upstream gradient g."""
    tau = torch.as_tensor(tau, dtype=e.dtype)
    s = torch.exp(tau / 2)
    raw = e.pow(2).sum(-1, keepdim=True).sqrt()
    nrm = raw.clamp_min(eps)
    nhat = e / nrm
    # d/de [e / max(|e|, eps)]: below eps the denominator is the constant eps
    proj = torch.where(raw > eps, (g * nhat).sum(-1, keepdim=True), torch.zeros_like(raw))
    de = s / nrm * (g - nhat * proj)
    dtau = 0.5 * (g * nhat * s).sum()
    return de, dtau


# ---------------------------------------------------------------------------------------------------------
# a2 / a3  CLIP InfoNCE.  ref: modules/training_utils.py:74-89 and :247-265
# ---------------------------------------------------------------------------------------------------------
def clip_loss(pep: torch.Tensor, recT: torch.Tensor) -> torch.Tensor:
    """Literal arithmetic of _contrastive_loss(pep [B,D], recT [D,B]) (naive exp, no max subtraction)."""
    s = pep @ recT                                   # :85
    e = s.exp()                                      # :86
    pos = s.diagonal().exp()
    l_r = -(pos / e.sum(dim=1)).log().mean()         # :87
    l_p = -(pos / e.sum(dim=0)).log().mean()         # :88
    return (l_r + l_p) * 0.5                         # :89


def clip_loss_gc(xs, ys) -> torch.Tensor:
    """_contrastive_loss_gc: cat_input_tensor concatenates list arguments, then the same math on x @ y.T."""
    x = torch.cat(list(xs), 0) if isinstance(xs, (list, tuple)) else xs
    y = torch.cat(list(ys), 0) if isinstance(ys, (list, tuple)) else ys
    return clip_loss(x, y.t())


def clip_loss_and_grads(x: torch.Tensor, y: torch.Tensor, chunk: int = 4096, grad_output: float = 1.0):
    """loss, dX, dY, row_lse, col_lse for S = x @ y.T, evaluated in row chunks (two passes: column sums first).

    Closed form of :85-89:  loss = 1/(2B) sum_i (log R_i + log C_i - 2 S_ii),
    dS = (E / R_i + E / C_j - 2 I) / (2B),  dX = dS y,  dY = dS^T x.
    """
    B = x.shape[0]
    assert y.shape[0] == B
    R = torch.empty(B, dtype=x.dtype)
    C = torch.zeros(B, dtype=x.dtype)
    diag = (x * y).sum(-1)
    for i0 in range(0, B, chunk):
        e = (x[i0:i0 + chunk] @ y.t()).exp()
        R[i0:i0 + chunk] = e.sum(1)
        C += e.sum(0)
    loss = (R.log() + C.log() - 2 * diag).sum() / (2 * B)
    dx = torch.empty_like(x)
    dy = torch.zeros_like(y)
    k = grad_output / (2 * B)
    for i0 in range(0, B, chunk):
        xs = x[i0:i0 + chunk]
        e = (xs @ y.t()).exp()
        g = e * (1 / R[i0:i0 + chunk, None] + 1 / C[None, :])
        idx = torch.arange(i0, min(i0 + chunk, B))
        g[idx - i0, idx] -= 2
        g *= k
        dx[i0:i0 + chunk] = g @ y
        dy += g.t() @ xs
    return loss, dx, dy, R.log(), C.log()


# ---------------------------------------------------------------------------------------------------------
# a5 / a6 / a7  scores, rank of the positive, top-k accuracy, cosine matrix
# ---------------------------------------------------------------------------------------------------------
def get_logits(x, y) -> torch.Tensor:
    """ref: modules/training_utils.py:157-172 -- returns exp(x @ y.T) (sic: the exponentiated logits)."""
    x = torch.cat(list(x), 0) if isinstance(x, (list, tuple)) else x
    y = torch.cat(list(y), 0) if isinstance(y, (list, tuple)) else y
    return (x @ y.t()).exp()


def rank_of_positive(scores: torch.Tensor, k: int) -> int:
    """ref: modules/training_utils.py:219-222 -- 1 + #{j : score_j > score_k}, strict compare."""
    return int((scores > scores[k]).sum().item()) + 1


def topk_accuracy_from_ranks(ranks, n: int = 256):
    """ref: topk.py:53-61 -- histogram of ranks, cumulative sum, divided by n.  (The reference indexes
    top_k[int(rank)] with a length-n list, so rank == n raises IndexError there; same here.)"""
    hist = [0] * n
    for r in ranks:
        hist[int(r)] += 1
    for i in range(1, n):
        hist[i] += hist[i - 1]
    return [h / n for h in hist]


def cosine_matrix(pep: torch.Tensor, rec: torch.Tensor, tau) -> torch.Tensor:
    """ref: modules/visualizations.py:51-52 -- mm(pep, rec.T) * exp(-tau) (undoes the two exp(tau/2))."""
    return pep @ rec.t() * torch.exp(-torch.as_tensor(tau, dtype=pep.dtype))


def topk_stable(q: torch.Tensor, bank: torch.Tensor, k: int, chunk: int = 1024):
    """Top-k by dot product, descending, exact ties broken by LOWER bank index (stable sort).

    The reference has no general top-k (only rank_of_positive); this is the oracle SURVEY.md 8(c) specifies."""
    idx = torch.empty(q.shape[0], k, dtype=torch.int64)
    val = torch.empty(q.shape[0], k, dtype=q.dtype)
    for i0 in range(0, q.shape[0], chunk):
        s = q[i0:i0 + chunk] @ bank.t()
        v, ix = torch.sort(s, dim=1, descending=True, stable=True)
        idx[i0:i0 + chunk] = ix[:, :k]
        val[i0:i0 + chunk] = v[:, :k]
    return idx, val


# ---------------------------------------------------------------------------------------------------------
# a8 - a11  FILIP token-wise late interaction
# ---------------------------------------------------------------------------------------------------------
def filip_normalize(h: torch.Tensor) -> torch.Tensor:
    """ref: modules/models.py:238-239 -- h / h.norm(dim=-1, keepdim=True), NO eps (pads included)."""
    return h / h.pow(2).sum(-1, keepdim=True).sqrt()


def filip_normalize_bwd(h, g):
    nrm = h.pow(2).sum(-1, keepdim=True).sqrt()
    n = h / nrm
    return (g - n * (g * n).sum(-1, keepdim=True)) / nrm


def masked_mean(t: torch.Tensor, mask: torch.Tensor, dim: int = 1, eps: float = 1e-6) -> torch.Tensor:
    """ref: modules/models.py:247-264 -- zero-fill invalid entries, sum, divide by clamp(count, eps)."""
    t = torch.where(mask, t, torch.zeros_like(t))
    return t.sum(dim=dim) / mask.sum(dim=dim).clamp(min=eps)


def filip_similarity(hA, hB, maskA, maskB, tau, want_arg: bool = False):
    """ref: modules/models.py:292-356 (_filip_similarity_score, include_group=False), one A-row at a time.

    S[a,b,t,s] = <hA[a,t], hB[b,s]> / tau; pads -> float32 finfo.min; simA = masked mean over t of max_s,
    simB = masked mean over s of max_t.  Returns (simA, simB[, argA, argB]) with arg* the FIRST maximiser."""
    maskA = maskA.bool()
    maskB = maskB.bool()
    Ba, TA, _ = hA.shape
    Bb, TB, _ = hB.shape
    tau = torch.as_tensor(tau, dtype=hA.dtype)
    simA = torch.empty(Ba, Bb, dtype=hA.dtype)
    simB = torch.empty(Ba, Bb, dtype=hA.dtype)
    argA = torch.empty(Ba, Bb, TA, dtype=torch.int64) if want_arg else None
    argB = torch.empty(Ba, Bb, TB, dtype=torch.int64) if want_arg else None
    fill = torch.tensor(F32_MIN, dtype=hA.dtype)
    for a in range(Ba):
        s = torch.einsum("td,bsd->bts", hA[a], hB) / tau                       # :332
        valid = maskA[a][None, :, None] & maskB[:, None, :]                      # :333-336
        s = torch.where(valid, s, fill)                                          # :338-340
        rmax, rarg = s.max(dim=2)                                                # :342  [Bb, TA]
        cmax, carg = s.max(dim=1)                                                # :343  [Bb, TB]
        simA[a] = masked_mean(rmax, maskA[a][None, :].expand(Bb, TA), dim=-1)    # :345-347
        simB[a] = masked_mean(cmax, maskB, dim=-1)                               # :349-351
        if want_arg:
            argA[a] = rarg
            argB[a] = carg
    if want_arg:
        return simA, simB, argA, argB
    return simA, simB


def filip_contrastive_loss(logits: torch.Tensor, use_dcl: bool = False) -> torch.Tensor:
    """ref: modules/training_utils_filip.py:4-23 -- row-wise InfoNCE with naive exp."""
    e = logits.exp()
    num = e.diagonal()
    if use_dcl:
        e = e.masked_fill(torch.eye(logits.shape[0], dtype=torch.bool), 0)
    return (-num.log() + e.sum(dim=1).log()).mean()


def filip_compute_loss(simA, simB) -> torch.Tensor:
    """ref: modules/training_utils_filip.py:26-39 -- both matrices row-wise (simB is NOT transposed)."""
    return (filip_contrastive_loss(simA) + filip_contrastive_loss(simB)) / 2


def filip_loss_and_grads(hA_raw, hB_raw, maskA, maskB, tau, normalize: bool = True):
    """Full FILIP step: (normalise ->) similarity -> loss, with closed-form gradients.

    Returns dict(loss, simA, simB, dhA, dhB, dtau).  The max is differentiated through its first maximiser
    (torch.amax splits evenly among exact ties; inputs used in tests have none)."""
    hA = filip_normalize(hA_raw) if normalize else hA_raw
    hB = filip_normalize(hB_raw) if normalize else hB_raw
    maskA = maskA.bool()
    maskB = maskB.bool()
    tau = torch.as_tensor(tau, dtype=hA.dtype)
    simA, simB, argA, argB = filip_similarity(hA, hB, maskA, maskB, tau, want_arg=True)
    loss = filip_compute_loss(simA, simB)
    B = simA.shape[0]
    eye = torch.eye(B, simA.shape[1], dtype=hA.dtype)
    gA = (torch.softmax(simA, dim=1) - eye) / (2 * B)
    gB = (torch.softmax(simB, dim=1) - eye) / (2 * B)
    nA = maskA.sum(1).clamp(min=1e-6).to(hA.dtype)
    nB = maskB.sum(1).clamp(min=1e-6).to(hA.dtype)
    dhA = torch.zeros_like(hA)
    dhB = torch.zeros_like(hB)
    Bb = hB.shape[0]
    for a in range(hA.shape[0]):
        for b in range(Bb):
            if not (maskA[a].any() and maskB[b].any()):
                continue
            wA = gA[a, b] / nA[a] / tau
            wB = gB[a, b] / nB[b] / tau
            tv = maskA[a].nonzero().flatten()
            sv = maskB[b].nonzero().flatten()
            ja = argA[a, b, tv]
            dhA[a].index_add_(0, tv, wA * hB[b, ja])
            dhB[b].index_add_(0, ja, wA * hA[a, tv])
            ib = argB[a, b, sv]
            dhB[b].index_add_(0, sv, wB * hA[a, ib])
            dhA[a].index_add_(0, ib, wB * hB[b, sv])
    dtau = -((gA * simA).sum() + (gB * simB).sum()) / tau
    if normalize:
        dhA = filip_normalize_bwd(hA_raw, dhA)
        dhB = filip_normalize_bwd(hB_raw, dhB)
    return dict(loss=loss, simA=simA, simB=simB, dhA=dhA, dhB=dhB, dtau=dtau, gA=gA, gB=gB,
                argA=argA, argB=argB)


# ---------------------------------------------------------------------------------------------------------
# encoder head after the ESM trunk
# ---------------------------------------------------------------------------------------------------------
def _ffn_apply(x, sd, name, ln_eps):
    """The nn.Sequential of models.py:66-73 from its state_dict entries, dropout inactive: blocks `<name>.<i>.0` (Linear),
    ReLU, `<name>.<i>.2` (LayerNorm), then the last plain Linear `<name>.<depth-1>`."""
    i = 0
    while f"{name}.{i}.0.weight" in sd:
        x = torch.relu(x @ sd[f"{name}.{i}.0.weight"].t() + sd[f"{name}.{i}.0.bias"])
        x = torch.nn.functional.layer_norm(x, (x.shape[-1],), sd[f"{name}.{i}.2.weight"], sd[f"{name}.{i}.2.bias"], ln_eps)
        i += 1
    return x @ sd[f"{name}.{i}.weight"].t() + sd[f"{name}.{i}.bias"]


def encoder_head(hidden, mask, sd, tau=None, pooled: bool = True, ln_eps: float = 1e-5):
    """Encoder.forward after the trunk (ref: modules/models.py:46-51; masked mean :75-90) when pooled, else
    FILIPEncoder.forward (:192-196).  `sd`: the encoder's state_dict (projection.*, amino_acid_ffn.*, embedding_ffn.*)."""
    h = _ffn_apply(hidden @ sd["projection.weight"].t() + sd["projection.bias"], sd, "amino_acid_ffn", ln_eps)
    if not pooled:
        return h
    mean = (h * mask.unsqueeze(-1)).sum(1) / mask.sum(1, keepdim=True).clamp(min=1)
    e = _ffn_apply(mean, sd, "embedding_ffn", ln_eps)
    return torch.nn.functional.normalize(e, dim=-1) * torch.exp(torch.as_tensor(tau) / 2)


def bf16_round(t: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest-even to bfloat16 and back (what the tensor-core path feeds the MMA)."""
    return t.to(torch.float32).to(torch.bfloat16).to(t.dtype)


def ln(x: float) -> float:
    return math.log(x)


# ---------------------------------------------------------------------------------------------------------
# bounded CPU-baseline samples (bench.py's cpu_baseline / --impl reference legs time these)
# ---------------------------------------------------------------------------------------------------------
def clip_rowblock_fwd_bwd(x: torch.Tensor, y: torch.Tensor, i0: int, n: int):
    """The work the reference arithmetic (modules/training_utils.py:85-89 + its autograd backward) spends on rows
    [i0, i0+n) of a B-pair batch: the [n, B] block of S and E (forward), and that block's share of dX and dY
    (backward: G block, G y, G^T x).  Column sums of a row block are partial, so the numbers are only a TIMING
    unit (n / B of one full fwd+bwd), not a loss value; the full-size value comes from clip_loss_and_grads."""
    xs = x[i0:i0 + n]
    s = xs @ y.t()                       # :85
    e = s.exp()                          # :86
    r = e.sum(1)                         # :87
    c = e.sum(0)                         # :88 (partial)
    g = e * (1 / r[:, None] + 1 / c[None, :])
    idx = torch.arange(i0, i0 + n)
    g[idx - i0, idx] -= 2
    dx = g @ y
    dy = g.t() @ xs
    return r, c, dx, dy


def topk_block(q: torch.Tensor, bank: torch.Tensor, k: int):
    """torch.topk over one query block (the CPU-baseline unit SURVEY.md 8(d) names for cfg 5)."""
    return torch.topk(q @ bank.t(), k, dim=1)
