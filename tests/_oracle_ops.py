"""CPU stand-in for the arithmetic entry points of pclip.ops, built on oracle/restate.py (TEST INFRASTRUCTURE).

Used ONLY by tests/offline_harness.py to run the reference's entry scripts against the drop-in `modules` package in a
container without a GPU: the host logic (modules/*, grad_cache shim, GradCache driver, rank cache, plotting) is the
product's, the numbers come from the oracle through torch autograd.  The product itself has no CPU path."""
import torch

from oracle import restate as R


def normalize_scale(e, tau, eps=1e-12):
    if tau is None:
        return R.filip_normalize(e)
    return R.normalize_scale(e, tau, eps)


def clip_loss(x, y, compute=None):
    xs = list(x) if isinstance(x, (list, tuple)) else [x]
    ys = list(y) if isinstance(y, (list, tuple)) else [y]
    return R.clip_loss_gc([t.float() for t in xs], [t.float() for t in ys])


def clip_loss_gc_direct(xs, ys, gscale=None):
    loss = clip_loss(list(xs), list(ys))
    (loss if gscale is None else loss * gscale.reshape(())).backward()
    return loss.detach()


def scores(q, bank, op="dot", neg_tau=None):
    s = q.detach().float() @ bank.detach().float().t()
    if op == "exp":
        s = s.exp()
    if neg_tau is not None:
        s = s * torch.exp(-neg_tau.detach().float())
    return s


def rank_of_positive(q, bank, pos):
    s = q.detach().float() @ bank.detach().float().t()
    return ((s > s.gather(1, pos.reshape(-1, 1).long())).sum(1) + 1).to(torch.int32)


def _as_mask(m, T):
    if m.dim() == 1:
        return (torch.arange(T, device=m.device)[None, :] < m[:, None]).long()
    return m


def filip_similarity(hA, hB, maskA, maskB, tau, compute=None):
    tau = torch.as_tensor(tau)
    return R.filip_similarity(hA, hB, _as_mask(maskA, hA.shape[1]), _as_mask(maskB, hB.shape[1]), tau)


def rowwise_infonce(logits, weight=1.0, diag_offset=0):
    assert diag_offset == 0
    return R.filip_contrastive_loss(logits) * weight


def filip_loss(simA, simB, diag_offset=0):
    assert diag_offset == 0
    return R.filip_compute_loss(simA, simB)


_NAMES = ("normalize_scale", "clip_loss", "clip_loss_gc_direct", "scores", "rank_of_positive", "filip_similarity",
          "rowwise_infonce", "filip_loss")


def install():
    """Swap the entry points of pclip.ops for the functions above; returns the undo function."""
    from pclip import ops
    saved = {n: getattr(ops, n) for n in _NAMES}
    for n in _NAMES:
        setattr(ops, n, globals()[n])

    def undo():
        for n, f in saved.items():
            setattr(ops, n, f)
    return undo
