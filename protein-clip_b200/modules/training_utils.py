"""Drop-in `modules.training_utils`: CLIP training / evaluation API with the loss, its backward, the exp-logits and the
rank-of-positive on libpclip.so.  Signatures and return values follow the reference's modules/training_utils.py
(train :7, evaluate :30, _process_batch :51, _contrastive_loss :74, train_gc :92, get_logits :157,
eval_gc_allrec_onepep :174, _call_model_gc :232, _contrastive_loss_gc :247) so main.py / main_2protein.py / topk.py
run unchanged.

What changed underneath:
  * `_contrastive_loss(pep, rec.t())` and `_contrastive_loss_gc(list, list)` are ONE fused op (pclip.ops.clip_loss):
    no torch.cat of the GradCache leaves, no B x B logits / exp(logits) tensors, gradients for every leaf produced in
    the same pass and scaled by autograd's grad_output (GradScaler's 65536) on the device.
  * under CUDA autocast (the GradCache path decorates its loss with autocast(), ref :247-249) the contraction runs on
    the bf16 tensor-core path, otherwise in fp32 -- the same split the reference gets from autocast's op lists.
  * `train_gc`: the frozen ESM trunk runs ONCE per sub-batch (its hidden states are parked in the
    sub-batch's dict and reused by GradCache's replay), and the big-batch loss + the gradients of all 32 leaves come
    from ONE kernel launch (pclip.ops.clip_loss_gc_direct) scaled by GradScaler's device-side scale.
  * rank evaluation: `build_bank` encodes every sequence once, `rank_all` ranks ALL pairs with one
    rank-kernel call, `topk_accuracy` is topk.py:53-61's cumulative histogram; `eval_gc_allrec_onepep` keeps its
    signature and return value as a thin wrapper over the cached ranks (the reference re-encodes the whole candidate
    set for each of its 256 calls, topk.py:47-51).
"""
from __future__ import annotations

import torch
from grad_cache.functional import cached, cat_input_tensor  # noqa: F401  (re-exported like the reference does)
from pclip import ops



def _encode(tokenizer, strings, device):
    return tokenizer(strings, return_tensors="pt", padding=True).to(device)


# ---------------------------------------------------------------------------------------------------------
# plain (non-cached) path
# ---------------------------------------------------------------------------------------------------------
def _contrastive_loss(pep_embedding, rec_embedding):
    """Symmetric InfoNCE of logits = pep_embedding @ rec_embedding, rec_embedding being the [D, B] transposed view the
    callers pass (`rec.t()`, ref :69).  The view is undone through its strides, never copied."""
    return ops.clip_loss(pep_embedding, rec_embedding.t())


def _process_batch(model, batch_data, tokenizer, device, compute_grad=False):
    """One batch: tokenise, encode both modalities, loss (+ backward when compute_grad).  Returns a Python float."""
    pep_strings, rec_strings = batch_data
    pep, rec = model(_encode(tokenizer, pep_strings, device), _encode(tokenizer, rec_strings, device))
    loss = _contrastive_loss(pep, rec.t())
    if compute_grad:
        loss.backward()
    return loss.item()


def train(model, data_loader, optimizer, tokenizer, device):
    """One epoch without gradient caching; mean loss over the loader's batches."""
    model.train()
    running = 0.0
    for batch in data_loader:
        optimizer.zero_grad()
        running += _process_batch(model, batch, tokenizer, device, compute_grad=True)
        optimizer.step()
    return running / len(data_loader)


def evaluate(model, data_loader, tokenizer, device):
    """Mean loss over the loader's batches, no gradients."""
    model.eval()
    running = 0.0
    with torch.no_grad():
        for batch in data_loader:
            running += _process_batch(model, batch, tokenizer, device)
    return running / len(data_loader)


# ---------------------------------------------------------------------------------------------------------
# GradCache path
# ---------------------------------------------------------------------------------------------------------
@cached
def _call_model_gc(model, input):
    """`cached` turns this into (detached leaf with requires_grad, closure that replays the forward with grad and
    back-propagates <output, leaf.grad>).  The forward runs under autocast as in the reference (ref :232-245)."""
    with torch.autocast("cuda", enabled=torch.cuda.is_available()):
        return model(input)


def _contrastive_loss_gc(x, y):
    """Loss over the concatenation of the cached sub-batch representations.  x / y: lists of [b_i, D] leaves (or
    tensors).  The lists go to the kernel as row-chunk descriptors; each leaf's .grad is filled by backward."""
    first = x[0] if isinstance(x, (list, tuple)) else x
    # the reference decorates this function with autocast(): half-precision GEMM, fp32 exp / sums
    compute = ops.BF16 if first.shape[-1] == 128 else ops.F32
    return ops.clip_loss(x, y, compute=compute)


def get_logits(x, y):
    """exp(x @ y.T) for lists of sub-batch representations (sic: the reference returns the exponentiated logits)."""
    xs = torch.cat(list(x), 0) if isinstance(x, (list, tuple)) else x
    ys = torch.cat(list(y), 0) if isinstance(y, (list, tuple)) else y
    return ops.scores(xs, ys, "exp")


class _Accumulator:
    """The four parallel lists train_gc keeps between optimizer steps."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.reps_x, self.reps_y, self.replay_x, self.replay_y = [], [], [], []

    def add(self, rx, cx, ry, cy):
        self.reps_x.append(rx)
        self.reps_y.append(ry)
        self.replay_x.append(cx)
        self.replay_y.append(cy)

    def replay(self):
        for closure, rep in zip(self.replay_x, self.reps_x):
            closure(rep)
        for closure, rep in zip(self.replay_y, self.reps_y):
            closure(rep)


def train_gc(model, data_loader, tokenizer, optimizer, scaler, device, accumulated_batches=1):
    """One epoch with gradient caching: every `accumulated_batches` sub-batches form one big contrastive batch.
    Mean of the big-batch losses; like the reference, a trailing incomplete group is dropped (and a loader shorter
    than one group divides by zero)."""
    model.train()
    acc = _Accumulator()
    loss_sum, groups = 0.0, 0
    for step, (pep_strings, rec_strings) in enumerate(data_loader):
        xx = _encode(tokenizer, pep_strings, device)
        yy = _encode(tokenizer, rec_strings, device)
        xx["temperature"] = yy["temperature"] = model.temperature
        xx["_cache_esm"] = yy["_cache_esm"] = True      # frozen trunk: the replay reuses the hidden states
        rx, cx = _call_model_gc(model.pep_encoder, xx)
        ry, cy = _call_model_gc(model.rec_encoder, yy)
        acc.add(rx, cx, ry, cy)
        if (step + 1) % accumulated_batches:
            continue
        groups += 1
        print(step)
        # loss = _contrastive_loss_gc(cache_x, cache_y); scaler.scale(loss).backward() (ref :137-139) as one launch:
        # the leaves' .grad are d(scale * loss) / d(rep), the scale being GradScaler's device tensor (no sync)
        gscale = scaler.scale(torch.ones((), dtype=torch.float32, device=rx.device))
        loss = ops.clip_loss_gc_direct(acc.reps_x, acc.reps_y, gscale)
        loss_sum += loss.item()
        acc.replay()                       # second pass of each sub-batch (head only), seeded by the leaf grads
        acc.reset()
        scaler.step(optimizer)
        scaler.update()
        optimizer.zero_grad()
    return loss_sum / groups


# ---------------------------------------------------------------------------------------------------------
# rank / top-k evaluation over an embedding bank built once (ref: :174-228, topk.py:47-61)
# ---------------------------------------------------------------------------------------------------------
def build_bank(model, data_loader, tokenizer, device, agg_batches=None):
    """Encode every peptide and every receptor of the loader's first `agg_batches` sub-batches (None: all) ONCE, the
    way the evaluation path encodes them (no grad, autocast as in _call_model_gc).  Returns (queries [N, D],
    bank [N, D]) fp32: row i of the queries is the peptide whose positive is row i of the bank."""
    model.eval()
    qs, bs = [], []
    with torch.no_grad():
        for step, (pep_strings, rec_strings) in enumerate(data_loader):
            if agg_batches is not None and step >= agg_batches:
                break
            xx = _encode(tokenizer, pep_strings, device)
            yy = _encode(tokenizer, rec_strings, device)
            xx["temperature"] = yy["temperature"] = model.temperature
            with torch.autocast("cuda", enabled=torch.cuda.is_available()):
                qs.append(model.pep_encoder(xx).float())
                bs.append(model.rec_encoder(yy).float())
    return torch.cat(qs, 0), torch.cat(bs, 0)


def rank_all(model, data_loader, device, tokenizer, agg_batches=2):
    """rank[k] = 1 + #{j : <pep_k, rec_j> > <pep_k, rec_k>} for EVERY pair k of the first `agg_batches` sub-batches
    (strict compare, ref :219-222; exp is monotone, so ranking the dot products is ranking exp(logits)): one bank
    build + ONE rank-kernel call instead of one pass over the loader per query.  int32 tensor [N] on the device."""
    q, bank = build_bank(model, data_loader, tokenizer, device, agg_batches)
    pos = torch.arange(q.shape[0], dtype=torch.int64, device=q.device)
    return ops.rank_of_positive(q, bank, pos)


def topk_accuracy(ranks, n=None):
    """topk.py:53-61: histogram of the ranks, cumulative sum, divided by the number of candidates: accs[k] is the share
    of queries whose positive ranks k or better (accs[0] is always 0: ranks start at 1).  A rank equal to the number of
    candidates falls off the end of the reference's list (IndexError there); here it is simply never counted."""
    ranks = torch.as_tensor(ranks).to(torch.int64).flatten().cpu()
    n = int(n if n is not None else ranks.numel())
    hist = torch.bincount(ranks.clamp(max=n), minlength=n + 1)[:n]
    return (torch.cumsum(hist, 0).double() / n).tolist()


_rank_cache = {}


def _weights_version(model):
    return sum(p._version for p in model.parameters())


def eval_gc_allrec_onepep(model, data_loader, device, tokenizer, agg_batches=2, k=0):
    """Rank of the k-th pair's receptor among the first `agg_batches` sub-batches of receptors for the k-th peptide
    (1 + number of candidates scoring strictly higher).  Returns [rank] after the first group, as the reference does.

    The reference encodes the whole candidate set again on every call (32 ESM forwards per query, topk.py:47-51 makes
    256 calls).  Here the first call for a (model weights, loader, agg_batches) encodes everything once and ranks every
    pair with one kernel call; later calls are a lookup.  The cache is dropped when a parameter of the model changes
    (in-place update counter) or another loader / model is passed."""
    key = (id(model), id(data_loader), int(agg_batches), str(device))
    ent = _rank_cache.get(key)
    ver = _weights_version(model)
    if ent is None or ent[0] != ver:
        _rank_cache.clear()
        ranks = rank_all(model, data_loader, device, tokenizer, agg_batches).cpu()
        ent = _rank_cache[key] = (ver, ranks)
    ranks = ent[1]
    if not 0 <= k < ranks.numel():
        raise IndexError(f"pair {k} is not among the first {ranks.numel()} candidates")
    return [int(ranks[k])]
