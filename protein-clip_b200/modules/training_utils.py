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
  * `eval_gc_allrec_onepep` keeps its loop protocol and return value but computes the rank with the rank kernel
    instead of materialising a 256 x 256 matrix of identical rows to read its diagonal.
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
        rx, cx = _call_model_gc(model.pep_encoder, xx)
        ry, cy = _call_model_gc(model.rec_encoder, yy)
        acc.add(rx, cx, ry, cy)
        if (step + 1) % accumulated_batches:
            continue
        groups += 1
        print(step)
        loss = _contrastive_loss_gc(acc.reps_x, acc.reps_y)
        loss_sum += loss.item()
        scaler.scale(loss).backward()      # leaves now hold d(scale * loss) / d(rep)
        acc.replay()                       # second forward of each sub-batch, with grad, seeded by the leaf grads
        acc.reset()
        scaler.step(optimizer)
        scaler.update()
        optimizer.zero_grad()
    return loss_sum / groups


def eval_gc_allrec_onepep(model, data_loader, device, tokenizer, agg_batches=2, k=0):
    """Rank of the k-th pair's receptor among the first `agg_batches` sub-batches of receptors for the k-th peptide
    (1 + number of candidates scoring strictly higher).  Returns [rank] after the first group, as the reference does."""
    model.eval()
    query = None
    for i, (pep_strings, rec_strings) in enumerate(data_loader):
        if k // len(pep_strings) == i:
            query = pep_strings[k % len(rec_strings)]
    bank, q_rep = [], None
    for step, (pep_strings, rec_strings) in enumerate(data_loader):
        yy = _encode(tokenizer, rec_strings, device)
        yy["temperature"] = model.temperature
        if q_rep is None:  # the reference re-encodes the same query for every candidate row; once is enough
            xx = _encode(tokenizer, (query,), device)
            xx["temperature"] = model.temperature
            q_rep, _ = _call_model_gc(model.pep_encoder, xx)
        ry, _ = _call_model_gc(model.rec_encoder, yy)
        bank.append(ry.detach())
        if (step + 1) % agg_batches == 0:
            cand = torch.cat(bank, 0).float()
            if not 0 <= k < cand.shape[0]:
                raise IndexError(f"pair {k} is not among the first {cand.shape[0]} candidates")
            pos = torch.tensor([k], dtype=torch.int64, device=cand.device)
            rank = ops.rank_of_positive(q_rep.detach().float()[:1], cand, pos)
            return [int(rank.item())]
