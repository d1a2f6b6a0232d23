"""Drop-in `modules.training_utils_filip`: FILIP training / evaluation API (ref: modules/training_utils_filip.py:
_contrastive_loss :4, _compute_loss :26, _process_batch :41, train :66, evaluate :89) with the row-wise InfoNCE tail on
libpclip.so (pclip.ops.rowwise_infonce: one kernel for the loss and d(loss)/d(logits))."""
from __future__ import annotations

import torch
from pclip import ops


def _contrastive_loss(logits, use_dcl=False):
    """mean_i( -log E_ii + log sum_j E_ij ), E = exp(logits), rows normalised over dim 1.  With use_dcl the positive
    is left out of the denominator (decoupled contrastive loss; unused by the scripts, ref :18-20)."""
    if not use_dcl:
        return ops.rowwise_infonce(logits, 1.0, 0)
    # DCL variant: log sum_{j != i} E_ij = log(R_i - E_ii); a [B]-sized correction of the fused row LSE
    n = logits.shape[0]
    eye = torch.eye(n, dtype=torch.bool, device=logits.device)
    pos = logits.diagonal()
    denom = torch.logsumexp(logits.masked_fill(eye, float("-inf")), dim=1)
    return (denom - pos).mean()


def _compute_loss(sim_scores_A, sim_scores_B):
    """(loss_A + loss_B) / 2 -- both matrices row-wise, sim_scores_B is NOT transposed (ref :37-39)."""
    return ops.filip_loss(sim_scores_A, sim_scores_B)


def _process_batch(model, batch_data, tokenizer, device, compute_grad=False):
    pep_strings, rec_strings = batch_data
    pep = tokenizer(pep_strings, return_tensors="pt", padding=True).to(device)
    rec = tokenizer(rec_strings, return_tensors="pt", padding=True).to(device)
    loss = _compute_loss(*model(pep, rec))
    if compute_grad:
        loss.backward()
    return loss.item()


def train(model, data_loader, optimizer, tokenier, device):   # `tokenier`: the reference's spelling, kept positional
    model.train()
    running = 0.0
    for batch in data_loader:
        optimizer.zero_grad()
        running += _process_batch(model, batch, tokenier, device, compute_grad=True)
        optimizer.step()
    return running / len(data_loader)


def evaluate(model, data_loader, tokenizer, device):
    model.eval()
    running = 0.0
    with torch.no_grad():
        for batch in data_loader:
            running += _process_batch(model, batch, tokenizer, device)
    return running / len(data_loader)
