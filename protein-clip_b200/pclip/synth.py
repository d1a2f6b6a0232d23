"""Seeded synthetic inputs for the contrastive hot path (there is no network: no Propedia / PDB / ESM weights).

Distributions follow SURVEY.md section 8(d): sequences are uniform over the 20 canonical residues with
Propedia-like peptide / receptor lengths (statistics of the 128-pair batch embedded in the reference's
notebooks/minimal_ram.ipynb) or PDB-like chain lengths (cap 2000, ref: modules/data_utils_2protein.py:88);
kernel-level tensors are N(0,1) embeddings, optionally passed through the encoder tail (ref: models.py:50-51).
Generation happens on the CPU generator so that the same seed gives the same tensors with or without a GPU.
"""
from __future__ import annotations

import math

import torch

RESIDUES = "ACDEFGHIKLMNPQRSTVWY"
SEED = 42  # ref: modules/seed.py:5


def _gen(cfg_id: int) -> torch.Generator:
    return torch.Generator().manual_seed(SEED + cfg_id)


def _lognormal_lengths(n, median, sigma, lo, hi, gen):
    z = torch.randn(n, generator=gen)
    return torch.clamp(torch.round(torch.exp(math.log(median) + sigma * z)), lo, hi).to(torch.int64)


def peptide_lengths(n, gen):
    return _lognormal_lengths(n, 11, 0.8, 2, 50, gen)


def receptor_lengths(n, gen, cap=1100):
    return _lognormal_lengths(n, 193, 0.55, 30, cap, gen)


def random_sequences(lengths, gen):
    out = []
    for L in lengths.tolist():
        ids = torch.randint(0, len(RESIDUES), (L,), generator=gen)
        out.append("".join(RESIDUES[i] for i in ids.tolist()))
    return out


def prefix_mask(lengths: torch.Tensor, T: int) -> torch.Tensor:
    """Right-padded attention mask [B, T] int64 (contiguous prefix of ones), as the ESM tokenizer produces."""
    return (torch.arange(T)[None, :] < lengths[:, None]).to(torch.int64)


def unit_rows(n, d, gen, scale=1.0):
    x = torch.randn(n, d, generator=gen)
    return x / x.norm(dim=-1, keepdim=True) * scale


def clip_raw_embeddings(B=16, D=128, cfg_id=1):
    """cfg 1: un-normalised encoder outputs e_pep, e_rec ~ N(0,1) [B, D], tau = 1."""
    g = _gen(cfg_id)
    return torch.randn(B, D, generator=g), torch.randn(B, D, generator=g), torch.tensor(1.0)


def clip_scaled_embeddings(B, D=128, cfg_id=4, tau=1.0, trained_like=False):
    """cfg 2 / 4: embeddings as they enter the loss: normalize(N(0,1)) * exp(tau/2).
    trained_like=True correlates the two modalities so the diagonal logits dominate (exercises exp range)."""
    g = _gen(cfg_id)
    s = math.exp(tau / 2)
    x = unit_rows(B, D, g)
    if trained_like:
        y = x + 0.35 * torch.randn(B, D, generator=g) / math.sqrt(D)
        y = y / y.norm(dim=-1, keepdim=True)
    else:
        y = unit_rows(B, D, g)
    return x * s, y * s


def gradcache_leaves(n_chunks=16, rows=16, D=128, cfg_id=2, tau=1.0):
    x, y = clip_scaled_embeddings(n_chunks * rows, D, cfg_id, tau)
    return list(x.split(rows)), list(y.split(rows))


def filip_tokens(B=256, TA=512, TB=512, D=128, cfg_id=3, ragged=False):
    """cfg 3: token embeddings hA, hB ~ N(0,1) and right-padded masks; ragged=True draws PDB-like lengths."""
    g = _gen(cfg_id)
    hA = torch.randn(B, TA, D, generator=g)
    hB = torch.randn(B, TB, D, generator=g)
    if ragged:
        lenA = torch.clamp(receptor_lengths(B, g, cap=2000) + 2, max=TA)
        lenB = torch.clamp(receptor_lengths(B, g, cap=2000) + 2, max=TB)
    else:
        lenA = torch.full((B,), TA, dtype=torch.int64)
        lenB = torch.full((B,), TB, dtype=torch.int64)
    return hA, hB, prefix_mask(lenA, TA), prefix_mask(lenB, TB), torch.tensor(1.0)


def retrieval_bank(N, Q, D=128, cfg_id=5, noise=0.5, device="cpu", dup_every=0):
    """cfg 5: unit-norm bank [N, D]; queries are noisy copies of known bank rows so the positive is known.
    dup_every > 0 duplicates bank rows (tie stress: identical scores, lower index must win)."""
    g = torch.Generator(device=device).manual_seed(SEED + cfg_id)
    bank = torch.randn(N, D, generator=g, device=device)
    bank = bank / bank.norm(dim=-1, keepdim=True)
    if dup_every:
        src = torch.arange(0, N - 1, dup_every, device=device)
        bank[src + 1] = bank[src]
    pos = torch.randint(0, N, (Q,), generator=g, device=device)
    q = bank[pos] + noise * torch.randn(Q, D, generator=g, device=device) / math.sqrt(D)
    q = q / q.norm(dim=-1, keepdim=True)
    return bank, q, pos
