"""Drop-in `modules.models` for protein-clip's entry scripts, with the contrastive hot path on libpclip.so.

Same public surface as the reference's modules/models.py (class names, constructor signatures, attribute names and
therefore state_dict keys: `pep_encoder.projection.weight`, `...amino_acid_ffn.0.0.weight`, `temperature`, ...), so
`main*.py`, `topk.py`, `cosine_similarity.py` and the authors' checkpoints work unchanged.  What differs is where the
hot-path arithmetic runs:

  * Encoder tail           ref: modules/models.py:50-51    -> pclip.ops.normalize_scale   (one fused kernel, fwd + bwd)
  * FILIP normalise        ref: modules/models.py:238-239  -> pclip.ops.normalize_scale(tau=None, eps=0)
  * token-wise similarity  ref: modules/models.py:292-356  -> pclip.ops.filip_similarity  (tcgen05 tiles; the 6-D score
                                                              tensor is never materialised)

The encoder trunk (frozen ESM -> Linear -> FFN -> masked mean -> FFN) is the CALLER of the hot path and stays stock
PyTorch.  There is no CPU fallback: on CPU tensors the pclip ops raise.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from pclip import ops


def _ffn(width: int, depth: int, p_drop: float) -> nn.Sequential:
    """depth-1 blocks of Linear/ReLU/LayerNorm/Dropout followed by a Linear (module indices match the reference, so
    checkpoints load: `<ffn>.<i>.0.weight` for the blocks, `<ffn>.<depth-1>.weight` for the head)."""
    blocks = []
    for _ in range(depth - 1):
        blocks.append(nn.Sequential(nn.Linear(width, width), nn.ReLU(inplace=True), nn.LayerNorm(width),
                                    nn.Dropout(p_drop)))
    blocks.append(nn.Linear(width, width))
    return nn.Sequential(*blocks)


class _Trunk(nn.Module):
    """Shared by Encoder and FILIPEncoder: ESM hidden states -> projection -> per-residue FFN."""

    def __init__(self, input_dim, embedding_dim, h1, h2, dropout_rate, esm_model):
        super().__init__()
        self.projection = nn.Linear(input_dim, embedding_dim)
        self.amino_acid_ffn = self._build_ffn(embedding_dim, h1, dropout_rate)
        self.embedding_ffn = self._build_ffn(embedding_dim, h2, dropout_rate)
        self.esm_model = esm_model

    def _build_ffn(self, embedding_dim, depth, dropout_rate):
        return _ffn(embedding_dim, depth, dropout_rate)

    def _esm_is_replayable(self):
        """True when a second forward of the ESM trunk on the same tokens must give the same hidden states and no
        gradient flows into it: every parameter frozen (ref: main.py:22-23) and no active dropout inside it."""
        esm = self.esm_model
        if any(p.requires_grad for p in esm.parameters()):
            return False
        return not (esm.training and any(isinstance(m, nn.Dropout) and m.p > 0 for m in esm.modules()))

    def _residues(self, seq):
        # GradCache runs every sub-batch twice (no-grad forward, then the closure's replay with grad, ref:
        # modules/training_utils.py:126-144).  With a frozen trunk the replay's ESM forward recomputes exactly what the
        # first one produced: train_gc asks (seq["_cache_esm"]) for the hidden states to be parked in the caller's dict,
        # and the replay -- which receives the same dict -- picks them up instead of running 30 transformer layers again.
        hidden = seq.get("_esm_hidden") if hasattr(seq, "get") else None
        if hidden is None:
            hidden = self.esm_model(input_ids=seq["input_ids"], attention_mask=seq["attention_mask"]).last_hidden_state
            if hasattr(seq, "get") and seq.get("_cache_esm") and self._esm_is_replayable():
                seq["_esm_hidden"] = hidden.detach()
        return self.amino_acid_ffn(self.projection(hidden))

    def _hidden(self, seq):
        """ESM hidden states of `seq` (the cached ones during a GradCache replay, see _residues)."""
        hidden = seq.get("_esm_hidden") if hasattr(seq, "get") else None
        if hidden is None:
            hidden = self.esm_model(input_ids=seq["input_ids"], attention_mask=seq["attention_mask"]).last_hidden_state
        return hidden

    def _fused_head_ok(self, pooled_ffn):
        """Inference (eval mode, no autograd graph wanted): the whole head runs as two fused launches
        (pclip.ops.encoder_head).  Training keeps the stock modules: their backward is autograd's.
        Under autocast the reference's Linears run in half precision; the fused head computes them in fp32."""
        if os.environ.get("PCLIP_FUSED_HEAD", "1") == "0" or torch.is_grad_enabled() or self.training:
            return False      # (training mode even without grad: GradCache's first pass must equal its replay)
        return self.projection.weight.is_cuda and ops.encoder_head_supported(
            torch.empty(0, self.projection.in_features, device=self.projection.weight.device), self.projection,
            self.amino_acid_ffn, pooled_ffn)


class Encoder(_Trunk):
    """Sequence encoder of ExtendedCLIP.  forward(seq) with seq = {'input_ids', 'attention_mask', 'temperature'} returns
    F.normalize(pooled) * exp(temperature / 2) as one fused kernel (ref: modules/models.py:32-52)."""

    def __init__(self, input_dim, embedding_dim, h1, h2, dropout_rate, esm_model):
        super().__init__(input_dim, embedding_dim, h1, h2, dropout_rate, esm_model)

    def forward(self, seq):
        mask = seq["attention_mask"]
        if self._fused_head_ok(self.embedding_ffn):
            return ops.encoder_head(self._hidden(seq), mask, self.projection, self.amino_acid_ffn, self.embedding_ffn,
                                    seq["temperature"], norm_eps=1e-12)
        pooled = self._masked_mean(self._residues(seq), mask)
        return ops.normalize_scale(self.embedding_ffn(pooled), seq["temperature"], eps=1e-12)

    @staticmethod
    def _masked_mean(h, mask):
        """Mean over the valid residues (count clamped to >= 1), ref: modules/models.py:75-90."""
        w = mask.unsqueeze(-1).to(h.dtype)
        return (h * w).sum(dim=1) / mask.sum(dim=1, keepdim=True).clamp(min=1)


class ExtendedCLIP(nn.Module):
    """Peptide / receptor twin encoders with a shared learnable temperature (ref: modules/models.py:92-130)."""

    def __init__(self, input_dim, embedding_dim, h1, h2, dropout, esm_model):
        super().__init__()
        self.pep_encoder = Encoder(input_dim, embedding_dim, h1, h2, dropout, esm_model)
        self.rec_encoder = Encoder(input_dim, embedding_dim, h1, h2, dropout, esm_model)
        self.temperature = nn.Parameter(torch.tensor(1.0))

    def forward(self, pep_seq, rec_seq):
        # the reference inserts the temperature into the callers' dicts; scripts rely on that side effect
        pep_seq["temperature"] = rec_seq["temperature"] = self.temperature
        return self.pep_encoder(pep_seq), self.rec_encoder(rec_seq)


class FILIPEncoder(_Trunk):
    """Token-level encoder of ExtendedFILIP: returns (per-residue embeddings [B,T,D], attention mask [B,T]); no pooling
    and no normalisation here (ref: modules/models.py:178-196).  `embedding_ffn` exists only for checkpoint parity."""

    def __init__(self, input_dim, embedding_dim, h1, h2, dropout_rate, esm_model):
        super().__init__(input_dim, embedding_dim, h1, h2, dropout_rate, esm_model)

    def forward(self, seq):
        if self._fused_head_ok(None):
            return ops.encoder_head(self._hidden(seq), seq["attention_mask"], self.projection,
                                    self.amino_acid_ffn), seq["attention_mask"]
        return self._residues(seq), seq["attention_mask"]


class ExtendedFILIP(nn.Module):
    """FILIP twin encoders; forward returns (simA, simB), both [B_pep, B_rec] (ref: modules/models.py:198-356)."""

    def __init__(self, input_dim, embedding_dim, h1, h2, dropout, esm_model):
        super().__init__()
        self.pep_encoder = FILIPEncoder(input_dim, embedding_dim, h1, h2, dropout, esm_model)
        self.rec_encoder = FILIPEncoder(input_dim, embedding_dim, h1, h2, dropout, esm_model)
        self.temperature = nn.Parameter(torch.tensor(1.0))

    def forward(self, pep_seq, rec_seq):
        pep_seq["temperature"] = rec_seq["temperature"] = self.temperature
        pep, pep_mask = self.pep_encoder(pep_seq)
        rec, rec_mask = self.rec_encoder(rec_seq)
        # h / ||h|| with NO eps, pads included (ref: :238-239): normalize_scale with tau=None, eps=0
        pep = ops.normalize_scale(pep, None, eps=0.0)
        rec = ops.normalize_scale(rec, None, eps=0.0)
        # tokenizer masks are right-padded (padding=True): lengths go straight to the kernel, no compaction pass
        return ops.filip_similarity(pep, rec, pep_mask, rec_mask, self.temperature)

    @staticmethod
    def _masked_mean(t, mask, dim=1, eps=1e-6):
        """Zero-fill the masked entries, sum over `dim`, divide by the clamped count (ref: :247-264).  Inside the
        similarity kernels this is the fused epilogue; the standalone form serves callers of the static method."""
        kept = torch.where(mask, t, torch.zeros((), dtype=t.dtype, device=t.device))
        return kept.sum(dim=dim) / mask.sum(dim=dim).clamp(min=eps)

    @staticmethod
    def _mean_average_similarity_score(hA, hB, maskA, maskB, temperature):
        """Mean-pooled baseline similarity (unused by the scripts; ref: :266-290): plain means over T of the masked
        tokens, then all-pairs dot products / temperature.  Differentiable like the reference's einsum (a [B, B] matmul
        off the hot path: stock torch)."""
        pa = (hA * maskA[..., None]).mean(dim=1)
        pb = (hB * maskB[..., None]).mean(dim=1)
        return pa @ pb.t() / temperature

    @staticmethod
    def _filip_similarity_score(hA, hB, maskA, maskB, temperature, include_group=False):
        """Token-wise late interaction (ref: :292-356): S = <hA[a,t], hB[b,s]> / temperature on the valid tokens;
        simA = masked mean over t of max_s, simB = masked mean over s of max_t.

        Any 0/1 mask is accepted, as in the reference: the valid tokens of every sample are moved to the front first
        (max and mean do not depend on token order; the gather is differentiable), so the kernels only ever see
        lengths.  include_group=True takes [m, b, t, d] / [n, b, t, d] inputs and returns [m, n, bA, bB] (:314-322): the
        group pairs are evaluated one after the other."""
        if include_group:
            rows_a, rows_b = [], []
            for m in range(hA.shape[0]):
                pa, pb = zip(*(ExtendedFILIP._filip_similarity_score(hA[m], hB[n], maskA[m], maskB[n], temperature)
                               for n in range(hB.shape[0])))
                rows_a.append(torch.stack(pa))
                rows_b.append(torch.stack(pb))
            return torch.stack(rows_a), torch.stack(rows_b)
        hA, lenA = ops.compact_tokens(hA, maskA)
        hB, lenB = ops.compact_tokens(hB, maskB)
        return ops.filip_similarity(hA, hB, lenA, lenB, temperature)
