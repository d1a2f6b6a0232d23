"""Offline stand-ins for the HF pieces the entry scripts download (TEST INFRASTRUCTURE): a 33-token ESM-style
tokenizer (vocab order of facebook/esm2_*: <cls> <pad> <eos> <unk> L A G V S E R T I D P K Q N F Y M H W C X B U Z O . -
<null_1> <mask>; adds <cls>/<eos>, right-pads with id 1) and a randomly initialised 2-layer EsmModel."""
import torch

VOCAB = ["<cls>", "<pad>", "<eos>", "<unk>"] + list("LAGVSERTIDPKQNFYMHWCXBUZO.-") + ["<null_1>", "<mask>"]
TOK = {t: i for i, t in enumerate(VOCAB)}


class Encoding(dict):
    def to(self, device):
        return Encoding({k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in self.items()})


class TinyTokenizer:
    def __call__(self, seqs, return_tensors="pt", padding=True):
        rows = [[0] + [TOK.get(c, 3) for c in s] + [2] for s in seqs]
        T = max(len(r) for r in rows)
        ids = torch.tensor([r + [1] * (T - len(r)) for r in rows], dtype=torch.int64)
        return Encoding(input_ids=ids, attention_mask=(ids != 1).to(torch.int64))


def tiny_esm(hidden=40, seed=0):
    from transformers import EsmConfig, EsmModel
    torch.manual_seed(seed)
    cfg = EsmConfig(vocab_size=33, hidden_size=hidden, num_hidden_layers=2, num_attention_heads=4,
                    intermediate_size=2 * hidden, position_embedding_type="rotary", token_dropout=True, pad_token_id=1,
                    mask_token_id=32, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = EsmModel(cfg, add_pooling_layer=False)
    for p in m.parameters():
        p.requires_grad_(False)     # ref: main.py:22-23 freezes the ESM trunk
    return m


def loader(n_batches, batch, seed=0, pep_len=(3, 12), rec_len=(8, 30)):
    from pclip import synth
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_batches):
        lp = torch.randint(pep_len[0], pep_len[1] + 1, (batch,), generator=g)
        lr = torch.randint(rec_len[0], rec_len[1] + 1, (batch,), generator=g)
        out.append((tuple(synth.random_sequences(lp, g)), tuple(synth.random_sequences(lr, g))))
    return out
