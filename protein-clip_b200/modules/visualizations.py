"""Drop-in for the two similarity-matrix functions of modules/visualizations.py that sit on the hot path
(_compute_embedding_cosine_similarities :38-52, _compute_embedding_cosine_similarities_filip :54-73).  The plotting
functions of the reference are matplotlib / pandas / BioPython CPU code and are out of scope;
`plot_embedding_cosine_similarities*` here compute the matrix and save it as .npy next to where the PNG would go."""
from __future__ import annotations

import os
import time

import numpy as np
import torch
from pclip import ops


def _compute_embedding_cosine_similarities(model, peptides, receptors):
    """cos[i, j] = <pep_i, rec_j> * exp(-temperature): undoes the two exp(temperature / 2) factors of the encoder tail."""
    pep, rec = model(peptides, receptors)
    return ops.scores(pep, rec, "dot", neg_tau=model.temperature)


def _compute_embedding_cosine_similarities_filip(model, peptides, receptors):
    sim_a, sim_b = model(peptides, receptors)
    print(sim_a)
    print(sim_b)
    return (sim_a + sim_b) / 2


def _save_plot(base_path, array=None, stem="figure"):
    out_dir = os.path.join(base_path, "figures")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, f"{stem}_{time.strftime('%Y%m%d_%H%M%S')}.npy")
    if array is not None:
        np.save(path, array)
    return path


def _first_batch(data_loader, tokenizer, device):
    pep, rec = next(iter(data_loader))
    return (tokenizer(pep, return_tensors="pt", padding=True).to(device),
            tokenizer(rec, return_tensors="pt", padding=True).to(device))


def plot_embedding_cosine_similarities(base_path, title, data_loader, tokenizer, model, device):
    with torch.no_grad():
        m = _compute_embedding_cosine_similarities(model, *_first_batch(data_loader, tokenizer, device))
    print(f"{title} matrix saved to {_save_plot(base_path, m.cpu().numpy(), 'cosine')}")


def plot_embedding_cosine_similarities_filip(base_path, title, data_loader, tokenizer, model, device):
    with torch.no_grad():
        m = _compute_embedding_cosine_similarities_filip(model, *_first_batch(data_loader, tokenizer, device))
    print(f"{title} matrix saved to {_save_plot(base_path, m.cpu().numpy(), 'filip_similarity')}")
