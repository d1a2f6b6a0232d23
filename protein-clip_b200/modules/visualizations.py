"""Drop-in for modules/visualizations.py of the reference.

On the hot path are the two similarity-matrix functions (_compute_embedding_cosine_similarities :38-52,
_compute_embedding_cosine_similarities_filip :54-73): they call the CUDA library.  Everything else here is the CPU
plotting the entry scripts call (main.py:41,55,96; main_2protein*.py:40-42,51,92; topk.py:87) with the reference's names,
arguments, printed messages and file naming ("figures/figure N.png", one shared counter), so that the scripts run
unchanged.  matplotlib and BioPython are optional: without matplotlib the numbers behind a figure are saved as
"figure N.npy" beside the path that is returned; FASTA files are read by a ten-line parser."""
from __future__ import annotations

import math
import os

import numpy as np
import torch
from pclip import ops

try:  # the reference imports pyplot unconditionally (visualizations.py:2); here it is optional
    import matplotlib
    matplotlib.use("Agg")
    import matplotlib.pyplot as plt
except Exception:  # pragma: no cover - depends on the image
    plt = None

_pending = {}   # data of the figure being built, saved by _save_plot when there is no matplotlib


def _compute_embedding_cosine_similarities(model, peptides, receptors):
    """cos[i, j] = <pep_i, rec_j> * exp(-temperature): undoes the two exp(temperature / 2) factors of the encoder tail
    (ref: visualizations.py:50-52)."""
    pep, rec = model(peptides, receptors)
    return ops.scores(pep, rec, "dot", neg_tau=model.temperature)


def _compute_embedding_cosine_similarities_filip(model, peptides, receptors):
    """(simA + simB) / 2, printing both matrices as the reference does (visualizations.py:68-73)."""
    sim_a, sim_b = model(peptides, receptors)
    print(sim_a)
    print(sim_b)
    return (sim_a + sim_b) / 2


def _save_plot(base_path, fig_num=[1]):
    """Save the current figure as <base_path>/figures/figure N.png and advance the shared counter
    (ref: visualizations.py:191-210; the mutable default IS the counter, as in the reference)."""
    folder = os.path.join(base_path, "figures")
    os.makedirs(folder, exist_ok=True)
    path = os.path.join(folder, f"figure {fig_num[0]}.png")
    if plt is not None:
        plt.tight_layout()
        plt.savefig(path)
        plt.clf()
    elif _pending:
        np.save(path[:-4] + ".npy", np.asarray(_pending.pop("data"), dtype=np.float64))
    _pending.clear()
    fig_num[0] += 1
    return path


def _first_batch(data_loader, tokenizer, device):
    pep, rec = next(iter(data_loader))
    return (tokenizer(pep, return_tensors="pt", padding=True).to(device),
            tokenizer(rec, return_tensors="pt", padding=True).to(device))


def _matrix_figure(base_path, title, matrix):
    m = matrix.cpu().detach().numpy()
    if plt is not None:
        plt.figure(figsize=(6, 4))
        plt.imshow(m, cmap="ocean", vmin=-1, vmax=1)
        plt.colorbar()
        plt.title(title)
        plt.xlabel("Receptor Protein")
        plt.ylabel("Peptide")
    else:
        _pending["data"] = m
    print(f"{title} plot saved to {_save_plot(base_path)}")


def plot_embedding_cosine_similarities(base_path, title, data_loader, tokenizer, model, device):
    """First batch of the loader -> 16 x 16 cosine matrix -> figure (ref: visualizations.py:7-36)."""
    _matrix_figure(base_path, title,
                   _compute_embedding_cosine_similarities(model, *_first_batch(data_loader, tokenizer, device)))


def plot_embedding_cosine_similarities_filip(base_path, title, data_loader, tokenizer, model, device):
    """ref: visualizations.py:76-105."""
    _matrix_figure(base_path, title,
                   _compute_embedding_cosine_similarities_filip(model, *_first_batch(data_loader, tokenizer, device)))


def plot_loss_curves(base_path, train_losses, val_losses, train_batch_size, val_batch_size):
    """Losses divided by ln(batch), the loss of a random model (ref: visualizations.py:108-130)."""
    title = "Training and Validation Loss Relative to Random"
    tr = [v / math.log(train_batch_size) for v in train_losses]
    va = [v / math.log(val_batch_size) for v in val_losses]
    if plt is not None:
        plt.plot(tr, label="Train Loss")
        plt.plot(va, label="Validation Loss")
        plt.title(title)
        plt.xlabel("Epochs")
        plt.ylabel("Loss Relative to Random")
        plt.legend()
    else:
        n = max(len(tr), len(va))
        _pending["data"] = [tr + [float("nan")] * (n - len(tr)), va + [float("nan")] * (n - len(va))]
    print(f"{title} plot saved to {_save_plot(base_path)}")


def _histogram(base_path, values, title, xlabel, ylabel):
    if plt is not None:
        plt.hist(values, bins=50, log=True)
        plt.xlabel(xlabel)
        plt.ylabel(ylabel)
        plt.title(title)
    else:
        _pending["data"] = list(values)
    print(f"{title} plot saved to {_save_plot(base_path)}")


def plot_clustering(base_path, data_path, prefix="protein2"):
    """Histogram of mmseqs cluster sizes from <prefix>DB_clustered.tsv: cluster id <TAB> member id
    (ref: visualizations.py:132-158)."""
    sizes = {}
    n = 0
    with open(f"{data_path}/{prefix}DB_clustered.tsv") as f:
        for line in f:
            if line.strip():
                cid = line.rstrip("\n").split("\t")[0]
                sizes[cid] = sizes.get(cid, 0) + 1
                n += 1
    title = f"MMSeqs Clustering for Protein-Protein. NSeq={n}, NClusters={len(sizes)}"
    _histogram(base_path, list(sizes.values()), title, "Number of Protein-Protein Sequences in Cluster",
               "Number of Clusters")


def _fasta_lengths(path):
    lengths, cur = [], None
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith(">"):
                if cur is not None:
                    lengths.append(cur)
                cur = 0
            elif cur is not None:
                cur += len(line)
    if cur is not None:
        lengths.append(cur)
    return lengths


def plot_protein_lengths(base_path, data_dir, prefix1="protein1", prefix2="protein2"):
    """Histogram of the sequence lengths of the two FASTA files (ref: visualizations.py:160-188)."""
    lengths = _fasta_lengths(os.path.join(str(data_dir), f"{prefix1}.fasta")) + \
        _fasta_lengths(os.path.join(str(data_dir), f"{prefix2}.fasta"))
    title = f"Distribution of Protein Sequences Lengths in Dataset. NSeq={len(lengths)}"
    _histogram(base_path, lengths, title, "Length of Protein Sequence", "Number of Sequences")
