"""Writes tests/golden/interface.json from the reference checkout (authoring container only; TEST INFRASTRUCTURE):
the public signatures of the hot-path modules and the state_dict keys of ExtendedCLIP / ExtendedFILIP built around a
2-layer random ESM, so that tests/test_interface.py can check the drop-in `modules` package without the reference."""
import inspect
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

from oracle import ref_loader  # noqa: E402


def tiny_esm():
    from transformers import EsmConfig, EsmModel
    cfg = EsmConfig(vocab_size=33, hidden_size=40, num_hidden_layers=2, num_attention_heads=4, intermediate_size=80,
                    position_embedding_type="rotary", token_dropout=True, pad_token_id=1, mask_token_id=32)
    return EsmModel(cfg, add_pooling_layer=False)


def signatures(mod, names):
    return {n: str(inspect.signature(getattr(mod, n))) for n in names}


def script_attrs():
    import re
    out = {}
    for script in ("main.py", "main_2protein.py", "main_2protein_filip.py", "topk.py", "cosine_similarity.py"):
        src = open(os.path.join(ref_loader.REFERENCE_ROOT, script)).read()
        for mod, attr in re.findall(r"\b(models|training_utils|training_utils_filip|visualizations|seed)\.(\w+)", src):
            out.setdefault(mod, set()).add(attr)
    return {m: sorted(v) for m, v in sorted(out.items())}


def main():
    ref = ref_loader.load()
    esm = tiny_esm()
    M = ref.models
    out = {
        "training_utils": signatures(ref.training_utils, ["train", "evaluate", "_process_batch", "_contrastive_loss",
                                                          "train_gc", "eval_gc_allrec_onepep"]),
        "training_utils_filip": signatures(ref.training_utils_filip, ["_contrastive_loss", "_compute_loss",
                                                                      "_process_batch", "train", "evaluate"]),
        "models": {
            "Encoder.__init__": str(inspect.signature(M.Encoder.__init__)),
            "ExtendedCLIP.__init__": str(inspect.signature(M.ExtendedCLIP.__init__)),
            "ExtendedCLIP.forward": str(inspect.signature(M.ExtendedCLIP.forward)),
            "FILIPEncoder.__init__": str(inspect.signature(M.FILIPEncoder.__init__)),
            "ExtendedFILIP.__init__": str(inspect.signature(M.ExtendedFILIP.__init__)),
            "ExtendedFILIP._masked_mean": str(inspect.signature(M.ExtendedFILIP._masked_mean)),
            "ExtendedFILIP._filip_similarity_score": str(inspect.signature(M.ExtendedFILIP._filip_similarity_score)),
        },
        "visualizations": signatures(ref.visualizations, ["_compute_embedding_cosine_similarities",
                                                          "_compute_embedding_cosine_similarities_filip",
                                                          "plot_embedding_cosine_similarities",
                                                          "plot_embedding_cosine_similarities_filip",
                                                          "plot_loss_curves", "plot_clustering",
                                                          "plot_protein_lengths", "_save_plot"]),
        # every attribute the entry scripts touch on the modules this repo replaces (the scripts must not
        # meet an AttributeError at start-up)
        "script_attrs": script_attrs(),
        "state_dict_keys": {
            "ExtendedCLIP": sorted(k for k in M.ExtendedCLIP(40, 128, 2, 2, 0.1, esm).state_dict()
                                   if "esm_model" not in k),
            "ExtendedFILIP": sorted(k for k in M.ExtendedFILIP(40, 128, 2, 2, 0.1, esm).state_dict()
                                    if "esm_model" not in k),
        },
    }
    path = os.path.join(REPO, "tests", "golden", "interface.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", path)


if __name__ == "__main__":
    main()
