"""`modules.seed.set_seed` (ref: modules/seed.py:5-23): seeds python / numpy / torch; like the reference it turns cuDNN
off globally on CUDA machines (the encoders use no convolutions, so nothing on the hot path is affected)."""
import random

import numpy as np
import torch


def set_seed(seed_value=42):
    for seeder in (random.seed, np.random.seed, torch.manual_seed):
        seeder(seed_value)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed_value)
        cudnn = torch.backends.cudnn
        cudnn.deterministic, cudnn.benchmark, cudnn.enabled = True, False, False
    return seed_value
