"""Pieces shared by make_golden.py (build container, imports the reference) and the tests
(anywhere): the case list and the tensor fingerprint.  Test infrastructure only."""
import numpy as np
import torch

CASES = [  # name, batch, H, W, steps, data seed
    ("b2_32x32", 2, 32, 32, 2, 123),
    ("b1_64x48", 1, 64, 48, 1, 7),
    ("b3_32x40", 3, 32, 40, 1, 11),
]


def fingerprint(t: torch.Tensor) -> np.ndarray:
    """[sum, abs-sum, sum of squares, first 13 values] -- small, order-sensitive enough."""
    f = t.detach().double().flatten()
    head = torch.zeros(13, dtype=torch.float64)
    head[: min(13, f.numel())] = f[:13]
    return torch.cat([torch.stack([f.sum(), f.abs().sum(), (f * f).sum()]), head]).numpy()

C_CASES = [  # compression front end: name, batch, H, W, data seed, train-mode BatchNorm
    ("b2_32x32", 2, 32, 32, 123, True),
    ("b1_64x48_eval", 1, 64, 48, 7, False),
    ("b3_32x40", 3, 32, 40, 11, True),
]

P3_CASES = [  # whole iteration with front end and phase 3: name, batch, H, W, steps, data seed
    ("b2_32x32", 2, 32, 32, 2, 123),
    ("b1_64x48", 1, 64, 48, 1, 7),
]

PSNR_CASES = [  # name, H, W, seed, kind
    ("noise_32x32", 32, 32, 1, "noise"), ("close_64x48", 64, 48, 2, "close"), ("out_of_range_40x40", 40, 40, 3, "wide"),
    ("identical_16x16", 16, 16, 4, "same"),
]


def psnr_inputs(h, w, seed, kind):
    """two images [1,3,H,W]; 'wide' leaves [0,1] so that the clip matters, 'same' gives mse = 0 (PSNR = inf)"""
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(1, 3, h, w, generator=g)
    if kind == "noise":
        b = torch.rand(1, 3, h, w, generator=g)
    elif kind == "close":
        b = a + 0.02 * torch.randn(1, 3, h, w, generator=g)
    elif kind == "wide":
        a = a * 1.6 - 0.3
        b = a + 0.1 * torch.randn(1, 3, h, w, generator=g)
    else:
        b = a.clone()
    return a, b

VGG_CASES = [  # perceptual loss alone: name, batch, H, W, data seed (H, W multiples of 16: four 2x2 poolings)
    ("b2_32x32", 2, 32, 32, 31),
    ("b1_64x48", 1, 64, 48, 32),
]
