"""Loss-curve envelope of the ORACLE for the 200-step tracking test (test infrastructure; build container, CPU).

north_star asks that the loss curves of the CUDA path track the reference over 200 steps.  GAN training amplifies
rounding differences, so two runs of the reference itself that differ by rounding-level noise separate after ~10
iterations; what "tracks" can mean is therefore "lies where runs of the reference lie".  This script makes that
region explicit: K runs of the oracle (the fp32 restatement of train.py:294-390, pinned to the unmodified reference by
tests/golden/ref_step_*.npz) on the SAME batches, run 0 from the stock initial weights and runs 1..K-1 from weights
perturbed by relative noise of bf16 rounding size (2^-9), at batch 8, 64x64 -- large enough that the window means
of the losses are tight.  The per-step losses of every run are stored; the test builds the per-window envelope from
them.  Usage: python oracle/make_envelope.py [K]   (about 2 minutes of CPU per run)
"""
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import pix2pix_oracle as orc  # noqa: E402

BATCH, SIZE, STEPS, DATA_SEED = 8, 64, 200, 3000
KEYS = ("loss_d", "loss_g", "loss_g_gan", "loss_feat", "loss_tv")


def perturbed(state, seed, rel=2.0 ** -9):
    if seed == 0:
        return state
    g = torch.Generator().manual_seed(seed)
    out = type(state)()
    for k, v in state.items():
        if v.dtype.is_floating_point and "running" not in k and not k.endswith(("weight_u", "weight_v")):
            out[k] = v * (1.0 + rel * torch.randn(v.shape, generator=g))
        else:
            out[k] = v.clone()
    return out


def run(k):
    ot = orc.OracleTrainer(perturbed(orc.init_g_state(123), 10 * k), perturbed(orc.init_d_state(124), 10 * k + 1))
    rec = np.zeros((STEPS, len(KEYS)))
    for it in range(STEPS):
        r = ot.step(*orc.synthetic_batch(BATCH, SIZE, SIZE, DATA_SEED + it))
        rec[it] = [r[key].item() for key in KEYS]
    return rec


if __name__ == "__main__":
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    torch.set_num_threads(os.cpu_count() or 1)
    curves = []
    for k in range(K):
        t0 = time.time()
        curves.append(run(k))
        print(f"run {k}: {time.time() - t0:.0f} s, mean loss_d {curves[-1][:, 0].mean():.4f} loss_g {curves[-1][:, 1].mean():.4f}",
              flush=True)
        np.savez_compressed(os.path.join(os.path.dirname(HERE), "tests", "golden", "oracle_envelope_b8_64x64_200.npz"),
                            curves=np.stack(curves), keys=np.array(KEYS), batch=BATCH, size=SIZE, data_seed=DATA_SEED)
