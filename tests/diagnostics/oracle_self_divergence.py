"""How far do two runs of the ORACLE drift apart over 200 steps when the initial weights differ by 1e-6
relative noise?  Calibrates the loss-curve tracking test (GAN dynamics amplify rounding differences)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from oracle import pix2pix_oracle as orc

torch.set_num_threads(4)
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
g0, d0 = orc.init_g_state(123), orc.init_d_state(124)
gen = torch.Generator().manual_seed(9)
g1 = {k: (v * (1 + 1e-6 * torch.randn(v.shape, generator=gen)) if v.is_floating_point() else v) for k, v in g0.items()}
d1 = {k: (v * (1 + 1e-6 * torch.randn(v.shape, generator=gen)) if v.is_floating_point() else v) for k, v in d0.items()}
t0, t1 = orc.OracleTrainer(g0, d0), orc.OracleTrainer(g1, d1)
c0, c1 = [], []
for it in range(steps):
    a, b, c = orc.synthetic_batch(2, 32, 32, 1000 + it)
    r0, r1 = t0.step(a, b, c), t1.step(a, b, c)
    c0.append([r0["loss_d"].item(), r0["loss_g"].item()]); c1.append([r1["loss_d"].item(), r1["loss_g"].item()])
c0, c1 = np.array(c0), np.array(c1)
np.savez("profiles/oracle_self_divergence_200.npz", base=c0, perturbed=c1)
for lo in range(0, steps, 20):
    print(lo, np.round(c0[lo:lo+20].mean(0), 4), np.round(c1[lo:lo+20].mean(0), 4))
