"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

Runs `/root/reference/networks.py`'s define_G / define_D / GANLoss through the iteration
body of `/root/reference/train.py:294-390` (VGG term off: its weights cannot be downloaded
offline, networks.py:35; `train.py` itself cannot be imported -- numpy.core.numeric.Inf at
train.py:7, skimage at train.py:19 -- so the loop body is re-typed here around the real
modules) on the weights / inputs that `oracle.pix2pix_oracle` generates, and stores what the
reference returns.  `/root/reference` does not exist on the GPU box; only the .npz files
travel.  Usage:  python oracle/make_golden.py
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")

import networks as ref  # noqa: E402  (the reference, unmodified)
from oracle import pix2pix_oracle as orc  # noqa: E402

from oracle.make_golden_util import CASES, fingerprint  # noqa: E402


def run_case(name, batch, h, w, steps, seed):
    torch.manual_seed(0)
    net_g = ref.define_G("normal", 0.02, gpu_id="cpu")
    net_d = ref.define_D(6, 64, gpu_id="cpu")
    net_g.load_state_dict(orc.init_g_state(123))
    net_d.load_state_dict(orc.init_d_state(124))
    crit_gan = ref.GANLoss(tensor=torch.FloatTensor)      # default tensor type needs CUDA (networks.py:810)
    crit_feat = nn.L1Loss()
    opt_g = torch.optim.Adam(net_g.parameters(), lr=2e-4, betas=(0.5, 0.999))
    opt_d = torch.optim.Adam(net_d.parameters(), lr=2e-4, betas=(0.5, 0.999))
    net_g.train(); net_d.train()
    out = {}
    for it in range(steps):
        real_a, real_b, compressed = orc.synthetic_batch(batch, h, w, seed + it)
        # ---- train.py:300-390, VGG term removed -------------------------------------
        fake_b = net_g(compressed.detach())
        fake_ab_pool = torch.cat((real_a, fake_b.detach()), 1)
        pred_fake = net_d.forward(fake_ab_pool)
        loss_d_fake = crit_gan(pred_fake, False)
        real_ab = torch.cat((real_a, real_b.detach()), 1)
        pred_real = net_d.forward(real_ab)
        loss_d_real = crit_gan(pred_real, True)
        loss_d = (loss_d_fake + loss_d_real) * 0.5
        fake_ab = torch.cat((real_a, fake_b), 1)
        pred_fake = net_d.forward(fake_ab)
        loss_g_gan = crit_gan(pred_fake, True)
        loss_feat = 0
        for i in range(3):
            for j in range(len(pred_fake[i]) - 1):
                loss_feat += (1.0 / 3) * (4.0 / 4) * crit_feat(pred_fake[i][j], pred_real[i][j].detach()) * 10.0
        loss_g = loss_g_gan + loss_feat
        tv = torch.mean(torch.abs(fake_b[:, :, :, :-1] - fake_b[:, :, :, 1:])) \
            + torch.mean(torch.abs(fake_b[:, :, :-1, :] - fake_b[:, :, 1:, :]))
        loss_g = loss_g + tv * 1.0
        opt_g.zero_grad(); loss_g.backward()
        g_grads = {k: p.grad.clone() for k, p in net_g.named_parameters() if p.grad is not None}
        opt_g.step()
        opt_d.zero_grad(); loss_d.backward()
        d_grads = {k: p.grad.clone() for k, p in net_d.named_parameters() if p.grad is not None}
        opt_d.step()
        # ------------------------------------------------------------------------------
        p = f"s{it}."
        out[p + "fake_b"] = fake_b.detach().numpy()
        for nm, v in (("loss_d", loss_d), ("loss_g_gan", loss_g_gan), ("loss_feat", loss_feat),
                      ("loss_tv", tv), ("loss_g", loss_g)):
            out[p + nm] = np.float64(v.item())
        for i in range(3):
            for j in range(5):
                out[p + f"pred_fake.{i}.{j}"] = fingerprint(pred_fake[i][j])
                out[p + f"pred_real.{i}.{j}"] = fingerprint(pred_real[i][j])
        for k, v in g_grads.items():
            out[p + "ggrad." + k] = fingerprint(v)
        for k, v in d_grads.items():
            out[p + "dgrad." + k] = fingerprint(v)
        for k, v in net_g.state_dict().items():
            out[p + "gstate." + k] = fingerprint(v)
        for k, v in net_d.state_dict().items():
            out[p + "dstate." + k] = fingerprint(v)
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", f"ref_step_{name}.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    torch.set_num_threads(8)
    for c in CASES:
        run_case(*c)
