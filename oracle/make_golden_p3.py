"""Generate tests/golden/ref_full_*.npz from the UNMODIFIED reference modules (build container only): the iteration
as train.py:294-402 executes it -- compression front end, G/D phases, phase 3 -- with the VGG terms off.  The loop
body is re-typed around the real define_G / define_D / define_C / GANLoss modules (train.py itself cannot be
imported, see make_golden.py).  Stored per iteration: the five losses, loss_c, fingerprints of fake_b, of the
generator input, of every BatchNorm buffer of G and of net_c, and whether any net_c parameter moved.
Usage:  python oracle/make_golden_p3.py
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
from oracle.make_golden_util import P3_CASES, fingerprint  # noqa: E402


def ref_compress(tensor, bit):                     # generate_dataset.py:29-34, verbatim arithmetic
    max_val = 2 ** bit - 1
    tensor = torch.clamp(tensor, 0.0, 1.0) * max_val
    tensor = torch.round(tensor)
    return tensor / max_val


def run_case(name, batch, h, w, steps, seed):
    torch.manual_seed(0)
    net_g = ref.define_G("normal", 0.02, gpu_id="cpu")
    net_d = ref.define_D(6, 64, gpu_id="cpu")
    net_c = ref.define_C("normal", 0.02, gpu_id="cpu")
    net_g.load_state_dict(orc.init_g_state(123))
    net_d.load_state_dict(orc.init_d_state(124))
    net_c.load_state_dict(orc.init_c_state(125))
    crit_gan = ref.GANLoss(tensor=torch.FloatTensor)
    crit_feat, mse_criterion = nn.L1Loss(), nn.MSELoss(reduction="mean")
    opt_g = torch.optim.Adam(net_g.parameters(), lr=2e-4, betas=(0.5, 0.999))
    opt_d = torch.optim.Adam(net_d.parameters(), lr=2e-4, betas=(0.5, 0.999))
    opt_c = torch.optim.Adam(net_c.parameters(), lr=2e-4, betas=(0.5, 0.999))
    net_g.train(); net_d.train(); net_c.train()
    c_before = {k: v.clone() for k, v in net_c.named_parameters()}
    out = {}
    for it in range(steps):
        real_a, real_b, _unused = orc.synthetic_batch(batch, h, w, seed + it)
        compressed = ref_compress(net_c(real_b), 3)                                   # train.py:297
        fake_b = net_g(compressed.detach())                                           # :300
        pred_fake = net_d.forward(torch.cat((real_a, fake_b.detach()), 1))
        loss_d_fake = crit_gan(pred_fake, False)
        pred_real = net_d.forward(torch.cat((real_a, real_b.detach()), 1))
        loss_d_real = crit_gan(pred_real, True)
        loss_d = (loss_d_fake + loss_d_real) * 0.5
        pred_fake = net_d.forward(torch.cat((real_a, fake_b), 1))
        loss_g_gan = crit_gan(pred_fake, True)
        loss_feat = 0
        for i in range(3):
            for j in range(len(pred_fake[i]) - 1):
                loss_feat += (1.0 / 3) * (4.0 / 4) * crit_feat(pred_fake[i][j], pred_real[i][j].detach()) * 10.0
        tv = torch.mean(torch.abs(fake_b[:, :, :, :-1] - fake_b[:, :, :, 1:])) \
            + torch.mean(torch.abs(fake_b[:, :, :-1, :] - fake_b[:, :, 1:, :]))
        loss_g = loss_g_gan + loss_feat + tv * 1.0
        opt_g.zero_grad(); loss_g.backward(); opt_g.step()
        opt_d.zero_grad(); loss_d.backward(); opt_d.step()
        fake_after = net_g(compressed)                                                # :392
        loss_c = mse_criterion(fake_after, real_b)                                    # :394
        opt_c.zero_grad(); loss_c.backward(); opt_c.step()                            # :400-402
        p = f"s{it}."
        for nm, v in (("loss_d", loss_d), ("loss_g_gan", loss_g_gan), ("loss_feat", loss_feat), ("loss_tv", tv),
                      ("loss_g", loss_g), ("loss_c", loss_c)):
            out[p + nm] = np.float64(v.item())
        out[p + "fake_b"] = fingerprint(fake_b)
        out[p + "compressed"] = fingerprint(compressed)
        out[p + "c_grad_max"] = np.float64(max(q.grad.abs().max().item() for q in net_c.parameters()))
        for k, v in list(net_g.state_dict().items()) + [("net_c." + k, v) for k, v in net_c.state_dict().items()]:
            if "running" in k or "num_batches" in k:
                out[p + "buf." + k] = fingerprint(v.double())
    out["c_params_moved"] = np.int64(sum(int(not torch.equal(c_before[k], v)) for k, v in net_c.named_parameters()))
    return out


if __name__ == "__main__":
    out_dir = os.path.join(os.path.dirname(HERE), "tests", "golden")
    for case in P3_CASES:
        res = run_case(*case)
        np.savez_compressed(os.path.join(out_dir, f"ref_full_{case[0]}.npz"), **res)
        print(case[0], {k: float(v) for k, v in res.items() if k.endswith(("loss_c", "loss_d", "c_grad_max"))},
              int(res["c_params_moved"]))
