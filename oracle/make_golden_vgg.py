"""Golden vectors for the VGG19 perceptual loss from the UNMODIFIED reference classes
`VGGLoss` / `Vgg19` (/root/reference/networks.py:18-62) -- build container only.

`Vgg19.__init__` calls `models.vgg19(pretrained=True)` (networks.py:35), a download that is impossible offline.  The
class itself is left untouched: `networks.models.vgg19` is replaced, for the duration of the construction, by a factory
that returns torchvision's own vgg19 architecture (`weights=None`) loaded with the seeded tensors of
`oracle.pix2pix_oracle.init_vgg_state`.  Everything after that -- the five slices, the 1/32 ... 1 weights, L1 against the
detached target, and the use inside the training iteration (train.py:377-380) -- is the reference's own code.

Writes tests/golden/ref_vgg_<case>.npz (loss, gradient w.r.t. x, fingerprints of the five taps) and
tests/golden/ref_stepvgg_b2_32x32.npz (one whole iteration with the perceptual term on).
Usage:  python oracle/make_golden_vgg.py
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn
import torchvision

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")

import networks as ref  # noqa: E402  (the reference, unmodified)
from oracle import pix2pix_oracle as orc  # noqa: E402
from oracle.make_golden_util import VGG_CASES, fingerprint  # noqa: E402


def reference_vgg_loss(state):
    """the reference's VGGLoss('cpu') holding the seeded weights"""
    stock = ref.models.vgg19              # (`ref.models` is the torchvision.models module itself)

    def factory(pretrained=True, **_kw):
        net = stock(weights=None)
        with torch.no_grad():
            for k, v in state.items():
                idx, leaf = k.split(".")[2:4]
                getattr(net.features[int(idx)], leaf).copy_(v)
        return net
    ref.models.vgg19 = factory
    try:
        crit = ref.VGGLoss("cpu")
    finally:
        ref.models.vgg19 = stock
    assert list(crit.state_dict().keys()) == list(state.keys()), "key naming differs from the reference module"
    for k, v in crit.state_dict().items():
        assert torch.equal(v, state[k]), k
    return crit


def loss_cases(crit):
    for name, batch, h, w, seed in VGG_CASES:
        a, b, _ = orc.synthetic_batch(batch, h, w, seed)
        x = torch.tanh(a * 1.5).requires_grad_(True)       # a generator-like output in (-1, 1)
        loss = crit(x, b)
        loss.backward()
        out = {"loss": np.float64(loss.item()), "grad_x": x.grad.numpy()}
        for i, t in enumerate(crit.vgg(x.detach())):
            out[f"tap{i}"] = fingerprint(t)
        path = os.path.join(os.path.dirname(HERE), "tests", "golden", f"ref_vgg_{name}.npz")
        np.savez_compressed(path, **out)
        print("wrote", path, os.path.getsize(path), "bytes, loss", loss.item())


def step_case(crit, name="b2_32x32", batch=2, h=32, w=32, seed=123):
    net_g = ref.define_G("normal", 0.02, gpu_id="cpu")
    net_d = ref.define_D(6, 64, gpu_id="cpu")
    net_g.load_state_dict(orc.init_g_state(123))
    net_d.load_state_dict(orc.init_d_state(124))
    crit_gan = ref.GANLoss(tensor=torch.FloatTensor)
    crit_feat = nn.L1Loss()
    opt_g = torch.optim.Adam(net_g.parameters(), lr=2e-4, betas=(0.5, 0.999))
    opt_d = torch.optim.Adam(net_d.parameters(), lr=2e-4, betas=(0.5, 0.999))
    net_g.train(); net_d.train()
    real_a, real_b, compressed = orc.synthetic_batch(batch, h, w, seed)
    # ---- train.py:300-390 with the perceptual term -------------------------------------------
    fake_b = net_g(compressed.detach())
    pred_fake = net_d.forward(torch.cat((real_a, fake_b.detach()), 1))
    loss_d_fake = crit_gan(pred_fake, False)
    pred_real = net_d.forward(torch.cat((real_a, real_b.detach()), 1))
    loss_d = (loss_d_fake + crit_gan(pred_real, True)) * 0.5
    pred_fake = net_d.forward(torch.cat((real_a, fake_b), 1))
    loss_g_gan = crit_gan(pred_fake, True)
    loss_feat = 0
    for i in range(3):
        for j in range(len(pred_fake[i]) - 1):
            loss_feat += (1.0 / 3) * (4.0 / 4) * crit_feat(pred_fake[i][j], pred_real[i][j].detach()) * 10.0
    loss_g = loss_g_gan + loss_feat
    content_loss = crit(fake_b, real_b) * 10.0                                              # train.py:377
    tv = torch.mean(torch.abs(fake_b[:, :, :, :-1] - fake_b[:, :, :, 1:])) \
        + torch.mean(torch.abs(fake_b[:, :, :-1, :] - fake_b[:, :, 1:, :]))
    loss_g = loss_g + content_loss + tv * 1.0                                               # :380
    opt_g.zero_grad(); loss_g.backward()
    g_grads = {k: p.grad.clone() for k, p in net_g.named_parameters() if p.grad is not None}
    opt_g.step()
    opt_d.zero_grad(); loss_d.backward(); opt_d.step()
    out = {"fake_b": fake_b.detach().numpy()}
    for nm, v in (("loss_d", loss_d), ("loss_g_gan", loss_g_gan), ("loss_feat", loss_feat), ("loss_tv", tv),
                  ("loss_vgg", content_loss), ("loss_g", loss_g)):
        out[nm] = np.float64(v.item())
    for k, v in g_grads.items():
        out["ggrad." + k] = fingerprint(v)
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", f"ref_stepvgg_{name}.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    torch.set_num_threads(8)
    torch.manual_seed(0)
    crit = reference_vgg_loss(orc.init_vgg_state(126))
    loss_cases(crit)
    step_case(crit)
