"""Per-stage error growth of the generator (forward activations and backward gradients) against an fp32 torch
graph on the same device.  Diagnostic only (not part of the product): python tests/diagnostics/debug_g.py [precision]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.nn.functional as F

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

import p2p_pytorch_b200 as P
from p2p_pytorch_b200 import engine as E
from oracle import pix2pix_oracle as orc

DEV = "cuda"


def rms(a, b):
    a, b = a.double(), b.double()
    return ((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt().clamp_min(1e-30)).item()


def mx(a, b):
    a, b = a.double(), b.double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def ref_forward(s, x):
    """orc.g_forward with every stage's post-activation output recorded (NCHW)"""
    acts = []
    a = s["relu.weight"]

    def stage(y, conv, bn, stride=1, up=0):
        if up:
            y = F.interpolate(y, scale_factor=up, mode="nearest")
        y = orc._reflect_conv(y, s[conv + ".conv2d.weight"], s[conv + ".conv2d.bias"], stride)
        return orc._bn(y, s, bn, True)

    y = F.interpolate(orc.pixel_unshuffle2(x), scale_factor=2, mode="nearest")
    y = F.prelu(stage(y, "conv1", "in1_e"), a); acts.append(("conv1", y))
    y = F.prelu(stage(y, "conv2", "in2_e", 2), a); acts.append(("conv2", y))
    y = F.prelu(stage(y, "conv3", "in3_e", 2), a); acts.append(("conv3", y))
    long_skip = y
    for r in range(1, 10):
        t = F.relu(stage(y, f"res{r}.conv1", f"res{r}.in1")); acts.append((f"res{r}.a", t))
        t = stage(t, f"res{r}.conv2", f"res{r}.in2")
        y = F.relu(t + y); acts.append((f"res{r}.b", y))
    y = F.leaky_relu(y + long_skip, 0.2); acts.append(("long_skip", y))
    y = F.prelu(stage(y, "deconv3", "in3_d", 1, 2), a); acts.append(("deconv3", y))
    y = F.prelu(stage(y, "deconv2", "in2_d", 1, 2), a); acts.append(("deconv2", y))
    y = torch.tanh(stage(y, "deconv1", "in1_d")); acts.append(("deconv1", y))
    for _, t in acts:
        t.retain_grad()
    return y, acts


def main():
    precision = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    B, S = 2, 32
    P.set_precision(precision)
    g = P.ExpandNetwork().to(DEV)
    g.load_state_dict(orc.init_g_state(123))
    g.train()
    _, _, c = orc.synthetic_batch(B, S, S, 5)
    c = c.to(DEV)
    s = {k: v.to(DEV) for k, v in orc.init_g_state(123).items()}
    for k in orc.g_trainable(s):
        s[k].requires_grad_(True)
    y_ref, acts = ref_forward(s, c)
    y = g(c)
    stages = y.grad_fn.stages
    print(f"== {precision}: forward, per stage: rms rel err / max err over max ref")
    for (name, ref), st in zip(acts, stages):
        if st.role == "long_skip" or name == "deconv1":
            continue
    # stage outputs live in the next stage's x_in (with halo / upsample): compare interiors
    for i, (name, ref) in enumerate(acts[:-1]):
        st = stages[i]
        nxt = stages[i + 1]
        buf = nxt.x_in if nxt.x_in is not None else nxt.r       # long_skip reads res9's compact output as r
        pad, up = st.pad, st.up
        t = buf.float()
        if pad:
            t = t[:, pad:-pad, pad:-pad]
        if up == 2:
            t = t[:, ::2, ::2]
        t = t[..., :ref.shape[1]].permute(0, 3, 1, 2)
        print(f"  {name:12s} rms {rms(t, ref):.3e}  max {mx(t, ref):.3e}")
    print(f"  {'output':12s} rms {rms(y, y_ref):.3e}  max {mx(y, y_ref):.3e}")
    torch.manual_seed(0)
    gy = torch.randn_like(y_ref)
    y_ref.backward(gy)
    y.backward(gy)
    print("== backward: parameter gradients, rms rel err")
    for name, p in g.named_parameters():
        print(f"  {name:28s} {rms(p.grad, s[name].grad):.3e}   |ref| {s[name].grad.abs().max().item():.3e}")


if __name__ == "__main__":
    main()
