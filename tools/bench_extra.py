"""Numbers for BASELINE.json configs 4 and 5 (not bench.py lines): 512x512 training step and generator-only
inference at 256x256, batch 64.  CUDA events, warm-up first, synthetic inputs resident on the device."""
import contextlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import p2p_pytorch_b200 as P

dev = torch.device("cuda", 0)
torch.manual_seed(123)
with contextlib.redirect_stdout(sys.stderr):
    g = P.define_G("normal", 0.02, gpu_id=dev)
    d = P.define_D(6, 64, gpu_id=dev)
gen = torch.Generator(device=dev).manual_seed(5)


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


out = {}
# config 5: generator inference (test.py path), eval-mode BatchNorm, no autograd
g.eval()
x = torch.rand(64, 3, 256, 256, device=dev, generator=gen) * 2 - 1
with torch.no_grad():
    ms = timed(lambda: g(x), 20)
out["infer_b64_256"] = {"ms_per_batch": ms, "images_per_s": 64 / ms * 1e3}
# config 4: 512x512 training step, 4 and 8 images per GPU
g.train()
for B in (4, 8):
    tr = P.Pix2PixTrainer(g, d, cuda_graph=True)
    a = torch.rand(B, 3, 512, 512, device=dev, generator=gen) * 2 - 1
    b = torch.rand(B, 3, 512, 512, device=dev, generator=gen) * 2 - 1
    c = torch.round((b * 0.5 + 0.5).clamp(0, 1) * 7) / 7
    ms = timed(lambda: tr.step(a, b, c), 10, warmup=4)
    out[f"train_b{B}_512"] = {"ms_per_step": ms, "images_per_s": B / ms * 1e3}
    del tr
# the iteration as the reference executes it: front end + G/D phases + phase 3, 16 x 256^2
with contextlib.redirect_stdout(sys.stderr):
    c = P.define_C("normal", 0.02, gpu_id=dev)
c.train()
tr = P.Pix2PixTrainer(g, d, cuda_graph=True, net_c=c)
a = torch.rand(16, 3, 256, 256, device=dev, generator=gen) * 2 - 1
b = torch.rand(16, 3, 256, 256, device=dev, generator=gen) * 2 - 1
ms = timed(lambda: tr.step_full(a, b), 20, warmup=4)
out["train_full_b16_256"] = {"ms_per_step": ms, "images_per_s": 16 / ms * 1e3}
ms = timed(lambda: c.front_end(b), 20)
out["front_end_b16_256"] = {"ms": ms}
print(json.dumps(out))
