"""Per-layer timing of the convolution kernels at the benchmark shapes (B=16, 256x256), CUDA events on the
launching stream, L2 flushed between launches.  Prints achieved TFLOP/s per layer and pass.
    python tools/bench_convs.py [fwd,dgrad,wgrad] [layer-substring]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import p2p_pytorch_b200 as P
from p2p_pytorch_b200 import _lib as L, engine as E

B = int(os.environ.get("B", "16"))
S = 256
# name, cin, cout, k, stride, pad, Hin (stored, incl. halo for G), count per step (fwd)
LAYERS = [
    ("g.conv1", 12, 32, 9, 1, 0, S + 8),
    ("g.conv2", 32, 64, 3, 2, 0, S + 2),
    ("g.conv3", 64, 128, 3, 2, 0, S // 2 + 2),
    ("g.res", 128, 128, 3, 1, 0, S // 4 + 2),
    ("g.deconv3", 128, 64, 3, 1, 0, S // 2 + 2),
    ("g.deconv2", 64, 32, 3, 1, 0, S + 2),
    ("g.deconv1", 32, 3, 9, 1, 0, S + 8),
    ("d0.l0", 6, 64, 4, 2, 2, S),
    ("d0.l1", 64, 128, 4, 2, 2, 129),
    ("d0.l2", 128, 256, 4, 2, 2, 65),
    ("d0.l3", 256, 512, 4, 1, 2, 33),
    ("d0.l4", 512, 1, 4, 1, 2, 34),
    ("d1.l3", 256, 512, 4, 1, 2, 17),
]


def timeit(fn, iters=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        fn()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def main():
    passes = (sys.argv[1] if len(sys.argv) > 1 else "fwd,dgrad,wgrad").split(",")
    sel = sys.argv[2] if len(sys.argv) > 2 else ""
    dt = torch.bfloat16
    print(f"B={B}  P2P_NO_HALO={os.environ.get('P2P_NO_HALO', '0')}")
    for name, cin, cout, k, s, pad, Hin in LAYERS:
        if sel not in name:
            continue
        Ci = E.pad16(cin)
        plan = E.ConvPlan(B, Hin, Hin, Ci, cin, cout, k, s, pad)
        x = torch.randn(B, Hin, Hin, Ci, device="cuda").to(dt)
        w = torch.randn(cout, cin, k, k, device="cuda") * 0.05
        plan.pack(w, None, dt)
        out = torch.empty(B, plan.Ho, plan.Wo, plan.Co, dtype=dt, device="cuda")
        go = torch.randn_like(out)
        din = torch.empty_like(x)
        stats = torch.zeros(2 * plan.Co, dtype=torch.float64, device="cuda")
        acc = torch.zeros(k, k, Ci, plan.Co, device="cuda")
        flops = 2.0 * B * plan.Ho * plan.Wo * cout * cin * k * k
        res = []
        if "fwd" in passes:
            ms = timeit(lambda: plan.fwd(x, None, out, stats if name.startswith("g.") else None))   # D has no BatchNorm
            res.append(f"fwd {ms * 1e3:8.1f} us {flops / ms / 1e9:7.1f} TF/s")
        if "dgrad" in passes:
            ms = timeit(lambda: plan.dgrad(go, din))
            res.append(f"dgrad {ms * 1e3:8.1f} us {flops / ms / 1e9:7.1f} TF/s")
        if "wgrad" in passes:
            ms = timeit(lambda: L.call("p2p_conv_wgrad", C.byref(plan.desc), x.data_ptr(), go.data_ptr(), acc.data_ptr(),
                                       L.BF16, L.IMPL_AUTO))
            res.append(f"wgrad {ms * 1e3:8.1f} us {flops / ms / 1e9:7.1f} TF/s")
        print(f"{name:10s} {cin:3d}->{cout:3d} k{k} s{s} out {plan.Ho:3d}x{plan.Wo:<3d} GF {flops / 1e9:7.2f} | " + " | ".join(res))


if __name__ == "__main__":
    main()
