"""Times the batched weight pack / gradient unpack launches at the discriminator's and generator's shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import p2p_pytorch_b200 as P
from p2p_pytorch_b200 import _lib as L, engine as E
from p2p_pytorch_b200._lib import call_multi, ptr

DEV = "cuda"
dt = torch.bfloat16


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


d_shapes = [(64, 6, 4), (128, 64, 4), (256, 128, 4), (512, 256, 4), (1, 512, 4)] * 3
g_shapes = [(32, 12, 9), (64, 32, 3), (128, 64, 3)] + [(128, 128, 3)] * 18 + [(64, 128, 3), (32, 64, 3), (3, 32, 9)]
for name, shapes in (("D (15 layers)", d_shapes), ("G (24 layers)", g_shapes)):
    pj, uj, keep = [], [], []
    for cout, cin, k in shapes:
        w = torch.randn(cout, cin, k, k, device=DEV) * 0.1
        Co, Ci = E.pad16(cout), E.pad16(cin)
        wp = torch.empty(Co, k, k, Ci, dtype=dt, device=DEV)
        wt = torch.empty(Ci, k, k, Co, dtype=dt, device=DEV)
        acc = torch.randn(k, k, Ci, Co, device=DEV)
        g = torch.zeros(cout, cin, k, k, device=DEV)
        dot = torch.zeros(1, device=DEV)
        sig = torch.ones(1, device=DEV)
        pj.append(L.PackJob(ptr(w), ptr(sig), ptr(wp), ptr(wt), cout, cin, Co, Ci, k, k))
        uj.append(L.UnpackJob(ptr(acc), ptr(g), ptr(w), ptr(sig), ptr(dot), cout, cin, Co, Ci, k, k))
        keep.append((w, wp, wt, acc, g, dot, sig))
    mb = sum(w.numel() * 4 + wp.numel() * 2 + wt.numel() * 2 for w, wp, wt, *_ in keep) / 1e6
    t_p = timeit(lambda: call_multi("p2p_pack_weights_multi", pj, L.BF16))
    t_u = timeit(lambda: call_multi("p2p_unpack_wgrad_multi", uj))
    print(f"{name}: pack {t_p:6.1f} us ({mb:.1f} MB moved), unpack {t_u:6.1f} us")
