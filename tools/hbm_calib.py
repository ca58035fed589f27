"""Calibration: what do plain torch elementwise kernels reach at the activation sizes of this network?"""
import torch
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def t(fn, cold, iters=10):
    for _ in range(2): fn()
    tot = 0.0
    for _ in range(iters):
        if cold: flush.zero_()
        else: torch.cuda._sleep(200000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize(); tot += e0.elapsed_time(e1)
    return tot / iters * 1e3
for mb in (16.8, 33.5, 67, 268):
    n = int(mb * 1e6 / 2)
    a = torch.randn(n, device=dev).bfloat16(); b = torch.randn(n, device=dev).bfloat16(); c = torch.empty_like(a); d = torch.randn(n, device=dev).bfloat16()
    for name, fn, streams in (("copy", lambda: c.copy_(a), 2), ("add3", lambda: torch.add(a, b, out=c), 3), ("addcmul4", lambda: torch.addcmul(a, b, d, out=c), 4)):
        for cold in (1, 0):
            us = t(fn, cold)
            print(f"{mb:6.1f} MB/tensor {name:9s} cold={cold}: {us:6.1f} us -> {streams * mb / us * 1e-3 * 1e3:6.0f} GB/s")
