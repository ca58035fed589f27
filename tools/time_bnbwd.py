import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import p2p_pytorch_b200 as P
from p2p_pytorch_b200 import _lib as L
B, H, W, C = 16, 64, 64, 128
dt = torch.bfloat16
dpad = torch.randn(B, H + 2, W + 2, C, device="cuda").to(dt)
g1 = torch.randn(B, H, W, C, device="cuda").to(dt)
r = torch.randn(B, H, W, C, device="cuda").to(dt)
skip = torch.randn(B, H + 2, W + 2, C, device="cuda").to(dt)
dz = torch.empty_like(r)
sb = torch.rand(4 * C, device="cuda")
red = torch.zeros(2 * C, dtype=torch.float64, device="cuda")
a = torch.tensor([0.25], device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run():
    L.call("p2p_bn_act_bwd_reduce", dpad.data_ptr(), g1.data_ptr(), None, None, r.data_ptr(), sb.data_ptr(), skip.data_ptr(), 1,
           a.data_ptr(), dz.data_ptr(), red.data_ptr(), None, B, H, W, C, C, 1, 1, 1, L.BF16)
for cold in (1, 0):
    for _ in range(3): run()
    tot = 0
    for _ in range(10):
        if cold: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); e1.synchronize(); tot += e0.elapsed_time(e1)
    mb = (dpad.numel() + g1.numel() + r.numel() + skip.numel() + dz.numel()) * 2 / 1e6
    print(f"cold={cold} no_atomics={os.environ.get('P2P_DEBUG_NO_RED_ATOMICS','0')}: {tot/10*1e3:.1f} us  ({mb:.0f} MB -> {mb/1e3/(tot/10):.0f} GB/s)")
