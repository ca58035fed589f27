"""Where does the operand-swapped halo convolution kernel wait?  Uses the library's debug hook
p2p_debug_halo_clocks (per-CTA clock64 sums of every mbarrier wait of the MMA issuer and the epilogue, and
%globaltimer stamps at kernel entry / MMA-loop start / MMA-loop end / epilogue end / exit)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import p2p_pytorch_b200 as P
from p2p_pytorch_b200 import _lib as L, engine as E

B, S = 16, 256
B = int(os.environ.get("B", "16"))
CASES = [("g.res fwd", 128, 128, 3, S // 4 + 2, "fwd"), ("g.res dgrad", 128, 128, 3, S // 4 + 2, "dgrad"),
         ("g.deconv3 dgrad", 128, 64, 3, S // 2 + 2, "dgrad")]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
lib = L.load()
lib.p2p_debug_halo_clocks.argtypes = [C.c_void_p]
lib.p2p_debug_halo_clocks.restype = None
clk = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
dt = torch.bfloat16
for name, cin, cout, k, Hin, what in CASES:
    plan = E.ConvPlan(B, Hin, Hin, cin, cin, cout, k, 1, 0)
    x = torch.randn(B, Hin, Hin, cin, device="cuda").to(dt)
    w = torch.randn(cout, cin, k, k, device="cuda") * 0.05
    plan.pack(w, None, dt)
    out = torch.empty(B, plan.Ho, plan.Wo, plan.Co, dtype=dt, device="cuda")
    go = torch.randn_like(out)
    din = torch.empty_like(x)
    stats = torch.zeros(2 * plan.Co, dtype=torch.float64, device="cuda")
    fn = (lambda: plan.fwd(x, None, out, stats)) if what == "fwd" else (lambda: plan.dgrad(go, din))
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    lib.p2p_debug_halo_clocks(clk.data_ptr())
    clk.zero_()
    if os.environ.get("COLD", "1") == "1":
        flush.zero_()            # cold L2, like bench.py's timing of the dominant kernel
    fn()
    torch.cuda.synchronize()
    lib.p2p_debug_halo_clocks(None)
    c = clk.view(148, 16).double().cpu()
    busy = c[:, 0] > 0
    if not bool(busy.any()):
        print(f"{name}: the swapped kernel did not run for this shape")
        continue
    m = c[busy].mean(0)
    mx = c[busy].max(0).values
    print(f"{name:16s} CTAs {int(busy.sum()):3d} | MMA loop {m[0]:8.0f} (max {mx[0]:8.0f}) cyc: wait t_empty {m[1]:7.0f} "
          f"a_full {m[2]:7.0f} b_full {m[3]:7.0f} | epilogue wait t_full {m[5]:7.0f} work {m[6]:7.0f}")
    g = c[busy][:, 8:13]
    t0 = g[:, 0].min()
    print(f"{'':16s} GPU span {(g[:, 4].max() - t0) / 1e3:6.1f} us; mean since first entry: MMA start "
          f"{(g[:, 1].mean() - t0) / 1e3:5.1f}, MMA end {(g[:, 2].mean() - t0) / 1e3:5.1f}, epilogue end "
          f"{(g[:, 3].mean() - t0) / 1e3:5.1f}, exit {(g[:, 4].mean() - t0) / 1e3:5.1f} (max {(g[:, 4].max() - t0) / 1e3:5.1f})")
