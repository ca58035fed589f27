import ctypes, os, sys
os.environ["P2P_HALO_DEBUG"] = str(32 | int(sys.argv[1]) if len(sys.argv) > 1 else 32)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import p2p_pytorch_b200 as P
from p2p_pytorch_b200 import _lib as L, engine as E
B = 16
plan = E.ConvPlan(B, 66, 66, 128, 128, 128, 3, 1, 0)
x = torch.randn(B, 66, 66, 128, device="cuda").bfloat16()
w = torch.randn(128, 128, 3, 3, device="cuda") * 0.05
plan.pack(w, None, torch.bfloat16)
out = torch.empty(B, 64, 64, 128, dtype=torch.bfloat16, device="cuda")
stats = torch.zeros(256, dtype=torch.float64, device="cuda")
for _ in range(3):
    plan.fwd(x, None, out, stats)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 64)()
lib = L.load()
lib.p2p_debug_halo_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
rc = lib.p2p_debug_halo_trace(buf, 64)
t = list(buf)
t0 = t[0]
names = {0: "start", 1: "setup done", 2: "mma: item0 A ready", 3: "mma: item0 issued", 4: "mma: item1 A ready", 5: "mma: item1 issued", 40: "end"}
for li in range(2):
    names[8 + li * 10] = f"epi: item{li} acc ready"
    for h in range(2):
        names[9 + li * 10 + h * 4] = f"epi: item{li} tile{h} tmem->stg done"
        names[10 + li * 10 + h * 4] = f"epi: item{li} tile{h} barrier"
        names[11 + li * 10 + h * 4] = f"epi: item{li} tile{h} stats done"
        names[12 + li * 10 + h * 4] = f"epi: item{li} tile{h} stores done"
for i in sorted(names, key=lambda i: t[i]):
    if t[i]:
        print(f"{t[i] - t0:8d} cyc  {names[i]}")
