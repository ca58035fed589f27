"""Runs the UMMA shifted-descriptor probe on the GPU and prints max errors per configuration."""
import ctypes, os, sys
import torch
here = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.path.join(here, "libprobe.so"))
lib.probe_run.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_void_p]
lib.probe_run.restype = ctypes.c_int
torch.manual_seed(0)
ROWS = 192
a = torch.randn(ROWS, 128, device="cuda").bfloat16()
out = torch.zeros(128, 16, device="cuda")
st = torch.cuda.current_stream().cuda_stream
print("mode 0 (K-major A, row-shifted start, 8-row groups grp rows apart)")
b = torch.randn(16, 64, device="cuda").bfloat16()
for grp in (8, 10, 18):
    for shift in (0, 1, 2, 3, 5, 8, 9, 13):
        idx = torch.tensor([shift + (i // 8) * grp + i % 8 for i in range(128)], device="cuda")
        if idx.max().item() >= ROWS:
            continue
        ref = a[idx, :64].float() @ b.float().t()
        rc = lib.probe_run(a.data_ptr(), b.data_ptr(), out.data_ptr(), 0, shift, grp, st)
        torch.cuda.synchronize()
        err = (out - ref).abs().max().item() / ref.abs().max().item()
        print(f"  grp {grp:2d} shift {shift:2d}: rc {rc} rel err {err:.3e} {'OK' if err < 1e-2 else 'MISMATCH'}")
print("mode 1 (MN-major A, K-row-shifted start)")
b = torch.randn(64, 16, device="cuda").bfloat16()
for shift in (0, 1, 2, 3, 7, 8, 9, 19, 40):
    ref = a[shift:shift + 64, :].float().t() @ b.float()
    rc = lib.probe_run(a.data_ptr(), b.data_ptr(), out.data_ptr(), 1, shift, 8, st)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item() / ref.abs().max().item()
    print(f"  shift {shift:2d}: rc {rc} rel err {err:.3e} {'OK' if err < 1e-2 else 'MISMATCH'}")

lib.probe2_run.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 4 + [ctypes.c_void_p]
lib.probe2_run.restype = ctypes.c_int
for rb in (32, 64):
    ch = rb // 2
    a2 = torch.randn(ROWS, ch, device="cuda").bfloat16()
    print(f"mode 2 (K-major, {rb}-byte rows)")
    b2 = torch.randn(16, ch, device="cuda").bfloat16()
    for grp in (8, 10, 16):
        for shift in (0, 1, 3, 8, 9):
            idx = torch.tensor([shift + (i // 8) * grp + i % 8 for i in range(128)], device="cuda")
            if idx.max().item() >= ROWS:
                continue
            ref = a2[idx].float() @ b2.float().t()
            rc = lib.probe2_run(a2.data_ptr(), b2.data_ptr(), out.data_ptr(), 2, shift, grp, rb, st)
            torch.cuda.synchronize()
            err = (out - ref).abs().max().item() / ref.abs().max().item()
            print(f"  grp {grp:2d} shift {shift:2d}: rc {rc} rel err {err:.3e} {'OK' if err < 1e-2 else 'MISMATCH'}")
    print(f"mode 3 (MN-major, {rb}-byte rows, overlapping atoms LBO = 1 row)")
    b3 = torch.randn(64, 16, device="cuda").bfloat16()
    ntap = 128 // ch
    for shift in (0, 1, 5, 8, 11):
        A = torch.stack([a2[shift + t: shift + t + 64].float() for t in range(ntap)], 0)   # [tap][k][c]
        ref = torch.einsum("tkc,kn->tcn", A, b3.float()).reshape(128, 16)
        rc = lib.probe2_run(a2.data_ptr(), b3.data_ptr(), out.data_ptr(), 3, shift, 8, rb, st)
        torch.cuda.synchronize()
        err = (out - ref).abs().max().item() / ref.abs().max().item()
        print(f"  shift {shift:2d}: rc {rc} rel err {err:.3e} {'OK' if err < 1e-2 else 'MISMATCH'}")
