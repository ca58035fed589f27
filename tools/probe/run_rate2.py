import ctypes, os, torch
here = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.path.join(here, "libprobe.so"))
lib.rate_run2.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 6
out = torch.zeros(2, dtype=torch.int64, device="cuda")
iters = 4000
for rb in (32, 64, 128):
    for sbo in (8, 10, 16, 24):
        for n in (16, 32, 128):
            rc = lib.rate_run2(out.data_ptr(), n, iters, 2, rb, sbo, 148)
            t = out.cpu().tolist()
            print(f"row bytes {rb:3d} group stride {sbo:2d} rows N {n:3d}: rc {rc} complete {t[1] / iters:6.1f} cyc/MMA")
