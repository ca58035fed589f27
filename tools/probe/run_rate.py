import ctypes, os, torch
here = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.path.join(here, "libprobe.so"))
lib.rate_run.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 5
out = torch.zeros(2, dtype=torch.int64, device="cuda")
iters = 4000
for grid in (1, 148):
    for mode in (0, 1):
        for n in (32, 64, 128, 256):
            for n_acc in (1, 2):
                if n * n_acc > 512:
                    continue
                rc = lib.rate_run(out.data_ptr(), n, iters, n_acc, mode, grid)
                t = out.cpu().tolist()
                print(f"grid {grid:3d} mode {mode} N {n:3d} acc {n_acc}: rc {rc} issue {t[0] / iters:6.1f} cyc/MMA, complete {t[1] / iters:6.1f} cyc/MMA"
                      f" (math floor {128 * n / 256:.0f})")
