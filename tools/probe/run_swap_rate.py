import ctypes, os, subprocess, torch
here = os.path.dirname(os.path.abspath(__file__))
so = os.path.join(here, "libswapprobe.so")
subprocess.run(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
                "--expt-relaxed-constexpr", "-shared", "-o", so, os.path.join(here, "umma_swap_rate_probe.cu"),
                os.path.join(here, "probe_stub.cu")], check=True)
lib = ctypes.CDLL(so)
lib.swap_rate_run.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 6
out = torch.zeros(8, dtype=torch.int64, device="cuda")
iters = 7200
for n_cols in (256, 176, 128):
    for sbo, shifts in ((8, 0), (10, 0), (10, 1)):
        for ldg in (0, 1, 2, 4):
            out.zero_()
            rc = lib.swap_rate_run(out.data_ptr(), iters, sbo, shifts, ldg, n_cols, 148)
            t = out.cpu().tolist()
            print(f"N {n_cols:3d} group stride {sbo:2d} rows, tap shifts {shifts}, TMEM reader groups {ldg}: rc {rc} "
                  f"{t[0] / iters:6.1f} cyc/MMA  (reader sweeps of 256 columns per warp: {t[1]})")
