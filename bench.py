#!/usr/bin/env python
"""bench.py -- pix2pix 256x256 training throughput (images/s) of the B200 path, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--size S]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one iteration of the reference's training loop body (train.py:294-390, VGG term off):
1 G forward+backward, 3 D forwards, D backward for loss_d, D data-gradient for loss_g, 2 Adam steps,
on one synthetic batch (16 images of 256x256 per GPU: BASELINE.json configs[1]; data parallel with a
bucketed NCCL gradient all-reduce for N > 1, weak scaling).  Rank 0 prints ONE JSON line.

--impl reference times the CPU oracle (the reference's computation restated in fp32 PyTorch; the
reference's own files cannot travel to the GPU box) on the host cores, batch 1 per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

RES_CONV_DRAM_BYTES = 18.19e6      # dram__bytes_read.sum + dram__bytes_write.sum of one residual-conv launch (profiles/r01d_ncu_res_fwd.md)
FLOP_PER_IMG_256 = 173.41e9          # necessary conv FLOPs of one training step per image
METRIC = "pix2pix 256x256 train images/sec"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p["hbm_gbs"], p["bf16_tflops"], p.get("bf16_tflops_sustained", p["bf16_tflops"]), "measured"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for nm, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args, rank, world):
    """CPU oracle, all host threads, batch 1 per step (a bounded sample of the 16-image workload)."""
    if rank != 0:
        return
    from oracle import pix2pix_oracle as orc
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    S = args.size
    tr = orc.OracleTrainer(orc.init_g_state(123), orc.init_d_state(124))
    batches = [orc.synthetic_batch(1, S, S, 500 + i) for i in range(2)]
    for i in range(args.warmup):
        tr.step(*batches[i % 2])
    t0 = time.perf_counter()
    for i in range(args.steps):
        tr.step(*batches[i % 2])
    dt = time.perf_counter() - t0
    ips = args.steps / dt
    line = {"metric": METRIC, "value": ips, "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"pix2pix {S}x{S} RGB->RGB training step (train.py:294-390, VGG term off), "
                                   f"{args.batch} images per GPU", "image": S, "batch_per_gpu": args.batch},
            "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                             "sample": f"batch 1 of the {args.batch}-image step, {args.steps} steps, "
                                       f"torch {torch.__version__} fp32 CPU, {cores} threads"},
            "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


_REAL_STDOUT = None


def divert_stdout():
    """Libraries (NCCL's version banner, the reference-style init prints) write to fd 1; the contract is ONE JSON
    line on stdout, so fd 1 is pointed at stderr for the duration of the run and the line is written to a saved
    duplicate of the original stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def time_dominant_kernel(P, B, S, iters=20):
    """The residual-block convolution (3x3, 128->128 on [B, S/4, S/4]) is 66% of G's FLOPs and the largest
    single kernel family.  Timed alone with CUDA events on the launching stream, L2 flushed between launches."""
    import ctypes as C
    from p2p_pytorch_b200 import _lib as L, engine as E
    dev = torch.device("cuda")
    h = S // 4
    plan = E.ConvPlan(B, h + 2, h + 2, 128, 128, 128, 3, 1, 0)
    x = torch.randn(B, h + 2, h + 2, 128, device=dev).bfloat16()
    w = torch.randn(128, 128, 3, 3, device=dev) * 0.03
    plan.pack(w, None, torch.bfloat16)
    out = torch.empty(B, h, h, 128, dtype=torch.bfloat16, device=dev)
    stats = torch.zeros(256, dtype=torch.float64, device=dev)
    bias = torch.zeros(128, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    args = (C.byref(plan.desc), x.data_ptr(), plan.wp.data_ptr(), bias.data_ptr(), out.data_ptr(), stats.data_ptr(),
            L.BF16, L.IMPL_TC)
    for _ in range(3):
        L.call("p2p_conv_fwd", *args)
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.call("p2p_conv_fwd", *args)
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / iters
    flops = 2.0 * B * h * h * 128 * 1152
    return ms, flops


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    import p2p_pytorch_b200 as P
    import contextlib
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    B, S = args.batch, args.size
    P.set_precision(args.precision)
    torch.manual_seed(123)                                   # the reference's default --seed (train.py:155)
    torch.cuda.manual_seed(123)
    with contextlib.redirect_stdout(sys.stderr):             # init_weights prints, like the reference's
        g = P.define_G("normal", 0.02, gpu_id=dev)
        d = P.define_D(6, 64, gpu_id=dev)
    g.train(); d.train()
    tr = P.Pix2PixTrainer(g, d, cuda_graph=not args.no_graph)
    tr.broadcast_parameters(0)

    # synthetic pairs generated on the device: real_a, real_b ~ U(-1,1), G input = 3-bit quantised real_b
    gen = torch.Generator(device=dev).manual_seed(123 + rank)
    n_sets = 2
    data = []
    for _ in range(n_sets):
        a = torch.rand(B, 3, S, S, device=dev, generator=gen) * 2 - 1
        b = torch.rand(B, 3, S, S, device=dev, generator=gen) * 2 - 1
        c = torch.round((b * 0.5 + 0.5).clamp(0, 1) * 7) / 7
        data.append((a, b, c))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    for i in range(args.warmup):
        tr.step(*data[i % n_sets])
    primed = 0
    while tr.cuda_graph and not tr._graphs:                  # --warmup below 3: record the graph outside the timed region
        tr.step(*data[primed % n_sets])
        primed += 1
    # ---- device-resident timing ---------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    l0 = P._lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        out = tr.step(*data[i % n_sets])
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = P._lib.launch_count() - l0
    if tr.cuda_graph:                                        # replayed launches do not pass through the C ABI's counter
        launches = tr.graph_launches * args.steps
    clocks = sampler.stop() if rank == 0 else None
    value = world * B * args.steps / (ms * 1e-3)

    # ---- end to end: pinned host inputs -> H2D -> step -> D2H of the losses, every step ---------
    host = [tuple(t.cpu().pin_memory() for t in s) for s in data]
    dbuf = tuple(torch.empty_like(t) for t in data[0])
    loss_host = torch.empty(2, dtype=torch.float32).pin_memory()
    def e2e_step(i):
        if tr.cuda_graph:                                    # step() itself copies host inputs into the graph's buffers
            return tr.step(*host[i % n_sets])
        for dst, src in zip(dbuf, host[i % n_sets]):
            dst.copy_(src, non_blocking=True)
        return tr.step(*dbuf)

    for i in range(2):
        e2e_step(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if tr.cuda_graph:
        tr.prefetch(*host[0])                                # inside the timed region, like every later H2D copy
    for i in range(args.steps):
        o = e2e_step(i)
        if tr.cuda_graph and i + 1 < args.steps:
            tr.prefetch(*host[(i + 1) % n_sets])             # the next batch travels while this iteration runs
        loss_host.copy_(torch.stack((o["loss_d"], o["loss_g"])), non_blocking=True)
        torch.cuda.current_stream().synchronize()       # the caller reads the losses every iteration
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    e2e = world * B * args.steps / (ms_e2e * 1e-3)
    h2d = sum(t.numel() * 4 for t in host[0])

    if rank != 0:
        return
    hbm, tf_burst, tf_sust, which = load_peaks()
    line = {"metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if args.precision.startswith("bf16") else "f32", "data": "synthetic",
            "config": {"workload": f"pix2pix {S}x{S} RGB->RGB training step (train.py:294-390, VGG term off), "
                                   f"{B} images per GPU", "image": S, "batch_per_gpu": B, "global_batch": B * world,
                       "parallelism": f"dp{world}", "precision": P.get_precision(),
                       "launch": (f"one CUDA graph replay per step ({tr.graph_launches} kernels of libp2p_b200.so recorded)"
                                  if tr.cuda_graph else "kernel-by-kernel launches"),
                       "l2": "per-step working set (>1 GB of activations) exceeds the 126 MB L2; no explicit flush"},
            "clocks": clocks,
            "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8},
            "gpu_launches": launches,
            "step_tensor_frac": value / world * FLOP_PER_IMG_256 * (S / 256.0) ** 2 / (tf_sust * 1e12),
            "final_losses": {"loss_d": out["loss_d"].item(), "loss_g": out["loss_g"].item()}}
    try:
        kms, kflops = time_dominant_kernel(P, B, S)
        ach = kflops / (kms * 1e-3) / 1e12
        line["roofline"] = {"bound": "tensor", "achieved": ach, "peak": tf_burst, "unit": "TFLOP/s",
                            "frac": ach / tf_burst,
                            # dram__bytes_read.sum + dram__bytes_write.sum of this launch, ncu --set full
                            # (profiles/r01c_ncu_dominant_kernels.md); valid for the default B=16, 256x256 shape only
                            "traffic": RES_CONV_DRAM_BYTES if (B == 16 and S == 256) else None, "peak_source": which,
                            "kernel": "conv_halo_swap_kernel<64> (3x3 128->128 residual conv fwd, bf16, fp32 accumulate)",
                            "ms_per_launch": kms}
    except Exception as e:                                   # e.g. fp32 verification mode: no tensor-core kernel
        line["roofline"] = {"bound": "tensor", "achieved": None, "peak": tf_burst, "unit": "TFLOP/s", "frac": None,
                            "traffic": None, "note": f"dominant-kernel timing unavailable: {e}"}
    if world == 1 and not args.no_cpu_baseline:
        from oracle import pix2pix_oracle as orc             # the CPU checker, timed as the reported baseline only
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        otr = orc.OracleTrainer(orc.init_g_state(123), orc.init_d_state(124))
        ob = orc.synthetic_batch(1, S, S, 77)
        for _ in range(2):
            otr.step(*ob)
        n = 10
        t0 = time.perf_counter()
        for _ in range(n):
            otr.step(*ob)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": n / dt, "unit": "images/s", "cores": cores, "kind": "port",
                                "sample": f"{n} steps at batch 1 ({S}x{S}) of the oracle (fp32 PyTorch restatement "
                                          f"of train.py:294-390), 2 warm-up steps, {cores} threads"}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=16, help="images per GPU")
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "bf16-simt", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch kernel by kernel instead of replaying a CUDA graph")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    divert_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
