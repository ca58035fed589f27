"""How long does the host take to enqueue one training iteration, against the GPU time of the same iteration?
If the two are close the step is launch-bound and a CUDA graph (or fewer launches) is the lever."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import p2p_pytorch_b200 as P

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = "cuda:0"
g = P.define_G("normal", 0.02, gpu_id=dev)
d = P.define_D(6, 64, gpu_id=dev)
tr = P.Pix2PixTrainer(g, d, cuda_graph=len(sys.argv) > 2)
gen = torch.Generator(device=dev).manual_seed(1)
a = torch.rand(B, 3, 256, 256, device=dev, generator=gen) * 2 - 1
b = torch.rand(B, 3, 256, 256, device=dev, generator=gen) * 2 - 1
c = torch.rand(B, 3, 256, 256, device=dev, generator=gen) * 2 - 1
for _ in range(3):
    tr.step(a, b, c)
torch.cuda.synchronize()
host, gpu = [], []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    n0 = P._lib.launch_count()
    tr.step(a, b, c)
    n1 = P._lib.launch_count()
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    host.append((t1 - t0) * 1e3)
    gpu.append(e0.elapsed_time(e1))
host.sort(); gpu.sort()
print(f"B={B}: host enqueue median {host[5]:.2f} ms (min {host[0]:.2f}), GPU span median {gpu[5]:.2f} ms (min {gpu[0]:.2f}), "
      f"{n1 - n0} own launches/step")
