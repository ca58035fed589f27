"""Parity at the shapes and in the precision the benchmark runs (BASELINE.json configs 2, 4 and 5), not only at the toy
shapes of test_gpu_ops.py / test_gpu_networks.py.

Which kernel serves a convolution is a function of its shape (csrc/conv_halo.cu:run_halo_problem picks resident or
streamed weights, the operand-swapped kernel, the rows per block; small maps fall back to conv_tc_kernel), so:

  * every convolution of the training step at B=16, 256x256 (24 generator layers = 7 distinct shapes, 15 discriminator
    layers, the re-expressed last layers), plus the 512x512 (B=4) and B=64 shapes, runs forward / data-gradient /
    weight-gradient through the C ABI with P2P_IMPL_AUTO and is compared with torch's fp32 convolution on the same
    bf16-rounded operands at the 1e-2 bar of test_gpu_ops.py;
  * the library records which kernel instantiation (and block geometry) served each call (p2p_trace_begin/_end); the
    last test records one whole B=16 256x256 training iteration and asserts that every instantiation it used was
    exercised by the per-layer comparisons above;
  * one whole B=16 256x256 iteration (eager and replayed from a CUDA graph, fp32 and bf16) is compared with the oracle.
"""
import math
import zlib

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import pix2pix_oracle as orc

DEV = "cuda"
P = L = E = None
SEEN = {}            # layer case id -> set of trace lines, filled by the per-layer tests


@pytest.fixture(scope="module", autouse=True)
def _pkg():
    global P, L, E
    import p2p_pytorch_b200 as P_
    from p2p_pytorch_b200 import _lib as L_, engine as E_
    P, L, E = P_, L_, E_
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    P.set_precision("bf16")


def to_stored(x_nchw, cpad, dt):
    B, Cc, H, W = x_nchw.shape
    out = torch.zeros(B, H, W, cpad, dtype=dt, device=x_nchw.device)
    out[..., :Cc] = x_nchw.permute(0, 2, 3, 1).to(dt)
    return out.contiguous()


def from_stored(t, c_real):
    return t[..., :c_real].permute(0, 3, 1, 2).float().contiguous()


def relerr(a, b):
    return (a.double() - b.double()).abs().max().item() / max(b.double().abs().max().item(), 1e-30)


def layer_cases(B, S, tag):
    """(id, B, cin, cout, stored input H = W, k, stride, pad) of every convolution of one iteration at S x S.
    G layers read pre-padded buffers (pad 0 here); D layers use zero padding 2 on maps 129/65/33/34/35 (and the two
    pooled scales)."""
    cs = [(f"{tag}.g.conv1", B, 12, 32, S + 8, 9, 1, 0), (f"{tag}.g.conv2", B, 32, 64, S + 2, 3, 2, 0),
          (f"{tag}.g.conv3", B, 64, 128, S // 2 + 2, 3, 2, 0), (f"{tag}.g.res", B, 128, 128, S // 4 + 2, 3, 1, 0),
          (f"{tag}.g.deconv3", B, 128, 64, S // 2 + 2, 3, 1, 0), (f"{tag}.g.deconv2", B, 64, 32, S + 2, 3, 1, 0),
          (f"{tag}.g.deconv1", B, 32, 3, S + 8, 9, 1, 0)]
    h = S
    for sc in range(3):
        hin = h
        for j, (cin, cout, stride, _sn) in enumerate(E_d_layers()):
            cs.append((f"{tag}.d{sc}.l{j}", B, cin, cout, hin, 4, stride, 2))
            hin = (hin + 4 - 4) // stride + 1
        h = (h - 1) // 2 + 1
    return cs


def E_d_layers():
    return [(6, 64, 2, False), (64, 128, 2, True), (128, 256, 2, True), (256, 512, 1, True), (512, 1, 1, False)]


def vgg_cases(B, S, tag):
    """the thirteen 3x3 pad-1 convolutions of vgg19.features[0:30] on the 2B images of cat(fake_b, real_b)
    (engine.VGGLossFunction): distinct shapes only"""
    return [(f"{tag}.vgg.c{cin}_{cout}_{h}", 2 * B, cin, cout, h, 3, 1, 1)
            for cin, cout, h in ((3, 64, S), (64, 64, S), (64, 128, S // 2), (128, 128, S // 2), (128, 256, S // 4),
                                 (256, 256, S // 4), (256, 512, S // 8), (512, 512, S // 8), (512, 512, S // 16))]


CASES = layer_cases(16, 256, "b16s256") + layer_cases(4, 512, "b4s512") + \
    [c for c in layer_cases(64, 256, "b64s256") if ".g." in c[0]] + vgg_cases(16, 256, "b16s256")


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_conv_at_bench_shape(case):
    name, B, cin, cout, H, k, s, pad = case
    dt = torch.bfloat16
    g = torch.Generator(device="cpu").manual_seed(zlib.crc32(name.encode()) % 100000)
    x = (torch.rand(B, cin, H, H, generator=g) * 2 - 1).to(DEV)
    w = (torch.randn(cout, cin, k, k, generator=g) / math.sqrt(cin * k * k)).to(DEV)
    bias = torch.randn(cout, generator=g).to(DEV)
    Ci, Co = E.pad16(cin), E.pad16(cout)
    xs = to_stored(x, Ci, dt)
    del x
    relu = ".vgg." in name                     # the VGG stack applies ReLU in the convolution's epilogue
    plan = E.ConvPlan(B, H, H, Ci, cin, cout, k, s, pad, L.ACT_RELU if relu else L.ACT_NONE)
    P.set_precision("bf16")
    L.trace_begin()
    try:
        plan.pack(w, None, dt)
        out = torch.empty(B, plan.Ho, plan.Wo, Co, dtype=dt, device=DEV)
        stats = torch.zeros(2 * Co, dtype=torch.float64, device=DEV)
        plan.fwd(xs, bias, out, None if relu else stats)      # (statistics are taken before any activation)
        out2 = torch.full_like(out, 3.0)
        plan.fwd(xs, bias, out2, None)          # the discriminator's call (one-channel layers: 1x1 + tap gather)
        torch.manual_seed(zlib.crc32(name.encode()) % 1000)
        gos = torch.zeros(B, plan.Ho, plan.Wo, Co, dtype=dt, device=DEV)
        gos[..., :cout] = torch.randn(B, plan.Ho, plan.Wo, cout, device=DEV).to(dt)
        din = torch.empty_like(xs)
        plan.dgrad(gos, din)
        gw = torch.zeros_like(w)
        plan.wgrad(xs, gos, gw)
        torch.cuda.synchronize()
    finally:
        SEEN[name] = set(L.trace_end())
    # reference: torch fp32 (TF32 off) on the same rounded operands
    xr = from_stored(xs, cin).requires_grad_(True)
    wr = w.to(dt).float().requires_grad_(True)
    pre = F.conv2d(xr, wr, bias, stride=s, padding=pad)
    ref = F.relu(pre) if relu else pre
    got = from_stored(out, cout)
    assert relerr(got, ref) <= 1e-2, f"{name} fwd {relerr(got, ref):.3e}"
    if Co > cout:
        assert out[..., cout:].abs().max().item() == 0
    if not relu:
        st_ref = torch.stack([got.double().sum((0, 2, 3)), (got.double() ** 2).sum((0, 2, 3))])
        assert relerr(stats.view(2, Co)[:, :cout], st_ref) <= 1e-6, f"{name} stats"     # sums of what was stored, in fp64
    assert relerr(from_stored(out2, cout), ref) <= 1e-2, f"{name} fwd without stats"
    pre.backward(from_stored(gos, cout))       # dgrad / wgrad take the gradient w.r.t. the convolution's own output
    assert relerr(from_stored(din, cin), xr.grad) <= 1e-2, f"{name} dgrad {relerr(from_stored(din, cin), xr.grad):.3e}"
    if Ci > cin:
        assert din[..., cin:].abs().max().item() == 0
    assert relerr(gw, wr.grad) <= 1e-2, f"{name} wgrad {relerr(gw, wr.grad):.3e}"


@pytest.mark.parametrize("B,S", [(16, 256), (4, 512)], ids=["b16s256", "b4s512"])
def test_fat_column_last_layer_at_bench_shape(B, S):
    """ExpandNetwork's last layer (9x9, 32 -> 3) as the engine runs it in bf16 mode: 9x1 convolution with 27 fat
    channels + horizontal gather, and the adjoint for the data-gradient."""
    P.set_precision("bf16")
    dt = torch.bfloat16
    torch.manual_seed(5)
    x = (torch.rand(B, 32, S + 8, S + 8, device=DEV) * 2 - 1)
    w = torch.randn(3, 32, 9, 9, device=DEV) / math.sqrt(32 * 81)
    bias = torch.randn(3, device=DEV)
    xs = to_stored(x, 32, dt)
    del x
    assert E.FatColumnConv.applies(3, 9, dt)
    fat = E.FatColumnConv(B, S + 8, S + 8, 32, 32, 3, 9)
    L.trace_begin()
    try:
        L.call_multi("p2p_pack_weights_multi", [fat.pack_job(w, dt)], L.BF16)
        out = torch.empty(B, S, S, 16, dtype=dt, device=DEV)
        fat.fwd(xs, bias, out)
        gos = torch.zeros(B, S, S, 16, dtype=dt, device=DEV)
        gos[..., :3] = torch.randn(B, S, S, 3, device=DEV).to(dt)
        din = torch.empty_like(xs)
        fat.dgrad(gos, din)
        torch.cuda.synchronize()
    finally:
        SEEN[f"fat.b{B}s{S}"] = set(L.trace_end())
    xr = from_stored(xs, 32).requires_grad_(True)
    ref = F.conv2d(xr, w.to(dt).float(), bias)
    # one extra bf16 rounding of the nine per-kx partial sums: 1.5e-2 instead of 1e-2
    assert relerr(from_stored(out, 3), ref) <= 1.5e-2
    ref.backward(from_stored(gos, 3))
    assert relerr(from_stored(din, 32), xr.grad) <= 1.5e-2


# ---- one whole iteration at the benchmark shape ---------------------------------------------------------------------------
_ORACLE = {}


def _oracle_step(B, S, seed):
    key = (B, S, seed)
    if key not in _ORACLE:
        torch.set_num_threads(max(1, torch.get_num_threads()))
        ot = orc.OracleTrainer(orc.init_g_state(123), orc.init_d_state(124))
        a, b, c = orc.synthetic_batch(B, S, S, seed)
        out = ot.step(a, b, c)
        _ORACLE[key] = ({k: v.clone() for k, v in out.items()}, ot.grads(), (a, b, c))
    return _ORACLE[key]


def _cos(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return (a.dot(b) / (a.norm() * b.norm()).clamp_min(1e-30)).item()


def _nets(precision):
    P.set_precision(precision)
    g = P.ExpandNetwork().to(DEV)
    d = P.MultiscaleDiscriminator(6, 64, getIntermFeat=True).to(DEV)
    g.load_state_dict(orc.init_g_state(123))
    d.load_state_dict(orc.init_d_state(124))
    g.train(); d.train()
    return g, d


@pytest.mark.parametrize("precision,graph,B,S", [("fp32", False, 16, 256), ("fp32", True, 16, 256), ("bf16", False, 16, 256),
                                                 ("bf16", True, 16, 256), ("bf16", True, 4, 512)],
                         ids=["fp32-eager-b16s256", "fp32-graph-b16s256", "bf16-eager-b16s256", "bf16-graph-b16s256",
                              "bf16-graph-b4s512"])
def test_train_step_at_bench_shape_matches_oracle(precision, graph, B, S):
    """B=16, 256x256 (BASELINE.json configs[1]) and B=4, 512x512 (configs[3] per GPU): generator output, the five losses
    and every gradient of one iteration against the oracle; with graph=True the iteration is recorded into a CUDA graph
    and REPLAYED on two streams (what bench.py times), else launched kernel by kernel in the reference's order."""
    ref, (gg, dg), (a, b, c) = _oracle_step(B, S, 4242)
    g, d = _nets(precision)
    # graph_warmup=0: the very first call captures and replays, so the compared iteration starts from the oracle's state
    tr = P.Pix2PixTrainer(g, d, cuda_graph=graph, graph_warmup=0)
    if graph:
        g0, d0 = _nets(precision)             # lazy one-time set-up (kernel attributes) happens outside the capture
        P.Pix2PixTrainer(g0, d0).step(a.to(DEV), b.to(DEV), c.to(DEV))
        torch.cuda.synchronize()
    out = tr.step(a.to(DEV), b.to(DEV), c.to(DEV))
    torch.cuda.synchronize()
    if graph:
        assert len(tr._graphs) == 1 and tr.graph_launches > 100
    diff = out["fake_b"].float().cpu().double() - ref["fake_b"].double()
    err_max = diff.abs().max().item() / ref["fake_b"].abs().max().item()
    err_rms = (diff.pow(2).mean().sqrt() / ref["fake_b"].double().pow(2).mean().sqrt()).item()
    tol_max, tol_rms, tol_loss = (2e-4, 1e-4, 2e-4) if precision == "fp32" else (0.2, 6e-2, 5e-2)
    print(f"[{precision} {'graph' if graph else 'eager'} B={B} {S}x{S}] fake_b max {err_max:.3e} rms {err_rms:.3e}")
    assert err_max <= tol_max and err_rms <= tol_rms, (err_max, err_rms)
    for nm in ("loss_d", "loss_g_gan", "loss_feat", "loss_tv", "loss_g"):
        rel = abs(out[nm].item() - ref[nm].item()) / abs(ref[nm].item())
        print(f"    {nm}: {out[nm].item():.6f} vs {ref[nm].item():.6f} ({rel:.2e})")
        assert rel <= tol_loss, (nm, out[nm].item(), ref[nm].item())
    # gradients (left in the flat buffers after the step: the optimizers zero them at the START of an iteration)
    worst = 1.0
    for net, refs in ((g, gg), (d, dg)):
        for name, p in net.named_parameters():
            if not p.requires_grad or name.endswith("conv2d.bias") or p.numel() == 1 or name not in refs:
                continue
            cs = _cos(p.grad, refs[name])
            worst = min(worst, cs)
            assert cs >= (0.995 if precision == "fp32" else 0.9), f"grad {name}: cos {cs:.4f}"
    print(f"    worst gradient cosine {worst:.5f}")


def test_bench_iteration_runs_on_the_instantiations_checked_above():
    """Every kernel instantiation (with its block geometry) that one B=16 256x256 bf16 iteration launches was compared
    with torch by a per-layer test of this file."""
    checked = set()
    for name, lines in SEEN.items():
        if name.startswith(("b16s256.", "fat.b16s256")):
            checked |= lines
    if not checked:
        pytest.skip("the per-layer tests of this file did not run in this session")
    g, d = _nets("bf16")
    tr = P.Pix2PixTrainer(g, d)
    a, b, c = (t.to(DEV) for t in orc.synthetic_batch(16, 256, 256, 1))
    tr.step(a, b, c)                        # lazy set-up
    L.trace_begin()
    try:
        tr.step(a, b, c)
        torch.cuda.synchronize()
    finally:
        used = L.trace_end()
    assert len(used) > 100
    missing = sorted(set(used) - checked)
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/bench_step_kernel_trace.txt", "w") as f:
        from collections import Counter
        for ln, n in sorted(Counter(used).items()):
            f.write(f"{n:4d}  {ln}\n")
    assert not missing, "instantiations used by the benchmark iteration but not compared per layer:\n" + "\n".join(missing)


def test_generator_inference_at_bench_shape_matches_oracle():
    """BASELINE.json configs[4]: net_g(input) in eval mode under no_grad at B=64, 256x256 (test.py:37-41), bf16 tcgen05
    kernels, against the oracle's eval-mode forward (running statistics)."""
    B, S = 64, 256
    g, _d = _nets("bf16")
    g.eval()
    _a, _b, c = orc.synthetic_batch(B, S, S, 99)
    with torch.no_grad():
        y = g(c.to(DEV))
    torch.set_num_threads(max(1, torch.get_num_threads()))
    with torch.no_grad():
        ref = orc.g_forward(orc.init_g_state(123), c, False)
    diff = y.float().cpu().double() - ref.double()
    err_max = diff.abs().max().item() / ref.abs().max().item()
    err_rms = (diff.pow(2).mean().sqrt() / ref.double().pow(2).mean().sqrt()).item()
    print(f"[infer B={B}] fake_b max {err_max:.3e} rms {err_rms:.3e}")
    assert y.shape == ref.shape and err_max <= 0.2 and err_rms <= 6e-2, (err_max, err_rms)
    assert int(g.state_dict()["in1_e.num_batches_tracked"]) == 0
    # the same batch split over CUDA streams (P.generate): images are independent in eval mode, so only the block
    # geometry of the kernels changes -- bf16 results agree to accumulation-order noise
    y3 = P.generate(g, c.to(DEV), streams=3)
    torch.cuda.synchronize()
    assert y3.shape == y.shape and (y3 - y).abs().max().item() <= 2e-2
    with pytest.raises(ValueError):
        g.train()
        P.generate(g, c.to(DEV)[:4])
