"""Per-kernel parity: every C-ABI entry point against plain fp32 PyTorch on the same seeded inputs.
fp32 kernels: max |err| <= 1e-5 of the output scale (the north star's fp32 verification bar);
bf16 / tcgen05 kernels: <= 1e-2 of the output scale against a reference fed the same bf16-rounded inputs."""
import ctypes as C
import math
import zlib

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

P = L = E = None


@pytest.fixture(scope="module", autouse=True)
def _pkg():
    global P, L, E
    import p2p_pytorch_b200 as P_
    from p2p_pytorch_b200 import _lib as L_, engine as E_
    P, L, E = P_, L_, E_
    # the torch references below must be true fp32 (cuDNN / cuBLAS default to TF32 for fp32 convs and matmuls)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield


DEV = "cuda"


def to_stored(x_nchw, cpad, dt):
    """NCHW fp32 -> stored NHWC with zero pad channels"""
    B, Cc, H, W = x_nchw.shape
    out = torch.zeros(B, H, W, cpad, dtype=dt, device=x_nchw.device)
    out[..., :Cc] = x_nchw.permute(0, 2, 3, 1).to(dt)
    return out.contiguous()


def from_stored(t, c_real):
    return t[..., :c_real].permute(0, 3, 1, 2).float().contiguous()


def relerr(a, b):
    return (a.double() - b.double()).abs().max().item() / max(b.double().abs().max().item(), 1e-30)


CONV_CASES = [
    # name,            B, cin, cout, H,  W,  k, s, pad
    ("g_res",          2, 128, 128, 16, 16, 3, 1, 0),
    ("g_conv2",        2, 32, 64, 34, 34, 3, 2, 0),
    ("g_conv1",        1, 12, 32, 40, 40, 9, 1, 0),
    ("g_deconv1",      1, 32, 3, 24, 24, 9, 1, 0),
    ("d_l0",           2, 6, 64, 32, 32, 4, 2, 2),
    ("d_l1_odd",       2, 64, 128, 17, 17, 4, 2, 2),
    ("d_l3",           1, 256, 512, 5, 5, 4, 1, 2),
    ("d_l4",           2, 512, 1, 6, 6, 4, 1, 2),
    ("d_l4_tiles",     2, 512, 1, 19, 23, 4, 1, 2),
    ("rect_odd",       3, 16, 32, 13, 21, 3, 1, 1),
]


def _conv_case(case, dt, impl, tol):
    name, B, cin, cout, H, W, k, s, pad = case
    g = torch.Generator(device="cpu").manual_seed(zlib.crc32(name.encode()) % 1000)
    x = torch.randn(B, cin, H, W, generator=g).to(DEV)
    w = (torch.randn(cout, cin, k, k, generator=g) / math.sqrt(cin * k * k)).to(DEV)
    bias = torch.randn(cout, generator=g).to(DEV)
    Ci, Co = E.pad16(cin), E.pad16(cout)
    xs = to_stored(x, Ci, dt)
    plan = E.ConvPlan(B, H, W, Ci, cin, cout, k, s, pad)
    old = dict(E._state)
    E._state.update(dtype=dt, impl=impl)
    try:
        plan.pack(w, None, dt)
        out = torch.empty(B, plan.Ho, plan.Wo, Co, dtype=dt, device=DEV)
        stats = torch.zeros(2 * Co, dtype=torch.float64, device=DEV)
        plan.fwd(xs, bias, out, stats)
        # reference on the same (rounded) operands
        # fp32 kernels are held to 1e-5 against an fp64 reference (two fp32 summation orders differ by more)
        rdt = torch.float64 if dt == torch.float32 else torch.float32
        xr = from_stored(xs, cin).to(rdt).requires_grad_(True)
        wr = w.to(dt).to(rdt).requires_grad_(True)
        ref = F.conv2d(xr, wr, bias.to(rdt), stride=s, padding=pad)
        got = from_stored(out, cout)
        assert relerr(got, ref) <= tol, f"{name} fwd"
        assert out[..., cout:].abs().max().item() == 0 if Co > cout else True
        # statistics of what was stored
        st_ref = torch.stack([got.sum((0, 2, 3)), (got * got).sum((0, 2, 3))]).double()
        st = stats.view(2, Co)[:, :cout]
        assert relerr(st, st_ref) <= max(tol, 1e-5) * 5, f"{name} stats"
        # without statistics (the discriminator's use): one-channel layers take the 1x1 + tap-gather route
        out2 = torch.full_like(out, 3.0)
        plan.fwd(xs, bias, out2, None)
        assert relerr(from_stored(out2, cout), ref) <= tol, f"{name} fwd without stats"
        assert out2[..., cout:].abs().max().item() == 0 if Co > cout else True
        # backward
        go = torch.randn(ref.shape, generator=g).to(DEV)
        gos = to_stored(go, Co, dt)
        gor = from_stored(gos, cout)
        ref.backward(gor.to(rdt))
        din = torch.empty_like(xs)
        plan.dgrad(gos, din)
        assert relerr(from_stored(din, cin), xr.grad) <= tol, f"{name} dgrad"
        if Ci > cin:
            assert din[..., cin:].abs().max().item() == 0
        gw = torch.zeros_like(w)
        plan.wgrad(xs, gos, gw)
        assert relerr(gw, wr.grad) <= tol, f"{name} wgrad"
    finally:
        E._state.update(old)


@pytest.mark.parametrize("case", CONV_CASES, ids=[c[0] for c in CONV_CASES])
def test_conv_simt_fp32(case):
    _conv_case(case, torch.float32, L.IMPL_SIMT, 1e-5)


@pytest.mark.parametrize("case", CONV_CASES, ids=[c[0] for c in CONV_CASES])
def test_conv_simt_bf16(case):
    _conv_case(case, torch.bfloat16, L.IMPL_SIMT, 1e-2)


@pytest.mark.parametrize("case", CONV_CASES, ids=[c[0] for c in CONV_CASES])
def test_conv_tcgen05_bf16(case):
    _conv_case(case, torch.bfloat16, L.IMPL_AUTO, 1e-2)


def test_conv_tcgen05_path_is_taken():
    """IMPL_TC must be accepted (not silently rerouted) for the flagship residual conv shape."""
    plan = E.ConvPlan(1, 18, 18, 128, 128, 128, 3, 1, 0)
    x = torch.randn(1, 18, 18, 128, device=DEV).bfloat16()
    w = torch.randn(128, 128, 3, 3, device=DEV) * 0.03
    plan.pack(w, None, torch.bfloat16)
    out = torch.empty(1, 16, 16, 128, dtype=torch.bfloat16, device=DEV)
    L.call("p2p_conv_fwd", C.byref(plan.desc), x.data_ptr(), plan.wp.data_ptr(), None, out.data_ptr(), None, L.BF16,
           L.IMPL_TC)
    ref = F.conv2d(x.permute(0, 3, 1, 2).float(), w.bfloat16().float())
    assert relerr(out.permute(0, 3, 1, 2).float(), ref) <= 1e-2


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_g_input_prep(dt):
    from oracle import pix2pix_oracle as orc
    x = torch.rand(2, 3, 16, 24, device=DEV)
    dst = torch.empty(2, 24, 32, 16, dtype=dt, device=DEV)
    L.call("p2p_g_input_prep", x.data_ptr(), dst.data_ptr(), 2, 16, 24, L.dtype_code(dt))
    ref = F.pad(F.interpolate(orc.pixel_unshuffle2(x), scale_factor=2, mode="nearest"), (4, 4, 4, 4), mode="reflect")
    assert relerr(from_stored(dst, 12), ref.to(dt).float()) == 0
    assert dst[..., 12:].abs().max().item() == 0


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("act,up,pad,skip", [(L_ACT, up, pad, skip) for L_ACT, up, pad, skip in
                                              [(1, 1, 1, True), (2, 1, 4, False), (3, 2, 1, True), (2, 2, 1, False),
                                               (0, 1, 0, False)]])
def test_bn_act_store_and_backward(dt, act, up, pad, skip):
    """train-mode BatchNorm + activation + (skip) + upsample + reflection pad, forward and both backward passes,
    against autograd on the equivalent torch graph."""
    torch.manual_seed(3)
    B, Cc, H, W = 2, 32, 6, 10
    tol = 1e-5 if dt == torch.float32 else 2e-2
    r = torch.randn(B, Cc, H, W, device=DEV) * 2 + 0.5
    rs = to_stored(r, Cc, dt)
    r_used = from_stored(rs, Cc).requires_grad_(True)
    gamma = (torch.rand(Cc, device=DEV) + 0.5).requires_grad_(True)
    beta = torch.randn(Cc, device=DEV).requires_grad_(True)
    a = torch.tensor([0.25], device=DEV, requires_grad=True)
    sk = torch.randn(B, Cc, H, W, device=DEV) if skip else None
    sks = None
    if skip:
        sks = torch.zeros(B, H + 2, W + 2, Cc, dtype=dt, device=DEV)
        sks[:, 1:-1, 1:-1] = sk.permute(0, 2, 3, 1).to(dt)
        sk_used = sks[:, 1:-1, 1:-1].permute(0, 3, 1, 2).float().requires_grad_(True)
    rm, rv = torch.zeros(Cc, device=DEV), torch.ones(Cc, device=DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    stats = torch.zeros(2 * Cc, dtype=torch.float64, device=DEV)
    dc = L.dtype_code(dt)
    L.call("p2p_channel_stats", rs.data_ptr(), stats.data_ptr(), B * H * W, Cc, dc)
    sb = torch.empty(4 * Cc, device=DEV)
    L.call("p2p_bn_finalize", stats.data_ptr(), B * H * W, gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
           rv.data_ptr(), nbt.data_ptr(), sb.data_ptr(), Cc, Cc, 1)
    dst = torch.empty(B, H * up + 2 * pad, W * up + 2 * pad, Cc, dtype=dt, device=DEV)
    L.call("p2p_bn_act_store", rs.data_ptr(), sb.data_ptr(), L.ptr(sks), 1, a.data_ptr(), dst.data_ptr(), None, B, H, W,
           Cc, Cc, act, up, pad, dc)
    # torch reference
    rm2, rv2 = torch.zeros(Cc, device=DEV), torch.ones(Cc, device=DEV)
    z = F.batch_norm(r_used, rm2, rv2, gamma, beta, True, 0.1, 1e-5)
    if skip:
        z = z + sk_used
    y = {0: lambda t: t, 1: F.relu, 2: lambda t: F.prelu(t, a), 3: lambda t: F.leaky_relu(t, 0.2), 4: torch.tanh}[act](z)
    if up == 2:
        y = F.interpolate(y, scale_factor=2, mode="nearest")
    if pad:
        y = F.pad(y, (pad,) * 4, mode="reflect")
    assert relerr(from_stored(dst, Cc), y) <= tol
    assert relerr(rm, rm2) <= 1e-5 and relerr(rv, rv2) <= 1e-5 and nbt.item() == 1
    # backward
    gy = torch.randn_like(y)
    gys = to_stored(gy, Cc, dt)
    y.backward(from_stored(gys, Cc))
    dz = torch.empty(B, H, W, Cc, dtype=dt, device=DEV)
    red = torch.zeros(2 * Cc, dtype=torch.float64, device=DEV)
    dpre = torch.zeros(1, dtype=torch.float64, device=DEV)
    L.call("p2p_bn_act_bwd_reduce", gys.data_ptr(), None, None, None, rs.data_ptr(), sb.data_ptr(), L.ptr(sks), 1,
           a.data_ptr(), dz.data_ptr(), red.data_ptr(), dpre.data_ptr(), B, H, W, Cc, Cc, act, up, pad, dc)
    dr = torch.empty_like(dz)
    dg, db_, dbias = (torch.zeros(Cc, device=DEV) for _ in range(3))
    L.call("p2p_bn_bwd_apply", dz.data_ptr(), rs.data_ptr(), sb.data_ptr(), red.data_ptr(), dr.data_ptr(), dg.data_ptr(),
           db_.data_ptr(), dbias.data_ptr(), B * H * W, Cc, Cc, dc)
    assert relerr(from_stored(dr, Cc), r_used.grad) <= tol * 2
    assert relerr(dg, gamma.grad) <= tol * 2 and relerr(db_, beta.grad) <= tol * 2
    if skip:
        assert relerr(from_stored(dz, Cc), sk_used.grad) <= tol * 2
    if act == 2:
        assert abs(dpre.item() - a.grad.item()) <= tol * 2 * max(abs(a.grad.item()), 1.0)


def test_avgpool_pyramid_and_input_grad():
    torch.manual_seed(0)
    B, H, W = 2, 17, 20
    a = torch.randn(B, 3, H, W, device=DEV)
    b = torch.randn(B, 3, H, W, device=DEV, requires_grad=True)
    x0 = torch.empty(B, H, W, 8, device=DEV)
    L.call("p2p_d_input_prep", a.data_ptr(), 3, b.data_ptr(), 3, x0.data_ptr(), B, H, W)
    cat = torch.cat((a, b), 1)
    assert torch.equal(x0[..., :6].permute(0, 3, 1, 2), cat)
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    x1 = torch.empty(B, Ho, Wo, 8, device=DEV)
    L.call("p2p_avgpool3s2_fwd", x0.data_ptr(), x1.data_ptr(), B, H, W, 8)
    ref = F.avg_pool2d(cat, 3, stride=2, padding=1, count_include_pad=False)
    assert relerr(x1[..., :6].permute(0, 3, 1, 2), ref) <= 1e-6
    g1 = torch.randn_like(ref)
    ref.backward(g1)
    g1s = torch.zeros(B, Ho, Wo, 8, device=DEV)
    g1s[..., :6] = g1.permute(0, 2, 3, 1)
    d0 = torch.zeros(B, H, W, 8, device=DEV)
    L.call("p2p_avgpool3s2_bwd_add", g1s.data_ptr(), d0.data_ptr(), B, H, W, 8)
    gb = torch.empty(B, 3, H, W, device=DEV)
    L.call("p2p_d_input_grad", d0.data_ptr(), gb.data_ptr(), 3, 3, B, H, W)
    assert relerr(gb, b.grad) <= 1e-6


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_losses(dt):
    torch.manual_seed(1)
    B, H, W, Cc, Cp = 2, 7, 9, 20, 32
    f = torch.randn(B, Cc, H, W, device=DEV)
    r = torch.randn(B, Cc, H, W, device=DEV)
    fs, rs = to_stored(f, Cp, dt).requires_grad_(True), to_stored(r, Cp, dt)
    fr = from_stored(fs.detach(), Cc).requires_grad_(True)
    rr = from_stored(rs, Cc)
    # LSGAN on 2 maps + feature-matching style L1, scaled
    loss = P.lsgan_loss([fs, fs], [Cc, Cc], 1.0) * 0.5
    ref = (((fr - 1.0) ** 2).mean() * 2) * 0.5
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    loss.backward(); ref.backward()
    tol = 1e-5 if dt == torch.float32 else 1e-2
    assert relerr(from_stored(fs.grad, Cc), fr.grad) <= tol
    assert fs.grad[..., Cc:].abs().max().item() == 0
    from p2p_pytorch_b200.losses import _MapLoss
    fs.grad = None; fr.grad = None
    l1 = _MapLoss.apply("l1", 0.0, [10.0 / 3], [Cc], 1, fs, rs)
    l1r = (10.0 / 3) * (fr - rr).abs().mean()
    assert abs(l1.item() - l1r.item()) <= 1e-5 * abs(l1r.item())
    l1.backward(); l1r.backward()
    assert relerr(from_stored(fs.grad, Cc), fr.grad) <= tol
    # TV and flat MSE on NCHW fp32
    x = torch.randn(2, 3, 12, 10, device=DEV, requires_grad=True)
    x2 = x.detach().clone().requires_grad_(True)
    from oracle import pix2pix_oracle as orc
    t, tr_ = P.tv_loss(x), orc.tv_loss(x2)
    assert abs(t.item() - tr_.item()) <= 1e-5 * abs(tr_.item())
    (t * 3).backward(); (tr_ * 3).backward()
    assert relerr(x.grad, x2.grad) <= 1e-5
    crit = P.GANLoss()
    p1 = torch.randn(2, 1, 5, 5, device=DEV, requires_grad=True)
    p2 = p1.detach().clone().requires_grad_(True)
    lg = crit([[p1]], True); lr_ = ((p2 - 1.0) ** 2).mean()
    lg.backward(); lr_.backward()
    assert abs(lg.item() - lr_.item()) <= 1e-6 and relerr(p1.grad, p2.grad) <= 1e-5


def test_fused_adam_matches_torch_adam():
    torch.manual_seed(2)
    shapes = [(1,), (32, 12, 9, 9), (32,), (128, 128, 3, 3), (7,)]
    ps = [torch.nn.Parameter(torch.randn(s, device=DEV)) for s in shapes]
    qs = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    mine = P.FusedAdam(ps, lr=2e-4, betas=(0.5, 0.999))
    ref = torch.optim.Adam(qs, lr=2e-4, betas=(0.5, 0.999))
    for it in range(5):
        mine.zero_grad(); ref.zero_grad()
        for p, q in zip(ps, qs):
            g = torch.randn_like(p) * (10.0 ** (-it))
            p.grad.add_(g); q.grad = g.clone()
        mine.step(); ref.step()
    for p, q in zip(ps, qs):
        assert relerr(p.data, q.data) <= 1e-6


def test_spectral_norm_power_iteration():
    from oracle import pix2pix_oracle as orc
    torch.manual_seed(4)
    ws = [torch.randn(128, 64, 4, 4, device=DEV) * 0.03, torch.randn(512, 256, 4, 4, device=DEV) * 0.02]
    us = [orc.l2normalize(torch.randn(w.shape[0], device=DEV)) for w in ws]
    vs = [orc.l2normalize(torch.randn(w[0].numel(), device=DEV)) for w in ws]
    us0 = [u.clone() for u in us]
    sig = E.spectral_power_iteration(ws, us, vs)
    for w, u0, u, v, s in zip(ws, us0, us, vs, sig):
        w2 = w.view(w.shape[0], -1)
        v_ref = orc.l2normalize(w2.t().mv(u0))
        u_ref = orc.l2normalize(w2.mv(v_ref))
        s_ref = u_ref.dot(w2.mv(v_ref))
        assert relerr(v, v_ref) <= 1e-5 and relerr(u, u_ref) <= 1e-5
        assert abs(s.item() - s_ref.item()) <= 1e-5 * abs(s_ref.item())


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_batched_pack_unpack_and_sigma_correction_match_single_calls(dt):
    """p2p_pack_weights_multi / p2p_unpack_wgrad_multi / p2p_sn_weight_grad_multi == the per-layer entry points."""
    from p2p_pytorch_b200._lib import call, call_multi, ptr
    torch.manual_seed(9)
    shapes = [(32, 12, 9), (128, 128, 3), (3, 32, 9), (64, 6, 4), (512, 256, 4), (1, 512, 4)]   # (cout, cin, k)
    dc = L.dtype_code(dt)
    pack_jobs, singles, keep = [], [], []
    for cout, cin, k in shapes:
        w = torch.randn(cout, cin, k, k, device=DEV) * 0.1
        sigma = torch.rand(1, device=DEV) + 0.5 if cout % 2 == 0 else None
        Co, Ci = E.pad16(cout), E.pad16(cin)
        a = [torch.full((Co, k, k, Ci), 7.0, dtype=dt, device=DEV), torch.full((Ci, k, k, Co), 7.0, dtype=dt, device=DEV)]
        b = [torch.full_like(a[0], 9.0), torch.full_like(a[1], 9.0)]
        want_wt = cout != 3
        pack_jobs.append(L.PackJob(ptr(w), ptr(sigma), ptr(a[0]), ptr(a[1]) if want_wt else None, cout, cin, Co, Ci, k, k))
        call("p2p_pack_weights", ptr(w), ptr(sigma), ptr(b[0]), ptr(b[1]) if want_wt else None, cout, cin, Co, Ci, k, k, dc)
        singles.append((a, b, want_wt))
        keep.append((w, sigma))
    call_multi("p2p_pack_weights_multi", pack_jobs, dc)
    for a, b, want_wt in singles:
        assert torch.equal(a[0], b[0])
        assert torch.equal(a[1], b[1]) if want_wt else bool((a[1] == 7.0).all())

    unpack_jobs, sn_jobs, checks = [], [], []
    for (cout, cin, k), (w, sigma) in zip(shapes, keep):
        Co, Ci = E.pad16(cout), E.pad16(cin)
        acc = torch.randn(k, k, Ci, Co, device=DEV)
        g0 = torch.randn(cout, cin, k, k, device=DEV)
        g_multi, g_single = g0.clone(), g0.clone()
        dot_m = torch.zeros(1, device=DEV) if sigma is not None else None
        dot_s = torch.zeros(1, device=DEV) if sigma is not None else None
        unpack_jobs.append(L.UnpackJob(ptr(acc), ptr(g_multi), ptr(w), ptr(sigma), ptr(dot_m), cout, cin, Co, Ci, k, k))
        call("p2p_unpack_wgrad", ptr(acc), ptr(g_single), ptr(w), ptr(sigma), ptr(dot_s), cout, cin, Co, Ci, k, k)
        u = v = None
        if sigma is not None:
            u, v = torch.randn(cout, device=DEV), torch.randn(cin * k * k, device=DEV)
            sn_jobs.append(L.SnGradJob(ptr(g_multi), ptr(dot_m), ptr(sigma), ptr(u), ptr(v), cout, cin * k * k))
        checks.append((acc, g_multi, g_single, dot_m, dot_s, sigma, u, v, cout, cin, k))
    call_multi("p2p_unpack_wgrad_multi", unpack_jobs)
    for acc, g_multi, g_single, dot_m, dot_s, sigma, u, v, cout, cin, k in checks:
        if sigma is None:
            assert torch.equal(g_multi, g_single)
        else:
            assert relerr(g_multi, g_single) <= 1e-6
            assert abs(dot_m.item() - dot_s.item()) <= 1e-4 * max(abs(dot_s.item()), 1.0)
            call("p2p_sn_weight_grad", ptr(g_single), ptr(dot_s), ptr(sigma), ptr(u), ptr(v), cout, cin * k * k)
    call_multi("p2p_sn_weight_grad_multi", sn_jobs)
    for acc, g_multi, g_single, dot_m, dot_s, sigma, u, v, cout, cin, k in checks:
        assert relerr(g_multi, g_single) <= 1e-4


@pytest.mark.parametrize("shape", [(2, 32, 3, 20, 28, 9), (1, 16, 2, 13, 300, 5)])
def test_fat_column_conv_matches_conv2d(shape):
    """(k x 1) convolution with k*cout fat channels + horizontal gather == the k x k convolution; its dgrad too."""
    B, cin, cout, H, W, k = shape
    dt = torch.bfloat16
    g = torch.Generator(device="cpu").manual_seed(21)
    x = torch.randn(B, cin, H + k - 1, W + k - 1, generator=g).to(DEV)
    w = (torch.randn(cout, cin, k, k, generator=g) / math.sqrt(cin * k * k)).to(DEV)
    bias = torch.randn(cout, generator=g).to(DEV)
    Ci, Co = E.pad16(cin), E.pad16(cout)
    xs = to_stored(x, Ci, dt)
    assert E.FatColumnConv.applies(cout, k, dt)
    fat = E.FatColumnConv(B, H + k - 1, W + k - 1, Ci, cin, cout, k)
    L.call_multi("p2p_pack_weights_multi", [fat.pack_job(w, dt)], L.BF16)
    out = torch.full((B, H, W, Co), 5.0, dtype=dt, device=DEV)
    fat.fwd(xs, bias, out)
    xr = from_stored(xs, cin).float().requires_grad_(True)
    wr = w.to(dt).float()
    ref = F.conv2d(xr, wr, bias)
    assert relerr(from_stored(out, cout), ref) <= 1.5e-2          # one extra bf16 rounding of the per-kx partial sums
    assert out[..., cout:].abs().max().item() == 0
    go = torch.randn(ref.shape, generator=g).to(DEV)
    gos = to_stored(go, Co, dt)
    ref.backward(from_stored(gos, cout).float())
    din = torch.full_like(xs, 5.0)
    fat.dgrad(gos, din)
    assert relerr(from_stored(din, cin), xr.grad) <= 1e-2
    if Ci > cin:
        assert din[..., cin:].abs().max().item() == 0


def test_decode_u8_is_bit_identical_to_totensor_normalize():
    """p2p_decode_u8 against the reference's input pipeline (dataset.py:19-20,30-40: ToTensor, Normalize(0.5, 0.5))
    and its quantiser (generate_dataset.py:29-34), all 256 byte values and a ragged length."""
    from oracle import pix2pix_oracle as orc
    g = torch.Generator().manual_seed(3)
    u8 = torch.cat([torch.arange(256, dtype=torch.uint8), torch.randint(0, 256, (3 * 37 * 41 + 5,), generator=g, dtype=torch.uint8)])
    t = u8.float().div(255)                       # ToTensor
    ref = t.clone().sub_(0.5).div_(0.5)           # Normalize (torchvision: tensor.sub_(mean).div_(std))
    ref_q = orc.compress(ref * 0.5 + 0.5, 3)
    src = u8.to(DEV)
    dst = torch.empty(u8.numel(), device=DEV)
    q = torch.empty(u8.numel(), device=DEV)
    L.call("p2p_decode_u8", src.data_ptr(), dst.data_ptr(), q.data_ptr(), u8.numel(), 3)
    assert torch.equal(dst.cpu(), ref)
    assert torch.equal(q.cpu(), ref_q)
    dst2 = torch.empty(u8.numel(), device=DEV)
    L.call("p2p_decode_u8", src.data_ptr(), dst2.data_ptr(), None, u8.numel(), 3)
    assert torch.equal(dst2, dst)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("training", [True, False])
def test_fused_bn_finalize_act_store_is_bit_identical(dt, training):
    """p2p_bn_finalize_act_store == p2p_bn_finalize followed by p2p_bn_act_store: stored map, `sb`, running buffers and
    the batch counter, bit for bit (skip connection, upsample and reflection pad on)."""
    torch.manual_seed(11)
    B, Cc, H, W, up, pad, act = 3, 64, 10, 14, 2, 1, L.ACT_PRELU
    dc = L.dtype_code(dt)
    r = (torch.randn(B, H, W, Cc, device=DEV) * 2 + 0.3).to(dt)
    skip = torch.randn(B, H + 2, W + 2, Cc, device=DEV).to(dt)
    gamma, beta = torch.rand(Cc, device=DEV) + 0.5, torch.randn(Cc, device=DEV)
    a = torch.tensor([0.25], device=DEV)
    stats = torch.zeros(2 * Cc, dtype=torch.float64, device=DEV)
    L.call("p2p_channel_stats", r.data_ptr(), stats.data_ptr(), B * H * W, Cc, dc)
    res = []
    for fused in (False, True):
        rm, rv = torch.full((Cc,), 0.1, device=DEV), torch.full((Cc,), 1.3, device=DEV)
        nbt = torch.zeros((), dtype=torch.int64, device=DEV)
        sb = torch.zeros(4 * Cc, device=DEV)
        dst = torch.zeros(B, H * up + 2 * pad, W * up + 2 * pad, Cc, dtype=dt, device=DEV)
        if fused:
            L.call("p2p_bn_finalize_act_store", stats.data_ptr(), B * H * W, gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                   rv.data_ptr(), nbt.data_ptr(), sb.data_ptr(), int(training), r.data_ptr(), skip.data_ptr(), 1,
                   a.data_ptr(), dst.data_ptr(), B, H, W, Cc, Cc, act, up, pad, dc)
        else:
            L.call("p2p_bn_finalize", stats.data_ptr(), B * H * W, gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                   rv.data_ptr(), nbt.data_ptr(), sb.data_ptr(), Cc, Cc, int(training))
            L.call("p2p_bn_act_store", r.data_ptr(), sb.data_ptr(), skip.data_ptr(), 1, a.data_ptr(), dst.data_ptr(), None,
                   B, H, W, Cc, Cc, act, up, pad, dc)
        torch.cuda.synchronize()
        res.append((dst, sb, rm, rv, nbt))
    for x, y in zip(*res):
        assert torch.equal(x, y)
    assert res[1][4].item() == (1 if training else 0)


def test_zero_fill():
    for n in (1 << 20, (1 << 22) + 7, 33_214_016):
        t = torch.full((n + 32,), 7, dtype=torch.uint8, device=DEV)
        L.call("p2p_zero_fill", t.data_ptr() + 16, n)
        torch.cuda.synchronize()
        assert t[16:16 + n].count_nonzero().item() == 0
        assert t[:16].eq(7).all().item() and t[16 + n:].eq(7).all().item()
