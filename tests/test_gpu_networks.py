"""Whole-network and whole-step parity of the CUDA path (through the drop-in nn.Module API, which calls the
C ABI) against the CPU oracle on the same seeded weights and inputs, plus the committed golden vectors that
were generated from the unmodified reference.

Tolerances, and where they come from (measured with tests/diagnostics/debug_g.py on a B200, batch 2, 32x32):
  * every kernel alone meets 1e-5 (fp32) / 1e-2 (bf16, same rounded operands): tests/test_gpu_ops.py;
  * through the generator rounding differences grow ~x10 (24 conv+BatchNorm stages on random-init weights):
    fp32 mode 1.2e-6 after conv1 -> 1.6e-5 rms / 4.5e-5 max at the output; bf16 mode 0.44% -> 4.1% rms,
    13% max.  The north star's "1e-2" example is therefore stated here per network as: fp32 mode
    <= 2e-4 (max), bf16 mode <= 6e-2 rms and <= 0.2 max for the generator output, losses <= 5e-2 relative;
  * whole-network GRADIENTS are compared loosely: a forward difference of 1e-5 flips the ReLU/PReLU mask of
    an isolated element, and at these tiny map sizes (8x8x128 per image) one flipped element changes every
    upstream gradient by ~1% rms (the jumps are visible per layer in tests/diagnostics/debug_g.py).  fp32: rms <= 5e-2;
    bf16 (where ~1% of the masks flip per layer): cosine similarity >= 0.85.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import pix2pix_oracle as orc
from oracle.make_golden_util import CASES

DEV = "cuda"
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def relerr(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)


def cosine(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return (a.dot(b) / (a.norm() * b.norm()).clamp_min(1e-30)).item()


def rms_relerr(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt().clamp_min(1e-30)).item()


def build(precision):
    import p2p_pytorch_b200 as P
    P.set_precision(precision)
    g = P.ExpandNetwork().to(DEV)
    d = P.MultiscaleDiscriminator(6, 64, getIntermFeat=True).to(DEV)
    g.load_state_dict(orc.init_g_state(123))
    d.load_state_dict(orc.init_d_state(124))
    g.train(); d.train()
    return P, g, d


@pytest.fixture(autouse=True)
def _restore_precision():
    yield
    import p2p_pytorch_b200 as P
    P.set_precision("bf16")


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("bf16-simt", 0.2), ("bf16", 0.2)])
def test_generator_forward_backward(precision, tol):
    P, g, _ = build(precision)
    a, b, c = orc.synthetic_batch(2, 32, 32, 5)
    og = orc.init_g_state(123)
    for k in orc.g_trainable(og):
        og[k].requires_grad_(True)
    y_ref = orc.g_forward(og, c, True)
    y = g(c.to(DEV))
    assert y.shape == y_ref.shape and y.dtype == torch.float32
    assert relerr(y, y_ref) <= tol, "generator output"
    assert rms_relerr(y, y_ref) <= (1e-4 if precision == "fp32" else 6e-2), "generator output (rms)"
    # BatchNorm running buffers follow the reference's update
    sd = g.state_dict()
    for k in ("in1_e.running_mean", "res5.in2.running_var", "in1_d.running_mean"):
        assert relerr(sd[k], og[k]) <= (1e-3 if precision == "fp32" else 5e-2), k
    assert sd["in2_d.num_batches_tracked"].item() == 1
    torch.manual_seed(0)
    gy = torch.randn_like(y_ref)
    y_ref.backward(gy)
    y.backward(gy.to(DEV))
    for name, p in g.named_parameters():
        ref = og[name].grad
        if name.endswith("conv2d.bias"):
            continue            # analytically zero (BatchNorm follows): float noise on both sides
        assert p.grad is not None, name
        if precision == "fp32":
            assert rms_relerr(p.grad, ref) <= 5e-2, f"grad {name}: {rms_relerr(p.grad, ref):.3e}"
        elif p.numel() > 1:
            assert cosine(p.grad, ref) >= 0.85, f"grad {name}: cos {cosine(p.grad, ref):.4f}"


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16-simt", 1.5e-2), ("bf16", 1.5e-2)])
def test_discriminator_forward_backward(precision, tol):
    P, _, d = build(precision)
    a, b, _ = orc.synthetic_batch(2, 32, 40, 9)
    od = orc.init_d_state(124)
    for k in orc.d_trainable(od):
        od[k].requires_grad_(True)
    xb = b.clone().requires_grad_(True)
    ref = orc.d_forward(od, torch.cat((a, xb), 1))
    xin = torch.cat((a, b), 1).to(DEV).requires_grad_(True)
    out = d(xin)                                        # drop-in API: list[3][5] of NCHW fp32
    assert len(out) == 3 and all(len(s) == 5 for s in out)
    for i in range(3):
        for j in range(5):
            assert out[i][j].shape == ref[i][j].shape, (i, j)
            assert relerr(out[i][j], ref[i][j]) <= tol, f"map {i}{j}"
    # u, v were advanced by the power iteration exactly like the reference's
    sd = d.state_dict()
    for k in ("scale2_layer1.0.module.weight_u", "scale0_layer3.0.module.weight_v"):
        assert relerr(sd[k], od[k]) <= 1e-4, k
    loss_ref = orc.gan_loss(ref, True) + orc.feature_matching_loss(ref, [[torch.zeros_like(t) for t in s] for s in ref])
    crit = P.GANLoss()
    loss = crit(out, True)
    for i in range(3):
        for j in range(4):
            loss = loss + (10.0 / 3) * out[i][j].abs().mean()
    assert abs(loss.item() - loss_ref.item()) <= max(tol, 1e-4) * abs(loss_ref.item())
    loss_ref.backward(); loss.backward()
    gtol = 2e-3 if precision == "fp32" else 6e-2
    for name, p in d.named_parameters():
        if not p.requires_grad:
            continue
        assert p.grad is not None, name
        assert rms_relerr(p.grad, od[name].grad) <= gtol, f"grad {name}: {rms_relerr(p.grad, od[name].grad):.3e}"
    assert rms_relerr(xin.grad[:, 3:], xb.grad) <= gtol, "input gradient"


def _run_trainer(precision, batch, h, w, seed, steps, snapshots=None):
    P, g, d = build(precision)
    tr = P.Pix2PixTrainer(g, d)
    outs = []
    for it in range(steps):
        a, b, c = orc.synthetic_batch(batch, h, w, seed + it)
        outs.append({k: v.float().cpu() for k, v in tr.step(a.to(DEV), b.to(DEV), c.to(DEV)).items()})
        if snapshots is not None:
            snapshots.append(({k: v.detach().float().cpu().clone() for k, v in g.state_dict().items()},
                              {k: v.detach().float().cpu().clone() for k, v in d.state_dict().items()}))
    return tr, g, d, outs


LR = 2e-4


def check_first_adam_update(pre, post, ref_post, ref_grad, what, min_frac=0.99):
    """The first Adam step moves every weight by -lr * sign(g) (m / sqrt(v) = +-1 wherever |g| >> eps): comparing
    post-step VALUES at a tolerance of ~lr cannot tell an update from no update, so compare the UPDATE: on the elements
    whose reference gradient is well clear of zero, the step must have the reference's sign and magnitude lr."""
    pre, post, ref_post, ref_grad = (np.asarray(t, dtype=np.float64).ravel() for t in (pre, post, ref_post, ref_grad))
    sel = np.abs(ref_grad) > max(1e-6, 1e-3 * np.abs(ref_grad).max())
    if sel.sum() == 0:
        return 0, 0
    d_mine, d_ref = (post - pre)[sel], (ref_post - pre)[sel]
    assert np.all(np.abs(np.abs(d_ref) - LR) <= 0.02 * LR), what      # the premise: the reference moved by lr
    ok = (np.sign(d_mine) == np.sign(d_ref)) & (np.abs(np.abs(d_mine) - LR) <= 0.02 * LR)
    assert ok.mean() >= min_frac, f"{what}: only {ok.mean():.4f} of {sel.sum()} updates have the reference's sign and size"
    return int(ok.sum()), int(sel.sum())


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_train_step_fp32_matches_reference_golden(case):
    """The committed golden vectors come from the unmodified reference (oracle/make_golden.py)."""
    from oracle.make_golden_util import fingerprint
    name, batch, h, w, steps, seed = case
    gold = np.load(os.path.join(GOLD, f"ref_step_{name}.npz"))
    snaps = []
    tr, g, d, outs = _run_trainer("fp32", batch, h, w, seed, steps, snaps)
    o = outs[0]
    assert relerr(o["fake_b"], torch.from_numpy(gold["s0.fake_b"])) <= 2e-4
    for nm in ("loss_d", "loss_g_gan", "loss_feat", "loss_tv", "loss_g"):
        assert abs(o[nm].item() - float(gold["s0." + nm])) <= 2e-4 * abs(float(gold["s0." + nm])), nm
    # weights after the two Adam steps, BatchNorm buffers, spectral-norm vectors
    init = (orc.init_g_state(123), orc.init_d_state(124))
    n_ok = n_sel = 0
    for which, pre in ((0, "s0.gstate."), (1, "s0.dstate.")):
        for k, v in snaps[0][which].items():
            if k.endswith("conv2d.bias") or k.endswith("num_batches_tracked"):
                continue
            fp, ref = fingerprint(v), gold[pre + k]
            rms = max(float(np.sqrt(ref[2] / max(v.numel(), 1))), 1e-12)
            assert np.abs(fp[3:] - ref[3:]).max() <= 5e-4 * rms + 4.1e-4, k      # 2*lr: Adam sign flips where |g|~eps
            assert abs(fp[1] - ref[1]) <= 2e-3 * abs(ref[1]) + 1e-6, k
            gk = pre.replace("state.", "grad.") + k
            if gk in gold.files:
                # the UPDATE of the first 13 elements of every trainable tensor: the reference's sign, magnitude lr
                n = min(13, v.numel())
                a, b = check_first_adam_update(init[which][k].flatten()[:n], v.flatten()[:n], ref[3:3 + n],
                                               gold[gk][3:3 + n], k, min_frac=0.0)
                n_ok, n_sel = n_ok + a, n_sel + b
    assert n_sel >= 500 and n_ok >= 0.99 * n_sel, (n_ok, n_sel)
    if steps > 1:
        # second iteration (m and v are no longer +-g): the update now depends on gradient magnitudes
        num = den = 0.0
        for which, pre in ((0, "gstate."), (1, "dstate.")):
            for k, v in snaps[1][which].items():
                gk = "s1." + pre.replace("state.", "grad.") + k
                if gk not in gold.files or k.endswith("conv2d.bias"):
                    continue
                n = min(13, v.numel())
                d_mine = (v.flatten()[:n] - snaps[0][which][k].flatten()[:n]).double().numpy()
                d_ref = gold["s1." + pre + k][3:3 + n] - gold["s0." + pre + k][3:3 + n]
                num += float(((d_mine - d_ref) ** 2).sum())
                den += float((d_ref ** 2).sum())
        # measured 0.145: fp32-mode gradients differ from the reference by up to 5e-2 rms (mask flips), which Adam's
        # m / sqrt(v) amplifies on small elements; "no second update" would give 1.0, a wrong sign 2.0
        assert den > 0 and (num / den) ** 0.5 <= 0.3, (num / den) ** 0.5


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("bf16", 0.2)])
def test_train_step_matches_oracle(precision, tol):
    snaps = []
    tr, g, d, outs = _run_trainer(precision, 2, 32, 32, 21, 1, snaps)
    ot = orc.OracleTrainer(orc.init_g_state(123), orc.init_d_state(124))
    a, b, c = orc.synthetic_batch(2, 32, 32, 21)
    ref = ot.step(a, b, c)
    assert relerr(outs[0]["fake_b"], ref["fake_b"]) <= tol
    ltol = 2e-4 if precision == "fp32" else 5e-2
    for nm in ("loss_d", "loss_g_gan", "loss_feat", "loss_tv", "loss_g"):
        assert abs(outs[0][nm].item() - ref[nm].item()) <= ltol * abs(ref[nm].item()), \
            f"{nm}: {outs[0][nm].item()} vs {ref[nm].item()}"
    gg, dg = ot.grads()
    for name, p in g.named_parameters():
        if name.endswith("conv2d.bias") or p.numel() == 1:
            continue
        if precision == "fp32":
            assert rms_relerr(p.grad, gg[name]) <= 5e-2, f"G grad {name}: {rms_relerr(p.grad, gg[name]):.3e}"
        else:
            assert cosine(p.grad, gg[name]) >= 0.85, f"G grad {name}: cos {cosine(p.grad, gg[name]):.4f}"
    for name, p in d.named_parameters():
        if p.requires_grad:
            if precision == "fp32":
                assert rms_relerr(p.grad, dg[name]) <= 1e-2, f"D grad {name}: {rms_relerr(p.grad, dg[name]):.3e}"
            else:
                assert cosine(p.grad, dg[name]) >= 0.9, f"D grad {name}: cos {cosine(p.grad, dg[name]):.4f}"
    # both optimizer steps happened, in the reference's direction, with the reference's size (whole tensors)
    init = (orc.init_g_state(123), orc.init_d_state(124))
    n_ok = n_sel = 0
    for which, (ref_state, ref_grads) in enumerate(((ot.g, gg), (ot.d, dg))):
        for name, gr in ref_grads.items():
            if name.endswith("conv2d.bias") or gr.numel() == 1:
                continue
            # bf16: ~1 % of the activation masks flip per layer, so a share of the small gradients changes sign
            # (the cosine bars above); per tensor at least half, over all weights at least 85 %
            a, b = check_first_adam_update(init[which][name], snaps[0][which][name], ref_state[name].detach(), gr, name,
                                           min_frac=0.99 if precision == "fp32" else 0.5)
            n_ok, n_sel = n_ok + a, n_sel + b
    assert n_sel > 1_000_000 and n_ok >= (0.99 if precision == "fp32" else 0.85) * n_sel, (n_ok, n_sel)


def test_drop_in_loop_equals_fused_trainer():
    """train.py's loop typed against the drop-in modules (torch.cat, GANLoss, nn.L1Loss, torch.optim.Adam)
    gives the same step as Pix2PixTrainer."""
    P, g, d = build("fp32")
    a, b, c = (t.to(DEV) for t in orc.synthetic_batch(2, 32, 32, 33))
    crit, feat = P.GANLoss(), torch.nn.L1Loss()
    opt_g = torch.optim.Adam(g.parameters(), lr=2e-4, betas=(0.5, 0.999))
    opt_d = torch.optim.Adam(d.parameters(), lr=2e-4, betas=(0.5, 0.999))
    fake_b = g(c)
    pred_fake = d.forward(torch.cat((a, fake_b.detach()), 1))
    loss_d_fake = crit(pred_fake, False)
    pred_real = d.forward(torch.cat((a, b), 1))
    loss_d = (loss_d_fake + crit(pred_real, True)) * 0.5
    pred_fake = d.forward(torch.cat((a, fake_b), 1))
    loss_g = crit(pred_fake, True)
    for i in range(3):
        for j in range(4):
            loss_g = loss_g + (1.0 / 3) * feat(pred_fake[i][j], pred_real[i][j].detach()) * 10.0
    loss_g = loss_g + P.tv_loss(fake_b)
    opt_g.zero_grad(); loss_g.backward(); opt_g.step()
    opt_d.zero_grad(); loss_d.backward(); opt_d.step()
    _, g2, d2, outs = _run_trainer("fp32", 2, 32, 32, 33, 1)
    assert relerr(fake_b, outs[0]["fake_b"]) <= 1e-5
    assert abs(loss_d.item() - outs[0]["loss_d"].item()) <= 1e-5 * abs(loss_d.item())
    assert abs(loss_g.item() - outs[0]["loss_g"].item()) <= 1e-5 * abs(loss_g.item())
    for (k, v), (_, v2) in zip(g.state_dict().items(), g2.state_dict().items()):
        if k.endswith("conv2d.bias"):
            continue
        assert (v.float() - v2.float()).abs().max().item() <= 4.1e-4, k
    for (k, v), (_, v2) in zip(d.state_dict().items(), d2.state_dict().items()):
        assert (v.float() - v2.float()).abs().max().item() <= 4.1e-4, k


def test_state_dict_interchange_and_eval_mode():
    P, g, d = build("fp32")
    assert list(g.state_dict().keys()) == list(orc.init_g_state().keys())
    assert list(d.state_dict().keys()) == list(orc.init_d_state().keys())
    g.eval()
    _, _, c = orc.synthetic_batch(1, 32, 32, 2)
    with torch.no_grad():
        y = g(c.to(DEV))
    og = orc.init_g_state(123)
    assert relerr(y, orc.g_forward(og, c, False)) <= 2e-4
    assert g.state_dict()["in1_e.num_batches_tracked"].item() == 0
    with pytest.raises(Exception):
        g(torch.zeros(1, 3, 30, 32, device=DEV))          # H not a multiple of 4
    with pytest.raises(Exception):
        g(torch.zeros(1, 3, 32, 32))                      # CPU tensor: no fallback path


@pytest.mark.parametrize("precision,graph", [("bf16", True), ("bf16", False), ("fp32", False)],
                         ids=["bf16-graph", "bf16-eager", "fp32-eager"])
def test_loss_curves_track_oracle_200_steps(precision, graph):
    """north star: loss curves track the reference over 200 steps -- in the precision and launch mode the benchmark
    runs (bf16, tcgen05 kernels, CUDA-graph replay), and in the fp32 verification mode.

    GAN training amplifies rounding noise: runs of the ORACLE itself whose initial weights differ by bf16-rounding-size
    noise separate after ~10 iterations.  oracle/make_envelope.py therefore ran the oracle six times at batch 8, 64x64
    (stock init + five perturbed inits, same 200 batches) and committed every per-step loss
    (tests/golden/oracle_envelope_b8_64x64_200.npz: 25-step window means of loss_d spread by 5-27 %, of loss_g by
    6-20 % between those runs).  Those six runs understate the tails: the fp32 verification mode -- exact arithmetic,
    losses equal to the oracle's to 1e-5 at step 0 -- leaves the one-range-widened [min, max] band in 2 of 16 windows
    (by up to 13 % of the window's median) and the bf16 path in 1 (a phase where D is ahead: loss_d -15 %, loss_g +29 %;
    profiles/r02c_loss_curves_200_*.npz).  Bounds, calibrated on that: (1) pointwise agreement with the unperturbed
    oracle run while the trajectories are still one (first 3 steps, 1e-3 relative; measured 1e-5 ... 4e-4);
    (2) every 25-step window mean of loss_d and loss_g within 35 % of the oracle's median for that window, and at least
    6 of the 8 windows of each loss inside the oracle's [min, max] widened by one range on each side; (3) the 200-step
    means within 10 % of the oracle's (its own runs differ by 3.5 %).  A path that trained 1.5x worse or better than the
    reference fails (2) in most windows and (3) outright."""
    env = np.load(os.path.join(GOLD, "oracle_envelope_b8_64x64_200.npz"))
    curves = env["curves"][:, :, :2]                      # [runs, steps, (loss_d, loss_g)]
    batch, size, seed0, steps = int(env["batch"]), int(env["size"]), int(env["data_seed"]), curves.shape[1]
    P, g, d = build(precision)
    tr = P.Pix2PixTrainer(g, d, cuda_graph=graph, graph_warmup=1)
    mine = []
    for it in range(steps):
        a, b, c = orc.synthetic_batch(batch, size, size, seed0 + it)
        o = tr.step(a.to(DEV), b.to(DEV), c.to(DEV))
        mine.append([o["loss_d"].item(), o["loss_g"].item()])
    mine = np.array(mine)
    os.makedirs("gpurun_out", exist_ok=True)
    np.savez(f"gpurun_out/loss_curves_200_{precision}_{'graph' if graph else 'eager'}.npz", mine=mine, oracle=curves)
    assert np.isfinite(mine).all()
    if graph:
        assert len(tr._graphs) == 1 and tr.graph_launches > 100
    assert np.all(np.abs(mine[:3] - curves[0, :3]) <= 1e-3 * np.abs(curves[0, :3])), (mine[:3], curves[0, :3])
    win = 25
    w_or = curves.reshape(curves.shape[0], steps // win, win, 2).mean(2)           # [runs, windows, 2]
    w_me = mine.reshape(steps // win, win, 2).mean(1)
    med, lo, hi = np.median(w_or, 0), w_or.min(0), w_or.max(0)
    rel = np.abs(w_me - med) / med
    assert np.all(rel <= 0.35), (np.round(rel, 3).tolist(), np.round(w_me, 3).tolist(), np.round(med, 3).tolist())
    inside = (w_me >= lo - (hi - lo)) & (w_me <= hi + (hi - lo))
    assert np.all(inside.sum(0) >= inside.shape[0] - 2), (np.round(w_me, 3).tolist(), np.round(lo, 3).tolist(),
                                                          np.round(hi, 3).tolist())
    m_or, m_me = curves.mean(1).mean(0), mine.mean(0)
    assert np.all(np.abs(m_me - m_or) <= 0.10 * m_or), (m_me, m_or)


# ---- compression front end -----------------------------------------------------------------------
from oracle.make_golden_util import C_CASES  # noqa: E402


@pytest.mark.parametrize("case", C_CASES, ids=[c[0] for c in C_CASES])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_compression_front_end_matches_reference_golden(case, precision):
    """define_C's forward and compress(net_c(real_b), 3) against the vectors from the unmodified reference."""
    import p2p_pytorch_b200 as P
    name, batch, h, w, seed, training = case
    gold = np.load(os.path.join(GOLD, f"ref_netc_{name}.npz"))
    P.set_precision(precision)
    net = P.CompressionNetwork().to(DEV)
    net.load_state_dict(orc.init_c_state(125))
    net.train(training)
    _a, real_b, _c = orc.synthetic_batch(batch, h, w, seed)
    y = net(real_b.to(DEV)).cpu().numpy()
    sd = {k: v.cpu() for k, v in net.state_dict().items()}
    net2 = P.CompressionNetwork().to(DEV)
    net2.load_state_dict(orc.init_c_state(125))
    net2.train(training)
    comp = net2.front_end(real_b.to(DEV)).cpu().numpy()
    diff = np.abs(comp - gold["compressed"])
    assert diff.max() <= 1.0 / 7 + 1e-6                      # never more than one quantisation level apart
    if precision == "fp32":
        tol = 2e-5                                           # exact arithmetic
        assert np.abs(y - gold["y"]).max() <= tol
        # levels may flip only where the reference value sits within `tol` of a rounding boundary
        frac = gold["y"].clip(0, 1) * 7
        near = np.abs(frac - np.floor(frac) - 0.5) <= tol * 7 + 1e-6
        assert not np.any((diff > 1e-6) & ~near)
    else:
        # bf16 convolutions, then F.normalize: the residual is a UNIT vector per pixel, so where the 12-channel conv
        # output is small its bf16 error is amplified without bound (measured: 4e-2 ... >0.1 max over the three
        # shapes, ~5e-3 rms on values of order 1): the bar is on the rms and on the 99.9th percentile
        err = np.abs(y - gold["y"])
        assert np.sqrt((err ** 2).mean()) <= 1.5e-2 and np.quantile(err, 0.999) <= 0.1 and err.max() <= 1.0, \
            (np.sqrt((err ** 2).mean()), np.quantile(err, 0.999), err.max())
        assert (diff > 1e-6).mean() <= 0.15                  # rms error 5e-3 against a level spacing of 1/7
    assert set(np.unique(np.round(comp * 7)).tolist()) <= set(range(8))
    if training:
        rtol = 1e-4 if precision == "fp32" else 2e-2
        assert np.allclose(sd["conv_block1.1.running_mean"].numpy(), gold["running_mean"], rtol=rtol, atol=rtol * 1e-1)
        assert np.allclose(sd["conv_block1.1.running_var"].numpy(), gold["running_var"], rtol=rtol, atol=rtol * 1e-1)
        assert int(sd["conv_block1.1.num_batches_tracked"]) == 1
    else:
        assert int(sd["conv_block1.1.num_batches_tracked"]) == 0


def test_define_c_is_a_drop_in():
    import p2p_pytorch_b200 as P
    net = P.define_C("normal", 0.02, gpu_id=DEV)
    assert list(net.state_dict().keys()) == list(orc.init_c_state(125).keys())
    assert sum(p.numel() for p in net.parameters()) == sum(v.numel() for k, v in orc.init_c_state(125).items()
                                                           if "running" not in k and "num_batches" not in k)
    x = torch.rand(2, 3, 64, 64, device=DEV) * 2 - 1
    out = net.front_end(x)
    assert out.shape == x.shape and out.dtype == torch.float32 and not out.requires_grad


# ---- the iteration as executed, with front end and phase 3 ------------------------------------------
from oracle.make_golden_util import P3_CASES, fingerprint as _fp  # noqa: E402


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "graph"])
@pytest.mark.parametrize("case", P3_CASES, ids=[c[0] for c in P3_CASES])
def test_full_iteration_with_phase3_matches_reference_golden(case, graph):
    """fp32 mode against vectors from the unmodified reference modules (oracle/make_golden_p3.py); with graph=True
    the second iteration is recorded into, and replayed from, a CUDA graph."""
    import p2p_pytorch_b200 as P
    name, batch, h, w, steps, seed = case
    gold = np.load(os.path.join(GOLD, f"ref_full_{name}.npz"))
    P, g, d = build("fp32")
    c = P.CompressionNetwork().to(DEV)
    c.load_state_dict(orc.init_c_state(125))
    c.train()
    c0 = {k: v.detach().clone() for k, v in c.named_parameters()}
    tr = P.Pix2PixTrainer(g, d, net_c=c, cuda_graph=graph, graph_warmup=1)
    for it in range(steps):
        a, b, _ = orc.synthetic_batch(batch, h, w, seed + it)
        out = tr.step_full(a.to(DEV), b.to(DEV))
        p = f"s{it}."
        scale = 25.0 if it else 1.0
        for k in ("loss_d", "loss_g_gan", "loss_feat", "loss_tv", "loss_g", "loss_c"):
            assert abs(out[k].item() - float(gold[p + k])) <= 2e-4 * scale * abs(float(gold[p + k])) + 1e-6, (it, k)
        fp_c = _fp(out["compressed"].cpu())[:3]
        assert np.allclose(fp_c, gold[p + "compressed"][:3], rtol=2e-3, atol=0.5)   # a few 1/7 level flips at most
        assert np.allclose(_fp(out["fake_b"].cpu())[:3], gold[p + "fake_b"][:3], rtol=5e-3 * scale, atol=1e-2)
        sd = {**{k: v for k, v in g.state_dict().items()}, **{"net_c." + k: v for k, v in c.state_dict().items()}}
        for k, v in sd.items():
            if "running" in k or "num_batches" in k:
                assert np.allclose(_fp(v.double().cpu())[:3], gold[p + "buf." + k][:3], rtol=5e-3 * scale, atol=2e-3), (it, k)
    assert all(torch.equal(c0[k], v.detach()) for k, v in c.named_parameters())     # quirk Q3: net_c never moves
    assert int(g.state_dict()["in1_e.num_batches_tracked"]) == 2 * steps
    assert int(c.state_dict()["conv_block1.1.num_batches_tracked"]) == steps


# ---- evaluation metric ---------------------------------------------------------------------------------
def test_psnr_on_device_is_exact():
    """p2p_psnr_sse against the oracle's numpy restatement and the vectors from the reference's own function."""
    import p2p_pytorch_b200 as P
    from oracle.make_golden_util import PSNR_CASES, psnr_inputs
    gold = np.load(os.path.join(GOLD, "ref_psnr.npz"))
    for name, h, w, seed, kind in PSNR_CASES:
        a, b = psnr_inputs(h, w, seed, kind)
        got = P.psnr(a.to(DEV), b.to(DEV)).item()
        want = float(gold[name])
        assert (np.isinf(got) and np.isinf(want)) or abs(got - want) <= 1e-12 * abs(want), (name, got, want)
    # a batch at once == image by image
    g = torch.Generator().manual_seed(9)
    a, b = torch.rand(5, 3, 24, 40, generator=g), torch.rand(5, 3, 24, 40, generator=g) * 1.2 - 0.1
    got = P.psnr(a.to(DEV), b.to(DEV)).cpu().numpy()
    want = np.array([orc.psnr(a[i:i + 1], b[i:i + 1]) for i in range(5)])
    assert np.allclose(got, want, rtol=1e-12, atol=0)


# ---- VGG19 perceptual loss --------------------------------------------------------------------------
from oracle.make_golden_util import VGG_CASES  # noqa: E402


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("case", VGG_CASES, ids=[c[0] for c in VGG_CASES])
def test_vgg_loss_matches_reference_golden(case, precision):
    """P.VGGLoss against vectors from the reference's own VGGLoss / Vgg19 classes holding the same seeded weights
    (oracle/make_golden_vgg.py): loss, gradient w.r.t. the generated image, the five tapped activations."""
    import p2p_pytorch_b200 as P
    from p2p_pytorch_b200 import engine as E
    name, batch, h, w, seed = case
    gold = np.load(os.path.join(GOLD, f"ref_vgg_{name}.npz"))
    P.set_precision(precision)
    crit = P.VGGLoss(DEV)
    crit.load_state_dict(orc.init_vgg_state(126))
    a, b, _ = orc.synthetic_batch(batch, h, w, seed)
    x = torch.tanh(a * 1.5).to(DEV).requires_grad_(True)
    loss = crit(x, b.to(DEV))
    loss.backward()
    # the oracle's restatement agrees with the reference vectors too (CPU side of the same pin)
    xo = torch.tanh(a * 1.5).requires_grad_(True)
    lo = orc.vgg_loss(orc.init_vgg_state(126), xo, b)
    lo.backward()
    assert abs(lo.item() - float(gold["loss"])) <= 1e-5 * abs(float(gold["loss"]))
    assert relerr(xo.grad, torch.from_numpy(gold["grad_x"])) <= 1e-4
    if precision == "fp32":
        assert abs(loss.item() - float(gold["loss"])) <= 1e-4 * abs(float(gold["loss"])), (loss.item(), float(gold["loss"]))
        assert relerr(x.grad, torch.from_numpy(gold["grad_x"])) <= 2e-3
        assert rms_relerr(x.grad, torch.from_numpy(gold["grad_x"])) <= 5e-4
    else:
        # bf16 operands through 13 convolutions: loss to 1e-2 (the north star's bf16 example).  The gradient image is
        # compared by direction: four max-pools route each gradient to the arg-max of a 2x2 window, and where the two
        # largest activations of a window differ by less than a bf16 rounding step (~1-2 % of the windows) the winner
        # changes, which moves that gradient to the neighbouring pixel -- measured 0.24 rms on the 32x32 case, all of it
        # such displaced entries (the fp32 mode meets 5e-4)
        assert abs(loss.item() - float(gold["loss"])) <= 1e-2 * abs(float(gold["loss"])), (loss.item(), float(gold["loss"]))
        assert rms_relerr(x.grad, torch.from_numpy(gold["grad_x"])) <= 0.35
        assert cosine(x.grad, torch.from_numpy(gold["grad_x"])) >= 0.93
    assert x.grad.shape == x.shape and x.grad.dtype == torch.float32
    # the target's activations computed ahead of time (what the trainer does beside the generator's forward pass)
    x2 = torch.tanh(a * 1.5).to(DEV).requires_grad_(True)
    loss2 = crit(x2, target_taps=crit.target_features(b.to(DEV)))
    loss2.backward()
    assert abs(loss2.item() - loss.item()) <= 1e-3 * abs(loss.item())
    assert rms_relerr(x2.grad, x.grad) <= (1e-4 if precision == "fp32" else 5e-2)
    with pytest.raises(ValueError):
        crit(x2)


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "graph"])
def test_train_step_with_vgg_matches_reference_golden(graph):
    """One whole iteration with the perceptual term on (train.py:377-380), fp32 mode, against the unmodified reference."""
    import p2p_pytorch_b200 as P
    from oracle.make_golden_util import fingerprint
    gold = np.load(os.path.join(GOLD, "ref_stepvgg_b2_32x32.npz"))
    P, g, d = build("fp32")
    crit = P.VGGLoss(DEV)
    crit.load_state_dict(orc.init_vgg_state(126))
    if graph:                                  # lazy one-time set-up outside the capture
        P2, g0, d0 = build("fp32")
        c0 = P.VGGLoss(DEV); c0.load_state_dict(orc.init_vgg_state(126))
        a, b, c = orc.synthetic_batch(2, 32, 32, 5)
        P.Pix2PixTrainer(g0, d0, vgg=c0).step(a.to(DEV), b.to(DEV), c.to(DEV))
    tr = P.Pix2PixTrainer(g, d, vgg=crit, cuda_graph=graph, graph_warmup=0)
    a, b, c = orc.synthetic_batch(2, 32, 32, 123)
    out = tr.step(a.to(DEV), b.to(DEV), c.to(DEV))
    assert relerr(out["fake_b"], torch.from_numpy(gold["fake_b"])) <= 2e-4
    for nm in ("loss_d", "loss_g_gan", "loss_feat", "loss_tv", "loss_vgg", "loss_g"):
        assert abs(out[nm].item() - float(gold[nm])) <= 2e-4 * abs(float(gold[nm])), (nm, out[nm].item(), float(gold[nm]))
    # generator gradients now carry the perceptual term: fingerprints of the reference's .grad tensors
    worst = 0.0
    for name, p in g.named_parameters():
        if name.endswith("conv2d.bias") or p.numel() == 1:
            continue
        fp, ref = fingerprint(p.grad.float().cpu()), gold["ggrad." + name]
        rel = abs(fp[2] - ref[2]) / max(ref[2], 1e-30)            # sum of squares
        worst = max(worst, rel)
        assert rel <= 0.1, (name, rel)
    # and it differs from the iteration without the term (the gradients really flow through the VGG stack)
    assert abs(out["loss_g"].item() - (out["loss_g_gan"] + out["loss_feat"] + out["loss_tv"]).item()) > 1e-3
