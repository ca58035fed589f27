"""CUDA-graph replay of the training iteration (Pix2PixTrainer(cuda_graph=True)) against the same iteration
launched kernel by kernel, and the device-step Adam entry point against the host-step one.

Trajectories cannot be compared: two EAGER runs from the same weights already drift apart (fp32 mode: losses
1e-4, generator output 0.25 max after seven iterations at batch 2, 32x32 -- tests/diagnostics/graph_vs_eager.py) because
the float atomics of the split-K weight gradients land in a different order each run and Adam's g/sqrt(v)
amplifies that.  So every iteration is compared from the SAME state: before each step the eager twin loads the
graphed trainer's weights, buffers and optimizer state, both take one step on the same batch, and outputs,
gradients and moments must agree to the single-step noise level."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import pix2pix_oracle as orc

DEV = "cuda"


@pytest.fixture(autouse=True)
def _restore_precision():
    yield
    import p2p_pytorch_b200 as P
    P.set_precision("bf16")


def _nets(P):
    g = P.ExpandNetwork().to(DEV)
    d = P.MultiscaleDiscriminator(6, 64, getIntermFeat=True).to(DEV)
    g.load_state_dict(orc.init_g_state(123))
    d.load_state_dict(orc.init_d_state(124))
    return g, d


def _rms_rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt().clamp_min(1e-30)).item()


@pytest.mark.parametrize("precision,tol_out,tol_grad", [("fp32", 1e-5, 1e-3), ("bf16", 1e-2, 5e-2)])
def test_graph_replay_matches_eager_step_by_step(precision, tol_out, tol_grad):
    import p2p_pytorch_b200 as P
    P.set_precision(precision)
    g1, d1 = _nets(P)
    g2, d2 = _nets(P)
    tr_graph = P.Pix2PixTrainer(g1, d1, cuda_graph=True, graph_warmup=1)
    tr_eager = P.Pix2PixTrainer(g2, d2)
    for it in range(6):
        g2.load_state_dict(g1.state_dict())
        d2.load_state_dict(d1.state_dict())
        tr_eager.optimizer_g.load_state_dict(tr_graph.optimizer_g.state_dict())
        tr_eager.optimizer_d.load_state_dict(tr_graph.optimizer_d.state_dict())
        a, b, c = orc.synthetic_batch(2, 32, 32, 50 + it)
        dev_in = [t.to(DEV) for t in (a, b, c)]
        # host tensors are legal inputs of the graphed step (copied into the graph's input buffers)
        graph_in = [t.pin_memory() for t in (a, b, c)] if it % 2 else dev_in
        out_g = {k: v.detach().float().clone() for k, v in tr_graph.step(*graph_in).items()}
        out_e = {k: v.detach().float() for k, v in tr_eager.step(*dev_in).items()}
        for k in ("loss_d", "loss_g", "loss_feat", "loss_tv", "loss_g_gan"):
            assert abs(out_g[k].item() - out_e[k].item()) <= tol_out * abs(out_e[k].item()), (it, k)
        assert (out_g["fake_b"] - out_e["fake_b"]).abs().max().item() <= tol_out, it
        for og, oe in ((tr_graph.optimizer_g, tr_eager.optimizer_g), (tr_graph.optimizer_d, tr_eager.optimizer_d)):
            assert _rms_rel(og.flat_g, oe.flat_g) <= tol_grad, (it, _rms_rel(og.flat_g, oe.flat_g))
            assert _rms_rel(og.exp_avg, oe.exp_avg) <= tol_grad, it
            assert _rms_rel(og.exp_avg_sq, oe.exp_avg_sq) <= 2 * tol_grad, it
            assert og.steps == oe.steps == it + 1
        for (n, b1), (_, b2) in zip(list(g1.named_buffers()) + list(d1.named_buffers()),
                                    list(g2.named_buffers()) + list(d2.named_buffers())):
            assert torch.allclose(b1.float(), b2.float(), rtol=20 * tol_out, atol=20 * tol_out), (it, n)
    assert len(tr_graph._graphs) == 1 and tr_graph.graph_launches > 100


def test_graph_recaptures_per_shape():
    import p2p_pytorch_b200 as P
    g, d = _nets(P)
    tr = P.Pix2PixTrainer(g, d, cuda_graph=True, graph_warmup=1)
    for it, (bsz, h, w) in enumerate([(2, 32, 32), (2, 32, 32), (1, 64, 48), (2, 32, 32), (1, 64, 48)]):
        a, b, c = orc.synthetic_batch(bsz, h, w, it)
        out = tr.step(a.to(DEV), b.to(DEV), c.to(DEV))
        assert out["fake_b"].shape == (bsz, 3, h, w)
        assert torch.isfinite(out["loss_g"]).item() and torch.isfinite(out["loss_d"]).item()
    assert len(tr._graphs) == 2 and tr.optimizer_g.steps == 5


def test_adam_device_step_matches_host_step():
    from p2p_pytorch_b200._lib import call, ptr
    gen = torch.Generator(device=DEV).manual_seed(5)
    n = 100_003 // 4 * 4 + 4
    p0 = torch.randn(n, device=DEV, generator=gen)
    state = [[p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)] for _ in range(2)]
    step_t = torch.zeros(1, dtype=torch.int32, device=DEV)
    for step in range(1, 8):
        grad = torch.randn(n, device=DEV, generator=gen)
        p, m, v = state[0]
        call("p2p_adam_step", ptr(p), ptr(grad), ptr(m), ptr(v), n, 2e-4, 0.5, 0.999, 1e-8, step, 0.5)
        p, m, v = state[1]
        step_t.add_(1)
        call("p2p_adam_step_dev", ptr(p), ptr(grad), ptr(m), ptr(v), n, 2e-4, 0.5, 0.999, 1e-8, ptr(step_t), 0.5)
    for x, y in zip(*state):
        assert torch.allclose(x, y, rtol=1e-6, atol=1e-9)


def test_prefetch_stages_the_next_batch():
    """step() after prefetch() of the same pinned tensors == step() that copies them itself (fp32 mode)."""
    import p2p_pytorch_b200 as P
    P.set_precision("fp32")
    batches = [tuple(t.pin_memory() for t in orc.synthetic_batch(2, 32, 32, 70 + i)) for i in range(5)]
    losses = {}
    for use_prefetch in (False, True):
        g, d = _nets(P)
        tr = P.Pix2PixTrainer(g, d, cuda_graph=True, graph_warmup=1)
        rec = []
        if use_prefetch:
            tr.prefetch(*batches[0])
        for i, bt in enumerate(batches):
            out = tr.step(*bt)
            if use_prefetch and i + 1 < len(batches):
                tr.prefetch(*batches[i + 1])
            rec.append((out["loss_d"].item(), out["loss_g"].item()))
        losses[use_prefetch] = rec
    for (d0, g0), (d1, g1) in zip(losses[False], losses[True]):
        assert abs(d0 - d1) <= 1e-3 * abs(d0) and abs(g0 - g1) <= 1e-3 * abs(g0)


def test_lr_scheduler_reaches_the_captured_adam_step():
    """The reference steps its LR schedulers every epoch (train.py:441-443; get_scheduler, networks.py:104-109 builds a
    LambdaLR).  Under cuda_graph=True the Adam launches are replayed, so lr must not be frozen at capture time: after
    the scheduler halves it, a replayed step moves the weights half as far, exactly like the eager trainer."""
    import p2p_pytorch_b200 as P
    P.set_precision("fp32")
    batches = [tuple(t.to(DEV) for t in orc.synthetic_batch(2, 32, 32, 90 + i)) for i in range(6)]
    deltas = {}
    for graph in (False, True):
        g, d = _nets(P)
        tr = P.Pix2PixTrainer(g, d, cuda_graph=graph, graph_warmup=1)
        scheds = [torch.optim.lr_scheduler.LambdaLR(o, lr_lambda=lambda e: 0.5 ** e)
                  for o in (tr.optimizer_g, tr.optimizer_d)]
        rec = []
        for i, bt in enumerate(batches):
            before = (tr.optimizer_g.flat_p.clone(), tr.optimizer_d.flat_p.clone())
            tr.step(*bt)
            rec.append(tuple((o.flat_p - b).abs().mean().item() for o, b in zip((tr.optimizer_g, tr.optimizer_d), before)))
            if i == 2:                      # "end of epoch": captured after step 1 (graph_warmup=1), so this is post-capture
                for s in scheds:
                    s.step()
                assert abs(tr.optimizer_g.param_groups[0]["lr"] - 1e-4) < 1e-12
        deltas[graph] = rec
        if graph:
            assert len(tr._graphs) == 1
    for graph in (False, True):
        for net in (0, 1):
            # Adam's per-step movement is ~lr while m / sqrt(v) stays of order one: halving lr halves it
            ratio = deltas[graph][3][net] / deltas[graph][2][net]
            assert 0.35 <= ratio <= 0.65, (graph, net, ratio, deltas[graph])
    for i in range(len(batches)):
        for net in (0, 1):
            assert abs(deltas[True][i][net] - deltas[False][i][net]) <= 0.05 * deltas[False][i][net], (i, net, deltas)


def test_adam_graph_entry_reads_lr_and_scale_from_device():
    from p2p_pytorch_b200._lib import call, ptr
    gen = torch.Generator(device=DEV).manual_seed(6)
    n = 50_000
    p0 = torch.randn(n, device=DEV, generator=gen)
    state = [[p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)] for _ in range(2)]
    step_t = torch.zeros(1, dtype=torch.int32, device=DEV)
    hyper = torch.zeros(2, device=DEV)
    for step in range(1, 6):
        lr, gs = 2e-4 * 0.7 ** step, 1.0 / step
        grad = torch.randn(n, device=DEV, generator=gen)
        p, m, v = state[0]
        call("p2p_adam_step", ptr(p), ptr(grad), ptr(m), ptr(v), n, lr, 0.5, 0.999, 1e-8, step, gs)
        p, m, v = state[1]
        step_t.add_(1)
        hyper.copy_(torch.tensor([lr, gs]))
        call("p2p_adam_step_graph", ptr(p), ptr(grad), ptr(m), ptr(v), n, 0.5, 0.999, 1e-8, ptr(step_t), ptr(hyper))
    for x, y in zip(*state):
        assert torch.allclose(x, y, rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "graph"])
def test_uint8_batches_equal_float_batches(graph):
    """step(uint8 a, uint8 b) == step(ToTensor+Normalize(a), ..(b), compress(b*0.5+0.5, 3)): the device-side decode
    is bit-identical to the reference's input pipeline, so the two iterations agree to the run-to-run noise of the
    split-K atomics (fp32 mode).  Also through prefetch() with pinned uint8 tensors."""
    import p2p_pytorch_b200 as P
    P.set_precision("fp32")
    gen = torch.Generator().manual_seed(11)
    rec = {}
    for mode in ("float", "u8"):
        g, d = _nets(P)
        tr = P.Pix2PixTrainer(g, d, cuda_graph=graph, graph_warmup=1)
        losses = []
        gen.manual_seed(11)
        batches = [(torch.randint(0, 256, (2, 3, 32, 32), generator=gen, dtype=torch.uint8),
                    torch.randint(0, 256, (2, 3, 32, 32), generator=gen, dtype=torch.uint8)) for _ in range(4)]
        if mode == "u8":
            batches = [tuple(t.pin_memory() for t in bt) for bt in batches]
        for i, (a8, b8) in enumerate(batches):
            if mode == "float":
                fa, fb = (t.float().div(255).sub_(0.5).div_(0.5) for t in (a8, b8))
                out = tr.step(fa.to(DEV), fb.to(DEV), orc.compress(fb * 0.5 + 0.5, 3).to(DEV))
            else:
                out = tr.step(a8, b8)
                if graph and i + 1 < len(batches):
                    tr.prefetch(*batches[i + 1])
            losses.append((out["loss_d"].item(), out["loss_g"].item(), out["fake_b"].float().abs().sum().item()))
        rec[mode] = losses
    for x, y in zip(rec["float"], rec["u8"]):
        for u, v in zip(x, y):
            assert abs(u - v) <= 1e-3 * abs(u), (rec["float"], rec["u8"])


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5), ("bf16", 1e-2)])
def test_two_stream_iteration_equals_single_stream(precision, tol):
    """Pix2PixTrainer(two_lanes=True) -- D's training on a second CUDA stream, its three power iterations hoisted to the
    front -- against the single-stream order of the reference, kernel by kernel launched, from the same state on the
    same batches: outputs, gradients and spectral-norm vectors agree to the run-to-run noise of the float atomics."""
    import p2p_pytorch_b200 as P
    P.set_precision(precision)
    g1, d1 = _nets(P)
    g2, d2 = _nets(P)
    tr2 = P.Pix2PixTrainer(g1, d1, two_lanes=True)
    tr1 = P.Pix2PixTrainer(g2, d2, two_lanes=False)
    assert tr2.two_lanes and not tr1.two_lanes
    for it in range(4):
        g2.load_state_dict(g1.state_dict())
        d2.load_state_dict(d1.state_dict())
        tr1.optimizer_g.load_state_dict(tr2.optimizer_g.state_dict())
        tr1.optimizer_d.load_state_dict(tr2.optimizer_d.state_dict())
        batch = [t.to(DEV) for t in orc.synthetic_batch(2, 32, 32, 150 + it)]
        o2 = {k: v.detach().float().clone() for k, v in tr2.step(*batch).items()}
        o1 = {k: v.detach().float() for k, v in tr1.step(*batch).items()}
        torch.cuda.synchronize()
        for k in ("loss_d", "loss_g", "loss_feat", "loss_tv", "loss_g_gan"):
            assert abs(o2[k].item() - o1[k].item()) <= tol * abs(o1[k].item()), (it, k, o2[k].item(), o1[k].item())
        assert (o2["fake_b"] - o1["fake_b"]).abs().max().item() <= tol, it
        for oa, ob in ((tr2.optimizer_g, tr1.optimizer_g), (tr2.optimizer_d, tr1.optimizer_d)):
            assert _rms_rel(oa.flat_g, ob.flat_g) <= (1e-3 if precision == "fp32" else 5e-2), (it, _rms_rel(oa.flat_g, ob.flat_g))
        for (n, b1), (_, b2) in zip(d1.named_parameters(), d2.named_parameters()):
            if n.endswith(("weight_u", "weight_v")):
                assert torch.allclose(b1, b2, rtol=1e-5, atol=1e-7), (it, n)


@pytest.mark.parametrize("with_vgg", [False, True], ids=["plain", "vgg"])
def test_multi_stream_graph_is_reproducible_run_to_run(with_vgg):
    """Race check for the streams inside the captured iteration (two lanes, priorities, weight-gradient and coarse-scale
    helper streams, the perceptual lanes): three trainers from the same state on the same six batches must produce the
    same losses at every iteration up to the order of the float atomics (fp32 mode: measured <=3e-4 after six
    iterations).  A missing dependency between streams shows up as a run that differs by per cents."""
    import p2p_pytorch_b200 as P
    P.set_precision("fp32")
    batches = [tuple(t.to(DEV) for t in orc.synthetic_batch(2, 32, 32, 170 + i)) for i in range(6)]
    runs = []
    for _r in range(3):
        g, d = _nets(P)
        vgg = None
        if with_vgg:
            vgg = P.VGGLoss(DEV)
            vgg.load_state_dict(orc.init_vgg_state(126))
        tr = P.Pix2PixTrainer(g, d, cuda_graph=True, graph_warmup=1, vgg=vgg)
        assert tr.two_lanes
        rec = []
        for bt in batches:
            o = tr.step(*bt)
            rec.append([o["loss_d"].item(), o["loss_g"].item()])
        runs.append(torch.tensor(rec, dtype=torch.float64))
    for other in runs[1:]:
        rel = ((other - runs[0]).abs() / runs[0].abs()).max().item()
        assert rel <= 2e-3, (rel, runs[0].tolist(), other.tolist())
