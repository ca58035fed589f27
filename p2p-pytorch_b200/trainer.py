"""The training iteration of the reference (train.py:294-390, VGG term off) on the fused kernels, with
optional batch data parallelism (one process per GPU, NCCL all-reduce of the flat gradient buffers).

Differences from running the reference loop verbatim on the drop-in modules (all result-preserving):
  * the three discriminator passes take (real_a, image) pairs -- no torch.cat copy (train.py:308,315,336);
  * the third pass treats D's weights as constants: the reference computes those weight gradients during
    `loss_g.backward()` and throws them away at `optimizer_d.zero_grad()` (train.py:385-388);
  * losses are the fused LSGAN / feature-matching / TV kernels on the stored maps;
  * no `.item()` host syncs: `step()` returns device scalars.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

from . import engine as E
from .losses import feature_matching_loss, lsgan_loss, tv_loss
from .optim import FusedAdam


class GradReducer:
    """Sum-all-reduce of a flat gradient buffer in buckets on the NCCL stream; the mean is folded into Adam
    (grad_scale = 1/world).  Buckets are slices of the flat buffer, launched asynchronously so that the other
    network's backward overlaps the transfer (D's gradients are reduced while G's backward runs)."""

    def __init__(self, group=None, bucket_bytes: int = 64 << 20):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.bucket = max(1, bucket_bytes // 4)
        self.pending = []

    def launch(self, flat: torch.Tensor):
        if self.world == 1:
            return
        n = flat.numel()
        # last layers' gradients are complete first and sit at the end of the flat buffer: send back to front
        end = n
        while end > 0:
            beg = max(0, end - self.bucket)
            self.pending.append(self.dist.all_reduce(flat[beg:end], op=self.dist.ReduceOp.SUM, group=self.group,
                                                     async_op=True))
            end = beg

    def wait(self):
        for h in self.pending:
            h.wait()
        self.pending = []


class Pix2PixTrainer:
    """optimizer_g / optimizer_d loop of train.py:241-242, :294-390."""

    def __init__(self, net_g, net_d, lr: float = 2e-4, beta1: float = 0.5, process_group=None,
                 bucket_bytes: int = 64 << 20, cuda_graph: bool = False, graph_warmup: int = 2, net_c=None, vgg=None,
                 two_lanes: Optional[bool] = None):
        self.net_g, self.net_d, self.net_c = net_g, net_d, net_c
        # vgg: a VGGLoss module (criterionVGG, train.py:193) -- adds content_loss = vgg(fake_b, real_b) * 10 to loss_g
        # (train.py:377-380) and, in step_full, c_percp = vgg(compressed, real_b) * 10 to the logged phase-3 loss (:395)
        self.vgg = vgg
        # cuda_graph=True: the first `graph_warmup` calls of step() run eagerly, the next one records the whole
        # iteration (~650 launches, both backward passes, the all-reduces and both Adam updates) into one CUDA
        # graph per input shape, and every later call is three input copies and one graph launch.  The host
        # needs ~10 ms to enqueue an iteration kernel by kernel, which bounds small batches; a replay needs
        # well under 1 ms.  The returned tensors are then the graph's own output buffers: each call overwrites
        # the values returned by the previous one (clone what must be kept).
        self.cuda_graph, self.graph_warmup = bool(cuda_graph), int(graph_warmup)
        self._graphs, self._eager_calls = {}, 0
        self._prefetched, self._copy_stream = None, None
        self.graph_launches = 0          # kernels of libp2p_b200.so recorded in the last captured graph
        self.optimizer_g = FusedAdam(net_g.parameters(), lr=lr, betas=(beta1, 0.999))
        self.optimizer_d = FusedAdam([p for p in net_d.parameters() if p.requires_grad], lr=lr, betas=(beta1, 0.999))
        if "P2P_BUCKET_MB" in os.environ:                # tuning knob, see DESIGN.md section 5
            bucket_bytes = int(os.environ["P2P_BUCKET_MB"]) << 20
        self.red_g = GradReducer(process_group, bucket_bytes)
        self.red_d = GradReducer(process_group, bucket_bytes)
        self.world = self.red_g.world
        # overlap=True (default): D's all-reduce starts right after loss_d.backward() and travels while G's backward runs;
        # P2P_DP_OVERLAP=0 starts both all-reduces after both backward passes.  Measured (ms/step, N=1 / N=2 / N=8):
        # one 8-GPU box overlapped 9.73 / 10.08 / 10.21 (profiles/r01e_bench_N*.json), not overlapped 9.75 / 10.17 /
        # 10.31 (profiles/r02g_bench_N*.json); a 2-GPU box overlapped 9.83 / 10.13, not overlapped 9.83 / 9.95
        # (profiles/r02f_scale2c.txt).  Overlapping hides the transfer, but NCCL's CTAs take SMs from the persistent
        # one-CTA-per-SM convolution kernels, whose 148-CTA grids then need a second wave; which effect wins depends on
        # the box, and the 8-GPU boxes favour overlapping.
        self.overlap = os.environ.get("P2P_DP_OVERLAP", "1") != "0"
        self.optimizer_g.grad_scale = self.optimizer_d.grad_scale = 1.0 / self.world
        self._c_reals = [c[1] for c in net_d._cfg]
        # two_lanes: the discriminator's own training (passes 1-2, loss_d, its backward, Adam D) runs on a second CUDA
        # stream beside the generator's (G forward, pass 3, loss_g, its backward, Adam G); see _step_two_lanes.
        # Default: on whenever the iteration is replayed from a CUDA graph (kernel-by-kernel launching is host-bound,
        # a second stream gains nothing there).  P2P_LANES=1 / 2 overrides.
        env_lanes = os.environ.get("P2P_LANES")
        self.two_lanes = (self.cuda_graph if two_lanes is None else bool(two_lanes)) if env_lanes is None else env_lanes == "2"
        self._side = self._side_vgg = self._side_vgg_lo = None
        if self.cuda_graph:
            self.optimizer_g.make_capturable()
            self.optimizer_d.make_capturable()

    def _sync_hyper(self):
        """learning rates changed by a scheduler since the last call reach the captured Adam launches (device memory)"""
        self.optimizer_g.sync_hyper()
        self.optimizer_d.sync_hyper()

    def broadcast_parameters(self, src: int = 0):
        """identical replicas at start (rank `src` wins), including BatchNorm buffers and SN u/v"""
        if self.world == 1:
            return
        import torch.distributed as dist
        dist.broadcast(self.optimizer_g.flat_p, src)
        dist.broadcast(self.optimizer_d.flat_p, src)
        for net in (self.net_g, self.net_d):
            for b in net.buffers():
                dist.broadcast(b, src)
            for p in net.parameters():
                if not p.requires_grad:
                    dist.broadcast(p.data, src)

    def step(self, real_a: torch.Tensor, real_b: torch.Tensor, g_input: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """One iteration (train.py:294-390, VGG term off).  Inputs, host (pinned) or device tensors:
          * fp32 [B,3,H,W] in [-1,1] (what the reference's DataLoader yields, dataset.py:30-40) plus the generator
            input `g_input` (train.py:297-300 feeds compress(net_c(real_b), 3); step_full() computes that);
          * or uint8 [B,3,H,W] images: ToTensor + Normalize run on the device (p2p_decode_u8, bit-identical to
            torchvision's fp32 arithmetic), a quarter of the host-to-device bytes; with g_input=None the generator
            input is compress(real_b * 0.5 + 0.5, 3) (generate_dataset.py:29-34), made in the same pass."""
        if real_a.dtype != real_b.dtype or real_a.dtype not in (torch.float32, torch.uint8):
            raise ValueError("real_a / real_b must both be float32 or both be uint8")
        if g_input is None and real_a.dtype != torch.uint8:
            raise ValueError("float inputs need g_input (step_full() derives it with the compression network)")
        args = (real_a, real_b) if g_input is None else (real_a, real_b, g_input)
        if not self.cuda_graph:
            return self._step_any(*[t.cuda(non_blocking=True) for t in args])
        key = (tuple((tuple(t.shape), t.dtype) for t in args), E.get_precision())
        rec = self._graphs.get(key)
        if rec is None:
            if self._eager_calls < self.graph_warmup:
                # lazily created workspaces, kernel attributes and NCCL communicators come to life here
                self._eager_calls += 1
                return self._step_any(*[t.cuda(non_blocking=True) for t in args])
            rec = self._capture(args)
            self._graphs[key] = rec
        pre = self._prefetched
        if pre is not None and pre["ids"] == tuple(id(t) for t in args):
            # the batch was staged by prefetch(): its host->device copy ran beside the previous iteration
            torch.cuda.current_stream().wait_event(pre["event"])
            srcs = pre["staged"]
        else:
            srcs = args
        self._prefetched = None
        for dst, src in zip(rec["inputs"], srcs):
            dst.copy_(src, non_blocking=True)
        if pre is not None and srcs is pre["staged"]:
            pre["free"].record(torch.cuda.current_stream())         # the staging slot may be overwritten from here on
        self._sync_hyper()
        rec["graph"].replay()
        return rec["outputs"]

    def _decode_u8(self, a8: torch.Tensor, b8: torch.Tensor, want_g_input: bool):
        from . import _lib as L
        a8, b8 = a8.contiguous(), b8.contiguous()
        ra = torch.empty(a8.shape, dtype=torch.float32, device=a8.device)
        rb = torch.empty(b8.shape, dtype=torch.float32, device=b8.device)
        gi = torch.empty(b8.shape, dtype=torch.float32, device=b8.device) if want_g_input else None
        L.call("p2p_decode_u8", L.ptr(a8), L.ptr(ra), None, a8.numel(), 3)
        L.call("p2p_decode_u8", L.ptr(b8), L.ptr(rb), L.ptr(gi), b8.numel(), 3)
        return ra, rb, gi

    def _step_any(self, real_a, real_b, g_input=None):
        """device tensors, fp32 or uint8 (see step())"""
        if real_a.dtype == torch.uint8:
            real_a, real_b, gi = self._decode_u8(real_a, real_b, g_input is None)
            g_input = gi if g_input is None else g_input
        return self._step_eager(real_a, real_b, g_input)

    def step_full(self, real_a: torch.Tensor, real_b: torch.Tensor, phase3: bool = True) -> Dict[str, torch.Tensor]:
        """The iteration as the reference executes it (train.py:294-402, VGG terms off): the generator input is
        compress(net_c(real_b), 3) (:297), then the G/D phases of step(), then phase 3 (:392-402).

        Phase 3 in the reference is a second train-mode forward of net_g (with the weights optimizer_g has just
        updated) on `compressed`, loss_c = MSE(., real_b), backward and optimizer_c.step().  Every gradient path
        from loss_c to net_c runs through torch.round inside compress(), whose gradient is zero, so net_c receives
        exact zeros and Adam leaves it unchanged (quirk Q3; pinned by tests/golden/ref_full_*.npz), and the
        gradients it leaves in net_g are discarded by the next optimizer_g.zero_grad().  What persists -- and what
        is done here, without the backward pass -- is the second update of every BatchNorm running buffer of G
        (num_batches_tracked advances by 2 per iteration) and the value of loss_c."""
        if self.net_c is None:
            raise ValueError("step_full needs the compression network: Pix2PixTrainer(..., net_c=define_C(...))")
        if not self.cuda_graph:
            return self._step_full_eager(real_a.cuda(non_blocking=True), real_b.cuda(non_blocking=True), phase3)
        key = ("full", tuple(real_a.shape), tuple(real_b.shape), bool(phase3), E.get_precision())
        rec = self._graphs.get(key)
        if rec is None:
            if self._eager_calls < self.graph_warmup:
                self._eager_calls += 1
                return self._step_full_eager(real_a.cuda(non_blocking=True), real_b.cuda(non_blocking=True), phase3)
            from . import _lib as L
            dev = self.optimizer_g.flat_p.device
            inputs = [torch.empty(t.shape, dtype=torch.float32, device=dev) for t in (real_a, real_b)]
            for dst, src in zip(inputs, (real_a, real_b)):
                dst.copy_(src, non_blocking=True)
            graph = torch.cuda.CUDAGraph()
            self._sync_hyper()
            torch.cuda.synchronize()
            before = L.launch_count()
            with torch.cuda.graph(graph):
                outputs = self._step_full_eager(inputs[0], inputs[1], phase3)
            self.graph_launches = L.launch_count() - before
            rec = self._graphs[key] = {"graph": graph, "inputs": inputs, "outputs": outputs}
        for dst, src in zip(rec["inputs"], (real_a, real_b)):
            dst.copy_(src, non_blocking=True)
        self._sync_hyper()
        rec["graph"].replay()
        return rec["outputs"]

    def _step_full_eager(self, real_a, real_b, phase3):
        from . import _lib as L
        compressed = self.net_c.front_end(real_b, 3)                                      # :297 (train-mode BatchNorm)
        out = dict(self._step_eager(real_a, real_b, compressed))                          # :300-390
        out["compressed"] = compressed
        if phase3:
            with torch.no_grad():
                fake_after = self.net_g(compressed)                                       # :392
                loss_c = torch.zeros((), dtype=torch.float32, device=compressed.device)
                L.call("p2p_mse_pair_fwd", L.ptr(fake_after), L.ptr(real_b.contiguous().float()), 1.0,
                       L.ptr(loss_c), fake_after.numel())                                 # :394
                if self.vgg is not None:
                    out["loss_c_percp"] = self.vgg(compressed, real_b) * 10.0             # :395 (its gradient is zero too)
            out["loss_c"] = loss_c
        return out

    def prefetch(self, real_a: torch.Tensor, real_b: torch.Tensor, g_input: Optional[torch.Tensor] = None) -> None:
        """Start copying the NEXT batch (pinned host tensors, fp32 or uint8) to the device on a side stream while the
        current iteration runs; the following step() called with the same tensor objects picks the staged copy up
        (what a DataLoader's pin_memory + non_blocking pipeline does for the reference loop, train.py:296-298)."""
        dev = self.optimizer_g.flat_p.device
        args = (real_a, real_b) if g_input is None else (real_a, real_b, g_input)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._staging = [None, None]
            self._slot_free = [torch.cuda.Event(), torch.cuda.Event()]
            for ev in self._slot_free:
                ev.record(torch.cuda.current_stream())
        slot = self._stage_slot = 1 - getattr(self, "_stage_slot", 1)
        sig = tuple((tuple(t.shape), t.dtype) for t in args)
        if self._staging[slot] is None or tuple((tuple(t.shape), t.dtype) for t in self._staging[slot]) != sig:
            self._staging[slot] = [torch.empty(sh, dtype=dt, device=dev) for sh, dt in sig]
        self._copy_stream.wait_event(self._slot_free[slot])            # the slot's previous reader is done with it
        with torch.cuda.stream(self._copy_stream):
            for dst, src in zip(self._staging[slot], args):
                dst.copy_(src, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self._prefetched = {"ids": tuple(id(t) for t in args), "staged": self._staging[slot], "event": ev,
                            "free": self._slot_free[slot]}

    def _capture(self, args):
        from . import _lib as L
        dev = self.optimizer_g.flat_p.device
        inputs = [torch.empty(t.shape, dtype=t.dtype, device=dev) for t in args]
        for dst, src in zip(inputs, args):
            dst.copy_(src, non_blocking=True)
        graph = torch.cuda.CUDAGraph()
        self._sync_hyper()
        torch.cuda.synchronize()
        before = L.launch_count()
        with torch.cuda.graph(graph):
            outputs = self._step_any(*inputs)
        self.graph_launches = L.launch_count() - before
        return {"graph": graph, "inputs": inputs, "outputs": outputs}

    def _step_two_lanes(self, real_a, real_b, g_input):
        """The same iteration on two CUDA streams.  The discriminator's training and the generator's are two chains of
        kernels that meet in three places only: D's passes over the fake pair need fake_b, the feature-matching loss
        needs pass 2's maps, and Adam may move D's weights only once pass 3 has read them.  Each chain alone leaves SMs
        idle (second waves of the persistent convolution kernels, small maps of the coarse D scales, memory-bound
        elementwise kernels); side by side they fill each other's gaps.

        Exactness: the three power iterations of the spectral-norm layers depend only on the weights and on their own
        previous u, v, so they run first, in the reference's order, and each pass gets its sigma; everything else is
        the eager order with independent work moved to the second stream (the first-layer sharing between passes 1
        and 3 is dropped: they now run at the same time)."""
        outer = torch.cuda.current_stream()
        if self._side is None:
            self._side = torch.cuda.Stream(device=real_a.device)
            # the generator's chain is the longer one: it runs on a HIGH-priority stream, so that whenever CTAs of both
            # chains are waiting for SMs the critical path goes first (P2P_G_PRIORITY=0: same priority)
            prio = -1 if os.environ.get("P2P_G_PRIORITY", "1") != "0" else 0
            self._g_lane = torch.cuda.Stream(device=real_a.device, priority=prio)
        side, main = self._side, self._g_lane
        main.wait_stream(outer)
        with torch.cuda.stream(main):
            out = self._two_lanes_body(real_a, real_b, g_input, main, side)
        outer.wait_stream(main)
        for t in out.values():
            t.record_stream(outer)
        return out

    def _two_lanes_body(self, real_a, real_b, g_input, main, side):
        net_g, net_d, cr = self.net_g, self.net_d, self._c_reals
        last = lambda maps: [s[-1] for s in maps]                                         # noqa: E731
        ones = [1] * 3
        sig1, sig2, sig3 = (net_d.power_iteration() for _ in range(3))                    # passes 1, 2, 3 (:309, :316, :337)
        vgg_taps = None
        if self.vgg is not None:
            # the target's VGG activations do not depend on the generator: low priority, beside G's forward pass
            if self._side_vgg is None:
                self._side_vgg = torch.cuda.Stream(device=real_a.device, priority=-1)
                self._side_vgg_lo = torch.cuda.Stream(device=real_a.device)
            self._side_vgg_lo.wait_stream(main)
            with torch.cuda.stream(self._side_vgg_lo):
                vgg_taps = self.vgg.target_features(real_b)
                ev_taps = torch.cuda.Event()
                ev_taps.record(self._side_vgg_lo)
        self.optimizer_g.zero_grad()                                                      # :384
        side.wait_stream(main)
        with torch.cuda.stream(side):                                                     # ---- D lane ----
            self.optimizer_d.zero_grad()                                                  # :388
            pred_real = net_d.forward_stored(real_a, real_b, sigma=sig2)                  # :315-316
            loss_d_real = lsgan_loss(last(pred_real), ones, 1.0)                          # :317
            ev_real = torch.cuda.Event()
            ev_real.record(side)
        fake_b = net_g(g_input)                                                           # :300 (beside pass 2)
        ev_fake = torch.cuda.Event()
        ev_fake.record(main)
        with torch.cuda.stream(side):
            side.wait_event(ev_fake)
            fake_b.record_stream(side)
            pred_fake_d = net_d.forward_stored(real_a, fake_b.detach(), sigma=sig1)       # :308-309
            loss_d_fake = lsgan_loss(last(pred_fake_d), ones, 0.0)                        # :310
            loss_d = (loss_d_fake + loss_d_real) * 0.5                                    # :320
        loss_vgg = None
        if self.vgg is not None:                           # ---- perceptual lane: beside pass 3 (its backward too) ----
            with torch.cuda.stream(self._side_vgg):
                self._side_vgg.wait_event(ev_fake)
                self._side_vgg.wait_event(ev_taps)
                fake_b.record_stream(self._side_vgg)
                for t in vgg_taps:
                    t.record_stream(self._side_vgg)
                loss_vgg = self.vgg(fake_b, target_taps=vgg_taps) * 10.0                  # :377
        # ---- G lane ----
        pred_fake = net_d.forward_stored(real_a, fake_b, frozen=True, sigma=sig3)         # :336-337
        ev_pass3 = torch.cuda.Event()
        ev_pass3.record(main)                              # D's weights have been read (packed) for the last time
        loss_g_gan = lsgan_loss(last(pred_fake), ones, 1.0)                               # :338
        main.wait_event(ev_real)
        for s in pred_real:
            for t in s:
                t.record_stream(main)
        loss_feat = feature_matching_loss(pred_fake, pred_real, cr)                       # :344-351
        loss_tv = tv_loss(fake_b)                                                         # :378
        loss_g = loss_g_gan + loss_feat + loss_tv                                         # :353,:380
        if loss_vgg is not None:
            main.wait_stream(self._side_vgg)
            loss_g = loss_g + loss_vgg
        with torch.cuda.stream(side):
            loss_d.backward()                                                             # :389
            self.red_d.launch(self.optimizer_d.flat_g)
            side.wait_event(ev_pass3)
            self.red_d.wait()
            self.optimizer_d.step()                                                       # :390
        loss_g.backward()                                                                 # :385
        self.red_g.launch(self.optimizer_g.flat_g)
        self.red_g.wait()
        self.optimizer_g.step()                                                           # :386
        main.wait_stream(side)
        out = {"fake_b": fake_b.detach(), "loss_d": loss_d.detach(), "loss_g_gan": loss_g_gan.detach(),
               "loss_feat": loss_feat.detach(), "loss_tv": loss_tv.detach(), "loss_g": loss_g.detach()}
        if loss_vgg is not None:
            out["loss_vgg"] = loss_vgg.detach()
        return out

    def _step_eager(self, real_a: torch.Tensor, real_b: torch.Tensor, g_input: torch.Tensor) -> Dict[str, torch.Tensor]:
        if self.two_lanes:
            return self._step_two_lanes(real_a, real_b, g_input)
        net_g, net_d, cr = self.net_g, self.net_d, self._c_reals
        last = lambda maps: [s[-1] for s in maps]                                         # noqa: E731
        ones = [1] * 3
        fake_b = net_g(g_input)                                                           # train.py:300
        share = {}                                         # passes 1 and 3 see the same pair: pyramid + layer 0 once
        pred_fake_d = net_d.forward_stored(real_a, fake_b.detach(), first_layer_cache=share)   # :308-309
        loss_d_fake = lsgan_loss(last(pred_fake_d), ones, 0.0)                            # :310
        pred_real = net_d.forward_stored(real_a, real_b)                                  # :315-316
        loss_d_real = lsgan_loss(last(pred_real), ones, 1.0)                              # :317
        loss_d = (loss_d_fake + loss_d_real) * 0.5                                        # :320
        pred_fake = net_d.forward_stored(real_a, fake_b, frozen=True, first_layer_cache=share)   # :336-337
        loss_g_gan = lsgan_loss(last(pred_fake), ones, 1.0)                               # :338
        loss_feat = feature_matching_loss(pred_fake, pred_real, cr)                       # :344-351
        loss_tv = tv_loss(fake_b)                                                         # :378
        loss_g = loss_g_gan + loss_feat + loss_tv                                         # :353,:380
        loss_vgg = None
        if self.vgg is not None:
            loss_vgg = self.vgg(fake_b, real_b) * 10.0                                    # :377
            loss_g = loss_g + loss_vgg
        self.optimizer_g.zero_grad()                                                      # :384
        self.optimizer_d.zero_grad()                                                      # :388 (hoisted: D grads untouched above)
        # The two backward passes are independent (loss_d sees fake_b detached, the third D pass is frozen), so the
        # reference's order (:385-390: G backward, G step, D backward, D step) can be permuted without changing a
        # bit: D's gradients (33 MB, the larger all-reduce) go first and travel while G's backward runs.
        loss_d.backward()                                                                 # :389
        if self.overlap:
            self.red_d.launch(self.optimizer_d.flat_g)    # overlaps with G's backward below
        loss_g.backward()                                                                 # :385
        if not self.overlap:
            self.red_d.launch(self.optimizer_d.flat_g)
        self.red_g.launch(self.optimizer_g.flat_g)
        self.red_d.wait()
        self.optimizer_d.step()                                                           # :390
        self.red_g.wait()
        self.optimizer_g.step()                                                           # :386
        out = {"fake_b": fake_b.detach(), "loss_d": loss_d.detach(), "loss_g_gan": loss_g_gan.detach(),
               "loss_feat": loss_feat.detach(), "loss_tv": loss_tv.detach(), "loss_g": loss_g.detach()}
        if loss_vgg is not None:
            out["loss_vgg"] = loss_vgg.detach()
        return out
