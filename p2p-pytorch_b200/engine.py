"""Kernel schedules of the generator and discriminator: what `ExpandNetwork.forward` and
`MultiscaleDiscriminator.forward` of the reference (networks.py:492-523, :734-755) and their
autograd backward become when every tensor op is one of the C-ABI kernels.

Each network is ONE torch.autograd.Function: forward enqueues the layer kernels and keeps the
intermediate NHWC buffers on the ctx, backward enqueues dgrad / wgrad / BatchNorm-backward kernels in
reverse order.  PyTorch is used for device memory, streams and the autograd edge only.

Parameter gradients: when a parameter already has a `.grad` tensor (FusedAdam keeps them as views of
one flat buffer) the kernels accumulate straight into it and autograd is handed None; otherwise a
fresh gradient tensor is returned the usual way.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import torch

from . import _lib as L
from ._lib import Conv, call, ptr

_state = {"dtype": torch.bfloat16, "impl": L.IMPL_AUTO}


def set_precision(mode: str) -> None:
    """'bf16' (default): bf16 activations/weights, fp32 accumulation, tcgen05 convolutions.
    'bf16-simt': same data types on the CUDA-core convolutions (debugging aid).
    'fp32': the verification mode -- fp32 activations and CUDA-core fp32 convolutions."""
    if mode == "bf16":
        _state.update(dtype=torch.bfloat16, impl=L.IMPL_AUTO)
    elif mode == "bf16-simt":
        _state.update(dtype=torch.bfloat16, impl=L.IMPL_SIMT)
    elif mode == "fp32":
        _state.update(dtype=torch.float32, impl=L.IMPL_SIMT)
    else:
        raise ValueError("precision must be 'bf16', 'bf16-simt' or 'fp32'")


def get_precision() -> str:
    if _state["dtype"] == torch.float32:
        return "fp32"
    return "bf16" if _state["impl"] == L.IMPL_AUTO else "bf16-simt"


_helper_streams = {}


def helper_stream(dev, which: str = "wgrad", same_priority: bool = False):
    """A second CUDA stream tied to the CURRENT one (one per (device, current stream, purpose)): work that nothing
    downstream waits for -- the weight gradients of a backward pass -- is forked onto it and joined at the end, so it
    runs beside the data-gradient chain (the persistent wgrad kernels beside the memory-bound BatchNorm-backward kernels:
    8.20 -> 7.79 ms/step).  Used whenever the backward pass runs on a stream of its own (the trainer's lanes, a graph
    capture); on the process's default stream it stays off -- with the lanes on the default stream one prefetch test
    differed from its un-prefetched twin at the second iteration, unexplained, and that configuration is avoided rather
    than trusted.  P2P_WGRAD_STREAM=0 / 1 forces it off / on."""
    import os
    mode = os.environ.get("P2P_WGRAD_STREAM", "auto")
    if which != "wgrad":
        mode = os.environ.get("P2P_SCALE_STREAMS", "auto")
    cur = torch.cuda.current_stream()
    if mode == "0" or (mode != "1" and cur == torch.cuda.default_stream(cur.device)):
        return None
    key = (str(dev), cur.cuda_stream, which)
    st = _helper_streams.get(key)
    if st is None:
        # weight gradients: lowest priority (nothing waits for them: 7.99 -> 7.79 ms/step against the lane's priority);
        # the coarse discriminator scales are part of the lane's chain and keep its priority
        st = _helper_streams[key] = torch.cuda.Stream(device=dev, priority=cur.priority if same_priority else 0)
    return st


def pad16(c: int) -> int:
    return (c + 15) // 16 * 16


def _empty(shape, dtype, dev):
    return torch.empty(shape, dtype=dtype, device=dev)


def _zeros(shape, dtype, dev):
    t = torch.empty(shape, dtype=dtype, device=dev)
    nbytes = t.numel() * t.element_size()
    if nbytes >= (1 << 20) and t.data_ptr() % 16 == 0:
        call("p2p_zero_fill", ptr(t), nbytes)          # large slabs: 16-byte stores (torch's fill runs at 0.5 TB/s here)
        return t
    return t.zero_()


def _grad_slot(p: torch.Tensor):
    """(tensor the kernels accumulate into, value to hand back to autograd)"""
    g = p.grad
    if g is not None and g.dtype == torch.float32 and g.is_contiguous() and g.shape == p.shape:
        return g, None
    z = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.contiguous_format)
    return z, z


class ConvPlan:
    """Geometry + packed weights of one convolution layer for one forward pass."""

    def __init__(self, N, Hi, Wi, Ci, cin_real, cout_real, k, stride, pad, act=L.ACT_NONE):
        self.k, self.stride, self.pad = k, stride, pad
        self.Ho = (Hi + 2 * pad - k) // stride + 1
        self.Wo = (Wi + 2 * pad - k) // stride + 1
        self.Co = pad16(cout_real)
        self.desc = Conv(N, Hi, Wi, Ci, self.Ho, self.Wo, self.Co, k, k, stride, pad, cout_real, cin_real, act)
        self.cin_real, self.cout_real, self.Ci, self.act = cin_real, cout_real, Ci, act
        self.wp = self.wt = self.w1x1 = None

    def one_channel(self, dt) -> bool:
        """D's last layer (512 -> 1): forward and data-gradient run as 1x1 convolutions over the filter taps."""
        return (self.cout_real == 1 and self.stride == 1 and self.k * self.k <= 16
                and _state["impl"] == L.IMPL_AUTO and dt == torch.bfloat16)

    def pack(self, weight, sigma, dt, want_wt=True):
        dev = weight.device
        self.wp = _empty((self.Co, self.k, self.k, self.Ci), dt, dev)
        self.wt = _empty((self.Ci, self.k, self.k, self.Co), dt, dev) if want_wt else None
        call("p2p_pack_weights", ptr(weight), ptr(sigma), ptr(self.wp), ptr(self.wt), self.cout_real, self.cin_real,
             self.Co, self.Ci, self.k, self.k, L.dtype_code(dt))
        for job in self.extra_pack_jobs(weight, sigma, dt):
            L.call_multi("p2p_pack_weights_multi", [job], L.dtype_code(dt))

    def pack_job(self, weight, sigma, dt, want_wt=True) -> L.PackJob:
        """Allocate the packed buffers and describe the packing for one batched p2p_pack_weights_multi launch."""
        dev = weight.device
        self.wp = _empty((self.Co, self.k, self.k, self.Ci), dt, dev)
        self.wt = _empty((self.Ci, self.k, self.k, self.Co), dt, dev) if want_wt else None
        return L.PackJob(ptr(weight), ptr(sigma), ptr(self.wp), ptr(self.wt), self.cout_real, self.cin_real, self.Co,
                         self.Ci, self.k, self.k)

    def extra_pack_jobs(self, weight, sigma, dt):
        """One-channel layers also keep the filter as a [Ci][taps -> 16] matrix: the PyTorch weight [1][ci][kh][kw]
        read as a 1x1 filter bank with ci outputs and kh*kw inputs (used by dgrad)."""
        if not self.one_channel(dt):
            return []
        self.w1x1 = _empty((self.Ci, 1, 1, 16), dt, weight.device)
        return [L.PackJob(ptr(weight), ptr(sigma), ptr(self.w1x1), None, self.cin_real, self.k * self.k, self.Ci, 16, 1, 1)]

    def fwd(self, x, bias, out, stats):
        d = self.desc
        if stats is None and self.one_channel(x.dtype):
            # one real output channel (D's last layer): a 1x1 convolution with one output column per filter tap -- the
            # [taps x Ci] filter row IS that 1x1 weight matrix -- reads the input once; then a 16-tap gather-sum
            t = _empty((d.N, d.Hi, d.Wi, 16), x.dtype, x.device)
            d1 = Conv(d.N, d.Hi, d.Wi, d.Ci, d.Hi, d.Wi, 16, 1, 1, 1, 0, 16, self.cin_real, L.ACT_NONE)
            call("p2p_conv_fwd", C.byref(d1), ptr(x), ptr(self.wp), None, ptr(t), None, L.dtype_code(x.dtype),
                 _state["impl"])
            call("p2p_tap_gather_sum", ptr(t), ptr(bias), ptr(out), d.N, d.Hi, d.Wi, d.Ho, d.Wo, self.k, self.k,
                 self.pad, self.act, L.dtype_code(x.dtype))
            return
        call("p2p_conv_fwd", C.byref(self.desc), ptr(x), ptr(self.wp), ptr(bias), ptr(out), ptr(stats),
             L.dtype_code(x.dtype), _state["impl"])

    def dgrad(self, dout, din):
        d = self.desc
        if self.w1x1 is not None and self.one_channel(dout.dtype):
            a = _empty((d.N, d.Hi, d.Wi, 16), dout.dtype, dout.device)
            call("p2p_tap_scatter", ptr(dout), ptr(a), d.N, d.Hi, d.Wi, d.Ho, d.Wo, self.k, self.k, self.pad,
                 L.dtype_code(dout.dtype))
            d1 = Conv(d.N, d.Hi, d.Wi, 16, d.Hi, d.Wi, d.Ci, 1, 1, 1, 0, self.cin_real, self.k * self.k, L.ACT_NONE)
            call("p2p_conv_fwd", C.byref(d1), ptr(a), ptr(self.w1x1), None, ptr(din), None, L.dtype_code(dout.dtype),
                 _state["impl"])
            return
        call("p2p_conv_dgrad", C.byref(self.desc), ptr(dout), ptr(self.wt), ptr(din), L.dtype_code(dout.dtype),
             _state["impl"])

    def acc_numel(self):
        return self.k * self.k * self.Ci * self.Co

    def wgrad(self, x, dout, grad, sigma=None, dot=None, w_for_dot=None, acc=None):
        """grad ([co][ci][kh][kw] fp32) += d loss / d weight  (scaled by 1/sigma if given).
        `acc`: zeroed fp32 scratch of acc_numel() elements (callers zero one slab for all their layers)."""
        if acc is None:
            acc = _zeros((self.acc_numel(),), torch.float32, x.device)
        call("p2p_conv_wgrad", C.byref(self.desc), ptr(x), ptr(dout), ptr(acc), L.dtype_code(x.dtype), _state["impl"])
        call("p2p_unpack_wgrad", ptr(acc), ptr(grad), ptr(w_for_dot), ptr(sigma), ptr(dot), self.cout_real,
             self.cin_real, self.Co, self.Ci, self.k, self.k)

    def wgrad_acc(self, x, dout, acc, grad, sigma=None, dot=None, w_for_dot=None) -> L.UnpackJob:
        """Accumulate d loss / d weight into the zeroed fp32 scratch `acc` and return the job that moves it into
        `grad` ([co][ci][kh][kw]); the caller runs all jobs of its pass in one p2p_unpack_wgrad_multi launch."""
        call("p2p_conv_wgrad", C.byref(self.desc), ptr(x), ptr(dout), ptr(acc), L.dtype_code(x.dtype), _state["impl"])
        return L.UnpackJob(ptr(acc), ptr(grad), ptr(w_for_dot), ptr(sigma), ptr(dot), self.cout_real, self.cin_real,
                           self.Co, self.Ci, self.k, self.k)


class FatColumnConv:
    """A stride-1 convolution with a wide filter and very few output channels (ExpandNetwork's last layer: 9x9,
    32 -> 3) as a (k x 1) convolution with k*cout "fat" output channels plus a horizontal gather: one MMA column
    per (kx, co) pair instead of 16 columns per filter tap.  Forward and data-gradient only (the weight gradient
    keeps the layer's ordinary plan)."""

    def __init__(self, N, Hi, Wi, Ci, cin_real, cout_real, k):
        self.k, self.cin_real, self.cout_real, self.Ci = k, cin_real, cout_real, Ci
        self.N, self.Hi, self.Wi = N, Hi, Wi
        self.Ho, self.Wo = Hi - k + 1, Wi - k + 1
        self.nfat = k * cout_real
        self.Cf = pad16(self.nfat)
        self.desc = Conv(N, Hi, Wi, Ci, self.Ho, Wi, self.Cf, k, 1, 1, 0, self.nfat, cin_real, L.ACT_NONE)
        self.wp = self.wt = self.w2 = None

    @staticmethod
    def applies(cout_real, k, dt) -> bool:
        return k >= 5 and cout_real <= 4 and dt == torch.bfloat16 and _state["impl"] == L.IMPL_AUTO

    def pack_job(self, weight, dt) -> L.PackJob:
        # W'[(kx*cout + co)][ci][ky][0] = w[co][ci][ky][kx]
        self.w2 = weight.detach().permute(3, 0, 1, 2).reshape(self.nfat, self.cin_real, self.k, 1).contiguous()
        dev = weight.device
        self.wp = _empty((self.Cf, self.k, 1, self.Ci), dt, dev)
        self.wt = _empty((self.Ci, self.k, 1, self.Cf), dt, dev)
        return L.PackJob(ptr(self.w2), None, ptr(self.wp), ptr(self.wt), self.nfat, self.cin_real, self.Cf, self.Ci,
                         self.k, 1)

    def fwd(self, x, bias, out):
        dc = L.dtype_code(x.dtype)
        t = _empty((self.N, self.Ho, self.Wi, self.Cf), x.dtype, x.device)
        call("p2p_conv_fwd", C.byref(self.desc), ptr(x), ptr(self.wp), None, ptr(t), None, dc, _state["impl"])
        call("p2p_kx_gather_sum", ptr(t), ptr(bias), ptr(out), self.N, self.Ho, self.Wi, self.Wo, self.k, self.cout_real,
             self.Cf, out.shape[3], dc)

    def dgrad(self, dout, din):
        dc = L.dtype_code(dout.dtype)
        a = _empty((self.N, self.Ho, self.Wi, self.Cf), dout.dtype, dout.device)
        call("p2p_kx_scatter", ptr(dout), ptr(a), self.N, self.Ho, self.Wo, self.Wi, self.k, self.cout_real,
             dout.shape[3], self.Cf, dc)
        call("p2p_conv_dgrad", C.byref(self.desc), ptr(a), ptr(self.wt), ptr(din), dc, _state["impl"])


# =====================================================================================================
# Generator
# =====================================================================================================
# (conv attr, bn attr, cout, k, stride, activation, role);  role: 'plain' | 'res_a' | 'res_b'
def _g_layers():
    t = [("conv1", "in1_e", 32, 9, 1, L.ACT_PRELU, "plain"),
         ("conv2", "in2_e", 64, 3, 2, L.ACT_PRELU, "plain"),
         ("conv3", "in3_e", 128, 3, 2, L.ACT_PRELU, "plain")]
    for r in range(1, 10):
        t.append((f"res{r}.conv1", f"res{r}.in1", 128, 3, 1, L.ACT_RELU, "res_a"))
        t.append((f"res{r}.conv2", f"res{r}.in2", 128, 3, 1, L.ACT_RELU, "res_b"))
    t += [("deconv3", "in3_d", 64, 3, 1, L.ACT_PRELU, "plain"),
          ("deconv2", "in2_d", 32, 3, 1, L.ACT_PRELU, "plain"),
          ("deconv1", "in1_d", 3, 9, 1, L.ACT_TANH, "plain")]
    return t


G_LAYERS = _g_layers()
G_N = len(G_LAYERS)  # 24 conv + BatchNorm stages


class _Stage:
    """One conv + BatchNorm + activation stage (or the long-skip add) and the buffers backward needs."""
    __slots__ = ("li", "plan", "x_in", "r", "sb", "skip", "skip_pad", "skip_src", "r_src", "act", "up", "pad",
                 "H", "W", "C", "c_real", "role", "fat")

    def __init__(self):
        self.skip = self.skip_src = self.r_src = self.plan = self.x_in = self.sb = self.fat = None
        self.skip_pad = 0
        self.li = -1


class GeneratorFunction(torch.autograd.Function):
    """y = ExpandNetwork(x) (networks.py:492-523).
    apply(x, prelu, aux, *params): `aux` = dict(buffers=[(running_mean, running_var, num_batches_tracked)] per
    layer, training=bool, param_refs=[Parameter objects in the order of `params`, prelu first]);
    params = per layer (conv.weight, conv.bias, bn.weight, bn.bias)."""

    @staticmethod
    def forward(ctx, x, prelu, aux, *params):
        L.require_cuda(x, "generator input")
        if x.dim() != 4 or x.shape[1] != 3:
            raise L.P2PError("generator input must be [B,3,H,W]")
        buffers, training = aux["buffers"], aux["training"]
        B, _, H, W = x.shape
        dt, dev = _state["dtype"], x.device
        dc = L.dtype_code(dt)
        x = x.detach().contiguous().float()
        stages: List[_Stage] = []
        cur = _empty((B, H + 8, W + 8, 16), dt, dev)
        call("p2p_g_input_prep", ptr(x), ptr(cur), B, H, W, dc)
        producer = None            # stage that wrote `cur`
        cin_real = 12
        block_in = block_in_src = long_skip = long_skip_src = None
        out = None
        # all 24 filter banks are packed (both layouts) by one launch, and all BatchNorm sums live in one zeroed slab
        packs, jobs, ci = [], [], 12
        for (_n, _bn, cout, k, _s, _a, _r) in G_LAYERS:
            pk = ConvPlan(1, k, k, pad16(ci), ci, cout, k, 1, 0)
            jobs.append(pk.pack_job(params[4 * len(packs)], None, dt))
            packs.append(pk)
            ci = cout
        fat = None
        _ln, _lbn, l_cout, l_k, _ls, _la, _lr = G_LAYERS[-1]
        if FatColumnConv.applies(l_cout, l_k, dt):
            # the last layer reads the H x W map padded by l_k // 2 (written by the stage before it)
            fat = FatColumnConv(B, H + 2 * (l_k // 2), W + 2 * (l_k // 2), packs[-1].Ci, packs[-1].cin_real, l_cout, l_k)
            jobs.append(fat.pack_job(params[4 * (G_N - 1)], dt))
        L.call_multi("p2p_pack_weights_multi", jobs, dc)
        stats_slab = _zeros((sum(2 * pk.Co for pk in packs),), torch.float64, dev) if training else None
        stats_off = 0
        for li, (name, _bn, cout, k, stride, act, role) in enumerate(G_LAYERS):
            _w, b, gamma, beta = params[4 * li: 4 * li + 4]
            rm, rv, nbt = buffers[li]
            plan = ConvPlan(B, cur.shape[1], cur.shape[2], cur.shape[3], cin_real, cout, k, stride, 0)
            plan.wp, plan.wt = packs[li].wp, packs[li].wt
            Ho, Wo, Co = plan.Ho, plan.Wo, plan.Co
            r = _empty((B, Ho, Wo, Co), dt, dev)
            stats = stats_slab[stats_off: stats_off + 2 * Co] if training else None
            stats_off += 2 * Co
            if fat is not None and li == G_N - 1:
                fat.fwd(cur, b, r)
                if stats is not None:
                    call("p2p_channel_stats", ptr(r), ptr(stats), B * Ho * Wo, Co, dc)
            else:
                plan.fwd(cur, b, r, stats)
            sb = _empty((4 * Co,), torch.float32, dev)
            # the BatchNorm bookkeeping rides in the launch that applies it (p2p_bn_finalize_act_store) wherever the
            # stage's output is a stored map; only the last stage (NCHW output) keeps the separate kernel
            fused_store = li != G_N - 1 and os.environ.get("P2P_FUSED_BN_STORE", "1") != "0"
            if not fused_store:
                call("p2p_bn_finalize", ptr(stats), B * Ho * Wo, ptr(gamma), ptr(beta), ptr(rm), ptr(rv), ptr(nbt), ptr(sb),
                     cout, Co, 1 if training else 0)

            def bn_store(skip, skip_pad, dst, up, pad):
                if fused_store:
                    call("p2p_bn_finalize_act_store", ptr(stats), B * Ho * Wo, ptr(gamma), ptr(beta), ptr(rm), ptr(rv),
                         ptr(nbt), ptr(sb), 1 if training else 0, ptr(r), ptr(skip), skip_pad, ptr(prelu), ptr(dst), B, Ho,
                         Wo, Co, cout, act, up, pad, dc)
                else:
                    call("p2p_bn_act_store", ptr(r), ptr(sb), ptr(skip), skip_pad, ptr(prelu), ptr(dst), None, B, Ho, Wo,
                         Co, cout, act, up, pad, dc)
            st = _Stage()
            st.li, st.plan, st.x_in, st.r, st.sb, st.act, st.role = li, plan, cur, r, sb, act, role
            st.fat = fat if li == G_N - 1 else None
            st.H, st.W, st.C, st.c_real = Ho, Wo, Co, cout
            if role == "res_a":
                block_in, block_in_src = cur, producer
            if role == "res_b":
                st.skip, st.skip_pad, st.skip_src = block_in, 1, block_in_src
            if li == G_N - 1:
                st.up, st.pad = 1, 0
                out = _empty((B, 3, H, W), torch.float32, dev)
                call("p2p_bn_act_store", ptr(r), ptr(sb), None, 0, ptr(prelu), None, ptr(out), B, Ho, Wo, Co, cout, act,
                     1, 0, dc)
                stages.append(st)
                break
            if name == "res9.conv2":
                # res9's output stays compact: its only consumer is the long-skip add, not a convolution
                st.up, st.pad = 1, 0
                y9 = _empty((B, Ho, Wo, Co), dt, dev)
                bn_store(st.skip, st.skip_pad, y9, 1, 0)
                stages.append(st)
                # y = LeakyReLU(res + residual) (networks.py:514-515), then nearest x2 + reflect 1 for deconv3
                ls = _Stage()
                ls.r, ls.act, ls.role, ls.r_src = y9, L.ACT_LRELU, "long_skip", st
                ls.H, ls.W, ls.C, ls.c_real = Ho, Wo, Co, cout
                ls.skip, ls.skip_pad, ls.skip_src, ls.up, ls.pad = long_skip, 1, long_skip_src, 2, 1
                cur = _empty((B, Ho * 2 + 2, Wo * 2 + 2, Co), dt, dev)
                call("p2p_bn_act_store", ptr(y9), None, ptr(long_skip), 1, None, ptr(cur), None, B, Ho, Wo, Co, cout,
                     L.ACT_LRELU, 2, 1, dc)
                stages.append(ls)
                producer = ls
            else:
                nxt_name, nxt_k = G_LAYERS[li + 1][0], G_LAYERS[li + 1][3]
                st.up, st.pad = (2 if nxt_name in ("deconv3", "deconv2") else 1), nxt_k // 2
                dst = _empty((B, Ho * st.up + 2 * st.pad, Wo * st.up + 2 * st.pad, Co), dt, dev)
                bn_store(st.skip, st.skip_pad, dst, st.up, st.pad)
                stages.append(st)
                cur, producer = dst, st
                if name == "conv3":
                    long_skip, long_skip_src = cur, st
            cin_real = cout
        ctx.set_materialize_grads(False)
        ctx.stages, ctx.prelu, ctx.param_refs = stages, prelu, aux["param_refs"]
        ctx.geom, ctx.training = (B, H, W), training
        return out

    @staticmethod
    def backward(ctx, g_out):
        if ctx.needs_input_grad[0]:
            raise L.P2PError("gradient w.r.t. the generator input is not implemented (the reference never uses it: "
                             "its only consumer is the quantiser, whose gradient is zero)")
        if not ctx.training:
            raise L.P2PError("backward through the generator needs training-mode BatchNorm statistics")
        if g_out is None:
            return (None,) * (3 + 4 * G_N)
        stages, prelu, refs = ctx.stages, ctx.prelu, ctx.param_refs
        B, _H, _W = ctx.geom
        dev, dt = g_out.device, stages[0].r.dtype
        dc = L.dtype_code(dt)
        g_out = g_out.contiguous().float()
        dprelu = _zeros((1,), torch.float64, dev)
        # one zeroed slab for all split-K weight-gradient accumulators and BatchNorm reductions of this pass
        acc_sizes = [st.plan.acc_numel() for st in stages if st.plan is not None]
        acc_slab = _zeros((sum(acc_sizes),), torch.float32, dev)
        acc_off = {}
        off = 0
        for st in stages:
            if st.plan is not None:
                acc_off[id(st)] = off
                off += st.plan.acc_numel()
        red_slab = _zeros((sum(2 * st.C for st in stages),), torch.float64, dev)
        red_off, off = {}, 0
        for st in stages:
            red_off[id(st)] = off
            off += 2 * st.C
        ret: List[Optional[torch.Tensor]] = [None] * (4 * G_N)
        unpack_jobs = []
        cur_stream = torch.cuda.current_stream()
        ws = helper_stream(dev)
        dpad = None            # gradient w.r.t. the (padded) buffer the stage being processed wrote
        extra = {}             # id(stage) -> compact gradients addressed to that stage's output (skip paths)
        for st in reversed(stages):
            ex = extra.pop(id(st), [])
            g1 = ex[0] if len(ex) > 0 else None
            g2 = ex[1] if len(ex) > 1 else None
            dz = _empty((B, st.H, st.W, st.C), dt, dev)
            if st.role == "long_skip":
                call("p2p_bn_act_bwd_reduce", ptr(dpad), None, None, None, ptr(st.r), None, ptr(st.skip), st.skip_pad,
                     None, ptr(dz), None, None, B, st.H, st.W, st.C, st.c_real, st.act, st.up, st.pad, dc)
                extra.setdefault(id(st.r_src), []).append(dz)      # res9 output
                extra.setdefault(id(st.skip_src), []).append(dz)   # conv3 output (the long residual)
                dpad = None
                continue
            li = st.li
            pw, pb, pg, pbeta = refs[1 + 4 * li: 5 + 4 * li]
            red = red_slab[red_off[id(st)]: red_off[id(st)] + 2 * st.C]
            call("p2p_bn_act_bwd_reduce", ptr(dpad), ptr(g1), ptr(g2), ptr(g_out) if li == G_N - 1 else None, ptr(st.r),
                 ptr(st.sb), ptr(st.skip), st.skip_pad, ptr(prelu), ptr(dz), ptr(red), ptr(dprelu), B, st.H, st.W, st.C,
                 st.c_real, st.act, st.up, st.pad, dc)
            if st.role == "res_b":
                extra.setdefault(id(st.skip_src), []).append(dz)   # z = bn + skip: the block input gets dz as is
            dr = _empty((B, st.H, st.W, st.C), dt, dev)
            gw, ret[4 * li] = _grad_slot(pw)
            gb, ret[4 * li + 1] = _grad_slot(pb)
            gg, ret[4 * li + 2] = _grad_slot(pg)
            gbe, ret[4 * li + 3] = _grad_slot(pbeta)
            call("p2p_bn_bwd_apply", ptr(dz), ptr(st.r), ptr(st.sb), ptr(red), ptr(dr), ptr(gg), ptr(gbe), ptr(gb),
                 B * st.H * st.W, st.C, st.c_real, dc)
            acc = acc_slab[acc_off[id(st)]: acc_off[id(st)] + st.plan.acc_numel()]
            if ws is None:
                unpack_jobs.append(st.plan.wgrad_acc(st.x_in, dr, acc, gw))
            else:
                # nothing downstream reads the weight gradient: it runs on the helper stream beside the dgrad chain
                ev = torch.cuda.Event()
                ev.record(cur_stream)
                with torch.cuda.stream(ws):
                    ws.wait_event(ev)
                    dr.record_stream(ws)
                    unpack_jobs.append(st.plan.wgrad_acc(st.x_in, dr, acc, gw))
            if li > 0:
                dpad = _empty(tuple(st.x_in.shape), dt, dev)
                (st.fat or st.plan).dgrad(dr, dpad)
            else:
                dpad = None
        if ws is not None:
            cur_stream.wait_stream(ws)
        L.call_multi("p2p_unpack_wgrad_multi", unpack_jobs)
        gp, ret_p = _grad_slot(refs[0])
        call("p2p_prelu_grad_finish", ptr(dprelu), ptr(gp))
        ctx.stages = None
        return (None, ret_p, None, *ret)


# =====================================================================================================
# Discriminator
# =====================================================================================================
def d_layers(input_nc=6, ndf=64):
    """(cin, cout, stride, spectral_norm) of one scale, networks.py:764-784"""
    return [(input_nc, ndf, 2, False), (ndf, ndf * 2, 2, True), (ndf * 2, ndf * 4, 2, True),
            (ndf * 4, min(ndf * 8, 512), 1, True), (min(ndf * 8, 512), 1, 1, False)]


NUM_D = 3


def spectral_power_iteration(wbars: Sequence[torch.Tensor], us, vs) -> torch.Tensor:
    """one batched power iteration (networks.py:537-547); returns sigma[n]; u, v are updated in place"""
    n = len(wbars)
    dev = wbars[0].device
    sigma = _empty((n,), torch.float32, dev)
    rows = [w.shape[0] for w in wbars]
    cols = [w[0].numel() for w in wbars]
    tmp = _empty((sum(rows) + sum(cols),), torch.float32, dev)
    PA, IA = C.c_void_p * n, C.c_int32 * n
    call("p2p_spectral_norm", n, PA(*[w.data_ptr() for w in wbars]), PA(*[u.data_ptr() for u in us]),
         PA(*[v.data_ptr() for v in vs]), PA(*[sigma.data_ptr() + 4 * i for i in range(n)]), IA(*rows), IA(*cols),
         ptr(tmp))
    return sigma


class DiscriminatorFunction(torch.autograd.Function):
    """All three scales of MultiscaleDiscriminator (networks.py:734-755) in one Function.

    apply(a, b, aux, *params):
      a [B,ca,H,W], b [B,cb,H,W] or None -- the pair is concatenated on the fly (train.py:308,315,336);
      aux = dict(layer_cfg=d_layers(..), sn_bufs=[(u, v)] for the spectral-norm layers in module order,
                 param_refs=[Parameter objects in the order of `params`]);
      params = per scale/layer (weight or weight_bar, bias) in module order scale0_layer0 .. scale2_layer4.
    returns 15 stored maps [B,Ho,Wo,pad16(C)] in the working dtype, in the reference's result order:
    full-resolution scale first (it runs the scale2_* modules, networks.py:749)."""

    @staticmethod
    def forward(ctx, a, b, aux, *params):
        L.require_cuda(a, "discriminator input")
        layer_cfg, sn_bufs = aux["layer_cfg"], aux["sn_bufs"]
        dt, dev = _state["dtype"], a.device
        dc = L.dtype_code(dt)
        # identity of the pair for the first-layer cache: same storage, not written to since
        pair_id = (a.data_ptr(), a._version, None if b is None else b.data_ptr(), None if b is None else b._version)
        a = a.detach().contiguous().float()
        ca, cb = a.shape[1], 0
        if b is not None:
            b = b.detach().contiguous().float()
            cb = b.shape[1]
        if ca + cb > 8:
            raise L.P2PError("the discriminator input may have at most 8 channels")
        B, _, H, W = a.shape
        nl = len(layer_cfg)
        sn_idx = [(sc, j) for sc in range(NUM_D) for j in range(nl) if layer_cfg[j][3]]
        sig_of = {}
        if sn_idx:
            sigma_all = aux.get("sigma")
            if sigma_all is None:
                wbars = [params[2 * (sc * nl + j)].detach() for sc, j in sn_idx]
                sigma_all = spectral_power_iteration(wbars, [u for u, _ in sn_bufs], [v for _, v in sn_bufs])
            # (else: the caller ran this pass's power iteration ahead of time, MultiscaleDiscriminator.power_iteration)
            sig_of = {key: sigma_all[i:i + 1] for i, key in enumerate(sn_idx)}
        # A second pass over the SAME pair within one iteration (train.py:308 and :336 both see (real_a, fake_b)) may
        # reuse the input pyramid and the first layer's outputs of the first one: that layer has no spectral norm, so
        # its weights -- unlike the later layers' W / sigma -- are identical in both passes.
        cache = aux.get("first_layer_cache")
        nl0 = len(layer_cfg)
        w0_id = tuple((params[2 * sc * nl0].data_ptr(), params[2 * sc * nl0]._version, params[2 * sc * nl0 + 1]._version)
                      for sc in range(NUM_D))
        cache_key = (tuple(a.shape), cb, dt, _state["impl"], pair_id, w0_id)
        reuse = bool(cache) and cache.get("key") == cache_key and not layer_cfg[0][3]
        pyr_shapes = cache["pyr_shapes"] if reuse else None
        xs32 = None
        if not reuse:
            # input pyramid in fp32 (8 stored channels), then cast to the working dtype (16 stored channels)
            xs32 = [_empty((B, H, W, 8), torch.float32, dev)]
            call("p2p_d_input_prep", ptr(a), ca, ptr(b), cb, ptr(xs32[0]), B, H, W)
            for _i in range(NUM_D - 1):
                h, w = xs32[-1].shape[1:3]
                nx = _empty((B, (h - 1) // 2 + 1, (w - 1) // 2 + 1, 8), torch.float32, dev)
                call("p2p_avgpool3s2_fwd", ptr(xs32[-1]), ptr(nx), B, h, w, 8)
                xs32.append(nx)
            pyr_shapes = [tuple(t.shape) for t in xs32]
        outs, plans, inputs = [], [], []
        # every filter bank of the three scales (divided by its sigma) is packed by one launch
        packs, jobs = {}, []
        for sc in range(NUM_D):
            ci = ca + cb
            for j, (_cin, cout, _stride, _sn) in enumerate(layer_cfg):
                pk = ConvPlan(1, 4, 4, pad16(ci), ci, cout, 4, 1, 2)
                jobs.append(pk.pack_job(params[2 * (sc * nl + j)], sig_of.get((sc, j)), dt))
                jobs += pk.extra_pack_jobs(params[2 * (sc * nl + j)], sig_of.get((sc, j)), dt)
                packs[(sc, j)] = pk
                ci = cout
        L.call_multi("p2p_pack_weights_multi", jobs, dc)
        # The three scales are independent chains of five convolutions; the two coarse ones (64x64 and 32x32 inputs
        # at 256x256) are short, latency-bound kernels.  Where this pass runs on a stream of its own they go to two
        # helper streams of the same priority and run beside the full-resolution scale.
        lane = torch.cuda.current_stream()
        scale_streams = [None] + [helper_stream(dev, f"scale{i}", same_priority=True) for i in range(1, NUM_D)]
        ev_ready = None
        if any(st is not None for st in scale_streams):
            ev_ready = torch.cuda.Event()
            ev_ready.record(lane)
        for i in range(NUM_D):
            sc = NUM_D - 1 - i
            h, w = pyr_shapes[i][1:3]
            st = scale_streams[i]
            if st is not None:
                st.wait_event(ev_ready)
            with torch.cuda.stream(st if st is not None else lane):
                if reuse:
                    cur = cache["inputs"][i]
                else:
                    cur = _empty((B, h, w, 16), dt, dev)
                    call("p2p_cast_pad_channels", ptr(xs32[i]), ptr(cur), B * h * w, 8, 16, dc)
                    if st is not None:
                        cur.record_stream(lane)
                cin_real = ca + cb
                for j, (_cin, cout, stride, _sn) in enumerate(layer_cfg):
                    bias = params[2 * (sc * nl + j) + 1]
                    act = L.ACT_LRELU if j != nl - 1 else L.ACT_NONE
                    plan = ConvPlan(B, cur.shape[1], cur.shape[2], cur.shape[3], cin_real, cout, 4, stride, 2, act)
                    plan.wp, plan.wt, plan.w1x1 = packs[(sc, j)].wp, packs[(sc, j)].wt, packs[(sc, j)].w1x1
                    if reuse and j == 0:
                        o = cache["l0"][i].detach()        # an alias of its own: each Function owns its output objects
                    else:
                        o = _empty((B, plan.Ho, plan.Wo, plan.Co), dt, dev)
                        plan.fwd(cur, bias, o, None)
                        if st is not None:
                            o.record_stream(lane)
                    plans.append(plan)
                    inputs.append(cur)
                    outs.append(o)
                    cur, cin_real = o, cout
        for st in scale_streams:
            if st is not None:
                lane.wait_stream(st)
        ctx.set_materialize_grads(False)
        ctx.plans, ctx.inputs, ctx.outs_ref = plans, inputs, outs
        ctx.sig_of, ctx.sn_bufs, ctx.sn_idx = sig_of, sn_bufs, sn_idx
        ctx.layer_cfg, ctx.param_refs = layer_cfg, aux["param_refs"]
        ctx.geom = (B, H, W, ca, cb)
        ctx.pyr_shapes = pyr_shapes
        if cache is not None and not reuse:
            cache.clear()
            cache.update(key=cache_key, pyr_shapes=pyr_shapes,
                         inputs=[inputs[i * nl] for i in range(NUM_D)], l0=[outs[i * nl] for i in range(NUM_D)])
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gouts):
        plans, inputs, outs = ctx.plans, ctx.inputs, ctx.outs_ref
        layer_cfg, refs = ctx.layer_cfg, ctx.param_refs
        B, H, W, ca, cb = ctx.geom
        nl = len(layer_cfg)
        dev, dt = outs[0].device, outs[0].dtype
        dc = L.dtype_code(dt)
        need_a = ctx.needs_input_grad[0]
        need_b = ctx.needs_input_grad[1] or need_a
        need_w = [ctx.needs_input_grad[3 + 2 * k] for k in range(NUM_D * nl)]
        need_bias = [ctx.needs_input_grad[4 + 2 * k] for k in range(NUM_D * nl)]
        ret: List[Optional[torch.Tensor]] = [None] * (2 * NUM_D * nl)
        sn_pos = {key: i for i, key in enumerate(ctx.sn_idx)}
        acc_slab, acc_offs = None, []
        if any(need_w):
            off = 0
            for pl in plans:
                acc_offs.append(off)
                off += pl.acc_numel()
            acc_slab = _zeros((off + len(sn_pos),), torch.float32, dev)   # + one dot product per spectral-norm layer
            dots = acc_slab[off:]
        unpack_jobs, sn_jobs = [], []
        lane = torch.cuda.current_stream()
        # as in forward: the coarse scales run on helper streams beside the full-resolution one
        scale_streams = [None] + [helper_stream(dev, f"scale{i}", same_priority=True) for i in range(1, NUM_D)]
        ev_ready = None
        if any(st is not None for st in scale_streams):
            ev_ready = torch.cuda.Event()
            ev_ready.record(lane)           # incoming gradients and the zeroed accumulators are ready
        dx32 = [None] * NUM_D       # fp32 gradient of each pyramid level (only when need_b)
        for i in reversed(range(NUM_D)):
            sc = NUM_D - 1 - i
            st = scale_streams[i]
            if st is not None:
                st.wait_event(ev_ready)
            with torch.cuda.stream(st if st is not None else lane):
                cur_stream = torch.cuda.current_stream()
                ws = helper_stream(dev) if any(need_w) else None      # weight gradients beside this scale's dgrad chain
                g_next = None           # gradient arriving from the layer above (its dgrad output)
                for j in reversed(range(nl)):
                    idx = i * nl + j
                    plan, x_in, o = plans[idx], inputs[idx], outs[idx]
                    gu = gouts[idx]
                    if gu is not None:
                        if gu.dtype != dt or tuple(gu.shape) != tuple(o.shape):
                            raise L.P2PError("gradient of a discriminator map must have the map's stored layout")
                        gu = gu.contiguous()
                        if st is not None:
                            gu.record_stream(st)
                    if gu is None and g_next is None:
                        continue        # nothing has reached this layer yet
                    k = sc * nl + j
                    cout = layer_cfg[j][1]
                    dz = _empty(tuple(o.shape), dt, dev)
                    gbias = None
                    if need_bias[k]:
                        gbias, ret[2 * k + 1] = _grad_slot(refs[2 * k + 1])
                    call("p2p_act_bwd", ptr(o), ptr(gu), ptr(g_next), ptr(dz), ptr(gbias), B * plan.Ho * plan.Wo, plan.Co,
                         cout, plan.act, dc)
                    if need_w[k]:
                        pw = refs[2 * k]
                        gw, ret[2 * k] = _grad_slot(pw)
                        acc = acc_slab[acc_offs[idx]: acc_offs[idx] + plan.acc_numel()]
                        if ws is not None:                        # the weight gradient runs beside the dgrad chain
                            ev = torch.cuda.Event()
                            ev.record(cur_stream)
                            ws.wait_event(ev)
                            dz.record_stream(ws)
                        with torch.cuda.stream(ws if ws is not None else cur_stream):
                            if layer_cfg[j][3]:
                                sigma = ctx.sig_of[(sc, j)]
                                dot = dots[sn_pos[(sc, j)]: sn_pos[(sc, j)] + 1]
                                # gW_bar = gW / sigma - (sum gW * W_bar) / sigma^2 * u v^T   (W = W_bar / sigma)
                                unpack_jobs.append(plan.wgrad_acc(x_in, dz, acc, gw, sigma=sigma, dot=dot,
                                                                  w_for_dot=pw.detach()))
                                u, v = ctx.sn_bufs[sn_pos[(sc, j)]]      # read at backward time: quirk Q1
                                sn_jobs.append(L.SnGradJob(ptr(gw), ptr(dot), ptr(sigma), ptr(u), ptr(v), pw.shape[0],
                                                           pw[0].numel()))
                            else:
                                unpack_jobs.append(plan.wgrad_acc(x_in, dz, acc, gw))
                    if j > 0 or need_b:
                        g_next = _empty(tuple(x_in.shape), dt, dev)
                        plan.dgrad(dz, g_next)
                    else:
                        g_next = None
                if need_b and g_next is not None:
                    h, w = ctx.pyr_shapes[i][1:3]
                    d32 = _zeros((B, h, w, 8), torch.float32, dev)
                    call("p2p_uncast_add", ptr(g_next), ptr(d32), B * h * w, 8, 16, dc)
                    if st is not None:
                        d32.record_stream(lane)
                    dx32[i] = d32
                if ws is not None:
                    cur_stream.wait_stream(ws)
        for st in scale_streams:
            if st is not None:
                lane.wait_stream(st)
        if unpack_jobs:
            L.call_multi("p2p_unpack_wgrad_multi", unpack_jobs)
        if sn_jobs:
            L.call_multi("p2p_sn_weight_grad_multi", sn_jobs)
        ga = gb = None
        if need_b:
            # fold the pyramid: d level i += avgpool_backward(d level i+1)
            for i in reversed(range(NUM_D - 1)):
                if dx32[i + 1] is None:
                    continue
                h, w = ctx.pyr_shapes[i][1:3]
                if dx32[i] is None:
                    dx32[i] = _zeros((B, h, w, 8), torch.float32, dev)
                call("p2p_avgpool3s2_bwd_add", ptr(dx32[i + 1]), ptr(dx32[i]), B, h, w, 8)
            if ctx.needs_input_grad[1]:
                gb = _zeros((B, cb, H, W), torch.float32, dev)
                if dx32[0] is not None:
                    call("p2p_d_input_grad", ptr(dx32[0]), ptr(gb), ca, cb, B, H, W)
            if need_a:
                ga = _zeros((B, ca, H, W), torch.float32, dev)
                if dx32[0] is not None:
                    call("p2p_d_input_grad", ptr(dx32[0]), ptr(ga), 0, ca, B, H, W)
        ctx.plans = ctx.inputs = ctx.outs_ref = None
        return (ga, gb, None, *ret)


class StoredToNCHW(torch.autograd.Function):
    """stored map [B,H,W,pad16(C)] (working dtype) -> NCHW fp32 [B,C,H,W]: the drop-in API edge."""

    @staticmethod
    def forward(ctx, t, c_real):
        B, H, W, Cs = t.shape
        out = _empty((B, c_real, H, W), torch.float32, t.device)
        call("p2p_nhwc_to_nchw", ptr(t), ptr(out), B, H, W, Cs, c_real, L.dtype_code(t.dtype))
        ctx.meta = (B, H, W, Cs, c_real, t.dtype)
        return out

    @staticmethod
    def backward(ctx, g):
        B, H, W, Cs, c_real, dt = ctx.meta
        g = g.contiguous().float()
        out = _empty((B, H, W, Cs), dt, g.device)
        call("p2p_nchw_to_nhwc", ptr(g), ptr(out), B, H, W, Cs, c_real, L.dtype_code(dt))
        return out, None


# =====================================================================================================
# Compression front end: CompressionNetwork.forward, reference networks.py:201-236
# =====================================================================================================
@torch.no_grad()
def compression_forward(x, convs, prelus, bn, training: bool, bits: Optional[int] = None, want_y: bool = True):
    """net_c(x) and/or compress(net_c(x), bits), forward only (the G/D iteration consumes it detached, train.py:300).
    convs = [(weight, bias)] x 3 (5x5 3->64, 3x3 64->64, 3x3 stride 2 64->12); prelus = the two PReLU scalars;
    bn = (gamma, beta, running_mean, running_var, num_batches_tracked) of the one BatchNorm.
    Returns (y or None, compressed or None), NCHW fp32."""
    L.require_cuda(x, "compression network input")
    if x.dim() != 4 or x.shape[1] != 3 or x.shape[2] % 2 or x.shape[3] % 2:
        raise L.P2PError("compression network input must be [B,3,H,W] with even H, W")
    B, _, H, W = x.shape
    dt, dev = _state["dtype"], x.device
    dc = L.dtype_code(dt)
    x = x.detach().contiguous().float()
    geo = [(3, 64, 5, 1), (64, 64, 3, 1), (64, 12, 3, 2)]
    plans, jobs = [], []
    hin, win = H, W
    for (cin, cout, k, stride), (w, _b) in zip(geo, convs):
        pl = ConvPlan(B, hin + 2 * (k // 2), win + 2 * (k // 2), pad16(cin), cin, cout, k, stride, 0)
        jobs.append(pl.pack_job(w, None, dt, want_wt=False))
        plans.append(pl)
        hin, win = pl.Ho, pl.Wo
    L.call_multi("p2p_pack_weights_multi", jobs, dc)
    # input: NCHW fp32 -> NHWC (16 stored channels) -> reflection pad 2
    x_nhwc = _empty((B, H, W, 16), dt, dev)
    call("p2p_nchw_to_nhwc", ptr(x), ptr(x_nhwc), B, H, W, 16, 3, dc)
    cur = _empty((B, H + 4, W + 4, 16), dt, dev)
    call("p2p_bn_act_store", ptr(x_nhwc), None, None, 0, None, ptr(cur), None, B, H, W, 16, 3, L.ACT_NONE, 1, 2, dc)
    # conv_input: 5x5 conv + PReLU, stored with the reflection pad 1 of the next conv
    r = _empty((B, H, W, 64), dt, dev)
    plans[0].fwd(cur, convs[0][1], r, None)
    cur = _empty((B, H + 2, W + 2, 64), dt, dev)
    call("p2p_bn_act_store", ptr(r), None, None, 0, ptr(prelus[0]), ptr(cur), None, B, H, W, 64, 64, L.ACT_PRELU, 1, 1, dc)
    # conv_block1: 3x3 conv + BatchNorm + PReLU
    gamma, beta, rm, rv, nbt = bn
    r = _empty((B, H, W, 64), dt, dev)
    stats = _zeros((128,), torch.float64, dev) if training else None
    plans[1].fwd(cur, convs[1][1], r, stats)
    sb = _empty((256,), torch.float32, dev)
    call("p2p_bn_finalize", ptr(stats), B * H * W, ptr(gamma), ptr(beta), ptr(rm), ptr(rv), ptr(nbt), ptr(sb), 64, 64,
         1 if training else 0)
    cur = _empty((B, H + 2, W + 2, 64), dt, dev)
    call("p2p_bn_act_store", ptr(r), ptr(sb), None, 0, ptr(prelus[1]), ptr(cur), None, B, H, W, 64, 64, L.ACT_PRELU, 1, 1,
         dc)
    # conv_block2: 3x3 stride-2 conv to 12 channels; PixelShuffle + normalize + identity (+ compress) in one kernel
    res = _empty((B, H // 2, W // 2, 16), dt, dev)
    plans[2].fwd(cur, convs[2][1], res, None)
    y = _empty((B, 3, H, W), torch.float32, dev) if want_y else None
    comp = _empty((B, 3, H, W), torch.float32, dev) if bits is not None else None
    call("p2p_c_output", ptr(res), ptr(x), ptr(y), ptr(comp), B, H, W, bits if bits is not None else 3, dc)
    return y, comp


# =====================================================================================================
# VGG19 perceptual loss: VGGLoss / Vgg19, reference networks.py:18-62
# =====================================================================================================
# torchvision vgg19().features[0:30]: output channels of the 3x3 pad-1 convolutions (each followed by ReLU), "M" = 2x2
# max-pool; Vgg19 taps the activation after features[1], [6], [11], [20], [29] (networks.py:42-51)
VGG_CFG = (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512)
VGG_CFG_CONV = tuple(v for v in VGG_CFG if v != "M")
VGG_TAP_AFTER_CONV = (0, 2, 4, 8, 12)          # index (0..12) of the convolution whose ReLU output is tapped
VGG_WEIGHTS = (1.0 / 32, 1.0 / 16, 1.0 / 8, 1.0 / 4, 1.0)      # networks.py:23


class VGGPack:
    """The thirteen frozen filter banks packed once for a (dtype, kernel family): forward and data-gradient layouts."""

    def __init__(self, weights, dt):
        self.key = (dt, _state["impl"], tuple(w._version for w in weights), tuple(w.data_ptr() for w in weights))
        self.plans, jobs = [], []
        dc = L.dtype_code(dt)
        for w in weights:
            cout, cin = w.shape[0], w.shape[1]
            pk = ConvPlan(1, 3, 3, pad16(cin), cin, cout, 3, 1, 1)
            jobs.append(pk.pack_job(w.detach(), None, dt))
            self.plans.append(pk)
        L.call_multi("p2p_pack_weights_multi", jobs, dc)


def vgg_stack(pack, biases, img, taps_only: bool):
    """vgg19.features[0:30] on an NCHW fp32 batch.  Returns (layers, taps): per convolution (plan, input, output,
    pooling input or None) -- empty when taps_only -- and the five tapped activations in stored layout."""
    N, _, H, W = img.shape
    dt, dev = _state["dtype"], img.device
    dc = L.dtype_code(dt)
    cur = _empty((N, H, W, 16), dt, dev)
    call("p2p_nchw_to_nhwc", ptr(img), ptr(cur), N, H, W, 16, 3, dc)
    layers, taps = [], []
    h, w, cin, ci = H, W, 3, 0
    pool_src = None
    for v in VGG_CFG:
        if v == "M":
            nxt = _empty((N, h // 2, w // 2, cur.shape[3]), dt, dev)
            call("p2p_maxpool2_fwd", ptr(cur), ptr(nxt), N, h, w, cur.shape[3], dc)
            pool_src = cur
            cur, h, w = nxt, h // 2, w // 2
            continue
        plan = ConvPlan(N, h, w, cur.shape[3], cin, v, 3, 1, 1, L.ACT_RELU)
        plan.wp, plan.wt = pack.plans[ci].wp, pack.plans[ci].wt
        out = _empty((N, h, w, plan.Co), dt, dev)
        plan.fwd(cur, biases[ci], out, None)
        if not taps_only:
            layers.append((plan, cur, out, pool_src))
        pool_src = None
        if ci in VGG_TAP_AFTER_CONV:
            taps.append(out)
        cur, cin, ci = out, v, ci + 1
    return layers, taps


class VGGLossFunction(torch.autograd.Function):
    """loss = sum_i w_i * mean|vgg(x)_i - vgg(y)_i| (VGGLoss.forward, networks.py:25-30), y detached.

    apply(x, y, pack, target_taps, *biases): x [B,3,H,W] fp32 NCHW (H, W multiples of 16); `pack` a VGGPack.  Either
    `y` (same shape: the feature stack then runs ONCE on the 2B images of cat(x, y)) or `target_taps` (the five tapped
    activations of the target, computed ahead of time by vgg_stack: the stack then runs on x alone).  Backward walks only
    the x half (frozen weights: data-gradients only)."""

    @staticmethod
    def forward(ctx, x, y, pack, target_taps, *biases):
        L.require_cuda(x, "VGGLoss input")
        if x.dim() != 4 or x.shape[1] != 3 or x.shape[2] % 16 or x.shape[3] % 16:
            raise L.P2PError("VGGLoss needs [B,3,H,W] batches with H, W multiples of 16")
        B, _, H, W = x.shape
        dt, dev = _state["dtype"], x.device
        dc = L.dtype_code(dt)
        if target_taps is None:
            L.require_cuda(y, "VGGLoss target")
            if y.shape != x.shape:
                raise L.P2PError("VGGLoss: input and target must have the same shape")
            img = torch.cat((x.detach().float(), y.detach().float()), 0).contiguous()
        else:
            img = x.detach().float().contiguous()
        layers, taps = vgg_stack(pack, biases, img, False)
        loss = torch.zeros((), dtype=torch.float32, device=dev)
        pairs = []
        for i, t in enumerate(taps):
            fake, real = (t[:B], t[B:]) if target_taps is None else (t, target_taps[i])
            if tuple(real.shape) != tuple(fake.shape) or real.dtype != fake.dtype:
                raise L.P2PError("VGGLoss: target activations do not match the input's")
            c_real = VGG_CFG_CONV[VGG_TAP_AFTER_CONV[i]]
            call("p2p_l1_fwd", ptr(fake), ptr(real), float(VGG_WEIGHTS[i]), ptr(loss), B * fake.shape[1] * fake.shape[2],
                 fake.shape[3], c_real, dc)
            pairs.append((fake, real))
        ctx.layers, ctx.pairs, ctx.geom = layers, pairs, (B, H, W)
        ctx.set_materialize_grads(False)
        return loss

    @staticmethod
    def backward(ctx, g):
        if ctx.needs_input_grad[1]:
            raise L.P2PError("VGGLoss: the target is detached in the reference (networks.py:29)")
        if g is None or not ctx.needs_input_grad[0]:
            return (None,) * (4 + 13)
        layers, pairs = ctx.layers, ctx.pairs
        B, H, W = ctx.geom
        dt, dev = layers[0][2].dtype, layers[0][2].device
        dc = L.dtype_code(dt)
        g = g.contiguous().float()
        g_next = None                    # gradient w.r.t. the OUTPUT of the layer being processed (x half only)
        for ci in reversed(range(len(layers))):
            plan, x_in, out, pool_src = layers[ci]
            d = plan.desc
            hB = _half_plan(plan, B)
            o_x = out[:B]
            gu = None
            if ci in VGG_TAP_AFTER_CONV:
                i = VGG_TAP_AFTER_CONV.index(ci)
                fake, real = pairs[i]
                gu = _empty(tuple(o_x.shape), dt, dev)
                call("p2p_l1_bwd", ptr(fake), ptr(real), float(VGG_WEIGHTS[i]), ptr(g), ptr(gu), B * d.Ho * d.Wo, plan.Co,
                     plan.cout_real, dc)
            if gu is None and g_next is None:
                continue
            dz = _empty(tuple(o_x.shape), dt, dev)
            call("p2p_act_bwd", ptr(o_x), ptr(gu), ptr(g_next), ptr(dz), None, B * d.Ho * d.Wo, plan.Co, plan.cout_real,
                 L.ACT_RELU, dc)
            din = _empty((B,) + tuple(x_in.shape[1:]), dt, dev)
            hB.dgrad(dz, din)
            if pool_src is not None:
                up = _empty((B,) + tuple(pool_src.shape[1:]), dt, dev)
                call("p2p_maxpool2_bwd", ptr(pool_src[:B]), ptr(din), ptr(up), B, pool_src.shape[1], pool_src.shape[2],
                     pool_src.shape[3], dc)
                din = up
            g_next = din
        gx = _empty((B, 3, H, W), torch.float32, dev)
        call("p2p_nhwc_to_nchw", ptr(g_next), ptr(gx), B, H, W, 16, 3, dc)
        ctx.layers = ctx.pairs = None
        return (gx, None, None, None) + (None,) * 13


def _half_plan(plan: ConvPlan, B: int) -> ConvPlan:
    """the same layer for the first B images of the batch it was built for"""
    d = plan.desc
    hp = ConvPlan(B, d.Hi, d.Wi, d.Ci, plan.cin_real, plan.cout_real, plan.k, plan.stride, plan.pad, plan.act)
    hp.wp, hp.wt, hp.w1x1 = plan.wp, plan.wt, plan.w1x1
    return hp
