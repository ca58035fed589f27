"""Fused objective kernels as autograd Functions: LSGAN (reference networks.py:840-850), the
discriminator-feature-matching L1 (train.py:344-351) and total variation (train.py:123-126).

"Stored" maps are what DiscriminatorFunction returns: NHWC, working dtype, channels padded to 16.
Every Function keeps only its inputs; the gradient is recomputed from them in one pass.
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from . import _lib as L
from ._lib import call, ptr


def _scalar(dev):
    return torch.zeros((), dtype=torch.float32, device=dev)


class _MapLoss(torch.autograd.Function):
    """sum_i weight_i * mean(f(p_i [, r_i])) over stored maps; kind 'mse' (vs constant) or 'l1' (vs r_i)."""

    @staticmethod
    def _jobs(kind, target, weights, c_reals, preds, refs, grads):
        jobs = []
        for i, p in enumerate(preds):
            B, H, W, C = p.shape
            jobs.append(L.LossJob(ptr(p), ptr(refs[i]) if kind == "l1" else None, ptr(grads[i]) if grads else None,
                                  float(target), float(weights[i]), B * H * W, C, c_reals[i], 1 if kind == "l1" else 0, 0))
        return jobs

    @staticmethod
    def forward(ctx, kind, target, weights, c_reals, n, *maps):
        preds, refs = maps[:n], maps[n:]
        loss = _scalar(preds[0].device)
        if any(p.dtype != preds[0].dtype for p in preds):
            raise L.P2PError("all maps of one loss must have the same dtype")
        # one launch for all maps (p2p_map_loss_multi_fwd), <= P2P_MULTI_MAX maps per launch
        L.call_multi("p2p_map_loss_multi_fwd", _MapLoss._jobs(kind, target, weights, c_reals, preds, refs, None), ptr(loss),
                     L.dtype_code(preds[0].dtype))
        ctx.kind, ctx.target, ctx.weights, ctx.c_reals, ctx.n = kind, target, weights, c_reals, n
        ctx.maps = maps
        return loss

    @staticmethod
    def backward(ctx, g):
        n, maps = ctx.n, ctx.maps
        preds, refs = maps[:n], maps[n:]
        g = g.contiguous().float()
        need = [i for i in range(n) if ctx.needs_input_grad[5 + i]]
        outs = [None] * n
        if need:
            for i in need:
                outs[i] = torch.empty_like(preds[i])
            jobs = _MapLoss._jobs(ctx.kind, ctx.target, [ctx.weights[i] for i in need], [ctx.c_reals[i] for i in need],
                                  [preds[i] for i in need], [refs[i] for i in need] if ctx.kind == "l1" else [],
                                  [outs[i] for i in need])
            L.call_multi("p2p_map_loss_multi_bwd", jobs, ptr(g), L.dtype_code(preds[0].dtype))
        return (None, None, None, None, None, *outs, *([None] * len(refs)))


def lsgan_loss(preds: Sequence[torch.Tensor], c_reals: Sequence[int], label: float) -> torch.Tensor:
    """sum over scales of mean((p - label)^2) on stored maps"""
    preds = [p.contiguous() for p in preds]
    return _MapLoss.apply("mse", label, [1.0] * len(preds), list(c_reals), len(preds), *preds)


def feature_matching_loss(fake: List[List[torch.Tensor]], real: List[List[torch.Tensor]],
                          c_reals: Sequence[int]) -> torch.Tensor:
    """train.py:344-351 on stored maps: sum_{i<3, j<4} (1/3) (4/4) 10 L1(fake_ij, real_ij.detach())"""
    num_d, n_layers_d = len(fake), 3
    w = (1.0 / num_d) * (4.0 / (n_layers_d + 1)) * 10.0
    f = [fake[i][j].contiguous() for i in range(num_d) for j in range(len(fake[i]) - 1)]
    r = [real[i][j].detach().contiguous() for i in range(num_d) for j in range(len(fake[i]) - 1)]
    cr = [c_reals[j] for i in range(num_d) for j in range(len(fake[i]) - 1)]
    return _MapLoss.apply("l1", 0.0, [w] * len(f), cr, len(f), *f, *r)


class _MseFlat(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p, target):
        L.require_cuda(p, "GANLoss input")
        p = p.contiguous().float()
        loss = _scalar(p.device)
        call("p2p_mse_flat_fwd", ptr(p), float(target), 1.0, ptr(loss), p.numel())
        ctx.p, ctx.target = p, target
        return loss

    @staticmethod
    def backward(ctx, g):
        gp = torch.empty_like(ctx.p)
        call("p2p_mse_flat_bwd", ptr(ctx.p), float(ctx.target), 1.0, ptr(g.contiguous().float()), ptr(gp), gp.numel())
        return gp, None


def mse_const_flat(p: torch.Tensor, target: float) -> torch.Tensor:
    """mean((p - target)^2) of any fp32 CUDA tensor (drop-in GANLoss path)"""
    return _MseFlat.apply(p, target)


class _TV(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        L.require_cuda(x, "tv_loss input")
        x = x.contiguous().float()
        B, C, H, W = x.shape
        loss = _scalar(x.device)
        call("p2p_tv_fwd", ptr(x), ptr(loss), B, C, H, W)
        ctx.x = x
        return loss

    @staticmethod
    def backward(ctx, g):
        x = ctx.x
        B, C, H, W = x.shape
        gx = torch.zeros_like(x)
        call("p2p_tv_bwd_add", ptr(x), ptr(g.contiguous().float()), ptr(gx), B, C, H, W)
        return gx


def tv_loss(x: torch.Tensor) -> torch.Tensor:
    """calc_tv_Loss of train.py:123-126 on an NCHW fp32 batch"""
    return _TV.apply(x)
