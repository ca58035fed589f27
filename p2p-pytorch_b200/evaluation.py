"""Evaluation metric and checkpoint format of the reference.

psnr(): train.py:60-65 on the device -- the images are quantised to uint8 exactly as tensor2img does
(train.py:32-42: x * 255 in float32, clip, truncate) and the squared differences are summed as integers by
p2p_psnr_sse, so the result equals the reference's for every image of a batch at once.
save_checkpoint() / load_checkpoint(): the dictionary train.py:514-526 writes (`epoch`, `state_dict_g`,
`state_dict_c`; the optimizer / scheduler entries are commented out in the reference and optional here).  The
modules of this package have the reference's state_dict keys, so files interchange in both directions.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _lib as L


def psnr(ground: torch.Tensor, compressed: torch.Tensor) -> torch.Tensor:
    """PSNR in dB of every image pair of two [B,3,H,W] (or [3,H,W]) fp32 CUDA tensors in [0,1] units; float64 [B].
    Identical images give +inf (the reference's loop replaces that by 60.0, train.py:482-484)."""
    L.require_cuda(ground, "psnr input")
    L.require_cuda(compressed, "psnr input")
    if ground.shape != compressed.shape:
        raise L.P2PError("psnr needs two tensors of the same shape")
    if ground.dim() == 3:
        ground, compressed = ground[None], compressed[None]
    B = ground.shape[0]
    per_image = ground[0].numel()
    a = ground.detach().contiguous().float()
    b = compressed.detach().contiguous().float()
    sse = torch.zeros(B, dtype=torch.int64, device=a.device)          # read by the kernel as uint64
    L.call("p2p_psnr_sse", L.ptr(a), L.ptr(b), L.ptr(sse), B, per_image)
    mse = sse.double() / float(per_image)
    return 10.0 * torch.log10(255.0 ** 2 / mse)


def save_checkpoint(path: str, epoch: int, net_g, net_c, **extra) -> None:
    """torch.save of the reference's state dictionary (train.py:514-526): epoch + 1, state_dict_g, state_dict_c."""
    state = {"epoch": epoch + 1, "state_dict_g": net_g.state_dict(), "state_dict_c": net_c.state_dict()}
    state.update(extra)
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    torch.save(state, path)


def load_checkpoint(path: str, net_g, net_c=None, net_d=None, map_location: Optional[str] = None,
                    trusted: bool = False) -> int:
    """Load a file written by the reference (or by save_checkpoint) and return its `epoch` (train.py:103-118 reads
    the same keys; entries the file does not hold are skipped).  The format holds an int and tensor state_dicts
    only, so the file is read with `weights_only=True`; pass `trusted=True` for a file of your own that carries
    other pickled objects (unpickling runs arbitrary code)."""
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    state = torch.load(path, map_location=map_location, weights_only=not trusted)
    net_g.load_state_dict(state["state_dict_g"])
    if net_c is not None and "state_dict_c" in state:
        net_c.load_state_dict(state["state_dict_c"])
    if net_d is not None and "state_dict_d" in state:
        net_d.load_state_dict(state["state_dict_d"])
    return int(state.get("epoch", 0))


_gen_streams = {}


@torch.no_grad()
def generate(net_g, x: torch.Tensor, streams: int = 2) -> torch.Tensor:
    """Generator inference as test.py:37-41 runs it (`net_g(input)`, no autograd) for a whole batch, with the batch
    split over `streams` CUDA streams: in eval mode BatchNorm uses its running statistics, so images are independent and
    the split is exact; the sub-batches' kernel chains fill each other's idle SMs (second waves of the persistent
    convolution kernels, memory-bound glue).  Returns fp32 [B,3,H,W]."""
    if net_g.training:
        raise ValueError("generate() needs net_g.eval(): train-mode BatchNorm statistics couple the images of a batch")
    B = x.shape[0]
    streams = max(1, min(int(streams), B))
    if streams == 1:
        return net_g(x)
    cur = torch.cuda.current_stream()
    key = (str(x.device), streams)
    if key not in _gen_streams:
        _gen_streams[key] = [torch.cuda.Stream(device=x.device) for _ in range(streams)]
    out = torch.empty((B, 3, x.shape[2], x.shape[3]), dtype=torch.float32, device=x.device)
    bounds = [B * i // streams for i in range(streams + 1)]
    for st, lo, hi in zip(_gen_streams[key], bounds[:-1], bounds[1:]):
        st.wait_stream(cur)
        with torch.cuda.stream(st):
            out[lo:hi].copy_(net_g(x[lo:hi]))
    for st in _gen_streams[key]:
        cur.wait_stream(st)
    return out
