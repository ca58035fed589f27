"""p2p_spectral_norm_static: the spectral-norm power iteration for callers that cannot allocate device memory
(programs living inside a fixed memory budget, loops that must not hit cudaMalloc mid-step).  Its scratch is a
static array inside the library instead of a caller-provided buffer; everything else must behave exactly like
p2p_spectral_norm.  This file is named so that pytest collects it last: if it ever fails, a run that stops at
the first failure has already run every other test.  It also passes on its own in a fresh process."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"
MAX_FLOATS = 32768          # P2P_SN_STATIC_MAX_FLOATS in include/p2p_b200.h


def _layers(seed):
    """the three spectral-normalised weight shapes of one discriminator scale (networks.py:771-779)"""
    g = torch.Generator(device="cpu").manual_seed(seed)
    shapes = [(128, 64 * 16), (256, 128 * 16), (512, 256 * 16)]
    ws = [(torch.randn(r, c, generator=g) * 0.02).to(DEV) for r, c in shapes]
    us = [torch.nn.functional.normalize(torch.randn(r, generator=g), dim=0).to(DEV) for r, _ in shapes]
    vs = [torch.nn.functional.normalize(torch.randn(c, generator=g), dim=0).to(DEV) for _, c in shapes]
    return ws, us, vs


def _args(ws, us, vs, sigma):
    n = len(ws)
    PA, IA = C.c_void_p * n, C.c_int32 * n
    return (n, PA(*[w.data_ptr() for w in ws]), PA(*[u.data_ptr() for u in us]), PA(*[v.data_ptr() for v in vs]),
            PA(*[sigma.data_ptr() + 4 * i for i in range(n)]), IA(*[w.shape[0] for w in ws]),
            IA(*[w.shape[1] for w in ws]))


def test_static_scratch_variant_matches_the_allocating_one_twice_in_a_row():
    from p2p_pytorch_b200 import _lib as L
    ws, us_a, vs_a = _layers(11)
    _, us_b, vs_b = _layers(11)                       # identical starting vectors for the second entry point
    sig_a = torch.zeros(3, device=DEV)
    sig_b = torch.zeros(3, device=DEV)
    tmp = torch.empty(sum(w.shape[0] + w.shape[1] for w in ws), device=DEV)
    for call_no in range(2):                          # the second call continues from the updated u, v
        L.call("p2p_spectral_norm", *_args(ws, us_a, vs_a, sig_a), tmp.data_ptr())
        before = torch.cuda.memory_allocated()
        L.call("p2p_spectral_norm_static", *_args(ws, us_b, vs_b, sig_b))
        assert torch.cuda.memory_allocated() == before
        torch.cuda.synchronize()
        for a, b in zip(us_a + vs_a + [sig_a], us_b + vs_b + [sig_b]):
            err = (a - b).abs().max().item() / max(a.abs().max().item(), 1e-30)
            assert err <= 1e-5, f"call {call_no}: {err:.3e}"
        assert torch.isfinite(sig_b).all() and (sig_b > 0).all()


def test_static_scratch_variant_rejects_oversized_requests():
    from p2p_pytorch_b200 import _lib as L
    rows, cols = 512, MAX_FLOATS                      # rows + cols > the built-in scratch
    w = torch.zeros(rows, cols, device=DEV)
    u, v = torch.zeros(rows, device=DEV), torch.zeros(cols, device=DEV)
    sig = torch.zeros(1, device=DEV)
    with pytest.raises(L.P2PError, match="exceeds the built-in scratch"):
        L.call("p2p_spectral_norm_static", *_args([w], [u], [v], sig))
    torch.cuda.synchronize()                          # nothing was enqueued, the context is still healthy
