"""CPU-side checks of the boundary: the library builds/loads, exports every symbol include/p2p_b200.h declares,
the ctypes binding covers all of them, and the product refuses to run without CUDA (no silent fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "p2p_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(p2p_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from p2p_pytorch_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib


def test_library_exports_every_declared_symbol(lib):
    dll = ctypes.CDLL(lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(dll, n), f"{n} declared in include/p2p_b200.h but not exported"


def test_binding_covers_header(lib):
    bound = set(lib.SIGNATURES) | set(lib.NOARG) | set(lib.MISC)
    assert bound == set(declared_symbols())


def test_version_and_error_string(lib):
    dll = lib.load()
    assert dll.p2p_version() == 100
    assert isinstance(lib.last_error(), str)


def test_argument_validation_needs_no_gpu(lib):
    """rejected arguments return a negative status before anything is enqueued"""
    dll = lib.load()
    rc = dll.p2p_adam_step(None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, 1, 1.0, None)
    assert rc < 0 and "p2p_adam_step" in lib.last_error()
    bad = lib.Conv(1, 8, 8, 12, 8, 8, 16, 3, 3, 1, 1, 16, 12, 0)       # Ci not a multiple of 16
    rc = dll.p2p_conv_fwd(ctypes.byref(bad), 1, 1, None, 1, None, lib.BF16, lib.IMPL_AUTO, None)
    assert rc < 0 and "multiples of 16" in lib.last_error()


def test_no_cpu_fallback():
    import p2p_pytorch_b200 as P
    g = P.ExpandNetwork()
    with pytest.raises(P._lib.P2PError):
        g(torch.zeros(1, 3, 32, 32))
    d = P.MultiscaleDiscriminator(6, 64, getIntermFeat=True)
    with pytest.raises(P._lib.P2PError):
        d(torch.zeros(1, 6, 32, 32))
    with pytest.raises(P._lib.P2PError):
        P.tv_loss(torch.zeros(1, 3, 8, 8))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "p2p-pytorch_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("oracle/", "").lower() or f == "__none__", f


def test_module_inventory_matches_reference():
    """same state_dict keys / shapes as the reference's define_G / define_D"""
    import p2p_pytorch_b200 as P
    from oracle import pix2pix_oracle as orc
    g, d = P.ExpandNetwork(), P.MultiscaleDiscriminator(6, 64, getIntermFeat=True)
    for net, ref in ((g, orc.init_g_state()), (d, orc.init_d_state())):
        sd = net.state_dict()
        assert list(sd) == list(ref)
        assert all(tuple(sd[k].shape) == tuple(ref[k].shape) for k in sd)
    assert sum(p.numel() for p in g.parameters()) == 2885290
    assert sum(p.numel() for p in d.parameters() if p.requires_grad) == 8303427
