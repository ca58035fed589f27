"""Seed-identical initialisation: after `torch.manual_seed(s)` this package's
define_G / define_D / define_C consume the RNG stream exactly like the reference's (networks.py:128-155 init_weights /
init_net, :157-171, :708-714), so every parameter and buffer is bit-identical -- including the spectral-norm wrapped
convolutions, which init_weights skips (their `weight` attribute is gone, networks.py:573) and which therefore keep
PyTorch's default initialisation, and the u / v vectors drawn at networks.py:566-577.

Runs against the reference's own module: oracle/_ref/networks.py (copied by oracle/build_ref.py in the build
container; skipped where it is absent)."""
import contextlib
import io

import pytest
import torch

from oracle.build_ref import load_reference_networks


@pytest.fixture(scope="module")
def ref():
    mod = load_reference_networks()
    if mod is None:
        pytest.skip("oracle/_ref/networks.py not built (no reference tree on this machine)")
    return mod


@pytest.mark.parametrize("seed", [123, 7])
def test_same_seed_gives_bit_identical_networks(ref, seed):
    import p2p_pytorch_b200 as P
    built = {}
    for name, mod in (("ref", ref), ("ours", P)):
        with contextlib.redirect_stdout(io.StringIO()):
            torch.manual_seed(seed)
            g = mod.define_G("normal", 0.02, gpu_id="cpu")
            d = mod.define_D(6, 64, gpu_id="cpu")
            c = mod.define_C("normal", 0.02, gpu_id="cpu")
        built[name] = (g.state_dict(), d.state_dict(), c.state_dict(), torch.rand(4))
    for a, b in zip(built["ref"][:3], built["ours"][:3]):
        assert list(a.keys()) == list(b.keys())
        for k in a:
            assert a[k].dtype == b[k].dtype and torch.equal(a[k], b[k]), k
    assert torch.equal(built["ref"][3], built["ours"][3])          # the RNG stream is left in the same state


def test_spectral_norm_layers_keep_default_init(ref):
    """quirk Q2: normal(0, 0.02) never reaches the SN-wrapped convolutions (|w| <= 1/sqrt(fan_in) instead)"""
    import p2p_pytorch_b200 as P
    with contextlib.redirect_stdout(io.StringIO()):
        torch.manual_seed(1)
        d = P.define_D(6, 64, gpu_id="cpu")
    sd = d.state_dict()
    w = sd["scale0_layer1.0.module.weight_bar"]
    assert abs(w.abs().max().item() - 1.0 / 1024 ** 0.5) < 1e-4      # kaiming-uniform bound of a 64x4x4 fan-in
    assert abs(sd["scale0_layer0.0.weight"].std().item() - 0.02) < 2e-3
