"""Drop-in replacement for the hot-path part of the reference's `networks.py`.

Same public names, signatures, `state_dict()` keys / shapes, parameter registration order (so the same
`torch.manual_seed` gives the same initial weights) and output structure as the reference:

    define_G(init_type, init_gain, gpu_id)                      reference networks.py:164-171
    define_D(input_nc, ndf, norm, use_sigmoid, init_type, ...)  reference networks.py:708-714
    GANLoss(use_lsgan, target_real_label, target_fake_label)    reference networks.py:808-850
    VGGLoss(device) / Vgg19()                                   reference networks.py:18-62
    init_weights / init_net                                     reference networks.py:128-155

The `nn.Conv2d` / `nn.BatchNorm2d` / `nn.PReLU` objects below are parameter containers only -- their
`forward` is never called.  `ExpandNetwork.forward` and `MultiscaleDiscriminator.forward` hand the
parameters to one autograd Function each (engine.py) whose forward/backward are sequences of the
sm_100a kernels in libp2p_b200.so.  There is no CPU path: inputs must be CUDA tensors.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
from torch.nn import Parameter, init

from . import _lib as L
from . import engine as E
from .losses import mse_const_flat


# -----------------------------------------------------------------------------------------------------
# initialisation (reference networks.py:128-155)
# -----------------------------------------------------------------------------------------------------
def init_weights(net, init_type="normal", gain=0.02):
    """Conv/Linear modules that still own a `weight`: chosen init + zero bias; BatchNorm2d: weight ~ N(1, gain),
    bias 0.  Spectral-norm wrapped convs have no `weight` attribute and keep their default init, as in the
    reference."""
    def visit(m):
        cls = type(m).__name__
        if hasattr(m, "weight") and ("Conv" in cls or "Linear" in cls):
            if init_type == "normal":
                init.normal_(m.weight.data, 0.0, gain)
            elif init_type == "xavier":
                init.xavier_normal_(m.weight.data, gain=gain)
            elif init_type == "kaiming":
                init.kaiming_normal_(m.weight.data, a=0, mode="fan_in")
            elif init_type == "orthogonal":
                init.orthogonal_(m.weight.data, gain=gain)
            else:
                raise NotImplementedError("initialization method [%s] is not implemented" % init_type)
            if getattr(m, "bias", None) is not None:
                init.constant_(m.bias.data, 0.0)
        elif "BatchNorm2d" in cls:
            init.normal_(m.weight.data, 1.0, gain)
            init.constant_(m.bias.data, 0.0)

    print("initialize network with %s" % init_type)
    net.apply(visit)


def init_net(net, init_type="normal", init_gain=0.02, gpu_id="cuda:0"):
    net.to(gpu_id)
    init_weights(net, init_type, gain=init_gain)
    return net


def define_G(init_type="normal", init_gain=0.02, gpu_id="cuda:0"):
    return init_net(ExpandNetwork(), init_type, init_gain, gpu_id)


def define_C(init_type="normal", init_gain=0.02, gpu_id="cuda:0"):
    """reference networks.py:157-162"""
    return init_net(CompressionNetwork(), init_type, init_gain, gpu_id)


def define_D(input_nc, ndf, norm="batch", use_sigmoid=False, init_type="normal", init_gain=0.02, gpu_id="cuda:0"):
    # `norm` is accepted and ignored, as in the reference (networks.py:712 passes norm_layer=None)
    net = MultiscaleDiscriminator(input_nc, ndf, n_layers=3, norm_layer=None, use_sigmoid=use_sigmoid, num_D=3,
                                  getIntermFeat=True)
    return init_net(net, init_type, init_gain, gpu_id)


# -----------------------------------------------------------------------------------------------------
# generator containers (reference networks.py:395-523)
# -----------------------------------------------------------------------------------------------------
class ConvLayer(nn.Module):
    """reflection pad k//2 + conv; holds `conv2d.{weight,bias}`"""

    def __init__(self, in_channels, out_channels, kernel_size, stride):
        super().__init__()
        self.reflection_pad = nn.ReflectionPad2d(kernel_size // 2)
        self.conv2d = nn.Conv2d(in_channels, out_channels, kernel_size, stride)


class UpsampleConvLayer(nn.Module):
    """[nearest upsample] + reflection pad + conv; holds `conv2d.{weight,bias}`"""

    def __init__(self, in_channels, out_channels, kernel_size, stride, upsample=None):
        super().__init__()
        self.upsample = nn.Upsample(scale_factor=upsample, mode="nearest") if upsample else None
        self.reflection_pad = nn.ReflectionPad2d(kernel_size // 2)
        self.conv2d = nn.Conv2d(in_channels, out_channels, kernel_size, stride)


class ResidualBlock(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv1 = ConvLayer(channels, channels, kernel_size=3, stride=1)
        self.in1 = nn.BatchNorm2d(channels, affine=True)
        self.relu = nn.ReLU()
        self.conv2 = ConvLayer(channels, channels, kernel_size=3, stride=1)
        self.in2 = nn.BatchNorm2d(channels, affine=True)


class PixelUnshuffle(nn.Module):
    def __init__(self, downscale_factor):
        super().__init__()
        self.downscale_factor = downscale_factor


class ExpandNetwork(nn.Module):
    """The generator `define_G` returns.  2,885,290 parameters, 169 state_dict keys."""

    def __init__(self):
        super().__init__()
        self.relu = nn.PReLU()
        self.leakyRelu = nn.LeakyReLU(0.2)
        self.tanh = nn.Tanh()
        self.inversePixel = PixelUnshuffle(2)
        self.upsclaing = nn.Upsample(scale_factor=2, mode="nearest")   # (sic) attribute name of the reference
        self.conv1 = ConvLayer(12, 32, kernel_size=9, stride=1)
        self.in1_e = nn.BatchNorm2d(32, affine=True)
        self.conv2 = ConvLayer(32, 64, kernel_size=3, stride=2)
        self.in2_e = nn.BatchNorm2d(64, affine=True)
        self.conv3 = ConvLayer(64, 128, kernel_size=3, stride=2)
        self.in3_e = nn.BatchNorm2d(128, affine=True)
        for r in range(1, 10):
            setattr(self, f"res{r}", ResidualBlock(128))
        self.deconv3 = UpsampleConvLayer(128, 64, kernel_size=3, stride=1, upsample=2)
        self.in3_d = nn.BatchNorm2d(64, affine=True)
        self.deconv2 = UpsampleConvLayer(64, 32, kernel_size=3, stride=1, upsample=2)
        self.in2_d = nn.BatchNorm2d(32, affine=True)
        self.deconv1 = UpsampleConvLayer(32, 3, kernel_size=9, stride=1)
        self.in1_d = nn.BatchNorm2d(3, affine=True)

    def _layer_modules(self):
        for conv_name, bn_name, *_ in E.G_LAYERS:
            conv, bn = self, self
            for part in conv_name.split("."):
                conv = getattr(conv, part)
            for part in bn_name.split("."):
                bn = getattr(bn, part)
            yield conv.conv2d, bn

    def forward(self, x):
        if x.shape[2] % 4 or x.shape[3] % 4:
            raise L.P2PError("generator input height and width must be multiples of 4")
        params, refs, buffers = [], [self.relu.weight], []
        for conv, bn in self._layer_modules():
            layer = [conv.weight, conv.bias, bn.weight, bn.bias]
            params += layer
            refs += layer
            buffers.append((bn.running_mean, bn.running_var, bn.num_batches_tracked))
        aux = dict(buffers=buffers, training=self.training, param_refs=refs)
        return E.GeneratorFunction.apply(x, self.relu.weight, aux, *params)


# -----------------------------------------------------------------------------------------------------
# discriminator containers (reference networks.py:525-806)
# -----------------------------------------------------------------------------------------------------
def l2normalize(v, eps=1e-12):
    return v / (v.norm() + eps)


class SpectralNorm(nn.Module):
    """Parameter container with the reference's registration order: the wrapped conv loses `weight` and gains
    `weight_u`, `weight_v` (requires_grad False) and `weight_bar` (networks.py:551-577).  The power iteration
    itself runs on the device, batched over all wrapped layers, inside DiscriminatorFunction."""

    def __init__(self, module, name="weight", power_iterations=1):
        super().__init__()
        if power_iterations != 1:
            raise NotImplementedError("the fused kernel does one power iteration per forward, like define_D")
        self.module, self.name, self.power_iterations = module, name, power_iterations
        if not hasattr(module, name + "_bar"):
            w = getattr(module, name)
            rows = w.data.shape[0]
            cols = w.data.view(rows, -1).shape[1]
            u = Parameter(w.data.new(rows).normal_(0, 1), requires_grad=False)
            v = Parameter(w.data.new(cols).normal_(0, 1), requires_grad=False)
            u.data = l2normalize(u.data)
            v.data = l2normalize(v.data)
            w_bar = Parameter(w.data)
            del module._parameters[name]
            module.register_parameter(name + "_u", u)
            module.register_parameter(name + "_v", v)
            module.register_parameter(name + "_bar", w_bar)


class NLayerDiscriminator(nn.Module):
    """One PatchGAN stack: k=4, pad=2, strides 2,2,2,1,1, LeakyReLU(0.2) after all but the last conv."""

    def __init__(self, input_nc, ndf=64, n_layers=3, norm_layer=None, use_sigmoid=False, getIntermFeat=False):
        super().__init__()
        if use_sigmoid or norm_layer is not None or not getIntermFeat or n_layers != 3:
            raise NotImplementedError("only the configuration define_D builds is implemented "
                                      "(n_layers=3, no norm layer, no sigmoid, getIntermFeat=True)")
        self.getIntermFeat, self.n_layers = getIntermFeat, n_layers
        kw = 4
        padw = int(np.ceil((kw - 1.0) / 2))
        for j, (cin, cout, stride, sn) in enumerate(E.d_layers(input_nc, ndf)):
            conv = nn.Conv2d(cin, cout, kernel_size=kw, stride=stride, padding=padw)
            mods = [SpectralNorm(conv) if sn else conv]
            if j != n_layers + 1:
                mods.append(nn.LeakyReLU(0.2, True))
            setattr(self, "model" + str(j), nn.Sequential(*mods))


class MultiscaleDiscriminator(nn.Module):
    """The discriminator `define_D` returns: three NLayerDiscriminators on a 1x, 1/2x, 1/4x pyramid.
    `forward(x)` returns list[3] of list[5] NCHW fp32 tensors like the reference; `forward_stored(a, b)`
    returns the same 15 maps in the kernels' own layout (NHWC, working dtype) for the fused loss kernels."""

    def __init__(self, input_nc, ndf=64, n_layers=3, norm_layer=None, use_sigmoid=False, num_D=3,
                 getIntermFeat=False):
        super().__init__()
        if num_D != 3:
            raise NotImplementedError("only num_D=3 (what define_D builds) is implemented")
        self.num_D, self.n_layers, self.getIntermFeat = num_D, n_layers, getIntermFeat
        self.input_nc, self.ndf = input_nc, ndf
        for i in range(num_D):
            net = NLayerDiscriminator(input_nc, ndf, n_layers, norm_layer, use_sigmoid, getIntermFeat)
            for j in range(n_layers + 2):
                setattr(self, "scale" + str(i) + "_layer" + str(j), getattr(net, "model" + str(j)))
        self.downsample = nn.AvgPool2d(3, stride=2, padding=[1, 1], count_include_pad=False)
        self._cfg = E.d_layers(input_nc, ndf)

    def _gather(self, frozen):
        params, sn_bufs = [], []
        for sc in range(self.num_D):
            for j, (_ci, _co, _s, sn) in enumerate(self._cfg):
                m = getattr(self, f"scale{sc}_layer{j}")[0]
                if sn:
                    conv = m.module
                    params += [conv.weight_bar, conv.bias]
                    sn_bufs.append((conv.weight_u, conv.weight_v))
                else:
                    params += [m.weight, m.bias]
        refs = list(params)
        if frozen:
            params = [p.detach() for p in params]
        return params, refs, sn_bufs

    def power_iteration(self):
        """One power iteration of every spectral-norm layer (what each forward pass starts with, networks.py:537-547):
        advances u, v in place and returns the nine sigmas.  u, v depend only on the weights and on their own previous
        value, so a caller that runs several passes between two optimizer steps may run their power iterations first,
        in order, and hand each pass its `sigma` -- the passes themselves then no longer depend on one another."""
        params, _refs, sn_bufs = self._gather(True)
        nl = len(self._cfg)
        sn_idx = [(sc, j) for sc in range(self.num_D) for j in range(nl) if self._cfg[j][3]]
        wbars = [params[2 * (sc * nl + j)] for sc, j in sn_idx]
        return E.spectral_power_iteration(wbars, [u for u, _ in sn_bufs], [v for _, v in sn_bufs])

    def forward_stored(self, a, b=None, frozen=False, first_layer_cache=None, sigma=None):
        """15 stored maps (scale-major, full resolution first).  `frozen=True` treats the weights as constants:
        gradients still reach the input, no weight gradients are computed (what `loss_g.backward()` needs;
        the reference computes and then discards them, train.py:385-388).
        `first_layer_cache`: a dict shared by two passes over the SAME image pair between two optimizer steps (the
        caller's promise): the first pass fills it, the second reuses the input pyramid and the outputs of the first
        layer (no spectral norm there, so they are bit-identical) instead of recomputing them."""
        params, refs, sn_bufs = self._gather(frozen)
        aux = dict(layer_cfg=self._cfg, sn_bufs=sn_bufs, param_refs=refs, first_layer_cache=first_layer_cache,
                   sigma=sigma)
        outs = E.DiscriminatorFunction.apply(a, b, aux, *params)
        nl = len(self._cfg)
        return [list(outs[i * nl:(i + 1) * nl]) for i in range(self.num_D)]

    def forward(self, input):
        stored = self.forward_stored(input, None)
        result = []
        for maps in stored:
            row = []
            for j, t in enumerate(maps):
                row.append(E.StoredToNCHW.apply(t, self._cfg[j][1]))
            result.append(row)
        return result


# -----------------------------------------------------------------------------------------------------
# compression front end (reference networks.py:201-236)
# -----------------------------------------------------------------------------------------------------
class CompressionNetwork(nn.Module):
    """The network `define_C` returns: same sub-module structure and state_dict keys as the reference.  FORWARD
    ONLY: the G/D iteration consumes `compress(net_c(real_b), 3).detach()` (train.py:297-300); training net_c
    itself is phase 3 of the reference loop (train.py:392-402), which is out of scope, so the output carries no
    grad_fn."""

    def __init__(self):
        super().__init__()
        self.conv_input = nn.Sequential(ConvLayer(3, 64, kernel_size=5, stride=1), nn.PReLU())
        self.conv_block1 = nn.Sequential(ConvLayer(64, 64, kernel_size=3, stride=1), nn.BatchNorm2d(64), nn.PReLU())
        self.conv_block2 = nn.Sequential(ConvLayer(64, 12, kernel_size=3, stride=2), nn.PixelShuffle(2))

    def _run(self, x, bits, want_y):
        c0, c1, c2 = self.conv_input[0].conv2d, self.conv_block1[0].conv2d, self.conv_block2[0].conv2d
        bn = self.conv_block1[1]
        return E.compression_forward(
            x, [(c0.weight, c0.bias), (c1.weight, c1.bias), (c2.weight, c2.bias)],
            [self.conv_input[1].weight, self.conv_block1[2].weight],
            (bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked), self.training, bits, want_y)

    def forward(self, x):
        """identity + normalize(pixel_shuffle(convs(x))) -- networks.py:222-236"""
        return self._run(x, None, True)[0]

    def front_end(self, real_b, bits=3):
        """compress(net_c(real_b), bits): the generator's input (train.py:297, :466) in one pass"""
        return self._run(real_b, bits, False)[1]


# -----------------------------------------------------------------------------------------------------
# GANLoss (reference networks.py:808-850)
# -----------------------------------------------------------------------------------------------------
class GANLoss(nn.Module):
    """LSGAN objective: sum over scales of MSE(last map, constant label).  `tensor` is accepted for signature
    compatibility; no label tensor is ever materialised (the constant is a kernel argument)."""

    def __init__(self, use_lsgan=True, target_real_label=1.0, target_fake_label=0.0, tensor=None):
        super().__init__()
        if not use_lsgan:
            raise NotImplementedError("only the LSGAN (MSE) variant the reference trains with is implemented")
        self.real_label, self.fake_label = target_real_label, target_fake_label
        self.Tensor = tensor

    def __call__(self, input, target_is_real):
        label = self.real_label if target_is_real else self.fake_label
        if isinstance(input[0], (list, tuple)):
            preds = [scale[-1] for scale in input]
        else:
            preds = [input[-1]]
        loss = 0
        for p in preds:
            loss = loss + mse_const_flat(p, label)
        return loss


# -----------------------------------------------------------------------------------------------------
# VGG19 perceptual loss (reference networks.py:18-62; train.py:193, :377, :395)
# -----------------------------------------------------------------------------------------------------
class Vgg19(nn.Module):
    """torchvision's vgg19().features[0:30] cut into the reference's five slices (networks.py:42-51); the modules
    are parameter containers with the reference's names (`slice2.5.weight` = features[5].weight), all frozen.

    The reference downloads the ImageNet weights in its constructor (networks.py:35).  Here construction never touches
    the network: pass `features_state_dict` (torchvision's `vgg19(...).features.state_dict()`, a whole-model
    `vgg19(...).state_dict()`, or this module's own keys) or call `load_pretrained` / `load_state_dict` later.
    Until then the weights are torchvision's random initialisation of the same architecture."""

    def __init__(self, requires_grad=False, features_state_dict=None):
        super().__init__()
        if requires_grad:
            raise NotImplementedError("the perceptual network is frozen in the reference (networks.py:52-54)")
        ends = (1, 6, 11, 20, 29)
        slices = [nn.Sequential() for _ in ends]
        idx, cin = 0, 3
        for v in E.VGG_CFG:
            k = next(i for i, e in enumerate(ends) if idx <= e)
            if v == "M":
                slices[k].add_module(str(idx), nn.MaxPool2d(kernel_size=2, stride=2))
                idx += 1
                continue
            conv = nn.Conv2d(cin, v, kernel_size=3, padding=1)
            init.kaiming_normal_(conv.weight, mode="fan_out", nonlinearity="relu")       # torchvision's vgg init
            init.constant_(conv.bias, 0.0)
            slices[k].add_module(str(idx), conv)
            k = next(i for i, e in enumerate(ends) if idx + 1 <= e)
            slices[k].add_module(str(idx + 1), nn.ReLU(inplace=True))
            idx, cin = idx + 2, v
        self.slice1, self.slice2, self.slice3, self.slice4, self.slice5 = slices
        for p in self.parameters():
            p.requires_grad = False
        if features_state_dict is not None:
            self.load_pretrained(features_state_dict)

    def load_pretrained(self, sd):
        """accepts `features.<i>.weight`, `<i>.weight` or `slice<k>.<i>.weight` keys"""
        own = self.state_dict()
        by_idx = {k.split(".")[1] + "." + k.split(".")[2]: k for k in own}
        new = {}
        for k, v in sd.items():
            parts = k.split(".")
            if parts[0] == "classifier":
                continue
            if parts[0] == "vgg":
                parts = parts[1:]
            if parts[0] == "features" or parts[0].startswith("slice"):
                parts = parts[1:]
            key = ".".join(parts[:2])
            if key in by_idx:
                new[by_idx[key]] = v
        missing = set(own) - set(new)
        if missing:
            raise KeyError(f"VGG19 weights missing for {sorted(missing)[:4]} ...")
        self.load_state_dict(new)

    def conv_params(self):
        ws, bs = [], []
        for sl in (self.slice1, self.slice2, self.slice3, self.slice4, self.slice5):
            for m in sl:
                if isinstance(m, nn.Conv2d):
                    ws.append(m.weight)
                    bs.append(m.bias)
        return ws, bs


class VGGLoss(nn.Module):
    """criterionVGG of the reference (networks.py:18-30): sum_i w_i L1(vgg(x)_i, vgg(y)_i.detach()), w = 1/32 ... 1.
    `VGGLoss(device)` as at train.py:193; `features_state_dict` supplies the pretrained weights (see Vgg19)."""

    def __init__(self, device=None, features_state_dict=None):
        super().__init__()
        self.vgg = Vgg19(features_state_dict=features_state_dict)
        self.criterion = nn.L1Loss()
        self.weights = list(E.VGG_WEIGHTS)
        self._pack = None
        if device is not None:
            self.to(device)

    def _packed(self):
        ws, bs = self.vgg.conv_params()
        L.require_cuda(ws[0], "VGG19 weights")
        dt = E._state["dtype"]
        key = (dt, E._state["impl"], tuple(w._version for w in ws), tuple(w.data_ptr() for w in ws))
        if self._pack is None or self._pack.key != key:
            self._pack = E.VGGPack(ws, dt)            # frozen banks: packed once, reused by every call
        return self._pack, [b.detach() for b in bs]

    @torch.no_grad()
    def target_features(self, y):
        """The five tapped activations of the (detached) target, for `forward(x, target_taps=...)`: they do not depend
        on the generator, so a training loop may compute them before -- or beside -- the generator's forward pass."""
        L.require_cuda(y, "VGGLoss target")
        pack, bs = self._packed()
        return E.vgg_stack(pack, bs, y.detach().float().contiguous(), True)[1]

    def forward(self, x, y=None, target_taps=None):
        if (y is None) == (target_taps is None):
            raise ValueError("VGGLoss needs the target image or its precomputed activations (exactly one of them)")
        pack, bs = self._packed()
        return E.VGGLossFunction.apply(x, y, pack, target_taps, *bs)
