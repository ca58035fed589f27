"""p2p_b200: the pix2pix G/D training-step hot path of Dev-Vault-Archived/p2p-pytorch on hand-written
sm_100a kernels, behind the reference's own `networks.py` API.  Import as `p2p_pytorch_b200`."""
from . import _lib
from .engine import get_precision, set_precision
from .networks import (CompressionNetwork, GANLoss, MultiscaleDiscriminator, ExpandNetwork, VGGLoss, Vgg19, define_C,
                       define_D, define_G, init_net, init_weights)
from .losses import feature_matching_loss, lsgan_loss, tv_loss
from .optim import FusedAdam
from .trainer import Pix2PixTrainer
from .evaluation import generate, load_checkpoint, psnr, save_checkpoint

_lib.load()   # fail at import time if the CUDA library has not been built: there is no fallback path

__all__ = ["generate", "VGGLoss", "Vgg19", "psnr", "save_checkpoint", "load_checkpoint", "define_G", "define_D", "define_C", "CompressionNetwork", "GANLoss", "init_net", "init_weights", "ExpandNetwork", "MultiscaleDiscriminator",
           "FusedAdam", "Pix2PixTrainer", "lsgan_loss", "feature_matching_loss", "tv_loss", "set_precision",
           "get_precision"]
