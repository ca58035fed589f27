"""Generate tests/golden/ref_netc_*.npz from the UNMODIFIED reference CompressionNetwork (build container only).

`/root/reference/networks.py:define_C()` (CompressionNetwork, :201-236) is loaded with the weights that
`oracle.pix2pix_oracle.init_c_state` generates and run in train mode on seeded inputs; stored are its output,
the generator input `compress(net_c(real_b), 3)` of train.py:297 (compress is re-typed from
generate_dataset.py:29-34 -- that file imports skimage and cannot be imported here), and the BatchNorm buffers
after the call.  Usage:  python oracle/make_golden_c.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")

import networks as ref  # noqa: E402  (the reference, unmodified)
from oracle import pix2pix_oracle as orc  # noqa: E402
from oracle.make_golden_util import C_CASES  # noqa: E402


def ref_compress(tensor, bit):                     # generate_dataset.py:29-34, verbatim arithmetic
    max_val = 2 ** bit - 1
    tensor = torch.clamp(tensor, 0.0, 1.0) * max_val
    tensor = torch.round(tensor)
    return tensor / max_val


def main():
    out_dir = os.path.join(os.path.dirname(HERE), "tests", "golden")
    for name, batch, h, w, seed, training in C_CASES:
        torch.manual_seed(0)
        net_c = ref.define_C("normal", 0.02, gpu_id="cpu")
        net_c.load_state_dict(orc.init_c_state(125))
        net_c.train(training)
        _a, real_b, _c = orc.synthetic_batch(batch, h, w, seed)
        with torch.no_grad():
            y = net_c(real_b)
            comp = ref_compress(y, 3)
        sd = net_c.state_dict()
        np.savez_compressed(os.path.join(out_dir, f"ref_netc_{name}.npz"), y=y.numpy(), compressed=comp.numpy(),
                            running_mean=sd["conv_block1.1.running_mean"].numpy(),
                            running_var=sd["conv_block1.1.running_var"].numpy(),
                            num_batches_tracked=sd["conv_block1.1.num_batches_tracked"].numpy())
        print(name, y.shape, float(y.abs().max()), float(comp.mean()))


if __name__ == "__main__":
    main()
