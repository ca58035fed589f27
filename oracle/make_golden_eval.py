"""Generate tests/golden/ref_psnr.npz by running the reference's OWN `tensor2img` and `psnr` (train.py:32-42,
:60-65) -- their source is cut out of /root/reference/train.py with `ast` and executed unmodified, because the
module itself cannot be imported (numpy.core.numeric.Inf, skimage).  Build container only.
Usage:  python oracle/make_golden_eval.py
"""
import ast
import os
import sys

import numpy as np
import torch
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle.make_golden_util import PSNR_CASES, psnr_inputs  # noqa: E402

src = open("/root/reference/train.py").read()
tree = ast.parse(src)
ns = {"np": np, "torch": torch, "Image": Image}
for node in tree.body:
    if isinstance(node, ast.FunctionDef) and node.name in ("tensor2img", "psnr"):
        exec(compile(ast.Module(body=[node], type_ignores=[]), "/root/reference/train.py", "exec"), ns)

out = {}
for name, h, w, seed, kind in PSNR_CASES:
    a, b = psnr_inputs(h, w, seed, kind)
    with np.errstate(divide="ignore"):
        out[name] = np.float64(ns["psnr"](a, b))
    print(name, out[name])
np.savez(os.path.join(os.path.dirname(HERE), "tests", "golden", "ref_psnr.npz"), **out)
