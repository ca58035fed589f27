"""Recipe for oracle/_ref/: the reference's own module, unmodified, next to the oracle (test infrastructure).

The reference (Dev-Vault-Archived/p2p-pytorch) is pure Python; the hot path lives in its `networks.py`, which imports
with torch + torchvision + numpy alone.  `/root/reference` exists only in the build container, so this script copies
that ONE file byte for byte to `oracle/_ref/networks.py` (git-ignored: reference sources never enter the history; not
gpurun-ignored: the copy travels to the GPU box like a built .so).  It is used by
  * `bench.py --impl reference` and bench.py's `cpu_baseline` leg: the reference's define_G / define_D / GANLoss run
    the loop body of train.py:294-390 on the host cores at the benchmark's batch size (`kind: "reference"`);
  * nothing else: parity tests use the oracle and the committed golden vectors.
`train.py` itself cannot be imported (numpy.core.numeric.Inf at train.py:7, skimage at train.py:19, a VGG19 download at
networks.py:35 / train.py:193), so its loop body is re-typed around the real modules in bench.py, VGG term off.
Usage: python oracle/build_ref.py        (called by __graft_entry__.build() when /root/reference is present)
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/networks.py"
DST_DIR = os.path.join(HERE, "_ref")
DST = os.path.join(DST_DIR, "networks.py")


def build_ref() -> bool:
    """copy the file if the reference tree is present; True if oracle/_ref/networks.py exists afterwards"""
    if os.path.isfile(SRC):
        os.makedirs(DST_DIR, exist_ok=True)
        shutil.copyfile(SRC, DST)
        with open(SRC, "rb") as f:
            digest = hashlib.sha256(f.read()).hexdigest()
        with open(os.path.join(DST_DIR, "SOURCE.txt"), "w") as f:
            f.write(f"{SRC} sha256 {digest}\n")
    return os.path.isfile(DST)


def load_reference_networks():
    """import oracle/_ref/networks.py as a module (None when it has not been built)"""
    if not os.path.isfile(DST):
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location("p2p_reference_networks", DST)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    ok = build_ref()
    print("oracle/_ref/networks.py", "present" if ok else "absent (no /root/reference here)")
    sys.exit(0 if ok else 1)
