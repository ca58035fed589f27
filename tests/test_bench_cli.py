"""bench.py's reference arm runs on the host cores (no GPU needed): one tiny step through it must print the one JSON
line the driver parses, with the keys of the contract; and the recipe that vendors the reference module is idempotent."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                         timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines                      # ONE line on stdout, everything else goes to stderr
    return json.loads(lines[0])


def test_reference_arm_prints_the_contract_line():
    d = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--batch", "1", "--size", "32")
    assert d["impl"] == "reference" and d["metric"] == "pix2pix 32x32 train images/sec" and d["unit"] == "images/s"
    assert d["higher_is_better"] is True and d["steps"] == 1 and d["n_gpus"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["batch_per_gpu"] == 1 and "workload" in d["config"]


def test_reference_arm_inference_mode():
    d = _run("--impl", "reference", "--mode", "infer", "--steps", "1", "--warmup", "0", "--batch", "2", "--size", "32")
    assert d["metric"] == "pix2pix 32x32 generator inference images/sec" and d["value"] > 0


def test_reference_arm_is_silent_on_other_ranks():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                          "--warmup", "0", "--batch", "1", "--size", "32"], capture_output=True, text=True, timeout=300,
                         cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_build_ref_is_idempotent():
    from oracle.build_ref import DST, build_ref
    had = os.path.isfile(DST)
    assert build_ref() == (had or os.path.isfile("/root/reference/networks.py"))
    if os.path.isfile(DST) and os.path.isfile("/root/reference/networks.py"):
        assert open(DST, "rb").read() == open("/root/reference/networks.py", "rb").read()     # byte for byte
