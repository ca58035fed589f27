"""Per-kernel SASS opcode summary of libp2p_b200.so: how many tcgen05 MMA (UTCHMMA), TMEM load (LDTM), TMA load
(UTMALDG), tensor-core barrier (UTCBAR) and related instructions each kernel instantiation contains.
    python tools/sass_summary.py > profiles/rNN_sass_opcodes.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "p2p-pytorch_b200", "csrc", "libp2p_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOMSWS", "SYNCS", "RED", "ATOM", "HMMA"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            for o in OPS:
                if op.startswith(o):
                    counts[cur][o] += 1
    names = subprocess.run(["cu++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
    print("# SASS opcode counts per kernel (cuobjdump -sass p2p-pytorch_b200/csrc/libp2p_b200.so, sm_100a)\n")
    print("UTCHMMA = tcgen05.mma (kind::f16), LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA load), "
          "UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, RED/ATOM = global reductions.\n")
    print("| kernel | instructions | " + " | ".join(OPS) + " |")
    print("|---|---|" + "---|" * len(OPS))
    tot = collections.Counter()
    for mangled, name in zip(counts, names):
        c = counts[mangled]
        name = name.replace("void ", "").replace("(int)", "").replace("(bool)", "")
        name = name.rsplit(">(", 1)[0] + ">" if ">(" in name else name.split("(")[0]
        print(f"| `{name}` | {c['_total']} | " + " | ".join(str(c[o]) if c[o] else "" for o in OPS) + " |")
        tot.update(c)
    print(f"| **total** | {tot['_total']} | " + " | ".join(str(tot[o]) for o in OPS) + " |")


if __name__ == "__main__":
    main()
