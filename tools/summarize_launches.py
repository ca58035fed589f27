"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: share of the step per kernel."""
import collections, csv, re, sys
path, nsteps = sys.argv[1], int(sys.argv[2])
rows = list(csv.reader(open(path)))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
kn, mv, gs = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hi + 1:]:
    try:
        t = float(r[mv].replace(",", ""))
    except Exception:
        continue
    name = re.sub(r"\(.*", "", r[kn]).replace("void ", "")
    agg[name][0] += 1
    agg[name][1] += t
tot = sum(v[1] for v in agg.values())
print(f"# {path}: {sum(v[0] for v in agg.values())} launches over {nsteps} steps, "
      f"{tot / 1e6 / nsteps:.3f} ms of kernel time per step (ncu: cold cache, serialised)")
print("| share | ms/step | launches/step | avg us | kernel |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if v[1] / tot < 0.0005:
        continue
    print(f"| {v[1] / tot * 100:.2f}% | {v[1] / 1e6 / nsteps:.3f} | {v[0] / nsteps:.1f} | {v[1] / v[0] / 1e3:.1f} | {k} |")
