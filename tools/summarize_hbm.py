"""Summarise an ncu --csv metrics pass (gpu__time_duration, dram__bytes_read/write, lts__t_bytes) per kernel and grid:
achieved DRAM and L2 bandwidth against the measured copy bandwidth of MEASURED_PEAKS.json.
    python tools/summarize_hbm.py gpurun_out/r02n_hbm.csv > profiles/r02n_hbm_kernels.md"""
import collections, csv, json, os, re, sys

path = sys.argv[1]
rows = list(csv.reader(open(path)))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
kn, gs, mn, mu, mv = (hdr.index(c) for c in ("Kernel Name", "Grid Size", "Metric Name", "Metric Unit", "Metric Value"))
idc = hdr.index("ID")
per = collections.OrderedDict()
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "%": 1, "usecond": 1, "nsecond": 1e-3}
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    key = r[idc]
    d = per.setdefault(key, {"name": re.sub(r"\(.*", "", r[kn]).replace("void ", "").replace("p2p::", ""), "grid": r[gs]})
    try:
        d[r[mn]] = float(r[mv].replace(",", "")) * scale.get(r[mu], 1)
    except ValueError:
        pass
agg = collections.OrderedDict()
for d in per.values():
    if "gpu__time_duration.sum" not in d:
        continue
    a = agg.setdefault((d["name"], d["grid"]), [0, 0.0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += d.get("dram__bytes_read.sum", 0.0)
    a[2] += d.get("dram__bytes_write.sum", 0.0)
    a[3] += d["gpu__time_duration.sum"]
    a[4] += d.get("lts__t_bytes.sum", 0.0)
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"]
print(f"# Bandwidth-bound kernels of one training iteration (B=16, 256x256, bf16), ncu metrics pass: {path}\n")
print("`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none`")
print("on `bench.py --steps 1 --warmup 1 --no-graph` (kernel by kernel, one stream; cold-cache, serialised launches).  DRAM")
print(f"GB/s = (read + write bytes) / duration against the measured copy bandwidth {peak:.0f} GB/s (MEASURED_PEAKS.json); written")
print("data that is still in the 126 MB L2 when the kernel ends does not show up as DRAM writes, so the L2 column (all")
print("bytes through the L2, reads and writes) is the better measure of what the kernel moves.  Rows with >= 1 MB moved.\n")
print("| kernel | grid | launches | avg us | DRAM read MB | DRAM write MB | DRAM GB/s | of peak | L2 MB | L2 GB/s |")
print("|---|---|---|---|---|---|---|---|---|---|")
tot_t = 0.0
for (name, grid), a in sorted(agg.items(), key=lambda kv: -kv[1][3]):
    n = a[0]
    rd, wr, t, l2 = a[1] / n, a[2] / n, a[3] / n, a[4] / n
    tot_t += a[3]
    if rd + wr + l2 < 1e6:
        continue
    gbs = (rd + wr) / t / 1e3
    print(f"| `{name}` | {grid} | {n} | {t:.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f} | {gbs:.0f} | {gbs / peak * 100:.0f} % | "
          f"{l2 / 1e6:.1f} | {l2 / t / 1e3:.0f} |")
print(f"\n{sum(a[0] for a in agg.values())} launches captured (the first 400 matching kernels of the run).")
