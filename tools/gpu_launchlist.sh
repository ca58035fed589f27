#!/usr/bin/env bash
# every launch of the benchmark iteration with its device time (kernel by kernel, one stream): ncu launch list + summary
set +e
cd "$(dirname "$0")/.."
TAG=${1:-r02z}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-graph"
$CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "[launch list] rc=$? lines=$(wc -l < gpurun_out/${TAG}_launches.csv)"
# bench.py runs warm-up 2 + (clock-sampler priming: a few) + timed 2 + end-to-end legs 2 x (2 + 2): count the Adam launches
python - <<PY
import csv
rows = list(csv.reader(open("gpurun_out/${TAG}_launches.csv")))
n = sum(1 for r in rows if any("adam_kernel" in c for c in r))
print("adam launches", n, "-> iterations", n // 2)
open("gpurun_out/${TAG}_iters.txt", "w").write(str(max(1, n // 2)))
PY
python tools/summarize_launches.py gpurun_out/${TAG}_launches.csv $(cat gpurun_out/${TAG}_iters.txt) > gpurun_out/${TAG}_launches.md
head -24 gpurun_out/${TAG}_launches.md
