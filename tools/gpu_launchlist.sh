#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-graph"
$CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "[launch list] rc=$? lines=$(wc -l < gpurun_out/launches.csv)"
python tools/summarize_launches.py gpurun_out/launches.csv 8 | head -45
