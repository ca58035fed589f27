#!/usr/bin/env bash
# On the GPU box: plain bench run, then (only if it exited 0) the ncu launch list and one full capture
# of the dominant kernel.  Outputs land in gpurun_out/.
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline"
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
echo "[bench default] rc=$? $(cut -c1-400 gpurun_out/bench_default.json)"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "[launch list] rc=$? lines=$(wc -l < gpurun_out/launches.csv)"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 60 -c 4 -o gpurun_out/prof_conv_tc $CMD > gpurun_out/ncu_full.log 2>&1
echo "[full capture] rc=$? $(ls -la gpurun_out/*.ncu-rep 2>/dev/null | wc -l) reports"
