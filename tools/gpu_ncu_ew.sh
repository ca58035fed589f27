#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:bn_act_bwd_reduce -s 33 -c 4 -o gpurun_out/prof_bnbwd $CMD > gpurun_out/ncu_a.log 2>&1; echo "[ncu bnbwd] rc=$?"
ls -la gpurun_out/*.ncu-rep
