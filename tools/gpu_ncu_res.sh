#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python tools/bench_convs.py fwd g.res"
$CMD > gpurun_out/plain_a.log 2>&1 || { echo plain failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:conv_halo -s 4 -c 2 -o gpurun_out/prof_halo_res $CMD > gpurun_out/ncu_a.log 2>&1; echo "[ncu halo] rc=$?"
P2P_NO_HALO=1 $CMD > gpurun_out/plain_b.log 2>&1 || { echo plain failed; exit 1; }
P2P_NO_HALO=1 ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 4 -c 2 -o gpurun_out/prof_tap_res $CMD > gpurun_out/ncu_b.log 2>&1; echo "[ncu tap] rc=$?"
ls -la gpurun_out/*.ncu-rep
