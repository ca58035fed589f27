#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/probe/run_probe.py > gpurun_out/probe.log 2>&1; echo "[probe] rc=$?"; cat gpurun_out/probe.log | tail -40
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:bn_act_bwd_reduce -s 25 -c 25 -o gpurun_out/prof_bnbwd $CMD > gpurun_out/ncu_a.log 2>&1; echo "[ncu bnbwd] rc=$?"
ncu --set full --clock-control none --import-source on -k regex:conv_wgrad_tc -s 54 -c 54 -o gpurun_out/prof_wgrad $CMD > gpurun_out/ncu_b.log 2>&1; echo "[ncu wgrad] rc=$?"
ncu --set full --clock-control none --import-source on -k "regex:conv_tc_kernel<128" -s 86 -c 40 -o gpurun_out/prof_conv128 $CMD > gpurun_out/ncu_c.log 2>&1; echo "[ncu conv128] rc=$?"
