#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "[bench] rc=$? $(cut -c1-160 gpurun_out/bench_default.json)"
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_reference.json 2>/dev/null; echo "[ref] rc=$? $(cut -c1-160 gpurun_out/bench_reference.json)"
python bench.py --batch 64 --steps 5 --no-cpu-baseline > gpurun_out/bench_b64.json 2>/dev/null; echo "[b64] rc=$? $(cut -c1-160 gpurun_out/bench_b64.json)"
CMD="python tools/bench_convs.py fwd g.res"
$CMD > gpurun_out/plain_a.log 2>&1 || { echo plain failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 4 -c 2 -o gpurun_out/prof_res_fwd $CMD > gpurun_out/ncu_a.log 2>&1; echo "[ncu res fwd] rc=$?"
CMD2="python tools/bench_convs.py wgrad g.res"
$CMD2 > gpurun_out/plain_b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:conv_wgrad_halo -s 4 -c 2 -o gpurun_out/prof_res_wgrad $CMD2 > gpurun_out/ncu_b.log 2>&1; echo "[ncu res wgrad] rc=$?"
ls -la gpurun_out/*.ncu-rep
