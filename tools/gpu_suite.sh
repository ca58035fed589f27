#!/usr/bin/env bash
# Runs on the GPU box (under gpurun): staged test suite with per-stage logs in gpurun_out/.
# Each tcgen05 case runs in its own process so that a faulting kernel cannot poison the other results.
set +e
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
stage() {  # name, timeout, command...
  local name=$1 to=$2; shift 2
  local t0=$(date +%s)
  timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s :: $(tail -n 1 gpurun_out/$name.log | cut -c1-200)"
}
stage ops_misc 600 python -m pytest tests/test_gpu_ops.py -q --maxfail=40 -k "not conv"
stage ops_conv_simt 600 python -m pytest tests/test_gpu_ops.py -q --maxfail=40 -k "conv_simt"
for c in g_res g_conv2 g_conv1 g_deconv1 d_l0 d_l1_odd d_l3 d_l4 rect_odd; do
  stage "tc_$c" 180 python -m pytest tests/test_gpu_ops.py -q -x -k "test_conv_tcgen05_bf16 and $c"
done
stage tc_taken 180 python -m pytest tests/test_gpu_ops.py -q -x -k "path_is_taken"
stage dbg_fp32 300 python tests/diagnostics/debug_g.py fp32
stage dbg_bf16 300 python tests/diagnostics/debug_g.py bf16
stage dbg_bf16simt 300 python tests/diagnostics/debug_g.py bf16-simt
stage net_fp32 900 python -m pytest tests/test_gpu_networks.py -q --maxfail=40 -k "fp32 or drop_in or state_dict"
stage net_bf16simt 600 python -m pytest tests/test_gpu_networks.py -q --maxfail=40 -k "bf16-simt"
stage net_bf16 600 python -m pytest tests/test_gpu_networks.py -q --maxfail=40 -k "bf16 and not simt"
stage curves 900 python -m pytest tests/test_gpu_networks.py -q -k "200_steps"
stage smoke 300 python -c "import __graft_entry__ as g; g.smoke()"
stage bench_fp32 600 python bench.py --steps 2 --warmup 1 --precision fp32 --no-cpu-baseline
stage bench_bf16 600 python bench.py --steps 3 --warmup 2
