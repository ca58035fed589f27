#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for mode in "" "P2P_FORCE_HALO=1" "P2P_NO_HALO=1"; do
  env $mode timeout 300 python -m pytest tests/test_gpu_ops.py -q -x -k "tcgen05" > gpurun_out/t_$mode.log 2>&1
  echo "[ops tcgen05 $mode] rc=$? $(tail -n 1 gpurun_out/t_$mode.log | cut -c1-150)"
done
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/all.log 2>&1; echo "[all] rc=$? $(tail -n 1 gpurun_out/all.log)"
timeout 300 python tools/bench_convs.py > gpurun_out/convs.log 2>&1; echo "[convs] rc=$?"; cat gpurun_out/convs.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "[bench] rc=$? $(cut -c1-220 gpurun_out/bench.json)"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "[smoke] rc=$?"; tail -n 4 gpurun_out/smoke.log
