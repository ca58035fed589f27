#!/usr/bin/env bash
set +e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/probe/run_probe.py > gpurun_out/probe2.log 2>&1; echo "[probe] rc=$?"; grep -c OK gpurun_out/probe2.log; grep MISMATCH gpurun_out/probe2.log | head -20
for c in g_res g_conv1 g_deconv1 d_l3 d_l4 rect_odd g_conv2 d_l1_odd d_l0; do
  timeout 120 python -m pytest tests/test_gpu_ops.py -q -x -k "test_conv_tcgen05_bf16 and $c" > gpurun_out/halo_$c.log 2>&1
  echo "[halo $c] rc=$? $(tail -n 1 gpurun_out/halo_$c.log | cut -c1-150)"
done
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/all.log 2>&1; echo "[all] rc=$? $(tail -n 1 gpurun_out/all.log)"
P2P_NO_HALO=1 timeout 300 python tools/bench_convs.py > gpurun_out/convs_nohalo.log 2>&1; echo "[convs nohalo] rc=$?"
timeout 300 python tools/bench_convs.py > gpurun_out/convs_halo.log 2>&1; echo "[convs halo] rc=$?"
paste -d'\n' gpurun_out/convs_nohalo.log gpurun_out/convs_halo.log | cut -c1-200
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "[bench] rc=$? $(cut -c1-200 gpurun_out/bench.json)"
