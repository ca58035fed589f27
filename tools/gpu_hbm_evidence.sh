#!/usr/bin/env bash
# ncu DRAM bytes + duration of the bandwidth-bound kernels of one benchmark iteration (kernel-by-kernel launch, one lane)
set +e
cd "$(dirname "$0")/.."
TAG=${1:-r02n}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph"
P2P_LANES=1 $CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
P2P_LANES=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:"adam_kernel|bn_act_bwd_reduce|bn_bwd_apply|bn_act_store|map_loss_fwd|map_loss_bwd|tv_fwd|tv_bwd|decode_u8|act_bwd_kernel|channel_stats|maxpool" \
    -c 400 --csv --log-file gpurun_out/${TAG}_hbm.csv $CMD > gpurun_out/ncu_hbm.log 2>&1
echo "[hbm kernels] rc=$? lines=$(wc -l < gpurun_out/${TAG}_hbm.csv)"
