#!/usr/bin/env bash
# Evidence set: launch list of the benchmark iteration, ncu --set full of the bandwidth-bound kernels and of the dominant conv
set +e
cd "$(dirname "$0")/.."
TAG=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-graph"
$CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "[launch list] rc=$? lines=$(wc -l < gpurun_out/${TAG}_launches.csv)"
python tools/summarize_launches.py gpurun_out/${TAG}_launches.csv 12 > gpurun_out/${TAG}_launches.md
head -30 gpurun_out/${TAG}_launches.md
CMD2="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph"
ncu --set full --clock-control none -k regex:"adam_kernel|bn_act_bwd_reduce|bn_bwd_apply|map_loss|bn_act_store_kernel|act_bwd_kernel|tv_fwd|unpack_wgrad_multi|pack_weights_multi" -s 60 -c 48 -o gpurun_out/${TAG}_hbm $CMD2 > gpurun_out/ncu_hbm.log 2>&1
echo "[hbm kernels] rc=$?"
python tools/bench_convs.py fwd g.res > gpurun_out/plain2.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:conv_halo_swap -s 3 -c 1 -o gpurun_out/${TAG}_res_fwd python tools/bench_convs.py fwd g.res > gpurun_out/ncu_res.log 2>&1
echo "[res conv] rc=$?"
ls -la gpurun_out/*.ncu-rep
