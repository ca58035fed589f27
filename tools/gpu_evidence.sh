#!/usr/bin/env bash
# Round-1 evidence run: default bench (both arms), per-layer conv table, ncu launch list, one ncu --set full
# capture of the dominant kernel.  Everything lands in gpurun_out/ and is copied into profiles/ by hand.
set +e
TAG=${TAG:-r01e}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "[bench] rc=$? $(cut -c1-160 gpurun_out/${TAG}_bench.json)"
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_ref.err; echo "[reference] rc=$? $(cut -c1-160 gpurun_out/${TAG}_bench_reference.json)"
python bench.py --no-graph --no-cpu-baseline > gpurun_out/${TAG}_bench_nograph.json 2>/dev/null; echo "[no-graph] rc=$? $(cut -c1-160 gpurun_out/${TAG}_bench_nograph.json)"
python bench.py --batch 64 --steps 10 --no-cpu-baseline > gpurun_out/${TAG}_bench_b64.json 2>/dev/null; echo "[b64] rc=$? $(cut -c1-160 gpurun_out/${TAG}_bench_b64.json)"
python tools/bench_convs.py > gpurun_out/${TAG}_convs.txt 2>&1; echo "[convs] rc=$?"
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-graph"
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "[launch list] rc=$? lines=$(wc -l < gpurun_out/${TAG}_launches.csv)"
python tools/summarize_launches.py gpurun_out/${TAG}_launches.csv 8 > gpurun_out/${TAG}_launches.md 2>/dev/null
python tools/bench_convs.py fwd g.res > /dev/null 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:conv_halo_swap -s 3 -c 1 -f -o gpurun_out/${TAG}_res_fwd python tools/bench_convs.py fwd g.res > gpurun_out/ncu_full.log 2>&1
echo "[ncu full] rc=$? $(ls -la gpurun_out/${TAG}_res_fwd.ncu-rep 2>/dev/null | awk '{print $5}') bytes"
