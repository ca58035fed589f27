#!/usr/bin/env bash
# BASELINE.json configs 3 and 4 on one 8-GPU box: global batch 128 on 1 / 2 / 8 GPUs (strong scaling), 512x512 global batch 32
set -u
mkdir -p gpurun_out
PORT=29600
trun() {  # name N args...
  local name=$1 N=$2; shift 2
  PORT=$((PORT + 7))
  local t0=$(date +%s)
  if [ $N -eq 1 ]; then timeout 240 python bench.py --gpus 1 --no-cpu-baseline "$@" 2>gpurun_out/cfg_err.txt > gpurun_out/$name.json
  else timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --no-cpu-baseline "$@" 2>gpurun_out/cfg_err.txt > gpurun_out/$name.json; fi
  local rc=$?
  python -c "import json; l=open('gpurun_out/$name.json').readline(); d=json.loads(l) if l.strip() else None; print('$name', (round(d['value'],1), 'img/s', round(d['ms_per_step'],3), 'ms/step; e2e', round(d['e2e']['value'],1)) if d else 'NO OUTPUT', 'rc=$rc wall', $(date +%s)-$t0, 's')"
  [ $rc -ne 0 ] && tail -3 gpurun_out/cfg_err.txt | cut -c1-300
}
trun r02h_cfg3_gb128_N1 1 --batch 128 --steps 10 --warmup 4
trun r02h_cfg3_gb128_N2 2 --batch 64 --steps 15 --warmup 4
trun r02h_cfg3_gb128_N8 8 --batch 16 --steps 30 --warmup 5
trun r02h_cfg4_512_gb32_N8 8 --size 512 --batch 4 --steps 30 --warmup 5
