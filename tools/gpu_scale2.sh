#!/usr/bin/env bash
# data-parallel runs at N GPUs of one box (each run bounded, own rendezvous port); reports whether the job EXITS
#   tools/gpu_scale2.sh "2 4 8" [bench.py args]
set -u
mkdir -p gpurun_out
PORT=29520
run() {
  local N=$1
  PORT=$((PORT + 7))
  local t0=$(date +%s)
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline "${@:2}" 2>gpurun_out/scale_err.txt > gpurun_out/scale_N$N.json
  local rc=$?
  python -c "import json,sys; l=open('gpurun_out/scale_N$N.json').readline(); d=json.loads(l) if l.strip() else None; print('N=$N', (round(d['value'],1), 'img/s', round(d['ms_per_step'],3), 'ms/step; e2e', round(d['e2e']['value'],1)) if d else 'NO OUTPUT', 'rc=$rc', 'wall', $(date +%s)-$t0, 's')"
  [ $rc -ne 0 ] && tail -3 gpurun_out/scale_err.txt | cut -c1-300
}
python bench.py --steps 30 --warmup 5 --no-cpu-baseline "${@:2}" 2>/dev/null > gpurun_out/scale_N1.json
python -c "import json; d=json.loads(open('gpurun_out/scale_N1.json').readline()); print('N=1', round(d['value'],1), 'img/s', round(d['ms_per_step'],3), 'ms/step; e2e', round(d['e2e']['value'],1))"
for n in $1; do run $n "${@:2}"; done
