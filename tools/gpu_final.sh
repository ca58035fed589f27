#!/usr/bin/env bash
# Round-end check on one GPU: full GPU test suite, smoke(), the benchmark's arms
set +e
cd "$(dirname "$0")/.."
TAG=${1:-r02z}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/${TAG}_pytest.log; tail -2 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^initialize" | tail -3 | tee gpurun_out/${TAG}_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; cut -c1-400 gpurun_out/${TAG}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err; cut -c1-200 gpurun_out/${TAG}_bench_reference.json
python bench.py --mode infer --batch 64 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_infer_b64.json 2>> gpurun_out/${TAG}_bench.err; cut -c1-250 gpurun_out/${TAG}_bench_infer_b64.json
python bench.py --with-vgg --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_with_vgg.json 2>> gpurun_out/${TAG}_bench.err; cut -c1-250 gpurun_out/${TAG}_bench_with_vgg.json
python bench.py --no-graph --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_nograph.json 2>> gpurun_out/${TAG}_bench.err; cut -c1-250 gpurun_out/${TAG}_bench_nograph.json
