#!/usr/bin/env bash
# First-contact script for a 1-GPU box: capabilities, smoke, GPU tests, 1-GPU bench (both arms).
# Everything is wrapped in `timeout`; outputs land in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
nvidia-smi topo -m >> gpurun_out/gpus.txt 2>&1
timeout 300 python - > gpurun_out/caps.json 2> gpurun_out/caps.err <<'PY'
import json
from pdcc_b200 import _native as N
print(json.dumps(N.probe(0), indent=1))
PY
echo "== caps"; cat gpurun_out/caps.json; tail -3 gpurun_out/caps.err
echo "== smoke"
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/smoke.log
echo "== pytest -m gpu"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/pytest_gpu.log
echo "== bench pdcc N=1"
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_pdcc_1.json 2> gpurun_out/bench_pdcc_1.err; echo "bench exit $?"; cat gpurun_out/bench_pdcc_1.json; tail -5 gpurun_out/bench_pdcc_1.err
echo "== bench reference N=1"
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_1.json 2> gpurun_out/bench_ref_1.err; echo "bench exit $?"; cat gpurun_out/bench_ref_1.json; tail -5 gpurun_out/bench_ref_1.err
