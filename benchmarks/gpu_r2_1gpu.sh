#!/usr/bin/env bash
# Round-2 single-GPU pass: GPU tests (emulation + world-1 + no-alloc), smoke, then
# optionally the 1-GPU bench. Everything under `timeout`; logs land in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r2}
echo "== pytest -m gpu"
timeout 900 python -m pytest tests -m gpu -x -q --timeout=300 > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu_$TAG.log
echo "== smoke"
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke_$TAG.log
if [ "${2:-}" = "bench" ]; then
  echo "== bench pdcc N=1"
  timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_pdcc_1_$TAG.json 2> gpurun_out/bench_pdcc_1_$TAG.err; echo "bench exit $?"; cat gpurun_out/bench_pdcc_1_$TAG.json; tail -5 gpurun_out/bench_pdcc_1_$TAG.err
fi
