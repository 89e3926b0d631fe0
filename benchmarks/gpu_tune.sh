#!/usr/bin/env bash
# Tuning sweep (run with `gpurun --gpus N`): correctness first, then the large-message
# variants at several CTA counts, then the small/medium range with graph-timed latency.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-tune_n$N}
rm -f gpurun_out/$TAG.jsonl
echo "== pytest gpu (emulation + multi)"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/pytest_$TAG.log
echo "== large all_reduce: variants x CTAs"
timeout 900 python benchmarks/sweep.py --gpus $N --colls all_reduce --algos all --nblocks 32,64,128 --min-bytes 16M --max-bytes 1G --factor 8 --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | tail -80
echo "== large rooted/movers: CTAs"
timeout 900 python benchmarks/sweep.py --gpus $N --colls reduce,broadcast,all_gather,scatter,gather --algos all --nblocks 32,64,128 --min-bytes 256M --max-bytes 256M --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#\|  ll  " | tail -80
echo "== mid range all_reduce"
timeout 900 python benchmarks/sweep.py --gpus $N --colls all_reduce --algos all --graph --min-bytes 64K --max-bytes 8M --factor 4 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -60
tail -5 gpurun_out/$TAG.err
