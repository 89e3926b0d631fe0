#!/usr/bin/env bash
# Round-2 starting point (run with `gpurun --gpus N --timeout 400 -- bash benchmarks/gpu_round2.sh`).
# Lessons from round 1, baked in: one job per question (process start-up dominates at 8
# GPUs), A/B planner knobs INSIDE a job (--env-sweep), every step under `timeout`, pytest with
# --timeout so a hung kernel costs seconds, not minutes. Budget: ~2.5 min of box time.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-r2_n$N}
rm -f gpurun_out/$TAG.jsonl
echo "== native correctness ($N GPUs)"
timeout 150 python -m pytest tests/test_gpu_multi.py -x -q -k "sweep or data_parallel" --timeout=120 --timeout-method=thread > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_$TAG.log
echo "== headline step, both arms"
timeout 120 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "pdcc exit $?"; cut -c1-160 gpurun_out/bench_pdcc_$TAG.json
timeout 120 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref exit $?"; grep '"metric"' gpurun_out/bench_ref_$TAG.json | cut -c1-160
echo "== tiles per CTA A/B on the three collectives that trail NCCL (one job)"
timeout 150 python benchmarks/sweep.py --gpus $N --colls all_reduce,reduce,broadcast --algos auto --sizes 4M,32M,256M --env-sweep PDCC_TILE_SPLIT=0,2,4,8 --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | tail -50
tail -3 gpurun_out/$TAG.err
