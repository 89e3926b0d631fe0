#!/usr/bin/env bash
# Short multi-GPU pass (run with `gpurun --gpus N`): native correctness sweep, the rooted
# collectives' variants, auto-plan spot checks against NCCL, bench.py for both arms.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-quick_n$N}
rm -f gpurun_out/$TAG.jsonl
echo "== pytest native sweep ($N GPUs)"
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q -k "sweep" > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_$TAG.log
echo "== rooted variants"
timeout 300 python benchmarks/sweep.py --gpus $N --colls reduce,broadcast --algos all --min-bytes 1M --max-bytes 256M --factor 16 --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#\|  ll  " | tail -40
echo "== auto plan vs nccl"
timeout 300 python benchmarks/sweep.py --gpus $N --colls all_reduce,all_gather,scatter,gather --algos auto --min-bytes 1M --max-bytes 1G --factor 32 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -40
echo "== bench pdcc N=$N"
timeout 300 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "bench exit $?"; cat gpurun_out/bench_pdcc_$TAG.json
echo "== bench reference N=$N"
timeout 300 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "bench exit $?"; grep '"metric"' gpurun_out/bench_ref_$TAG.json
tail -3 gpurun_out/$TAG.err
