#!/usr/bin/env bash
# Multi-GPU check (run with `gpurun --gpus N`): native multi-process tests, a short sweep of
# every collective against NCCL, and bench.py for both arms at N = all visible GPUs.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-n$N}
nvidia-smi topo -m > gpurun_out/topo_$TAG.txt 2>&1
echo "== pytest multi ($N GPUs)"
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi_$TAG.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/pytest_multi_$TAG.log
echo "== sweep all_reduce variants"
rm -f gpurun_out/sweep_$TAG.jsonl
timeout 600 python benchmarks/sweep.py --gpus $N --colls all_reduce --algos all --graph --min-bytes 8 --max-bytes 1G --factor 8 --out gpurun_out/sweep_$TAG.jsonl 2> gpurun_out/sweep_$TAG.err | tail -60; tail -5 gpurun_out/sweep_$TAG.err
echo "== sweep other collectives"
timeout 600 python benchmarks/sweep.py --gpus $N --colls reduce,broadcast,scatter,gather,all_gather --algos all --graph --min-bytes 64 --max-bytes 256M --factor 16 --out gpurun_out/sweep_$TAG.jsonl 2>> gpurun_out/sweep_$TAG.err | tail -80; tail -5 gpurun_out/sweep_$TAG.err
echo "== bench pdcc N=$N"
timeout 600 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "bench exit $?"; cat gpurun_out/bench_pdcc_$TAG.json; grep -v Warning gpurun_out/bench_pdcc_$TAG.err | tail -5
echo "== bench reference N=$N"
timeout 600 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "bench exit $?"; cat gpurun_out/bench_ref_$TAG.json; grep -v -i warning gpurun_out/bench_ref_$TAG.err | tail -5
