#!/usr/bin/env bash
# Full multi-GPU pass (run with `gpurun --gpus N`, N = 4 or 8): correctness on real GPUs,
# variant sweeps against NCCL, bench.py for both arms. Outputs land in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-full_n$N}
rm -f gpurun_out/$TAG.jsonl
nvidia-smi topo -m > gpurun_out/topo_$TAG.txt 2>&1
echo "== pytest multi ($N GPUs)"
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "sweep or contract or do_scatter" > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/pytest_$TAG.log
echo "== large all_reduce: variants x CTAs"
timeout 600 python benchmarks/sweep.py --gpus $N --colls all_reduce --algos all --nblocks 64,128 --min-bytes 16M --max-bytes 1G --factor 8 --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | tail -60
echo "== mid range all_reduce"
timeout 600 python benchmarks/sweep.py --gpus $N --colls all_reduce --algos all --graph --min-bytes 8 --max-bytes 8M --factor 8 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -70
echo "== other collectives, large"
timeout 600 python benchmarks/sweep.py --gpus $N --colls reduce,broadcast,all_gather,scatter,gather --algos all --nblocks 64,128 --min-bytes 256M --max-bytes 256M --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#\|  ll  " | tail -60
echo "== other collectives, small/mid"
timeout 600 python benchmarks/sweep.py --gpus $N --colls reduce,broadcast,all_gather,scatter,gather --algos all --graph --min-bytes 64 --max-bytes 4M --factor 16 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -100
echo "== bench pdcc N=$N"
timeout 600 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "bench exit $?"; cat gpurun_out/bench_pdcc_$TAG.json
echo "== bench reference N=$N"
NCCL_DEBUG=INFO timeout 600 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "bench exit $?"; grep '"metric"' gpurun_out/bench_ref_$TAG.json
grep -i -m5 "nvls" gpurun_out/bench_ref_$TAG.err | cut -c1-200
grep '"metric"' gpurun_out/bench_ref_$TAG.json > gpurun_out/bench_ref_$TAG.line.json
tail -3 gpurun_out/$TAG.err
