#!/usr/bin/env bash
# Final multi-GPU pass (run with `gpurun --gpus N`): native correctness sweep, an A/B of the
# tile split on the large messages, bench.py for both arms.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-final_n$N}
rm -f gpurun_out/$TAG.jsonl
echo "== pytest native sweep ($N GPUs)"
timeout 200 python -m pytest tests/test_gpu_multi.py -x -q -k "sweep" --timeout=150 --timeout-method=thread > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_$TAG.log
for SP in 0 1 4; do
echo "== TILE_SPLIT=$SP"
PDCC_TILE_SPLIT=$SP timeout 200 python benchmarks/sweep.py --gpus $N --impls pdcc --colls all_reduce,reduce,broadcast --algos auto --min-bytes 32M --max-bytes 256M --factor 8 --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | tail -8
done
echo "== bench pdcc N=$N"
timeout 300 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "bench exit $?"; cat gpurun_out/bench_pdcc_$TAG.json
echo "== bench reference N=$N"
timeout 300 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "bench exit $?"; grep '"metric"' gpurun_out/bench_ref_$TAG.json
tail -3 gpurun_out/$TAG.err
