#!/usr/bin/env bash
# Experiment: CTA count x staggered lookahead on the large-message collectives.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-exp_n$N}
rm -f gpurun_out/$TAG.jsonl
echo "== pytest 1-GPU emulation subset + native sweep"
timeout 300 python -m pytest tests/test_gpu_multi.py tests/test_gpu_emulation.py -x -q -k "sweep or stress or auto_multi" > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_$TAG.log
for ST in 1 0; do
echo "== STAGGER=$ST"
PDCC_STAGGER=$ST timeout 300 python benchmarks/sweep.py --gpus $N --impls pdcc --colls reduce,broadcast --algos nvls,twoshot_nvls --nblocks 128,256,296 --min-bytes 256M --max-bytes 256M --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | tail -40
PDCC_STAGGER=$ST timeout 300 python benchmarks/sweep.py --gpus $N --impls pdcc --colls all_reduce --algos twoshot_nvls,twoshot_p2p --nblocks 128,256,296 --min-bytes 32M --max-bytes 256M --factor 8 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -40
done
echo "== nccl reference points"
timeout 300 python benchmarks/sweep.py --gpus $N --impls nccl --colls reduce,broadcast,all_reduce --min-bytes 32M --max-bytes 256M --factor 8 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -12
tail -3 gpurun_out/$TAG.err
