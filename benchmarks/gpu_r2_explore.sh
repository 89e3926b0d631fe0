#!/usr/bin/env bash
# Exploration pass (gpurun --gpus N): ring size x tile size x variant on the three large
# collectives that trail NCCL, NCCL rows once, then in-kernel timelines of the chosen variants.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-explore_n$N}
rm -f gpurun_out/$TAG.jsonl gpurun_out/trace_$TAG.jsonl
timeout 100 python benchmarks/sweep.py --gpus $N --colls all_reduce,reduce,broadcast --impls nccl --sizes 256M --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | cut -c1-140
for SB in 262144 524288; do
echo "== PDCC_SLOT_BYTES=$SB"
PDCC_SLOT_BYTES=$SB timeout 200 python benchmarks/sweep.py --gpus $N --colls all_reduce,reduce,broadcast --algos twoshot_p2p,twoshot_nvls,nvls --impls pdcc --sizes 256M --env-sweep PDCC_TILE_BYTES=65536,131072,262144 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | cut -c1-140
done
export PDCC_SLOT_BYTES=524288 PDCC_TILE_BYTES=131072
for CA in "all_reduce twoshot_nvls" "reduce twoshot_p2p" "broadcast twoshot_p2p"; do
set -- $CA
timeout 100 python benchmarks/trace_timeline.py --gpus $N --coll $1 --size 256M --algo $2 --out gpurun_out/trace_$TAG.jsonl 2>> gpurun_out/$TAG.err | grep "^{" | cut -c1-900
done
tail -3 gpurun_out/$TAG.err
