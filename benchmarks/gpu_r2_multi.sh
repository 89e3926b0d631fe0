#!/usr/bin/env bash
# Round-2 multi-GPU pass (gpurun --gpus N): native correctness, both bench arms, and the
# tile-size / variant A/B on the large messages. One job per question, every step under
# `timeout`, logs in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-r2_n$N}
MODE=${2:-all}
rm -f gpurun_out/$TAG.jsonl
if [ "$MODE" = "all" ] || [ "$MODE" = "test" ]; then
echo "== native correctness ($N GPUs)"
timeout 420 python -m pytest tests/test_gpu_multi.py -x -q -k "sweep or data_parallel or contract or do_scatter" --timeout=300 --timeout-method=thread > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_$TAG.log
fi
echo "== headline step, both arms"
timeout 200 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "pdcc exit $?"; cat gpurun_out/bench_pdcc_$TAG.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['verified'], d['small_msg_p50_us'], {k:v['ms'] for k,v in d['per_collective'].items()})"; tail -3 gpurun_out/bench_pdcc_$TAG.err
timeout 200 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref exit $?"; grep '"metric"' gpurun_out/bench_ref_$TAG.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['verified'], d['small_msg_p50_us'], {k:v['ms'] for k,v in d['per_collective'].items()})"
if [ "$MODE" = "all" ] || [ "$MODE" = "sweep" ]; then
echo "== tile size x variant on the three collectives that trailed NCCL"
timeout 240 python benchmarks/sweep.py --gpus $N --colls all_reduce,reduce,broadcast --algos auto,twoshot_p2p,twoshot_nvls,nvls --sizes 32M,256M --env-sweep PDCC_TILE_BYTES=32768,65536,131072,262144 --out gpurun_out/$TAG.jsonl 2> gpurun_out/$TAG.err | grep -v "^#" | tail -100
tail -3 gpurun_out/$TAG.err
echo "== the other three, auto"
timeout 120 python benchmarks/sweep.py --gpus $N --colls scatter,gather,all_gather --algos auto --sizes 32M,256M --env-sweep PDCC_TILE_BYTES=32768,65536,131072 --out gpurun_out/$TAG.jsonl 2>> gpurun_out/$TAG.err | grep -v "^#" | tail -30
fi
