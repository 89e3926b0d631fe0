#!/usr/bin/env bash
# Round-2 validation pass on N GPUs (gpurun --gpus N): the whole multi-GPU test file, both
# bench arms, the DP MLP step (BASELINE.json config 5) with the DDP baseline.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-val_n$N}
echo "== pytest tests/test_gpu_multi.py ($N GPUs)"
timeout 600 python -m pytest tests/test_gpu_multi.py -q --timeout=300 --timeout-method=thread > gpurun_out/pytest_multi_$TAG.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_multi_$TAG.log
echo "== bench, both arms"
timeout 240 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "pdcc exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_pdcc_$TAG.json").read().strip().splitlines()[-1])
print("pdcc", d["value"], d["ms_per_step"], "verified", d["verified"], "small", d["small_msg_p50_us"], {k:v["ms"] for k,v in d["per_collective"].items()})
print("sym ", d.get("symmetric_memory"))
print("e2e ", d["e2e"], d["clocks"])
PY
tail -3 gpurun_out/bench_pdcc_$TAG.err
timeout 240 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref exit $?"
python - <<PY
import json
d=[l for l in open("gpurun_out/bench_ref_$TAG.json").read().splitlines() if '"metric"' in l][-1]
d=json.loads(d)
print("ref ", d["value"], d["ms_per_step"], "verified", d["verified"], "small", d["small_msg_p50_us"], {k:v["ms"] for k,v in d["per_collective"].items()})
PY
if [ "$N" -gt 1 ]; then
echo "== DP MLP (hidden 8192, 4 layers, bf16)"
timeout 300 python benchmarks/dp_mlp.py --gpus $N --steps 10 --warmup 3 --mode all > gpurun_out/dp_mlp_$TAG.json 2> gpurun_out/dp_mlp_$TAG.err; echo "dp exit $?"; grep '"config"' gpurun_out/dp_mlp_$TAG.json; tail -3 gpurun_out/dp_mlp_$TAG.err
fi
