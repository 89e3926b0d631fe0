#!/usr/bin/env bash
# Round-2 8-GPU pass: correctness (native sweep incl. NVLS, zero-copy, contract, DP), both
# bench arms, the DP MLP step against DDP, and the sweeps BASELINE.json configs 2-4 name.
# One process per question; everything under `timeout`.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TAG=${1:-val_n$N}
echo "== pytest ($N GPUs)"
timeout 420 python -m pytest tests/test_gpu_multi.py -q -k "sweep or symmetric or contract or data_parallel or do_scatter" --timeout=300 --timeout-method=thread > gpurun_out/pytest_multi_$TAG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_multi_$TAG.log
echo "== bench, both arms"
timeout 240 python bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_pdcc_$TAG.json 2> gpurun_out/bench_pdcc_$TAG.err; echo "pdcc exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_pdcc_$TAG.json").read().strip().splitlines()[-1])
print("pdcc", d["value"], d["ms_per_step"], "verified", d["verified"], "small", d["small_msg_p50_us"], {k:v["ms"] for k,v in d["per_collective"].items()})
print("sym ", {k:(v["ms"] if isinstance(v,dict) else v) for k,v in (d.get("symmetric_memory") or {}).items()})
print("e2e ", d["e2e"], d["clocks"])
PY
tail -2 gpurun_out/bench_pdcc_$TAG.err
timeout 240 python bench.py --gpus $N --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref exit $?"
python - <<PY
import json
d=[l for l in open("gpurun_out/bench_ref_$TAG.json").read().splitlines() if '"metric"' in l][-1]
d=json.loads(d)
print("ref ", d["value"], d["ms_per_step"], "verified", d["verified"], "small", d["small_msg_p50_us"], {k:v["ms"] for k,v in d["per_collective"].items()})
PY
echo "== DP MLP (hidden 8192, 4 layers, bf16)"
timeout 200 python benchmarks/dp_mlp.py --gpus $N --steps 10 --warmup 3 --mode all --comm-ctas 8,16 > gpurun_out/dp_mlp_$TAG.json 2> gpurun_out/dp_mlp_$TAG.err; echo "dp exit $?"; grep '"config"' gpurun_out/dp_mlp_$TAG.json; tail -2 gpurun_out/dp_mlp_$TAG.err
for CFG in config2_all_reduce config3_rooted config4_all_gather; do
echo "== sweep $CFG"
timeout 300 python benchmarks/sweep.py --gpus $N --jobs-file benchmarks/jobs_$CFG.json --out gpurun_out/${CFG}_$TAG.jsonl 2> gpurun_out/${CFG}_$TAG.err | grep -v "^#" | cut -c1-132; tail -2 gpurun_out/${CFG}_$TAG.err
done
