"""K9 local copy: chunk size x stages x CTAs per SM of the TMA bulk-copy kernel against
torch's device-to-device copy, 256 MiB fp32 (larger than L2), CUDA events, best of 3 x 20."""
import itertools
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdcc_b200 import _native as N
from pdcc_b200.ops import collectives as K

torch.cuda.set_device(0)
n = 64 << 20
src = torch.ones(n, device="cuda")
dst = torch.empty_like(src)


def timed(fn, iters=20, reps=3):
    best = 1e9
    for _ in range(reps):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters)
    return best


rows = [{"impl": "torch", "ms": round(timed(lambda: dst.copy_(src)), 4)}]
for chunk, stages, per_sm in itertools.product((8, 16, 32, 64), (3, 4, 6, 8, 12), (1, 2, 3, 4)):
    if chunk * stages * per_sm > 220 or chunk * stages > 200:
        continue
    os.environ.update(PDCC_COPY_CHUNK_KB=str(chunk), PDCC_COPY_STAGES=str(stages),
                      PDCC_COPY_CTAS_PER_SM=str(per_sm))
    N.reload_env()
    ms = timed(lambda: K.local_copy(dst, src), iters=10, reps=2)
    rows.append({"impl": "tma", "chunk_kb": chunk, "stages": stages, "ctas_per_sm": per_sm,
                 "ms": round(ms, 4), "TBps": round(2 * n * 4 / ms / 1e9, 3)})
assert torch.equal(dst, src)
rows.sort(key=lambda r: r["ms"])
for r in rows[:12]:
    print(json.dumps(r))
print(json.dumps([r for r in rows if r["impl"] == "torch"][0]))
with open("gpurun_out/copy_sweep.jsonl", "w") as f:
    for r in rows:
        f.write(json.dumps(r) + "\n")
