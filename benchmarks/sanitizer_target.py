"""Small single-GPU program for `compute-sanitizer --tool memcheck|racecheck` (SURVEY.md §4.4,
§5.2): every collective of the emulated 4-rank world (the real kernels and flag protocol in
one cooperative launch), several tiles per ring so that slots are reused, long device-side
wait budgets because the sanitizer slows the kernels down by an order of magnitude."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdcc_b200.ops import collectives as K

torch.cuda.set_device(0)
n, count = 4, 70001
emu = K.Emulator(n, 0, slot_bytes=16 << 10, max_blocks=4, device_timeout_ms=300000)
bad = []
for algo in ("ll", "oneshot_p2p", "twoshot_p2p"):
    xs = [torch.full((count,), float(r + 1), device="cuda") for r in range(n)]
    emu.all_reduce(xs, K.SUM, algo=algo)
    emu.check()
    if not all(bool((x == 10.0).all()) for x in xs):
        bad.append(("all_reduce", algo))
for algo in ("p2p", "twoshot_p2p"):
    xs = [torch.full((count,), float(r + 1), device="cuda") for r in range(n)]
    emu.reduce(xs, 1, K.MAX, algo=algo)
    emu.check()
    if not bool((xs[1] == 4.0).all()) or not bool((xs[0] == 1.0).all()):
        bad.append(("reduce", algo))
    emu.broadcast(xs, 2, algo=algo)
    emu.check()
    if not all(bool((x == 3.0).all()) for x in xs):
        bad.append(("broadcast", algo))
src = [torch.full((count,), float(10 * r), device="cuda") for r in range(n)]
outs = [torch.zeros(count, device="cuda") for _ in range(n)]
emu.scatter(outs, src, 0)
dst = [torch.zeros(count, device="cuda") for _ in range(n)]
emu.gather(outs, dst, 3)
lists = [[torch.zeros(count, device="cuda") for _ in range(n)] for _ in range(n)]
emu.all_gather(lists, outs)
emu.check()
if not all(bool((outs[r] == 10.0 * r).all()) and bool((dst[r] == 10.0 * r).all()) for r in range(n)):
    bad.append(("scatter/gather",))
if not all(bool((lists[r][p] == 10.0 * p).all()) for r in range(n) for p in range(n)):
    bad.append(("all_gather",))
emu.destroy()
torch.cuda.synchronize()
print("sanitizer target:", "FAILED " + str(bad) if bad else "all results correct")
sys.exit(1 if bad else 0)
