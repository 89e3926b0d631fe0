"""Small single-GPU program for `ncu --set full` (B200_PROFILING.md: with fewer GPUs than ranks,
emulate the ranks as ONE kernel over all ranks' data): four emulated ranks x 32 CTAs in one
cooperative launch, the three pipeline kernels of the bench path that cross "links" (two-shot
all_reduce, rooted two-shot reduce, two-phase broadcast), each launched twice so that
`-k regex:'reduce_kernel|move_kernel' -s 3 -c 3` captures the warm second round, plus the K9
local copy. Sizes are kept small: ncu saves and restores device memory around ~40 replays."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdcc_b200.ops import collectives as K

torch.cuda.set_device(0)
n = 4
emu = K.Emulator(n, 0, slot_bytes=256 << 10, max_blocks=32)
xs = [torch.full((4 << 20,), float(r + 1), device="cuda") for r in range(n)]  # 16 MiB each
for _ in range(2):
    emu.all_reduce(xs, K.SUM, algo="twoshot_p2p")
    emu.reduce(xs, 0, K.SUM, algo="twoshot_p2p")
    emu.broadcast(xs, 0, algo="twoshot_p2p")
emu.check()
src = torch.ones(16 << 20, device="cuda")
dst = torch.empty_like(src)
for _ in range(2):
    K.local_copy(dst, src)
torch.cuda.synchronize()
print("ok", float(dst[-1]), float(xs[0][0]), float(xs[3][-1]))
