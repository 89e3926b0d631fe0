"""Small single-GPU program for `ncu --set full`: the K9 copy kernel of the 1-GPU bench, the
world-size-1 all_reduce (tile pipeline with self-signalled flags) and one 4-rank emulated
two-shot all_reduce (one cooperative kernel over all ranks' data, as B200_PROFILING.md asks
for when there are fewer GPUs than ranks)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdcc_b200.ops import collectives as K

torch.cuda.set_device(0)
n = 64 << 20
src = torch.ones(n, device="cuda")
dst = torch.empty_like(src)
for _ in range(4):
    K.local_copy(dst, src)
comm = K.NativeComm(None, "ncu", 0, 1, 0)
x = torch.ones(16 << 20, device="cuda")
for _ in range(3):
    comm.all_reduce(x, K.SUM, algo="twoshot_p2p")
comm.check()
emu = K.Emulator(4, 0, slot_bytes=1 << 20, max_blocks=32)
xs = [torch.ones(8 << 20, device="cuda") for _ in range(4)]
for _ in range(3):
    emu.all_reduce(xs, K.SUM, algo="twoshot_p2p")
emu.check()
torch.cuda.synchronize()
print("ok", float(dst[-1]), float(x[0]), float(xs[0][0]))
