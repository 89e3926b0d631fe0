"""The allocation-free entry point, pdcc_emu2_all_reduce_noalloc, called straight through
the C ABI (ctypes, no Python wrapper in between).

It exists for hosts that cannot allocate device memory at the moment they need a
collective — for example a plugin inside a server that has already taken all free device
memory into its own pool, where a cudaMalloc would fail or fragment that pool, or a loop that
cannot afford cudaMalloc's device-wide synchronisation. Its scratch is a static device array,
so the only thing to verify is that it computes what the allocating entry point computes.

Oracle: pdcc_emu_collective on a 2-rank emulator (the existing, allocating entry point) fed
the same inputs. Two calls in a row, because the static scratch keeps its ring counters and
flags from one call to the next and the second call has to cope with what the first left.
This file is collected last on purpose (see its name): a run that stops at the first failure
has then already run every other test.
"""
import ctypes as C

import pytest
import torch

from pdcc_b200 import _native as N
from pdcc_b200.ops import collectives as K

pytestmark = pytest.mark.gpu


def _check():
    """Waits for the device; raises if a wait inside a no-alloc kernel ran out of time."""
    rc = N.lib().pdcc_emu2_noalloc_check(0)
    if rc != 0:
        raise N.PdccError(rc, N.lib().pdcc_last_error().decode())


def _no_alloc(bufs, redop=K.SUM, dtype_code=None):
    fn = N.lib().pdcc_emu2_all_reduce_noalloc
    if dtype_code is None:
        dtype_code = K.dtype_code(bufs[0].dtype, True)
    rc = fn(0, bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[0].numel(),
            dtype_code, redop, torch.cuda.current_stream().cuda_stream)
    if rc != 0:
        raise N.PdccError(rc, N.lib().pdcc_last_error().decode())


@pytest.mark.parametrize("dtype,count", [(torch.float32, 1), (torch.float32, 4099),
                                         (torch.float32, 1 << 20), (torch.bfloat16, 300001)])
def test_matches_allocating_entry_point_twice(dtype, count):
    torch.cuda.set_device(0)
    emu = K.Emulator(2, 0, slot_bytes=256 << 10, max_blocks=4)
    g = torch.Generator(device="cpu").manual_seed(count)
    tol = {torch.float32: 2e-6, torch.bfloat16: 1.6e-2}[dtype]
    try:
        for call in range(2):  # two calls in a row on the shared static scratch
            src = [(torch.rand(count, generator=g) * 2 - 1).to(dtype).cuda() for _ in range(2)]
            want = [s.clone() for s in src]
            emu.all_reduce(want, K.SUM)
            emu.check()
            got = [s.clone() for s in src]
            _no_alloc(got)
            _check()
            for r in range(2):
                torch.testing.assert_close(got[r].float(), want[r].float(), rtol=tol, atol=tol)
            assert torch.equal(got[0], got[1]), "both ranks must hold the identical result"
    finally:
        emu.destroy()


def test_other_ops_and_second_use_of_the_scratch():
    for redop, ref in ((K.MAX, torch.maximum), (K.MIN, torch.minimum), (K.PRODUCT, torch.mul)):
        a = torch.randn(70001, device="cuda")
        b = torch.randn(70001, device="cuda")
        want = ref(a, b)
        _no_alloc([a, b], redop)
        _check()
        torch.testing.assert_close(a, want, rtol=1e-6, atol=1e-6)
        assert torch.equal(a, b)


def test_rejects_sizes_beyond_the_stated_maximum():
    limit = 64 << 20  # PDCC_NOALLOC_MAX_BYTES in include/pdcc.h
    assert "#define PDCC_NOALLOC_MAX_BYTES (64ull << 20)" in (
        N.HERE.parent.parent / "include" / "pdcc.h").read_text()
    a = torch.zeros(limit // 4 + 1, device="cuda")
    b = torch.zeros_like(a)
    with pytest.raises(N.PdccError) as e:
        _no_alloc([a, b])
    assert e.value.code == N.EINVAL and "limited to" in str(e.value)
    # exactly at the limit is accepted
    a = torch.ones(limit // 4, device="cuda")
    b = torch.ones_like(a)
    _no_alloc([a, b])
    _check()
    assert float(a[0]) == 2.0 and float(b[-1]) == 2.0


def test_refuses_bad_arguments_before_enqueuing_and_keeps_the_current_device():
    a = torch.ones(16, dtype=torch.int8, device="cuda")  # int8 has no reduction kernel
    with pytest.raises(N.PdccError) as e:
        _no_alloc([a, a.clone()], dtype_code=N.I8)
    assert e.value.code == N.EUNSUPPORTED
    host = torch.ones(16).pin_memory()                   # not device memory of device 0
    dev = torch.ones(16, device="cuda")
    with pytest.raises(N.PdccError) as e:
        _no_alloc([dev, host])
    assert e.value.code == N.EINVAL and "not device memory" in str(e.value)
    assert float(dev[0]) == 1.0
    assert torch.cuda.current_device() == 0
    _check()
