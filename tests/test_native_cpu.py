"""The native library must load, and fail loudly rather than fall back, without a GPU."""
import ctypes as C

import pytest
import torch

from pdcc_b200 import _native as N
from pdcc_b200.ops import collectives as K


def test_library_loads_in_tree():
    L = N.lib()
    assert N.LIB_PATH.exists() and N.LIB_PATH.parent.name == "_native"
    assert b"sm_100a" in L.pdcc_version()
    assert N.launch_count() >= 0


def test_header_and_binding_agree():
    text = (N.HERE.parent.parent / "include" / "pdcc.h").read_text()
    assert f"#define PDCC_MAX_RANKS {N.MAX_RANKS}" in text
    assert f"#define PDCC_MAX_BLOCKS {N.MAX_BLOCKS}" in text
    assert f"#define PDCC_RING_DEPTH {N.RING_DEPTH}" in text
    for name in ("pdcc_all_reduce", "pdcc_reduce", "pdcc_broadcast", "pdcc_scatter", "pdcc_gather",
                 "pdcc_all_gather", "pdcc_collective", "pdcc_emu_collective", "pdcc_sym_alloc",
                 "pdcc_sym_query", "pdcc_comm_read_trace", "pdcc_reload_env",
                 "pdcc_emu2_all_reduce_noalloc", "pdcc_emu2_noalloc_check"):
        assert name in text and hasattr(N.lib(), name)
    assert "PDCC_ALGO_SYM = 9" in text and N.ALGO_NAMES["sym"] == 9
    assert C.sizeof(N.Op) == 4 * 8 + 8 + 8 + 8 + 8 * N.MAX_RANKS


def test_planner_knobs_are_cached_until_reloaded(monkeypatch):
    """The planner reads PDCC_* once; pdcc_reload_env re-reads (benchmarks A/B a knob in one job)."""
    base = N.plan_dry(8, N.ALL_REDUCE, 64 << 20, N.F32)
    monkeypatch.setenv("PDCC_TILE_BYTES", "32768")
    assert N.plan_dry(8, N.ALL_REDUCE, 64 << 20, N.F32)["tile_bytes"] == base["tile_bytes"]
    N.reload_env()
    try:
        assert N.plan_dry(8, N.ALL_REDUCE, 64 << 20, N.F32)["tile_bytes"] == 32768
    finally:
        monkeypatch.delenv("PDCC_TILE_BYTES")
        N.reload_env()
    assert N.plan_dry(8, N.ALL_REDUCE, 64 << 20, N.F32) == base


def test_dtype_codes():
    assert K.dtype_code(torch.float32) == N.F32 and K.dtype_code(torch.bfloat16) == N.BF16
    with pytest.raises(TypeError):
        K.dtype_code(torch.int8, reduce=True)
    with pytest.raises(TypeError):
        K.dtype_code(torch.complex64)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_cuda_paths_fail_loudly_without_gpu():
    with pytest.raises(N.PdccError) as e:
        N.probe(0)
    assert e.value.code == N.ECUDA
    with pytest.raises(N.PdccError):
        K.NativeComm(None, "t", 0, 1, 0)
    with pytest.raises(N.PdccError):
        K.Emulator(2, 0)
