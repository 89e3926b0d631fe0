"""The size -> (algorithm, CTAs, tile) table of csrc/launch.cu, checked on the CPU through
pdcc_plan_dry: selection follows the measured crossovers recorded in BASELINE.md, and every
plan respects the slot layout the kernels assume."""
import itertools

import pytest

from pdcc_b200 import _native as N

SLOT = 256 << 10   # a ring is 4 x this (1 MiB); every call cuts it into `depth` slots
SIZES = [1, 8, 100, 4096, 4097, 64 << 10, (64 << 10) + 4, 300_001, 1 << 20, (1 << 20) + 16,
         3 << 20, 32 << 20, (256 << 20) + 12, 1 << 30]
COLLS = {"all_reduce": N.ALL_REDUCE, "reduce": N.REDUCE, "broadcast": N.BROADCAST,
         "scatter": N.SCATTER, "gather": N.GATHER, "all_gather": N.ALL_GATHER}


def _plan(n, coll, nbytes, **kw):
    dtype = N.F32 if coll in (N.ALL_REDUCE, N.REDUCE) else N.U8
    count = max(1, nbytes // 4) if dtype == N.F32 else nbytes
    return N.plan_dry(n, coll, count, dtype, **kw), count * (4 if dtype == N.F32 else 1)


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("mc", [True, False])
def test_every_auto_plan_fits_the_slot_layout(n, mc):
    for name, coll in COLLS.items():
        for nbytes in SIZES:
            p, real = _plan(n, coll, nbytes, has_multicast=mc, slot_bytes=SLOT)
            algo, nb, tile, nt = p["algo"], p["nblocks"], p["tile_bytes"], p["ntiles"]
            assert 1 <= nb <= N.MAX_BLOCKS and nt >= 1 and p["threads"] == 512
            assert nb * nt * tile >= real, (name, n, nbytes, p)          # the tiles cover the message
            assert (nt - 1) * nb * tile < real or nt == 1, (name, n, nbytes, p)  # no idle rounds
            if algo == "ll":
                assert tile == 512 * 16 and nb <= 16 and real <= (512 << 10) // max(n, 1)
                continue
            assert tile % 16 == 0
            if "nvls" in algo:
                assert mc and n >= 2
            rooted_2p = algo == "twoshot_p2p" and name in ("reduce", "broadcast")
            owners = n - 1 if rooted_2p else n
            if algo in ("twoshot_p2p", "twoshot_nvls"):
                assert tile % (16 * owners) == 0, (name, n, nbytes, p)
            # the call's slots fit its ring, and a tile fits one slot of this variant's layout
            slot, depth = p["slot_stride"], p["depth"]
            assert 1 <= depth <= 16 and slot * depth <= 4 * SLOT and slot % 256 == 0
            if name in ("all_reduce", "reduce") and algo in ("oneshot_p2p", "p2p"):
                assert tile <= slot // n                       # one sub-slot per source rank
            elif name in ("all_reduce", "reduce") and algo == "twoshot_p2p":
                assert tile // owners * (n + owners) <= slot   # n partials + reduced chunks
            elif name == "gather" or (name == "all_gather" and algo != "pull"):
                assert tile <= slot // n
            else:
                assert tile <= slot


def test_selection_follows_the_measured_crossovers():
    ar = lambda n, b, **kw: _plan(n, N.ALL_REDUCE, b, **kw)[0]["algo"]  # noqa: E731
    assert ar(8, 4096) == "ll" and ar(2, 200 << 10) == "ll"
    assert ar(8, 128 << 10) == "oneshot_nvls"           # above 512 KiB / n, below 1 MiB
    assert ar(8, 128 << 10, has_multicast=False) == "oneshot_p2p"
    assert ar(3, 512 << 10) == "oneshot_p2p"            # NVLS only beyond 3 ranks (measured)
    assert ar(2, 1 << 30) == "oneshot_p2p"              # n = 2: one-shot moves no more bytes
    assert ar(4, 256 << 20) == "twoshot_nvls"           # measured in round 2: 680 us vs 707 us push
    assert ar(3, 256 << 20) == "twoshot_p2p"
    assert ar(8, 256 << 20) == "twoshot_nvls"
    assert ar(8, 256 << 20, has_multicast=False) == "twoshot_p2p"
    assert ar(8, 256 << 20, redop=2) == "twoshot_p2p"   # PRODUCT: the switch cannot multiply
    rd = lambda n, b, **kw: _plan(n, N.REDUCE, b, **kw)[0]["algo"]  # noqa: E731
    assert rd(8, 256 << 20) == "twoshot_nvls" and rd(2, 256 << 20) == "p2p"
    assert rd(8, 256 << 20, has_multicast=False) == "twoshot_p2p"
    assert rd(8, 256 << 20, redop=2) == "twoshot_p2p"   # PRODUCT: the switch cannot multiply
    assert rd(8, 1 << 20) == "nvls" and rd(8, 1 << 20, has_multicast=False) == "p2p"
    bc = lambda n, b, **kw: _plan(n, N.BROADCAST, b, **kw)[0]["algo"]  # noqa: E731
    assert bc(2, 256 << 20) == "p2p" and bc(4, 256 << 20) == "twoshot_p2p" and bc(4, 1 << 20) == "nvls"
    assert bc(8, 256 << 20) == "twoshot_p2p" and bc(8, 1 << 20) == "nvls"
    ag = lambda n, b, **kw: _plan(n, N.ALL_GATHER, b, **kw)[0]["algo"]  # noqa: E731
    assert ag(2, 256 << 20) == "p2p" and ag(8, 256 << 20) == "nvls"
    assert _plan(8, N.SCATTER, 256 << 20)[0]["algo"] == "p2p"


def test_big_messages_use_every_sm_and_explicit_requests_are_honoured():
    p, _ = _plan(8, N.ALL_REDUCE, 256 << 20)
    assert p["nblocks"] == N.MAX_BLOCKS
    p, _ = _plan(8, N.ALL_REDUCE, 256 << 20, nblocks=32, algo=N.ALGO_NAMES["twoshot_p2p"])
    assert p["nblocks"] == 32 and p["algo"] == "twoshot_p2p"
    with pytest.raises(N.PdccError):
        _plan(8, N.ALL_REDUCE, 1 << 20, has_multicast=False, algo=N.ALGO_NAMES["twoshot_nvls"])
    with pytest.raises(N.PdccError):
        _plan(8, N.ALL_REDUCE, 1 << 20, nblocks=N.MAX_BLOCKS + 1)
