"""Single-GPU degenerate case: world size 1 on the native backend, where
every collective is a copy or a no-op, plus the local K9 kernels."""
import os

import pytest
import torch

import pdcc_b200 as dist
from pdcc_b200 import _native as N
from pdcc_b200.ops import collectives as K
import _workers

pytestmark = pytest.mark.gpu


def test_probe_reports_b200_capabilities():
    caps = N.probe(0)
    assert caps["sm_count"] > 0 and caps["cooperative_launch"] == 1
    print(caps)


def test_local_copy_and_scale_cast():
    # >= 1 MiB, 16-byte aligned sizes take the TMA bulk-copy kernel (multi-chunk, ring wrap)
    for nbytes in (1, 15, 16, 4097, 1 << 20, (1 << 20) + 16, (1 << 24) + 3, 48 << 20, 256 << 20):
        src = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device="cuda")
        dst = torch.zeros_like(src)
        K.local_copy(dst, src)
        assert torch.equal(dst, src)
    off = torch.randint(0, 255, (5000,), dtype=torch.uint8, device="cuda")
    d2 = torch.zeros(5001, dtype=torch.uint8, device="cuda")
    K.local_copy(d2[1:], off)
    assert torch.equal(d2[1:], off)
    x = torch.randn(100003, device="cuda").bfloat16()
    o32 = torch.empty(100003, device="cuda")
    K.local_scale_cast(o32, x, 0.25)
    torch.testing.assert_close(o32, x.float() * 0.25, rtol=1e-6, atol=1e-6)
    o16 = torch.empty_like(x)
    K.local_scale_cast(o16, x, 0.5)
    torch.testing.assert_close(o16.float(), x.float() * 0.5, rtol=1e-2, atol=1e-2)


def test_world1_native_comm_runs_every_kernel():
    """nranks == 1 through the real communicator object (self-signalling flags)."""
    comm = K.NativeComm(None, "w1", 0, 1, 0, slot_bytes=64 << 10, max_blocks=8)
    assert comm.handle_kind == "local" and not comm.has_multicast
    before = N.launch_count()
    x = torch.randn(100001, device="cuda")
    ref = x.clone()
    comm.all_reduce(x, K.SUM, algo="oneshot_p2p")
    comm.all_reduce(x, K.MAX, algo="twoshot_p2p")
    comm.reduce(x, 0, K.SUM)
    comm.broadcast(x, 0)
    out = torch.zeros_like(x)
    comm.scatter(out, [x], 0)
    lst = [torch.zeros_like(x)]
    comm.gather(x, lst, 0)
    lst2 = [torch.zeros_like(x)]
    comm.all_gather(lst2, x)
    comm.barrier()
    comm.check()
    assert N.launch_count() - before == 8
    for t in (x, out, lst[0], lst2[0]):
        assert torch.equal(t, ref)
    comm.destroy()


def test_cuda_graph_replay_world1():
    """Launches carry no host-side epoch (ring counters live on the device), so a captured
    sequence of collectives can be replayed any number of times."""
    comm = K.NativeComm(None, "w1g", 0, 1, 0, slot_bytes=64 << 10, max_blocks=8)
    x = torch.zeros(300001, device="cuda")
    y = torch.zeros(1000, device="cuda")
    out, out2 = [torch.zeros_like(x)], torch.zeros_like(y)
    comm.all_gather(out, x)
    comm.scatter(out2, [y], 0)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            comm.all_gather(out, x)     # multi-tile, wraps the ring
            comm.scatter(out2, [y], 0)
            comm.all_reduce(y, K.SUM)
    for i in range(1, 6):
        x.fill_(float(i))
        y.fill_(float(-i))
        g.replay()
        torch.cuda.synchronize()
        assert float(out[0][0]) == i and float(out[0][-1]) == i and float(out2[7]) == -i
    comm.check()
    comm.destroy()


@pytest.mark.parametrize("name", ["do_reduce", "do_all_reduce", "do_scatter", "do_gather",
                                  "do_all_gather", "do_broadcast", "hello_world"])
def test_demos_world1_b200(name, tmp_path):
    """The reference's demos with size=1 on cuda:0 through init_process(..., 'b200')."""
    dist.launch(_workers.run_demo, size=1, backend="b200", timeout=300,
                env={"PDCC_TEST_OUT": str(tmp_path), "PDCC_TEST_DEMO": name})
    res = torch.load(os.path.join(tmp_path, f"{name}_0.pt"))
    want = {"do_reduce": 1.0, "do_all_reduce": 1.0, "do_scatter": 1.0, "do_broadcast": 0.0}
    if name in want:
        assert res.item() == want[name]
    elif name in ("do_gather", "do_all_gather"):
        assert [t.item() for t in res] == [0.0]


def test_named_c_abi_entry_points_once_each():
    """pdcc_all_reduce ... pdcc_barrier (include/pdcc.h) called straight through ctypes on a
    one-rank communicator: the call shapes of main.py:14,23,37,52,68,81 as a C host program
    would make them."""
    import ctypes as C

    L = N.lib()
    comm = K.NativeComm(None, "cabi", 0, 1, 0)
    h, st = comm._h, torch.cuda.current_stream().cuda_stream
    n = 70001
    try:
        x = torch.arange(n, dtype=torch.float32, device="cuda")
        want = x.clone()
        N.check(L.pdcc_all_reduce(h, x.data_ptr(), n, N.F32, K.SUM, st))
        N.check(L.pdcc_reduce(h, x.data_ptr(), n, N.F32, K.MAX, 0, st))
        N.check(L.pdcc_broadcast(h, x.data_ptr(), n, N.F32, 0, st))
        out = torch.zeros(n, device="cuda")
        N.check(L.pdcc_all_reduce_scaled(h, x.data_ptr(), out.data_ptr(), n, N.F32, N.F32, 0.5, st))
        lst = (C.c_void_p * 1)(x.data_ptr())
        sc = torch.zeros(n, device="cuda")
        N.check(L.pdcc_scatter(h, sc.data_ptr(), lst, n, N.F32, 0, st))
        ga = torch.zeros(n, device="cuda")
        glst = (C.c_void_p * 1)(ga.data_ptr())
        N.check(L.pdcc_gather(h, x.data_ptr(), glst, n, N.F32, 0, st))
        ag = torch.zeros(n, device="cuda")
        alst = (C.c_void_p * 1)(ag.data_ptr())
        N.check(L.pdcc_all_gather(h, alst, x.data_ptr(), n, N.F32, st))
        N.check(L.pdcc_barrier(h, st))
        comm.check()
        assert torch.equal(x, want) and torch.equal(sc, want) and torch.equal(ga, want)
        assert torch.equal(ag, want)
        torch.testing.assert_close(out, want * 0.5)
    finally:
        comm.destroy()
