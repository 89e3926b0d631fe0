"""GPU correctness on ONE device: N virtual ranks, one cooperative launch
per collective, the real kernels and the real cross-rank flag protocol (P2P variants).

Oracle: the inputs gathered on the host and combined by torch on the CPU (fp64 for
SUM/PRODUCT, exact for MAX/MIN and for every data-movement collective)."""
import pytest
import torch

from pdcc_b200.ops import collectives as K

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def emus():
    made = {}

    def get(n):
        if n not in made:
            # small slots + few rings so that modest sizes already span many tiles/rings
            made[n] = K.Emulator(n, 0, slot_bytes=64 << 10, max_blocks=8, device_timeout_ms=4000)
        return made[n]

    yield get
    for e in made.values():
        e.destroy()


def _inputs(n, count, dtype, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed + count)
    if dtype.is_floating_point:
        xs = [(torch.rand(count, generator=g, dtype=torch.float64) * 2 - 1) for _ in range(n)]
    elif dtype == torch.uint8:
        xs = [torch.randint(0, 3, (count,), generator=g).double() for _ in range(n)]
    else:
        xs = [torch.randint(-3, 4, (count,), generator=g).double() for _ in range(n)]
    return [x.to(dtype).to(DEV) for x in xs]


def _oracle(xs, op):
    st = torch.stack([x.cpu().double() for x in xs])
    if op == K.SUM:
        return st.sum(0)
    if op == K.PRODUCT:
        return st.prod(0)
    if op == K.MIN:
        return st.min(0).values
    return st.max(0).values


def _tol(dtype, op, n):
    if not dtype.is_floating_point or op in (K.MIN, K.MAX):
        return 0.0, 0.0
    eps = {torch.float32: 1e-6, torch.float64: 1e-14, torch.bfloat16: 8e-3, torch.float16: 1e-3}[dtype]
    return eps * n, eps * n


SIZES = [1, 3, 4, 7, 1000, 4099, 65536 + 5, 300000]


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("algo", ["oneshot_p2p", "twoshot_p2p", "ll"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_all_reduce_sum_sizes(emus, n, algo, dtype):
    e = emus(n)
    for count in SIZES:
        xs = _inputs(n, count, dtype)
        want = _oracle(xs, K.SUM)
        e.all_reduce(xs, K.SUM, algo=algo)
        e.check()
        rtol, atol = _tol(dtype, K.SUM, n)
        for r in range(n):
            torch.testing.assert_close(xs[r].cpu().double(), want, rtol=rtol, atol=atol)
            assert torch.equal(xs[r], xs[0]), "all_reduce must be bitwise identical across ranks"


@pytest.mark.parametrize("op", [K.SUM, K.PRODUCT, K.MIN, K.MAX])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16, torch.float64,
                                   torch.int32, torch.int64, torch.uint8])
def test_all_reduce_ops_dtypes(emus, op, dtype):
    n = 4
    e = emus(n)
    for algo in ("oneshot_p2p", "twoshot_p2p", "ll"):
        xs = _inputs(n, 5003, dtype, seed=op)
        want = _oracle(xs, op)
        e.all_reduce(xs, op, algo=algo)
        e.check()
        rtol, atol = _tol(dtype, op, n)
        for r in range(n):
            torch.testing.assert_close(xs[r].cpu().double(), want, rtol=rtol, atol=atol)


def test_all_reduce_auto_multi_tile(emus):
    """Auto plan, a message that needs several tiles per ring and wraps the ring."""
    n = 4
    e = emus(n)
    count = (8 * (64 << 10) * 6) // 4 + 13  # > rings * slot * depth
    xs = _inputs(n, count, torch.float32)
    want = _oracle(xs, K.SUM)
    e.all_reduce(xs, K.SUM)
    e.check()
    for r in range(n):
        torch.testing.assert_close(xs[r].cpu().double(), want, rtol=4e-6, atol=4e-6)


def test_all_reduce_fused_scale_and_cast(emus):
    """K2: bf16 in -> fp32 accumulate -> x 1/world -> fp32 (or bf16) out, out of place."""
    n = 4
    e = emus(n)
    for algo in ("oneshot_p2p", "twoshot_p2p", "ll"):
        xs = _inputs(n, 70001, torch.bfloat16)
        want = _oracle(xs, K.SUM) / n
        outs = [torch.empty(70001, dtype=torch.float32, device=DEV) for _ in range(n)]
        e.all_reduce(xs, K.SUM, outs=outs, scale=1.0 / n, algo=algo)
        e.check()
        tol = 8e-3 if algo == "twoshot_p2p" else 1e-6  # two-shot ships bf16 partials
        for r in range(n):
            torch.testing.assert_close(outs[r].cpu().double(), want, rtol=tol, atol=tol)
        ys = _inputs(n, 4097, torch.bfloat16, seed=3)
        want = _oracle(ys, K.SUM) / n
        e.all_reduce(ys, K.SUM, scale=1.0 / n, algo=algo)
        e.check()
        for r in range(n):
            torch.testing.assert_close(ys[r].cpu().double(), want, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("n", [2, 4, 8])
def test_reduce_leaves_non_root_untouched(emus, n):
    e = emus(n)
    for root, algo in ((0, "p2p"), (n - 1, "p2p"), (0, "ll"), (n - 1, "ll"), (0, "twoshot_p2p"),
                       (n - 1, "twoshot_p2p")):
        for op in (K.SUM, K.MAX):
            xs = _inputs(n, 33333, torch.float32, seed=root)
            before = [x.clone() for x in xs]
            want = _oracle(xs, op)
            e.reduce(xs, root, op, algo=algo)
            e.check()
            torch.testing.assert_close(xs[root].cpu().double(), want, rtol=1e-5, atol=1e-5)
            for r in range(n):
                if r != root:
                    assert torch.equal(xs[r], before[r])


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.uint8])
@pytest.mark.parametrize("algo", ["p2p", "ll", "twoshot_p2p"])
def test_broadcast(emus, n, dtype, algo):
    e = emus(n)
    for count in SIZES:
        for root in (0, n - 1):
            xs = _inputs(n, count, dtype, seed=root)
            want = xs[root].clone()
            e.broadcast(xs, root, algo=algo)
            e.check()
            for r in range(n):
                assert torch.equal(xs[r], want)


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("algo", ["p2p", "ll", "pull"])
def test_scatter_gather_all_gather(emus, n, algo):
    e = emus(n)
    for count in SIZES:
        for root in (0, n // 2):
            src = _inputs(n, count, torch.float32, seed=root)  # list on the root
            outs = [torch.zeros(count, device=DEV) for _ in range(n)]
            e.scatter(outs, src, root, algo=None if algo == "pull" else algo)
            e.check()
            for r in range(n):
                assert torch.equal(outs[r], src[r])

            ins = _inputs(n, count, torch.float32, seed=10 + root)
            dst = [torch.zeros(count, device=DEV) for _ in range(n)]
            e.gather(ins, dst, root, algo=None if algo == "pull" else algo)
            e.check()
            for r in range(n):
                assert torch.equal(dst[r], ins[r])

        ins = _inputs(n, count, torch.bfloat16, seed=20)
        lists = [[torch.zeros(count, dtype=torch.bfloat16, device=DEV) for _ in range(n)]
                 for _ in range(n)]
        e.all_gather(lists, ins, algo=algo)
        e.check()
        for r in range(n):
            for p in range(n):
                assert torch.equal(lists[r][p], ins[p])


def test_misaligned_views(emus):
    """User tensors that are not 16-byte aligned take the scalar paths."""
    n = 4
    e = emus(n)
    base = [torch.zeros(10007, device=DEV) for _ in range(n)]
    xs = _inputs(n, 10000, torch.float32)
    views = []
    for r in range(n):
        v = base[r][3:10003]
        v.copy_(xs[r])
        views.append(v)
    want = _oracle(xs, K.SUM)
    for algo in ("oneshot_p2p", "twoshot_p2p"):
        cur = [v.clone() for v in views]
        vv = []
        for r in range(n):
            b = torch.zeros(10007, device=DEV)
            w = b[1:10001]
            w.copy_(cur[r])
            vv.append(w)
        e.all_reduce(vv, K.SUM, algo=algo)
        e.check()
        for r in range(n):
            torch.testing.assert_close(vv[r].cpu().double(), want, rtol=4e-6, atol=4e-6)
    u8 = [torch.arange(1001, device=DEV, dtype=torch.uint8)[1:] + r for r in range(n)]
    want = u8[2].clone()
    e.broadcast(u8, 2)
    e.check()
    for r in range(n):
        assert torch.equal(u8[r], want)


def test_stress_mixed_sequence_no_host_sync(emus):
    """back-to-back collectives of alternating kinds and sizes with no host
    sync in between; catches flag/ring reuse bugs across calls."""
    n = 4
    e = emus(n)
    torch.manual_seed(0)
    sizes = [1, 17, 4096, 70001, 3, 250000, 129]
    acc_want = []
    acc_got = []
    for it in range(60):
        count = sizes[it % len(sizes)]
        kind = it % 4
        if kind == 0:
            xs = _inputs(n, count, torch.float32, seed=it)
            acc_want.append(_oracle(xs, K.SUM).float())
            e.all_reduce(xs, K.SUM)
            acc_got.append(xs[it % n])
        elif kind == 1:
            xs = _inputs(n, count, torch.float32, seed=it)
            acc_want.append(xs[it % n].clone().cpu())
            e.broadcast(xs, it % n)
            acc_got.append(xs[(it + 1) % n])
        elif kind == 2:
            ins = _inputs(n, count, torch.float32, seed=it)
            lists = [[torch.zeros(count, device=DEV) for _ in range(n)] for _ in range(n)]
            acc_want.append(torch.cat([x.cpu() for x in ins]))
            e.all_gather(lists, ins)
            acc_got.append(lists[it % n])
        else:
            xs = _inputs(n, count, torch.float32, seed=it)
            acc_want.append(_oracle(xs, K.MAX).float())
            e.reduce(xs, it % n, K.MAX)
            acc_got.append(xs[it % n])
    e.check()
    for want, got in zip(acc_want, acc_got):
        got = torch.cat([g.cpu() for g in got]) if isinstance(got, list) else got.cpu()
        torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5)


def test_barrier(emus):
    e = emus(2)
    e.barrier()
    e.check()
