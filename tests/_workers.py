"""Rank functions run by the multi-process tests (must be importable by spawned children)."""
import os
import warnings

import torch

import pdcc_b200 as dist
from pdcc_b200.models import demos


def _save(rank, name, obj):
    d = os.environ["PDCC_TEST_OUT"]
    torch.save(obj, os.path.join(d, f"{name}_{rank}.pt"))


def run_demo(rank, size):
    name = os.environ["PDCC_TEST_DEMO"]
    out = demos.DEMOS[name](rank, size)
    if isinstance(out, list):
        out = [t.cpu() for t in out]
    elif isinstance(out, torch.Tensor):
        out = out.cpu()
    _save(rank, name, out)


def contract_checks(rank, size):
    """Everything SURVEY.md §2.2 'semantics the new API must preserve', in one world of 4."""
    dev = demos.device()
    res = {}

    # ReduceOp set with inputs r+1: SUM=10 PRODUCT=24 MAX=4 MIN=1
    world = dist.new_group(list(range(size)))
    for op in (dist.ReduceOp.SUM, dist.ReduceOp.PRODUCT, dist.ReduceOp.MAX, dist.ReduceOp.MIN):
        t = torch.tensor([float(rank + 1)], device=dev)
        dist.all_reduce(t, op=op, group=world)
        res[f"all_reduce_{op.name}"] = t.item()
        t = torch.tensor([float(rank + 1)], device=dev)
        dist.reduce(t, dst=0, op=op, group=world)
        res[f"reduce_{op.name}"] = t.item()

    # sub-group [3, 1] is sorted to [1, 3]; global 3 has group rank 1; src is a GLOBAL rank
    sub = dist.new_group([3, 1])
    if rank in (1, 3):
        res["sub_ranks"] = list(sub.ranks)
        res["sub_group_rank"] = dist.get_rank(sub)
        t = torch.tensor([float(rank)], device=dev)
        dist.broadcast(t, src=3, group=sub)
        res["sub_bcast"] = t.item()
        t = torch.tensor([float(rank)], device=dev)
        dist.all_reduce(t, group=sub)
        res["sub_sum"] = t.item()
    else:
        res["non_member_sentinel"] = sub == dist.GroupMember.NON_GROUP_MEMBER
        t = torch.tensor([123.0], device=dev)
        with warnings.catch_warnings(record=True) as w:
            warnings.simplefilter("always")
            ret = dist.all_reduce(t, group=sub)
        res["non_member_ret"] = ret
        res["non_member_warned"] = len(w) == 1
        res["non_member_tensor"] = t.item()

    # halves: new_group is collective over the world, so everyone creates both, in order
    lower = dist.new_group([0, 1])
    upper = dist.new_group([2, 3])
    mine = lower if rank < 2 else upper
    lst = [torch.zeros(1, device=dev) for _ in range(2)]
    dist.all_gather(lst, torch.tensor([float(rank)], device=dev), group=mine)
    res["half_all_gather"] = [x.item() for x in lst]

    # list-on-root-only errors
    t = torch.zeros(1, device=dev)
    errs = {}
    try:
        if rank == 0:
            dist.scatter(t, scatter_list=[], src=0, group=world)
        else:
            dist.scatter(t, scatter_list=[torch.zeros(1, device=dev)] * size, src=0, group=world)
    except ValueError as e:
        errs["scatter"] = str(e)
    try:
        if rank == 0:
            dist.gather(t, gather_list=[], dst=0, group=world)
        else:
            dist.gather(t, gather_list=[torch.zeros(1, device=dev)] * size, dst=0, group=world)
    except ValueError as e:
        errs["gather"] = str(e)
    try:
        dist.all_gather([torch.zeros(1, dtype=torch.float64, device=dev)] * size, t, group=world)
    except ValueError as e:
        errs["dtype"] = str(e)
    try:
        dist.all_reduce([t], group=world)
    except TypeError as e:
        errs["type"] = str(e)
    try:
        dist.new_group([0, size])
    except ValueError as e:
        errs["range"] = str(e)
    res["errors"] = errs

    # async_op returns a handle; fused average
    t = torch.tensor([float(rank)], device=dev)
    w = dist.all_reduce(t, group=world, async_op=True)
    w.wait()
    res["async_sum"] = t.item()
    res["async_done"] = w.is_completed()
    t = torch.tensor([float(rank)], device=dev)
    dist.all_reduce_avg(t, group=world)
    res["avg"] = t.item()
    t = torch.tensor([float(rank)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.AVG, group=world)
    res["avg_op"] = t.item()

    # gather / scatter with a non-zero root
    lst = [torch.zeros(2, device=dev) for _ in range(size)] if rank == 2 else []
    dist.gather(torch.full((2,), float(rank), device=dev), gather_list=lst, dst=2, group=world)
    if rank == 2:
        res["gather_root2"] = [x.tolist() for x in lst]
    out = torch.zeros(3, device=dev)
    sl = [torch.full((3,), float(10 * i), device=dev) for i in range(size)] if rank == 1 else []
    dist.scatter(out, scatter_list=sl, src=1, group=world)
    res["scatter_root1"] = out.tolist()
    dist.barrier(group=world)
    _save(rank, "contract", res)


def crash_rank1(rank, size):
    if rank == 1:
        raise SystemExit(3)
    dist.barrier()


def _det_input(rank, count, dtype, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(1000 * seed + rank)
    if dtype.is_floating_point:
        x = torch.rand(count, generator=g, dtype=torch.float64) * 2 - 1
    else:
        x = torch.randint(-3, 4, (count,), generator=g).double()
    return x.to(dtype).to(dev)


def native_sweep(rank, size):
    """Every collective x dtype x op x size x algorithm variant on real GPUs, against an
    oracle each rank rebuilds locally from the seeds. Large sizes only for
    the SUM paths so that the CPU oracle does not dominate the GPU time."""
    from pdcc_b200.ops import collectives as K

    torch.set_num_threads(4)
    dev = demos.device()
    world = dist.new_group(list(range(size)))
    comm = dist.native_comm(world)
    failures, checked = [], 0
    small_sizes = [1, 5, 1024, 4099, 70001]
    big_sizes = [262144 + 3, (1 << 20) + 1]
    ar_algos = ["ll", "oneshot_p2p", "twoshot_p2p"]
    if comm.has_multicast:
        ar_algos += ["oneshot_nvls", "twoshot_nvls"]
    ops = {"SUM": K.SUM, "PRODUCT": K.PRODUCT, "MIN": K.MIN, "MAX": K.MAX}

    def oracle(count, dtype, seed, op):
        st = torch.stack([_det_input(r, count, dtype, seed, "cpu").double() for r in range(size)])
        return {"SUM": st.sum(0), "PRODUCT": st.prod(0), "MIN": st.min(0).values,
                "MAX": st.max(0).values}[op]

    seed = 0
    for dtype in (torch.float32, torch.bfloat16, torch.float16, torch.int32):
        for opname, op in ops.items():
            for algo in ar_algos:
                if "nvls" in algo and (opname == "PRODUCT" or
                                       (dtype == torch.float32 and opname != "SUM")):
                    continue
                sizes = list(small_sizes)
                if opname == "SUM" and dtype in (torch.float32, torch.bfloat16) and algo != "ll":
                    sizes += big_sizes
                for count in sizes:
                    seed += 1
                    x = _det_input(rank, count, dtype, seed, dev)
                    comm.all_reduce(x, op, algo=algo)
                    want = oracle(count, dtype, seed, opname)
                    tol = 0.0
                    if dtype.is_floating_point and opname in ("SUM", "PRODUCT"):
                        tol = {torch.float32: 1e-5, torch.bfloat16: 4e-2, torch.float16: 5e-3}[dtype]
                    got = x.cpu().double()
                    checked += 1
                    if not torch.allclose(got, want, rtol=tol, atol=tol):
                        failures.append(("all_reduce", str(dtype), opname, algo, count,
                                         float((got - want).abs().max())))
    mc = comm.has_multicast
    red_algos = ["ll", "p2p", "twoshot_p2p"] + (["nvls", "twoshot_nvls"] if mc else [])
    bc_algos = ["ll", "p2p", "twoshot_p2p"] + (["nvls", "twoshot_nvls"] if mc else [])
    ag_algos = ["ll", "p2p", "pull"] + (["nvls"] if mc else [])
    for count in small_sizes + big_sizes:
        for root in (0, size - 1):
            for algo in red_algos:
                seed += 1
                x = _det_input(rank, count, torch.float32, seed, dev)
                keep = x.clone()
                comm.reduce(x, root, K.SUM, algo=algo)
                checked += 1
                if rank == root:
                    if not torch.allclose(x.cpu().double(), oracle(count, torch.float32, seed, "SUM"),
                                          rtol=1e-5, atol=1e-5):
                        failures.append(("reduce", algo, count, root))
                elif not torch.equal(x, keep):
                    failures.append(("reduce-nonroot-touched", algo, count, root))
            for algo in bc_algos:
                seed += 1
                x = _det_input(rank, count, torch.float32, seed, dev)
                comm.broadcast(x, root, algo=algo)
                checked += 1
                if not torch.equal(x.cpu(), _det_input(root, count, torch.float32, seed, "cpu")):
                    failures.append(("broadcast", algo, count, root))
            for algo in ("ll", "p2p"):
                seed += 1
                out = torch.zeros(count, device=dev)
                sl = [_det_input(100 + j, count, torch.float32, seed, dev) for j in range(size)] \
                    if rank == root else None
                comm.scatter(out, sl, root, algo=algo)
                checked += 1
                if not torch.equal(out.cpu(), _det_input(100 + rank, count, torch.float32, seed, "cpu")):
                    failures.append(("scatter", algo, count, root))
                seed += 1
                x = _det_input(rank, count, torch.float32, seed, dev)
                gl = [torch.zeros(count, device=dev) for _ in range(size)] if rank == root else None
                comm.gather(x, gl, root, algo=algo)
                checked += 1
                if rank == root:
                    for p in range(size):
                        if not torch.equal(gl[p].cpu(), _det_input(p, count, torch.float32, seed, "cpu")):
                            failures.append(("gather", algo, count, root, p))
                            break
        for algo in ag_algos:
            seed += 1
            x = _det_input(rank, count, torch.bfloat16, seed, dev)
            lst = [torch.zeros(count, dtype=torch.bfloat16, device=dev) for _ in range(size)]
            comm.all_gather(lst, x, algo=algo)
            checked += 1
            for p in range(size):
                if not torch.equal(lst[p].cpu(), _det_input(p, count, torch.bfloat16, seed, "cpu")):
                    failures.append(("all_gather", algo, count, p))
                    break
    sizes = small_sizes
    # stress: no host sync between 300 mixed calls
    xs = []
    for it in range(300):
        count = sizes[it % 4]
        x = _det_input(rank, count, torch.float32, 5000 + it, dev)
        if it % 3 == 0:
            comm.all_reduce(x, K.SUM)
        elif it % 3 == 1:
            comm.broadcast(x, it % size)
        else:
            comm.reduce(x, it % size, K.MAX)
        xs.append((it, count, x))
    for it, count, x in xs:
        checked += 1
        if it % 3 == 0:
            ok = torch.allclose(x.cpu().double(), oracle(count, torch.float32, 5000 + it, "SUM"),
                                rtol=1e-5, atol=1e-5)
        elif it % 3 == 1:
            ok = torch.equal(x.cpu(), _det_input(it % size, count, torch.float32, 5000 + it, "cpu"))
        else:
            ok = rank != it % size or torch.equal(
                x.cpu().double(), oracle(count, torch.float32, 5000 + it, "MAX"))
        if not ok:
            failures.append(("stress", it, count))
    # CUDA-graph replay of a mixed sequence (LL + tile-pipelined + rooted)
    small = torch.zeros(1000, device=dev)
    big = torch.zeros(400003, device=dev)
    comm.all_reduce(small, K.SUM)
    comm.all_reduce(big, K.SUM)
    comm.broadcast(big, size - 1)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            comm.all_reduce(small, K.SUM)
            comm.all_reduce(big, K.SUM)
            comm.broadcast(small, size - 1)
    for i in range(1, 5):
        small.fill_(float(rank + i))
        big.fill_(float(i))
        graph.replay()
        torch.cuda.synchronize()
        checked += 1
        want_small = float(sum(r + i for r in range(size)))
        if float(small[0]) != want_small or float(small[-1]) != want_small or \
                float(big[0]) != float(i * size) or float(big[-1]) != float(i * size):
            failures.append(("graph", i, float(small[0]), float(big[-1])))
    try:
        comm.check()
    except Exception as e:  # device-side timeout
        failures.append(("device", str(e)))
    _save(rank, "sweep", {"failures": failures, "checked": checked,
                          "multicast": comm.has_multicast, "handle": comm.handle_kind})


def symmetric_checks(rank, size):
    """Zero-copy collectives on tensors from dist.empty_symmetric (csrc/sym.cuh) against the
    same oracle as native_sweep, AUTO routing and forced, interleaved with staged calls."""
    from pdcc_b200.ops import collectives as K

    torch.set_num_threads(4)
    dev = demos.device()
    world = dist.new_group(list(range(size)))
    comm = dist.native_comm(world)
    res = {"multicast": comm.has_multicast, "failures": [], "checked": 0}
    if not comm.has_multicast:
        _save(rank, "sym", res)
        return
    fails = res["failures"]

    def oracle(count, dtype, seed, op):
        st = torch.stack([_det_input(r, count, dtype, seed, "cpu").double() for r in range(size)])
        return {"SUM": st.sum(0), "MIN": st.min(0).values, "MAX": st.max(0).values}[op]

    seed = 9000
    sizes = [16, 16 * 1024 + 16, 300 * 1024, (2 << 20) + 64, (8 << 20) + 16 * size * 3]  # elements
    for dtype, ops in ((torch.float32, ("SUM",)), (torch.bfloat16, ("SUM", "MAX", "MIN"))):
        k = 16 // torch.empty(0, dtype=dtype).element_size()
        for opname in ops:
            op = {"SUM": K.SUM, "MIN": K.MIN, "MAX": K.MAX}[opname]
            for count in sizes:
                count = count // k * k  # the zero-copy kernels work on whole 16-byte units
                for algo in (None, "sym"):
                    seed += 1
                    x = dist.empty_symmetric(count, dtype=dtype, group=world)
                    assert dist.is_symmetric(x, world)
                    x.copy_(_det_input(rank, count, dtype, seed, dev))
                    comm.all_reduce(x, op, algo=algo)
                    want = oracle(count, dtype, seed, opname)
                    tol = 0.0 if opname != "SUM" else {torch.float32: 1e-5, torch.bfloat16: 4e-2}[dtype]
                    res["checked"] += 1
                    if not torch.allclose(x.cpu().double(), want, rtol=tol, atol=tol):
                        fails.append(("all_reduce", str(dtype), opname, count, algo))
    # fused average, reduce to both ends, broadcast (direct and two-phase), all_gather
    for count in (4096, (4 << 20) + 16):
        seed += 1
        x = dist.empty_symmetric(count, group=world)
        x.copy_(_det_input(rank, count, torch.float32, seed, dev))
        dist.all_reduce_avg(x, group=world)
        res["checked"] += 1
        if not torch.allclose(x.cpu().double(), oracle(count, torch.float32, seed, "SUM") / size,
                              rtol=1e-5, atol=1e-5):
            fails.append(("avg", count))
        for root in (0, size - 1):
            seed += 1
            x.copy_(_det_input(rank, count, torch.float32, seed, dev))
            keep = x.clone()
            comm.reduce(x, root, K.SUM, algo="sym")
            res["checked"] += 1
            if rank == root:
                if not torch.allclose(x.cpu().double(), oracle(count, torch.float32, seed, "SUM"),
                                      rtol=1e-5, atol=1e-5):
                    fails.append(("reduce", count, root))
            elif not torch.equal(x, keep):
                fails.append(("reduce-nonroot-touched", count, root))
            for algo in (None, "sym", "nvls"):
                seed += 1
                x.copy_(_det_input(rank, count, torch.float32, seed, dev))
                comm.broadcast(x, root, algo=algo)
                res["checked"] += 1
                if not torch.equal(x.cpu(), _det_input(root, count, torch.float32, seed, "cpu")):
                    fails.append(("broadcast", count, root, algo))
        seed += 1
        lst = [dist.empty_symmetric(count, dtype=torch.bfloat16, group=world) for _ in range(size)]
        xin = _det_input(rank, count, torch.bfloat16, seed, dev)
        for t in lst:
            t.zero_()
        comm.all_gather(lst, xin, algo="sym")
        res["checked"] += 1
        for p in range(size):
            if not torch.equal(lst[p].cpu(), _det_input(p, count, torch.bfloat16, seed, "cpu")):
                fails.append(("all_gather", count, p))
                break
    # interleave zero-copy and staged calls with no host sync in between
    a = dist.empty_symmetric(1 << 20, group=world)
    b = torch.empty(1 << 20, device=dev)
    outs = []
    for it in range(40):
        t = a if it % 2 == 0 else b
        t.fill_(float(rank + it))
        comm.all_reduce(t, K.SUM)
        outs.append((it, t[:4].clone(), t[-4:].clone()))
    for it, h, tl in outs:
        want = float(sum(r + it for r in range(size)))
        res["checked"] += 1
        if not (torch.all(h == want) and torch.all(tl == want)):
            fails.append(("interleave", it, float(h[0]), want))
    # graph capture of a zero-copy call
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            comm.all_reduce(a, K.SUM)
    for i in range(1, 4):
        a.fill_(float(rank + i))
        graph.replay()
        torch.cuda.synchronize()
        res["checked"] += 1
        want = float(sum(r + i for r in range(size)))
        if float(a[0]) != want or float(a[-1]) != want:
            fails.append(("graph", i, float(a[0]), want))
    try:
        comm.check()
    except Exception as e:
        fails.append(("device", str(e)))
    _save(rank, "sym", res)


def dp_mlp_check(rank, size):
    """D1: bucketed gradient averaging == one big-batch step (README.md:5), any backend."""
    from pdcc_b200.models.mlp import MLP, DataParallelTrainer

    dev = demos.device()
    dtype = torch.float32
    torch.manual_seed(1234 + rank)  # ranks start with DIFFERENT params: broadcast must fix it
    # initialise on the CPU generator (the oracle below does the same), then move
    model = MLP(hidden=32, layers=3, dtype=dtype).to(dev)
    tr = DataParallelTrainer(model, lr=0.1, bucket_bytes=3000)  # several buckets
    g = torch.Generator(device="cpu").manual_seed(7)
    xs = torch.randn(size, 8, 32, generator=g)
    ys = torch.randn(size, 8, 32, generator=g)
    losses = []
    for _ in range(3):
        losses.append(float(tr.step(xs[rank].to(dev), ys[rank].to(dev))))
    if dev.type == "cuda":
        torch.cuda.synchronize()

    # single-process oracle: rank 0's initial parameters, the concatenated batch
    torch.manual_seed(1234)
    ref = MLP(hidden=32, layers=3, dtype=dtype)
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    for _ in range(3):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(ref(xs.reshape(-1, 32)), ys.reshape(-1, 32))
        loss.backward()
        opt.step()
    err = max(float((p.detach().cpu() - q.detach()).abs().max())
              for p, q in zip(model.parameters(), ref.parameters()))
    _save(rank, "dp", {"err": err, "losses": losses, "buckets": len(tr.avg.buckets)})


def api_misc(rank, size):
    """Lifecycle and bookkeeping of the API layer on the gloo path."""
    res = {}
    res["rank"], res["world"], res["backend"] = dist.get_rank(), dist.get_world_size(), dist.get_backend()
    res["initialized"] = dist.is_initialized()
    try:
        dist.init_process_group("gloo", rank=rank, world_size=size)
    except ValueError as e:
        res["double_init"] = str(e)
    g = dist.new_group([0, size - 1])
    if g != dist.GroupMember.NON_GROUP_MEMBER:
        res["group_size"] = dist.get_world_size(g)
        res["group_rank"] = dist.get_rank(g)
        res["global_of_1"] = dist.get_global_rank(g, 1)
        res["group_of_last"] = dist.get_group_rank(g, size - 1)
        try:
            dist.broadcast(torch.zeros(1), src=1 if size > 2 else 5, group=g)
        except ValueError as e:
            res["bad_src"] = str(e)
        # group_src / group_dst spellings
        t = torch.tensor([float(rank)])
        dist.broadcast(t, group=g, group_src=1)
        res["bcast_group_src"] = t.item()
        t = torch.tensor([1.0])
        dist.reduce(t, group=g, group_dst=0)
        res["reduce_group_dst"] = t.item()
        w = dist.barrier(group=g, async_op=True)
        w.wait()
        res["barrier_done"] = w.is_completed()
        dist.destroy_process_group(g)
    else:
        res["sentinel_rank"] = dist.get_rank(g)
        res["sentinel_size"] = dist.get_world_size(g)
    # symmetric memory degrades to ordinary tensors where there is nothing to gain (gloo, CPU)
    sym = dist.empty_symmetric(3, 5, dtype=torch.float64)
    res["sym_fallback"] = (tuple(sym.shape), str(sym.dtype), dist.is_symmetric(sym))
    try:
        dist.all_reduce(torch.zeros(1), op=17)
    except ValueError as e:
        res["bad_op"] = str(e)
    try:
        dist.all_reduce(torch.zeros(1), group="nope")
    except TypeError as e:
        res["bad_group"] = str(e)
    if torch.cuda.is_available() is False:
        try:
            # CPU-only box: a "CUDA" request cannot be served by the gloo backend
            dist.native_comm()
        except RuntimeError as e:
            res["native_on_gloo"] = str(e)
    dist.barrier()
    _save(rank, "api", res)
