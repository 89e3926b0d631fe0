#!/usr/bin/env python
"""Message-size sweeps for BASELINE.json configs 2-4: every collective x size x algorithm
variant of this library against the same call on torch's NCCL backend, in ONE job so both
see the same box, clocks and rank placement.

  torchrun --nproc-per-node 8 benchmarks/sweep.py --colls all_reduce --min 1K --max 1G
  python benchmarks/sweep.py --gpus 2 --colls all_reduce,all_gather --algos auto,oneshot_p2p

Rows (JSON lines, rank 0) go to --out: collective, impl, algo, bytes, us (device time, max
over ranks), algbw, busbw (nccl-tests factors) and fraction of the NVLink roofline
(B200_PROFILING.md: 770 GB/s measured per direction per GPU).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

NVLINK_GBPS = 770.0
COLL_ID = {"all_reduce": 0, "reduce": 1, "broadcast": 2, "scatter": 3, "gather": 4, "all_gather": 5}


def parse_size(s: str) -> int:
    s = s.strip().upper()
    mult = {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30}
    if s[-1] in mult:
        return int(float(s[:-1]) * mult[s[-1]])
    return int(s)


BUS = {"all_reduce": lambda n: 2.0 * (n - 1) / n, "all_gather": lambda n: (n - 1) / n,
       "scatter": lambda n: (n - 1) / n, "gather": lambda n: (n - 1) / n,
       "reduce": lambda n: 1.0, "broadcast": lambda n: 1.0}

# bytes that must cross the busiest NVLink direction of the busiest GPU, per message byte
# (the lower bound any algorithm has without in-switch reduction / multicast)
LINK_BYTES = {"all_reduce": lambda n: 2.0 * (n - 1) / n, "all_gather": lambda n: (n - 1) / n,
              "scatter": lambda n: (n - 1) / n, "gather": lambda n: (n - 1) / n,
              "reduce": lambda n: (n - 1) / n, "broadcast": lambda n: (n - 1) / n}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--colls", default="all_reduce,reduce,broadcast,scatter,gather,all_gather")
    ap.add_argument("--min-bytes", dest="min", default="1K")
    ap.add_argument("--max-bytes", dest="max", default="1G")
    ap.add_argument("--factor", type=int, default=4)
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--op", default="SUM")
    ap.add_argument("--algos", default="auto", help="comma list; 'all' = every variant")
    ap.add_argument("--impls", default="pdcc,nccl")
    ap.add_argument("--nblocks", default="0", help="comma list of CTA counts (0 = auto)")
    ap.add_argument("--out", default="gpurun_out/sweep.jsonl")
    ap.add_argument("--subgroup", default="", help="comma list of global ranks, e.g. 1,3")
    ap.add_argument("--sizes", default="", help="explicit comma list of sizes (overrides min/max/factor)")
    ap.add_argument("--env-sweep", default="",
                    help="A/B a planner knob inside ONE job, e.g. PDCC_TILE_BYTES=32768,65536 "
                         "(the native planner reads its knobs from the environment per call)")
    ap.add_argument("--jobs-file", default="",
                    help="JSON list of dicts, each overriding options (colls, sizes, dtype, op, "
                         "algos, impls, subgroup, symmetric, graph, factor, min, max ...): all of "
                         "them run in THIS process, one after the other (process start-up costs "
                         "more than most sweeps on an 8-GPU box)")
    ap.add_argument("--symmetric", action="store_true",
                    help="pdcc buffers come from dist.empty_symmetric (zero-copy kernels); the NCCL "
                         "rows keep ordinary tensors")
    ap.add_argument("--graph", action="store_true",
                    help="also time small messages captured in a CUDA graph (device-only latency)")
    ap.add_argument("--graph-max", type=parse_size, default=1 << 20)
    args = ap.parse_args()

    if args.gpus > 1 and "RANK" not in os.environ:
        from pdcc_b200.launcher import find_free_port

        os.execvp(sys.executable, [
            sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
            f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
            str(find_free_port()), os.path.abspath(__file__), *sys.argv[1:]])

    import torch
    import torch.distributed as tdist

    import pdcc_b200 as dist
    from pdcc_b200 import _native as N
    from pdcc_b200.ops import collectives as K

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)) % torch.cuda.device_count())
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("b200", rank=rank, world_size=world)
    import copy

    jobs = [args]
    if args.jobs_file:
        jobs = []
        for over in json.load(open(args.jobs_file)):
            j = copy.copy(args)
            for k, v in over.items():
                setattr(j, k.replace("-", "_"), v)
            jobs.append(j)
    cache = {}  # rank tuple -> (group, nccl group, comm): groups are created once per process

    def run_job(args):
        ranks = [int(x) for x in args.subgroup.split(",")] if args.subgroup else list(range(world))
        impls = args.impls.split(",")
        member = rank in ranks
        n = len(ranks)
        root = ranks[0]
        if tuple(ranks) not in cache:
            g = dist.new_group(ranks)
            cache[tuple(ranks)] = [g, None, dist.native_comm(g) if member and n > 1 else None]
        ent = cache[tuple(ranks)]
        if "nccl" in impls and ent[1] is None:
            ent[1] = tdist.new_group(ranks, backend="nccl")
        group, nccl, comm = ent
        dtype = getattr(torch, args.dtype)
        esz = torch.empty(0, dtype=dtype).element_size()
        op_p = getattr(dist.ReduceOp, args.op)
        op_t = getattr(tdist.ReduceOp, args.op)
        kop = int(op_p)

        sizes = []
        if args.sizes:
            sizes = [parse_size(x) for x in args.sizes.split(",")]
        else:
            s = parse_size(args.min)
            while s <= parse_size(args.max):
                sizes.append(s)
                s *= args.factor
        env_name, env_values = None, [None]
        if args.env_sweep:
            env_name, vals = args.env_sweep.split("=", 1)
            env_values = vals.split(",")

        def timed(fn, nbytes: int) -> float:
            iters = 200 if nbytes <= (1 << 16) else 50 if nbytes <= (1 << 22) else 10 if nbytes <= (1 << 28) else 5
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            tdist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], dtype=torch.float64)
            tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
            return float(t.item())

        def timed_graph(fn, reps: int = 20, replays: int = 10) -> float:
            """us per call with the launches captured in a CUDA graph: device-side latency with
            the Python/launch overhead of the caller taken out."""
            side = torch.cuda.Stream()
            fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    for _ in range(reps):
                        fn()
            g.replay()
            torch.cuda.synchronize()
            tdist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(replays):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / (reps * replays) * 1e3], dtype=torch.float64)
            tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
            return float(t.item())

        rows = []
        out = open(args.out, "a") if rank == 0 else None
        info = {"world": world, "group": ranks, "multicast": bool(comm and comm.has_multicast),
                "handle": comm.handle_kind if comm else None, "dtype": args.dtype, "op": args.op}
        if rank == 0:
            print("# " + json.dumps(info), flush=True)
            out.write(json.dumps({"info": info}) + "\n")

        for coll in args.colls.split(","):
            algos = args.algos.split(",")
            if algos == ["all"]:
                if coll == "all_reduce":
                    algos = ["auto", "ll", "oneshot_p2p", "twoshot_p2p"]
                    if comm and comm.has_multicast and not (args.op == "PRODUCT" or
                                                            (dtype == torch.float32 and args.op != "SUM")):
                        algos += ["oneshot_nvls", "twoshot_nvls"]
                elif coll in ("reduce", "broadcast", "all_gather"):
                    algos = ["auto", "ll", "p2p"] + (["nvls"] if comm and comm.has_multicast else [])
                    algos += {"reduce": ["twoshot_p2p"], "all_gather": ["pull"]}.get(coll, [])
                else:
                    algos = ["auto", "ll", "p2p"]
            for nbytes in sizes:
                count = max(n, nbytes // esz // n * n)
                shard = count // n
                real = count * esz
                # input magnitudes keep repeated in-place SUMs finite
                x = torch.full((count,), 1e-20, dtype=dtype, device=dev) if dtype.is_floating_point \
                    else torch.ones(count, dtype=dtype, device=dev)
                xs = x[:shard].clone()
                lst = [torch.empty(shard, dtype=dtype, device=dev) for _ in range(n)]
                px, plst = x, lst  # what the pdcc rows use
                if args.symmetric and member:
                    px = dist.empty_symmetric(count, dtype=dtype, group=group)
                    px.copy_(x)
                    plst = [dist.empty_symmetric(shard, dtype=dtype, group=group) for _ in range(n)]
                is_root = rank == root

                def make(impl, algo, nb):
                    if impl == "nccl":
                        g = nccl
                        return {
                            "all_reduce": lambda: tdist.all_reduce(x, op=op_t, group=g),
                            "reduce": lambda: tdist.reduce(x, dst=root, op=op_t, group=g),
                            "broadcast": lambda: tdist.broadcast(x, src=root, group=g),
                            "scatter": lambda: tdist.scatter(xs, scatter_list=lst if is_root else [], src=root, group=g),
                            "gather": lambda: tdist.gather(xs, gather_list=lst if is_root else [], dst=root, group=g),
                            "all_gather": lambda: tdist.all_gather(lst, xs, group=g),
                        }[coll]
                    a = None if algo == "auto" else algo
                    gr = group.to_group_rank(root)
                    return {
                        "all_reduce": lambda: comm.all_reduce(px, kop, algo=a, nblocks=nb),
                        "reduce": lambda: comm.reduce(px, gr, kop, algo=a, nblocks=nb),
                        "broadcast": lambda: comm.broadcast(px, gr, algo=a, nblocks=nb),
                        "scatter": lambda: comm.scatter(xs, lst if is_root else None, gr, algo=a, nblocks=nb),
                        "gather": lambda: comm.gather(xs, lst if is_root else None, gr, algo=a, nblocks=nb),
                        "all_gather": lambda: comm.all_gather(plst, xs, algo=a, nblocks=nb),
                    }[coll]

                for impl in impls:
                    variants = [("nccl", 0)] if impl == "nccl" else \
                        [(a, int(nb)) for a in algos for nb in args.nblocks.split(",")]
                    variants = [(a, nb, ev) for (a, nb) in variants
                                for ev in (env_values if impl == "pdcc" else [None])]
                    for algo, nb, ev in variants:
                        if env_name is not None and ev is not None:
                            # same value on every rank: same plan. Several knobs at once are
                            # written NAME=a+OTHER=b inside one value.
                            for kv in (f"{env_name}={ev}").split("+"):
                                kk, vv = kv.split("=", 1)
                                os.environ[kk] = vv
                            N.reload_env()
                        if algo == "ll" and real > (1 << 20):
                            continue  # LL doubles the wire bytes: a small-message variant only
                        if not member or (impl == "pdcc" and comm is None):
                            # non-members still take part in the host-side timing collectives
                            us = timed(lambda: None, real)
                            continue
                        if impl == "pdcc":
                            try:  # a variant this collective / group cannot run: same answer on every rank
                                comm.plan(COLL_ID[coll], count if coll in ("all_reduce", "reduce") else
                                          real if coll == "broadcast" else shard * esz,
                                          torch.uint8 if coll not in ("all_reduce", "reduce") else dtype,
                                          redop=kop, algo=None if algo == "auto" else algo, nblocks=nb)
                            except N.PdccError:
                                continue
                        us = timed(make(impl, algo, nb), real)
                        graph_us = None
                        if args.graph and real <= args.graph_max:
                            try:
                                graph_us = round(timed_graph(make(impl, algo, nb)), 2)
                            except Exception as ex:  # capture unsupported for this call
                                graph_us = f"capture failed: {type(ex).__name__}"
                        algbw = real / (us * 1e-6) / 1e9
                        roof_us = real * LINK_BYTES[coll](n) / (NVLINK_GBPS * 1e9) * 1e6
                        row = {"coll": coll, "impl": impl, "algo": algo, "nblocks": nb, "bytes": real,
                               "us": round(us, 2), "algbw_GBps": round(algbw, 2),
                               "busbw_GBps": round(algbw * BUS[coll](n), 2),
                               "roofline_frac": round(roof_us / us, 3) if us > 0 else None, "n": n,
                               "graph_us": graph_us}
                        if env_name is not None and impl == "pdcc":
                            row["env"] = {env_name: ev}
                        if impl == "pdcc" and args.symmetric:
                            row["symmetric"] = True
                        if impl == "pdcc":
                            row["plan"] = comm.plan(COLL_ID[coll],
                                                    count if coll in ("all_reduce", "reduce") else
                                                    real if coll == "broadcast" else shard * esz,
                                                    torch.uint8 if coll not in ("all_reduce", "reduce") else dtype,
                                                    redop=kop, algo=None if algo == "auto" else algo, nblocks=nb)
                        rows.append(row)
                        if rank == 0:
                            out.write(json.dumps(row) + "\n")
                            out.flush()
                            tag = f" {env_name}={ev}" if env_name is not None and impl == "pdcc" else ""
                            print(f"{coll:11s} {impl:5s} {algo:13s} nb={nb:<3d}{tag} {real:>11d} B  {us:10.2f} us  "
                                  f"alg {algbw:8.2f}  bus {row['busbw_GBps']:8.2f} GB/s  roof {row['roofline_frac']}"
                                  + (f"  graph {graph_us} us" if graph_us is not None else ""),
                                  flush=True)
                del x, xs, lst, px, plst
        if comm is not None:
            comm.check()

    for job in jobs:
        run_job(job)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
