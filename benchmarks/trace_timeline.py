#!/usr/bin/env python
"""Where does the time of one collective go?  Runs one call with PDCC_TRACE=1 and prints,
per rank and per pipeline stage of CTA 0, how long the workers waited for admission and how
long they worked, plus the publisher's fences (count, mean latency). NVLink counters are not
reachable through ncu on these boxes; this timeline, taken on the real multi-GPU run, is the
evidence used to size tiles, rings and warp groups.

  torchrun --nproc-per-node 4 benchmarks/trace_timeline.py --coll reduce --size 256M
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--coll", default="all_reduce")
    ap.add_argument("--size", default="256M")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--out", default="gpurun_out/trace.jsonl")
    args = ap.parse_args()
    if args.gpus > 1 and "RANK" not in os.environ:
        from pdcc_b200.launcher import find_free_port

        os.execvp(sys.executable, [
            sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
            f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
            str(find_free_port()), os.path.abspath(__file__), *sys.argv[1:]])
    os.environ["PDCC_TRACE"] = "1"
    import torch
    import torch.distributed as tdist

    import pdcc_b200 as dist
    from benchmarks.sweep import parse_size

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29513")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)) % torch.cuda.device_count())
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("b200", rank=rank, world_size=world)
    comm = dist.native_comm()
    n = world
    nbytes = parse_size(args.size)
    count = nbytes // 4 // n * n
    x = torch.full((count,), 1e-20, device=dev)
    xs = x[:count // n].clone()
    lst = [torch.empty(count // n, device=dev) for _ in range(n)]
    a = None if args.algo == "auto" else args.algo
    fn = {"all_reduce": lambda: comm.all_reduce(x, 0, algo=a),
          "reduce": lambda: comm.reduce(x, 0, 0, algo=a),
          "broadcast": lambda: comm.broadcast(x, 0, algo=a),
          "scatter": lambda: comm.scatter(xs, lst if rank == 0 else None, 0, algo=a),
          "gather": lambda: comm.gather(xs, lst if rank == 0 else None, 0, algo=a),
          "all_gather": lambda: comm.all_gather(lst, xs, algo=a)}[args.coll]
    for _ in range(3):
        fn()
    comm.read_trace()  # drop the warm-up events
    tdist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    ev = comm.read_trace()
    us = e0.elapsed_time(e1) * 1e3
    t_min = min(e["begin_ns"] for e in ev) if ev else 0
    t_max = max(e["end_ns"] for e in ev) if ev else 0
    rep = {"rank": rank, "coll": args.coll, "bytes": count * 4, "algo": args.algo, "kernel_us": round(us, 1),
           "span_us": round((t_max - t_min) / 1e3, 1), "events": len(ev),
           "plan": comm.plan({"all_reduce": 0, "reduce": 1, "broadcast": 2, "scatter": 3, "gather": 4,
                              "all_gather": 5}[args.coll],
                             count if args.coll in ("all_reduce", "reduce") else
                             count * 4 if args.coll == "broadcast" else count // n * 4,
                             torch.float32 if args.coll in ("all_reduce", "reduce") else torch.uint8, algo=a)}
    for s, name in enumerate(("IN", "MID", "OUT")):
        w = [e for e in ev if e["kind"] == "worker" and e["stage"] == s]
        if not w:
            continue
        warps = sorted({e["warp"] for e in w})
        one = [e for e in w if e["warp"] == warps[0]]
        wait = sum(e["aux"] for e in one) / 1e3
        total = sum(e["end_ns"] - e["begin_ns"] for e in one) / 1e3
        rep[name] = {"warps": len(warps), "tiles": len(one), "wait_us": round(wait, 1),
                     "work_us": round(total - wait, 1),
                     "first_start_us": round((min(e["begin_ns"] for e in one) - t_min) / 1e3, 1),
                     "last_end_us": round((max(e["end_ns"] for e in one) - t_min) / 1e3, 1)}
    f = [e for e in ev if e["kind"] == "fence"]
    if f:
        lat = [(e["end_ns"] - e["begin_ns"]) / 1e3 for e in f]
        rep["fences"] = {"count": len(f), "mean_us": round(sum(lat) / len(lat), 2), "max_us": round(max(lat), 2)}
    for r in range(world):
        tdist.barrier()
        if r == rank and rank in (0, 1, world - 1):
            print(json.dumps(rep), flush=True)
            with open(args.out, "a") as fh:
                fh.write(json.dumps(rep) + "\n")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
