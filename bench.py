#!/usr/bin/env python
"""Headline benchmark: the reference's six collectives as one timed step.

BASELINE.json's metric is "bus GB/s per collective vs msg size; p50 small-msg latency".
One step = the six calls of main.py (reduce, all_reduce, scatter, gather, all_gather,
broadcast: main.py:14,23,37,52,68,81) with the reference's call shape, on fp32 CUDA
tensors of `--bytes` total each (nccl-tests size convention: the list collectives move
`bytes` in total, i.e. lists of N tensors of bytes/N).

  value            = 6 * bytes / step time, in GB/s  (aggregate algorithmic bandwidth)
  per_collective   = nccl-tests bus bandwidth of each collective, timed on its own
  small_msg_p50_us = median device time of an 8-byte all_reduce
  e2e              = the same step through the public API, plus per-step H2D copies of every
                     input from pinned host memory and a D2H read of one result element per
                     collective

  python bench.py                       # 1 GPU (degenerate: list ops are copies, in-place ops no-ops)
  torchrun --nproc-per-node 8 bench.py --gpus 8
  python bench.py --impl reference      # the unmodified reference's init_process on NCCL

Timing: CUDA events on the launching stream, barrier + synchronize on both sides, max over
ranks. Every tensor is larger than L2 (126 MB) at the default size, so no flush is needed.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="pdcc", choices=["pdcc", "reference"])
    ap.add_argument("--bytes", type=int, default=256 << 20,
                    help="message size of every collective (total bytes, fp32)")
    ap.add_argument("--detail-iters", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def _unavailable(why: str) -> None:
    print(json.dumps({"impl": "reference", "unavailable": why}))
    sys.exit(0)


class Suite:
    """Buffers + the six-collective step for one API namespace `D` (pdcc_b200 or
    torch.distributed: same call shape, which is the point of the reference)."""

    def __init__(self, D, rank: int, size: int, nbytes: int, dev):
        import torch

        self.D, self.rank, self.size, self.dev = D, rank, size, dev
        self.count = max(size, nbytes // 4 // size * size)
        self.shard = self.count // size
        self.nbytes = self.count * 4
        self.group = D.new_group(list(range(size)))
        tiny = 1e-20  # SUM grows by x size per step when inputs are not refreshed
        mk = lambda n: torch.full((n,), tiny, dtype=torch.float32, device=dev)  # noqa: E731
        self.t_red, self.t_ar, self.t_bc = mk(self.count), mk(self.count), mk(self.count)
        self.sc_out, self.ga_in, self.ag_in = mk(self.shard), mk(self.shard), mk(self.shard)
        self.ag_list = [mk(self.shard) for _ in range(size)]
        self.sc_list = [mk(self.shard) for _ in range(size)] if rank == 0 else []
        self.ga_list = [mk(self.shard) for _ in range(size)] if rank == 0 else []

    # the reference's six call sites, in main.py order
    def c_reduce(self):
        self.D.reduce(self.t_red, dst=0, op=self.D.ReduceOp.SUM, group=self.group)

    def c_all_reduce(self):
        self.D.all_reduce(self.t_ar, op=self.D.ReduceOp.SUM, group=self.group)

    def c_scatter(self):
        self.D.scatter(self.sc_out, scatter_list=self.sc_list, src=0, group=self.group)

    def c_gather(self):
        self.D.gather(self.ga_in, gather_list=self.ga_list, dst=0, group=self.group)

    def c_all_gather(self):
        self.D.all_gather(self.ag_list, self.ag_in, group=self.group)

    def c_broadcast(self):
        self.D.broadcast(self.t_bc, src=0, group=self.group)

    COLLS = ("reduce", "all_reduce", "scatter", "gather", "all_gather", "broadcast")

    def step(self):
        self.c_reduce()
        self.c_all_reduce()
        self.c_scatter()
        self.c_gather()
        self.c_all_gather()
        self.c_broadcast()

    # ---- correctness of the very buffers that were timed -------------------------------
    def verify(self) -> dict:
        """Each of the six collectives once more, on rank- and position-dependent inputs
        (x_r[i] = i % 251 + r + 1, exact in fp32), checked element by element against the
        closed form on every rank. Returns {collective: bool} for this rank."""
        import torch

        n, r, dev = self.size, self.rank, self.dev
        pat = lambda m: (torch.arange(m, device=dev) % 251).to(torch.float32)  # noqa: E731
        pc, ps = pat(self.count), pat(self.shard)
        tri = n * (n + 1) / 2.0
        ok = {}
        self.t_red.copy_(pc + (r + 1))
        self.c_reduce()
        ok["reduce"] = bool(torch.equal(self.t_red, n * pc + tri if r == 0 else pc + (r + 1)))
        self.t_ar.copy_(pc + (r + 1))
        self.c_all_reduce()
        ok["all_reduce"] = bool(torch.equal(self.t_ar, n * pc + tri))
        if r == 0:
            for j, t in enumerate(self.sc_list):
                t.copy_(ps + 1000.0 * (j + 1))
        self.sc_out.zero_()
        self.c_scatter()
        ok["scatter"] = bool(torch.equal(self.sc_out, ps + 1000.0 * (r + 1)))
        self.ga_in.copy_(ps + 1000.0 * (r + 1))
        for t in self.ga_list:
            t.zero_()
        self.c_gather()
        ok["gather"] = all(bool(torch.equal(t, ps + 1000.0 * (j + 1)))
                           for j, t in enumerate(self.ga_list))
        self.ag_in.copy_(ps + 1000.0 * (r + 1))
        for t in self.ag_list:
            t.zero_()
        self.c_all_gather()
        ok["all_gather"] = all(bool(torch.equal(t, ps + 1000.0 * (j + 1)))
                               for j, t in enumerate(self.ag_list))
        if r == 0:
            self.t_bc.copy_(pc + 7.0)
        else:
            self.t_bc.zero_()
        self.c_broadcast()
        ok["broadcast"] = bool(torch.equal(self.t_bc, pc + 7.0))
        torch.cuda.synchronize()
        return ok

    # ---- end to end: inputs come from pinned host memory, results go back ------------
    def make_host(self):
        import torch

        pin = lambda n: torch.full((n,), 1e-20, dtype=torch.float32).pin_memory()  # noqa: E731
        self.h_red, self.h_ar = pin(self.count), pin(self.count)
        self.h_ga, self.h_ag = pin(self.shard), pin(self.shard)
        self.h_bc = pin(self.count) if self.rank == 0 else None
        self.h_sc = pin(self.count) if self.rank == 0 else None
        self.h_res = torch.zeros(6, dtype=torch.float32).pin_memory()
        h2d = 2 * self.count + 2 * self.shard + (2 * self.count if self.rank == 0 else 0)
        return h2d * 4, 6 * 4

    def e2e_step(self):
        import torch

        self.t_red.copy_(self.h_red, non_blocking=True)
        self.c_reduce()
        self.t_ar.copy_(self.h_ar, non_blocking=True)
        self.c_all_reduce()
        if self.rank == 0:
            for j, t in enumerate(self.sc_list):
                t.copy_(self.h_sc[j * self.shard:(j + 1) * self.shard], non_blocking=True)
        self.c_scatter()
        self.ga_in.copy_(self.h_ga, non_blocking=True)
        self.c_gather()
        self.ag_in.copy_(self.h_ag, non_blocking=True)
        self.c_all_gather()
        if self.rank == 0:
            self.t_bc.copy_(self.h_bc, non_blocking=True)
        self.c_broadcast()
        outs = (self.t_red, self.t_ar, self.sc_out,
                self.ga_list[-1] if self.rank == 0 else self.ga_in, self.ag_list[-1], self.t_bc)
        for i, t in enumerate(outs):
            self.h_res[i:i + 1].copy_(t[0:1], non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the host reads the step's result
        return float(self.h_res.sum())


BUS_FACTOR = {  # nccl-tests conventions
    "all_reduce": lambda n: 2.0 * (n - 1) / n, "all_gather": lambda n: (n - 1) / n,
    "scatter": lambda n: (n - 1) / n, "gather": lambda n: (n - 1) / n,
    "reduce": lambda n: 1.0, "broadcast": lambda n: 1.0,
}


def run(D, rank: int, size: int, args, launch_count) -> dict:
    import torch
    import torch.distributed as tdist

    from pdcc_b200.utils.clocks import ClockSampler

    dev = torch.device("cuda", torch.cuda.current_device())
    suite = Suite(D, rank, size, args.bytes, dev)

    def sync_all():
        torch.cuda.synchronize()
        tdist.barrier()  # host barrier (gloo for pdcc, nccl for the reference arm)
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        t = torch.tensor([ms], dtype=torch.float64)
        if tdist.get_backend() == "nccl":
            t = t.to(dev)
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, iters: int) -> float:
        """ms per call of fn, device-timed, max over ranks"""
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        sync_all()
        return max_over_ranks(ms)

    for _ in range(max(args.warmup, 3)):
        suite.step()
    sync_all()

    # ---- headline: exactly K steps --------------------------------------------------
    l0 = launch_count()
    with ClockSampler(torch.cuda.current_device()) as cs:
        ms_step = timed(suite.step, args.steps)
    launches = launch_count() - l0
    clocks = cs.summary()

    # ---- per collective + small-message latency (explanatory, outside the K steps) ---
    per = {}
    for name in Suite.COLLS:
        ms = timed(getattr(suite, "c_" + name), args.detail_iters)
        algbw = suite.nbytes / (ms * 1e-3) / 1e9
        per[name] = {"ms": round(ms, 4), "algbw_GBps": round(algbw, 1),
                     "busbw_GBps": round(algbw * BUS_FACTOR[name](size), 1)}
    small = torch.zeros(2, dtype=torch.float32, device=dev)
    lat = []
    for _ in range(5):
        D.all_reduce(small, op=D.ReduceOp.SUM, group=suite.group)
    for _ in range(9):
        lat.append(timed(lambda: D.all_reduce(small, op=D.ReduceOp.SUM, group=suite.group), 200))
    small_p50 = statistics.median(lat) * 1e3

    # ---- the same collectives on tensors from the library's symmetric memory (zero-copy
    # kernels; explanatory, not part of `value`: the reference has no such allocator) -------
    symmetric = None
    if args.impl == "pdcc" and size > 1 and hasattr(D, "empty_symmetric"):
        t_sym = D.empty_symmetric(suite.count, dtype=torch.float32, group=suite.group)
        if D.is_symmetric(t_sym, suite.group):
            t_sym.fill_(1e-20)
            l_sym = [D.empty_symmetric(suite.shard, dtype=torch.float32, group=suite.group)
                     for _ in range(size)]
            calls = {
                "reduce": lambda: D.reduce(t_sym, dst=0, op=D.ReduceOp.SUM, group=suite.group),
                "all_reduce": lambda: D.all_reduce(t_sym, op=D.ReduceOp.SUM, group=suite.group),
                "all_gather": lambda: D.all_gather(l_sym, suite.ag_in, group=suite.group),
                "broadcast": lambda: D.broadcast(t_sym, src=0, group=suite.group),
            }
            symmetric = {}
            for name, fn in calls.items():
                fn()
                ms = timed(fn, args.detail_iters)
                algbw = suite.nbytes / (ms * 1e-3) / 1e9
                symmetric[name] = {"ms": round(ms, 4), "algbw_GBps": round(algbw, 1),
                                   "busbw_GBps": round(algbw * BUS_FACTOR[name](size), 1)}
            # and one correctness check of the zero-copy all_reduce on the timed buffer
            t_sym.copy_((torch.arange(suite.count, device=dev) % 251).to(torch.float32) + (rank + 1))
            D.all_reduce(t_sym, op=D.ReduceOp.SUM, group=suite.group)
            want = size * (torch.arange(suite.count, device=dev) % 251).to(torch.float32) + size * (size + 1) / 2.0
            symmetric["all_reduce_verified"] = bool(torch.equal(t_sym, want))
            del want

    # ---- verification: the timed call sites, on checkable inputs, on every rank ----------
    mine = suite.verify()
    bad = [k for k, v in mine.items() if not v]
    flag = torch.tensor([0.0 if bad else 1.0], dtype=torch.float64)
    if tdist.get_backend() == "nccl":
        flag = flag.to(dev)
    tdist.all_reduce(flag, op=tdist.ReduceOp.MIN)
    verified = bool(flag.item() == 1.0)
    if bad:
        print(f"[rank {rank}] verification FAILED for: {bad}", file=sys.stderr, flush=True)
    sync_all()

    # ---- end to end ------------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        h2d, d2h = suite.make_host()
        for _ in range(2):
            suite.e2e_step()
        sync_all()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            suite.e2e_step()
        e1.record()
        torch.cuda.synchronize()
        wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        ms_e2e = max_over_ranks(max(e0.elapsed_time(e1) / args.steps, wall_ms))
        h2d_max = int(max_over_ranks(float(h2d)))
        e2e = {"value": round(6 * suite.nbytes / (ms_e2e * 1e-3) / 1e9, 2), "unit": "GB/s",
               "ms_per_step": round(ms_e2e, 3), "h2d_bytes_per_step": h2d_max,
               "d2h_bytes_per_step": d2h}
    sync_all()

    value = 6 * suite.nbytes / (ms_step * 1e-3) / 1e9
    return {
        "metric": "six_collective_algbw_GBps", "value": round(value, 2), "unit": "GB/s",
        "n_gpus": size, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": args.impl,
        "config": {
            "model": "main.py six-collective step (reduce, all_reduce, scatter, gather, "
                     "all_gather, broadcast), root 0, op SUM",
            "bytes_per_collective": suite.nbytes, "global_batch": None, "seq_len": None,
            "parallelism": f"world{size}",
            "l2": "every tensor is larger than the 126 MB L2 (no flush needed)"
                  if suite.nbytes > (126 << 20) else "inputs smaller than L2",
            "value_definition": "6 x bytes_per_collective / step time; busbw per collective "
                                "(nccl-tests factors) is in per_collective",
        },
        "per_collective": per, "small_msg_p50_us": round(small_p50, 2),
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "verified": verified,
        "symmetric_memory": symmetric,
    }


def main() -> None:
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1 and "RANK" not in os.environ:
        # convenience: re-exec under torchrun the way the driver would
        from pdcc_b200.launcher import find_free_port

        os.execvp(sys.executable, [
            sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
            f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
            str(find_free_port()), os.path.abspath(__file__), *sys.argv[1:]])
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if "MASTER_PORT" not in os.environ:
        from pdcc_b200.launcher import find_free_port

        os.environ["MASTER_PORT"] = str(find_free_port())

    import torch

    if not torch.cuda.is_available():
        if args.impl == "reference":
            _unavailable("no CUDA device visible")
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local_rank % torch.cuda.device_count())

    if args.impl == "pdcc":
        import pdcc_b200 as D
        from pdcc_b200 import _native as N

        D.init_process_group("b200", rank=rank, world_size=world)
        result = run(D, rank, world, args, N.launch_count)
        if rank == 0:
            print(json.dumps(result), flush=True)
        D.destroy_process_group()
        if not result["verified"]:
            raise SystemExit(3)
        return

    # ---- reference arm: the UNMODIFIED reference's own init_process, on NCCL ------------
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "main.py")):
        _unavailable("baseline/_ref/main.py missing (run baseline/install_reference.sh)")
    sys.path.insert(0, ref_dir)
    import main as ref  # the reference module, byte-identical to /root/reference/main.py
    import torch.distributed as tdist

    # main.py:92-93 hard-codes MASTER_PORT=29500. Under torchrun the workers would try to
    # reach the *agent's* store on that port; unless that is where the agent listens, let
    # rank 0 host the store itself, as it does when the reference is run as shipped.
    if os.environ.get("MASTER_PORT") != "29500":
        os.environ.pop("TORCHELASTIC_USE_AGENT_STORE", None)
    out = {}

    def fn(r: int, size: int) -> None:
        # reference call shape; the only change is where the tensors live
        out["result"] = run(tdist, r, size, args, lambda: 0)

    ref.init_process(rank, world, fn, backend="nccl")
    if rank == 0:
        out["result"]["gpu_launches"] = None
        print(json.dumps(out["result"]), flush=True)
    tdist.destroy_process_group()
    if not out["result"]["verified"]:
        raise SystemExit(3)


if __name__ == "__main__":
    main()
