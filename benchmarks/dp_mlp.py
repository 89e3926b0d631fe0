#!/usr/bin/env python
"""BASELINE.json config 5: data-parallel 4-layer MLP (hidden 8192, bf16), gradients averaged
by the fused all_reduce x 1/world kernel, bucket by bucket, overlapped with backward.

  torchrun --nproc-per-node 8 benchmarks/dp_mlp.py --steps 20
  python benchmarks/dp_mlp.py --gpus 2 --mode all

Modes: `overlap` (side stream, buckets in symmetric memory: zero-copy fused average),
`staged` (same, buckets in ordinary memory), `serial` (symmetric buckets, no overlap), `ddp`
(torch.nn.parallel.DistributedDataParallel on NCCL: bucketed and overlapped, the real
baseline) and `nccl` (README.md:5 done naively: all_reduce(SUM) per gradient on NCCL, then
divide). Reports ms/step (device-timed, max over ranks) and the exposed communication
time = step time minus the step time of the same model with averaging off.
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--batch", type=int, default=4096, help="per-GPU batch (weak scaling)")
    ap.add_argument("--mode", default="all", choices=["overlap", "staged", "serial", "ddp", "nccl", "all"])
    ap.add_argument("--comm-ctas", default="0", help="comma list: CTAs per averaging kernel in the "
                    "overlap mode (0 = library default); more than one value adds overlap_c<N> rows")
    ap.add_argument("--bucket-mb", type=int, default=128)
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
    from pdcc_b200.models.mlp import MLP, DataParallelTrainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)) % torch.cuda.device_count())
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("b200", rank=rank, world_size=world)
    nccl = tdist.new_group(list(range(world)), backend="nccl") if world > 1 else None

    x = torch.randn(args.batch, args.hidden, device=dev, dtype=torch.bfloat16)
    y = torch.randn(args.batch, args.hidden, device=dev, dtype=torch.bfloat16)

    def timed(step) -> float:
        for _ in range(args.warmup):
            step()
        torch.cuda.synchronize()
        tdist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64)
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        return float(t.item())

    results = {}
    modes = ["overlap", "staged", "serial", "ddp", "nccl"] if args.mode == "all" else [args.mode]
    from pdcc_b200.utils.clocks import ClockSampler

    sampler = ClockSampler(torch.cuda.current_device())
    sampler.__enter__()
    # compute-only reference: same model, averaging switched off
    torch.manual_seed(0)
    base = MLP(args.hidden, args.layers, device=dev)
    opt = torch.optim.SGD(base.parameters(), lr=1e-3)

    def compute_only():
        opt.zero_grad(set_to_none=False)
        torch.nn.functional.mse_loss(base(x).float(), y.float()).backward()
        opt.step()

    results["compute_only_ms"] = timed(compute_only)
    del base, opt
    ctas_list = [int(v) for v in args.comm_ctas.split(",")]
    expanded = []
    for mode in modes:
        if mode == "overlap" and len(ctas_list) > 1:
            expanded += [f"overlap_c{c}" for c in ctas_list]
        else:
            expanded.append(mode)
    for mode in expanded:
        torch.manual_seed(0)
        model = MLP(args.hidden, args.layers, device=dev)
        if mode == "nccl":
            opt = torch.optim.SGD(model.parameters(), lr=1e-3)

            def step():
                opt.zero_grad(set_to_none=False)
                torch.nn.functional.mse_loss(model(x).float(), y.float()).backward()
                if nccl is not None:
                    for p in model.parameters():
                        tdist.all_reduce(p.grad, op=tdist.ReduceOp.SUM, group=nccl)
                        p.grad.div_(world)
                opt.step()
        elif mode == "ddp":
            if nccl is None:
                continue
            from torch.nn.parallel import DistributedDataParallel as DDP

            ddp = DDP(model, device_ids=[dev.index], process_group=nccl, bucket_cap_mb=128,
                      gradient_as_bucket_view=True)
            opt = torch.optim.SGD(ddp.parameters(), lr=1e-3)

            def step():
                opt.zero_grad(set_to_none=False)
                torch.nn.functional.mse_loss(ddp(x).float(), y.float()).backward()
                opt.step()
        else:
            ctas = int(mode.split("_c")[1]) if "_c" in mode else ctas_list[0]
            tr = DataParallelTrainer(model, overlap=(mode != "serial"), symmetric=(mode != "staged"),
                                     comm_ctas=ctas, bucket_bytes=args.bucket_mb << 20)

            def step():
                tr.step(x, y)
        results[mode + "_ms"] = timed(step)
        results[mode + "_exposed_comm_ms"] = results[mode + "_ms"] - results["compute_only_ms"]
    sampler.__exit__(None, None, None)
    results["clocks"] = sampler.summary()
    if rank == 0:
        grad_bytes = args.layers * (args.hidden * args.hidden + args.hidden) * 2
        print(json.dumps({"config": {"hidden": args.hidden, "layers": args.layers,
                                     "batch_per_gpu": args.batch, "world": world,
                                     "grad_bytes": grad_bytes, "dtype": "bf16"}, **results}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
