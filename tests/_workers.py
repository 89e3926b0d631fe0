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
