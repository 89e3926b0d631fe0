"""The reference's seven demo functions (main.py:9-87), same names, same call shape,
same printed lines — importing this library instead of torch.distributed.

Each `fn(rank, size)` builds its tensors on `device()`: CPU under backend "gloo" (what
the reference does), `cuda:<current>` under backend "b200", so one set of demos is the
CPU plumbing oracle and the GPU smoke test. Expected output: README.md:76-81, 105-110,
140-145, 175-180, 211-213, 245-250, 279-284 (re-observed on gloo).
"""
from __future__ import annotations

import torch

from .. import distributed as dist


def device() -> torch.device:
    if dist.is_initialized() and dist.get_backend() == "b200" and torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def do_reduce(rank: int, size: int):
    """main.py:9-17 — only rank 0 is defined to hold `size` afterwards."""
    group = dist.new_group(list(range(size)))
    tensor = torch.ones(1, device=device())
    dist.reduce(tensor, dst=0, op=dist.ReduceOp.SUM, group=group)
    print(f"[{rank}] data = {tensor[0]}")
    return tensor


def do_all_reduce(rank: int, size: int):
    """main.py:19-26 — every rank prints `size`."""
    group = dist.new_group(list(range(size)))
    tensor = torch.ones(1, device=device())
    dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)
    print(f"[{rank}] data = {tensor[0]}")
    return tensor


def do_scatter(rank: int, size: int):
    """main.py:29-41 — rank r ends with r + 1."""
    group = dist.new_group(list(range(size)))
    tensor = torch.empty(1, device=device())
    if rank == 0:
        tensor_list = [torch.tensor([i + 1], dtype=torch.float32, device=device())
                       for i in range(size)]
        dist.scatter(tensor, scatter_list=tensor_list, src=0, group=group)
    else:
        dist.scatter(tensor, scatter_list=[], src=0, group=group)
    print(f"[{rank}] data = {tensor[0]}")
    return tensor


def do_gather(rank: int, size: int):
    """main.py:44-58 — only rank 0 prints [0., 1., ..., size-1]."""
    group = dist.new_group(list(range(size)))
    tensor = torch.tensor([rank], dtype=torch.float32, device=device())
    tensor_list = None
    if rank == 0:
        tensor_list = [torch.empty(1, device=device()) for _ in range(size)]
        dist.gather(tensor, gather_list=tensor_list, dst=0, group=group)
    else:
        dist.gather(tensor, gather_list=[], dst=0, group=group)
    if rank == 0:
        print(f"[{rank}] data = {tensor_list}")
    return tensor_list


def do_all_gather(rank: int, size: int):
    """main.py:61-70 — every rank prints [0., 1., ..., size-1]."""
    group = dist.new_group(list(range(size)))
    tensor = torch.tensor([rank], dtype=torch.float32, device=device())
    tensor_list = [torch.empty(1, device=device()) for _ in range(size)]
    dist.all_gather(tensor_list, tensor, group=group)
    print(f"[{rank}] data = {tensor_list}")
    return tensor_list


def do_broadcast(rank: int, size: int):
    """main.py:73-83 — every rank prints rank 0's tensor([0.])."""
    group = dist.new_group(list(range(size)))
    if rank == 0:
        tensor = torch.tensor([rank], dtype=torch.float32, device=device())
    else:
        tensor = torch.empty(1, device=device())
    dist.broadcast(tensor, src=0, group=group)
    print(f"[{rank}] data = {tensor}")
    return tensor


def hello_world(rank: int, size: int):
    """main.py:86-87 — proves the launcher without any collective."""
    print(f"[{rank}] say hi!")


DEMOS = {f.__name__: f for f in (do_reduce, do_all_reduce, do_scatter, do_gather, do_all_gather,
                                 do_broadcast, hello_world)}
