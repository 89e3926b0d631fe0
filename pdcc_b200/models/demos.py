"""The reference's seven demo entry points, rebuilt on this library.

Same names, same `fn(rank, size)` shape and the same printed lines as `main.py:9-87`, so
`init_process(rank, size, do_scatter)` keeps working for a user of the reference; the bodies
are written against a few local helpers instead of being seven hand-unrolled scripts.

Every demo builds its tensors on `device()`: CPU under backend "gloo" (what the reference
does), the current CUDA device under backend "b200", so one set of demos is both the CPU
plumbing oracle and the GPU smoke test. Expected output: README.md:76-81, 105-110, 140-145,
175-180, 211-213, 245-250, 279-284 (re-observed on gloo).
"""
from __future__ import annotations

from typing import Callable, Dict, List

import torch

from .. import distributed as dist

ROOT = 0  # every rooted demo of the reference uses rank 0 (main.py:14,37,52,81)


def device() -> torch.device:
    """Where demo tensors live: the GPU of this rank on the native backend, else the CPU."""
    on_gpu = dist.is_initialized() and dist.get_backend() == "b200" and torch.cuda.is_available()
    return torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")


def _everyone(size: int):
    """The full-world sub-group each demo creates first (main.py:11,21,31,46,63,75)."""
    return dist.new_group(list(range(size)))


def _scalar(value: float) -> torch.Tensor:
    return torch.full((1,), float(value), dtype=torch.float32, device=device())


def _blank() -> torch.Tensor:
    return torch.empty(1, dtype=torch.float32, device=device())


def _blanks(count: int) -> List[torch.Tensor]:
    return [_blank() for _ in range(count)]


def _report(rank: int, what) -> None:
    print(f"[{rank}] data = {what}")


def do_reduce(rank: int, size: int):
    """main.py:9-17: ones summed onto rank 0; only rank 0's value (= size) is a contract."""
    everyone = _everyone(size)
    value = _scalar(1)
    dist.reduce(value, dst=ROOT, op=dist.ReduceOp.SUM, group=everyone)
    _report(rank, value[0])
    return value


def do_all_reduce(rank: int, size: int):
    """main.py:19-26: ones summed in place everywhere; every rank prints `size`."""
    everyone = _everyone(size)
    value = _scalar(1)
    dist.all_reduce(value, op=dist.ReduceOp.SUM, group=everyone)
    _report(rank, value[0])
    return value


def do_scatter(rank: int, size: int):
    """main.py:29-41: rank 0 deals 1..size, rank r ends up holding r + 1."""
    everyone = _everyone(size)
    received = _blank()
    deck = [_scalar(i + 1) for i in range(size)] if rank == ROOT else []
    dist.scatter(received, scatter_list=deck, src=ROOT, group=everyone)
    _report(rank, received[0])
    return received


def do_gather(rank: int, size: int):
    """main.py:44-58: every rank sends its number; only rank 0 collects and prints."""
    everyone = _everyone(size)
    mine = _scalar(rank)
    collected = _blanks(size) if rank == ROOT else []
    dist.gather(mine, gather_list=collected, dst=ROOT, group=everyone)
    if rank == ROOT:
        _report(rank, collected)
        return collected
    return None


def do_all_gather(rank: int, size: int):
    """main.py:61-70: every rank ends with [0., 1., ..., size-1] and prints it."""
    everyone = _everyone(size)
    collected = _blanks(size)
    dist.all_gather(collected, _scalar(rank), group=everyone)
    _report(rank, collected)
    return collected


def do_broadcast(rank: int, size: int):
    """main.py:73-83: rank 0's tensor([0.]) lands on every rank."""
    everyone = _everyone(size)
    value = _scalar(rank) if rank == ROOT else _blank()
    dist.broadcast(value, src=ROOT, group=everyone)
    _report(rank, value)
    return value


def hello_world(rank: int, size: int):
    """main.py:86-87: proves the launcher without any collective."""
    print(f"[{rank}] say hi!")


DEMOS: Dict[str, Callable[[int, int], object]] = {
    fn.__name__: fn
    for fn in (do_reduce, do_all_reduce, do_scatter, do_gather, do_all_gather, do_broadcast,
               hello_world)
}
