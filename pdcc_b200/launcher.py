"""L3/L4 of the reference: process bootstrap and the spawn-per-rank launcher.

`init_process(rank, size, fn, backend)` keeps the reference's signature (main.py:90-95);
`launch(fn, size, backend)` is its `__main__` block (main.py:98-108) turned into a
function that also checks exit codes, kills the siblings of a crashed rank, and can pick
a free port (the reference hard-codes 29500, so two concurrent runs collide).
"""
from __future__ import annotations

import os
import socket
import sys
import time
from typing import Callable, Optional, Sequence

import torch.multiprocessing as mp

from . import distributed as dist


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def init_process(rank: int, size: int, fn: Callable[[int, int], None], backend: str = "gloo") -> None:
    """Initialize the distributed environment, then run `fn(rank, size)` (main.py:90-95).

    MASTER_ADDR / MASTER_PORT default to the reference's 127.0.0.1:29500 but an existing
    environment value wins. With backend "b200" the rank also selects its GPU
    (rank modulo the visible device count), which the reference never does.
    """
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if backend.lower() in dist.NATIVE_BACKENDS:
        import torch

        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)) % torch.cuda.device_count())
    dist.init_process_group(backend, rank=rank, world_size=size)
    try:
        fn(rank, size)
    finally:
        dist.destroy_process_group()


def _entry(rank: int, size: int, fn, backend: str, env: dict) -> None:
    os.environ.update(env)
    init_process(rank, size, fn, backend)


def launch(fn: Callable[[int, int], None], size: int = 4, backend: str = "gloo",
           master_port: Optional[int] = None, timeout: Optional[float] = None,
           env: Optional[dict] = None) -> Sequence[int]:
    """Spawns `size` processes running `init_process(rank, size, fn, backend)`.

    Returns the exit codes (all zero) or raises RuntimeError naming the first rank that
    failed, after terminating the others (the reference joins blindly, main.py:107-108).
    """
    ctx = mp.get_context("spawn")
    child_env = {"MASTER_ADDR": "127.0.0.1",
                 "MASTER_PORT": str(master_port if master_port is not None else find_free_port())}
    child_env.update(env or {})
    procs = []
    for rank in range(size):
        p = ctx.Process(target=_entry, args=(rank, size, fn, backend, child_env))
        p.start()
        procs.append(p)
    deadline = None if timeout is None else time.monotonic() + timeout
    timed_out = False
    while any(p.is_alive() for p in procs):
        if any(p.exitcode not in (0, None) for p in procs):
            break
        if deadline is not None and time.monotonic() > deadline:
            timed_out = True
            break
        time.sleep(0.05)
    if timed_out or any(p.exitcode not in (0, None) for p in procs):
        # A dead rank takes its peers down with connection errors a moment later; give
        # them that moment so the report names every rank that failed, first cause included.
        time.sleep(0.5)
        failed = {r: p.exitcode for r, p in enumerate(procs) if p.exitcode not in (0, None)}
        for p in procs:
            if p.is_alive():
                p.terminate()
        for p in procs:
            p.join(5)
            if p.is_alive():
                p.kill()
        if timed_out and not failed:
            raise RuntimeError(f"launch timed out after {timeout}s")
        raise RuntimeError("; ".join(f"rank {r} failed with exit code {c}"
                                     for r, c in sorted(failed.items())))
    for p in procs:
        p.join()
    return [p.exitcode for p in procs]


def main(argv: Optional[Sequence[str]] = None) -> int:
    """`python -m pdcc_b200.launcher --fn do_scatter --size 4 --backend gloo`"""
    import argparse

    from .models import demos

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--fn", default="do_scatter", choices=sorted(demos.DEMOS))
    ap.add_argument("--size", type=int, default=4)
    ap.add_argument("--backend", default="gloo")
    ap.add_argument("--port", type=int, default=None)
    args = ap.parse_args(argv)
    launch(demos.DEMOS[args.fn], args.size, args.backend, master_port=args.port)
    return 0


if __name__ == "__main__":
    sys.exit(main())
