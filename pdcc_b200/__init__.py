"""pdcc_b200 — B200-native collective communication with the torch.distributed call shape
of FrancescoSaverioZuppichini/pytorch-distributed-collective-communication (main.py).

    import pdcc_b200 as dist
    dist.init_process(rank, size, fn, backend="b200")      # main.py:90-95
    group = dist.new_group(list(range(size)))              # main.py:11
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)  # main.py:23

Layout: `distributed` (API, validation, routing), `ops` (native kernels' wrappers),
`parallel` (bucketed data-parallel gradient averaging), `models` (the seven demos and the
DP MLP), `utils` (timing, clocks), `_native` (libpdcc.so + ctypes binding).
"""
from .distributed import (  # noqa: F401
    Group, GroupMember, ReduceOp, Work, all_gather, all_reduce, all_reduce_avg, barrier,
    broadcast, destroy_process_group, gather, get_backend, get_global_rank, get_group_rank,
    get_rank, get_world_size, init_process_group, is_initialized, native_comm, new_group, reduce,
    scatter, empty_symmetric, is_symmetric,
)
from .launcher import find_free_port, init_process, launch  # noqa: F401

__version__ = "0.1.0"
