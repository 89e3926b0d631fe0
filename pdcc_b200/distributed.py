"""Public API: the `torch.distributed` call shape of the reference, B200-native underneath.

    import pdcc_b200 as dist
    dist.init_process_group("b200", rank=rank, world_size=size)
    group = dist.new_group(list(range(size)))
    dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)

Routing is a plain branch, not a backend registry: a CUDA tensor goes to the native
sm_100a kernels (pdcc_b200/ops/collectives.py -> libpdcc.so); a CPU tensor goes to
torch's gloo process group, which is kept as the plumbing/oracle path.

Semantics preserved from `[torch] distributed/distributed_c10d.py` (the only layer the
reference talks to, main.py:11-81), each verified against gloo in tests/:
  * src/dst are GLOBAL ranks, translated to group ranks (:1183-1210)
  * new_group sorts ranks, rejects out-of-range ones and hands non-members the
    NON_GROUP_MEMBER sentinel (:5507, :5516-5521, :1988-1997)
  * a collective called by a non-member warns and returns None, tensor untouched
    (:1036-1049)
  * all tensors of one call share a dtype, else ValueError (:1221-1240)
  * scatter_list/gather_list must be given on the root and must not be given elsewhere
    (:4488-4499, :4317-4326)
  * synchronous by default, `async_op=True` returns a work handle (:3070-3075)
"""
from __future__ import annotations

import enum
import os
import warnings
from typing import List, Optional, Sequence

import torch
import torch.distributed as tdist

from . import _native as N
from .ops import collectives as K

__all__ = [
    "ReduceOp", "GroupMember", "Group", "Work", "init_process_group", "destroy_process_group",
    "is_initialized", "new_group", "get_rank", "get_world_size", "get_backend", "get_group_rank",
    "get_global_rank", "reduce", "all_reduce", "all_reduce_avg", "scatter", "gather", "all_gather",
    "broadcast", "barrier", "native_comm", "empty_symmetric", "is_symmetric",
]

NATIVE_BACKENDS = ("b200", "native", "nvlink")


class ReduceOp(enum.IntEnum):
    """Numbering of c10d::ReduceOp ([torch] .../c10d/Types.hpp:37-48). The reference names
    SUM, PRODUCT, MAX, MIN (main.py:14-15); AVG backs the fused gradient average."""
    SUM = 0
    AVG = 1
    PRODUCT = 2
    MIN = 3
    MAX = 4


class GroupMember:
    WORLD: Optional["Group"] = None
    NON_GROUP_MEMBER = -100


class Group:
    """A set of global ranks with this process's position in it."""

    def __init__(self, gid: int, ranks: Sequence[int], my_global_rank: int, torch_pg):
        self.gid = gid
        self.ranks = tuple(ranks)
        self._index = {g: i for i, g in enumerate(self.ranks)}
        self.global_rank = my_global_rank
        self.group_rank = self._index.get(my_global_rank, -1)
        self.torch_pg = torch_pg  # gloo group for CPU tensors
        self.comm: Optional[K.NativeComm] = None  # created on first CUDA collective

    def size(self) -> int:
        return len(self.ranks)

    def rank(self) -> int:
        return self.group_rank

    def to_group_rank(self, global_rank: int) -> int:
        try:
            return self._index[global_rank]
        except KeyError:
            raise ValueError(f"Global rank {global_rank} is not part of group {self.ranks}") from None

    def __repr__(self) -> str:
        return f"<pdcc_b200.Group gid={self.gid} ranks={list(self.ranks)}>"


class Work:
    """Completion handle of an `async_op=True` collective.

    Native collectives are stream ordered: the handle owns a CUDA event recorded right
    after the kernel, and `wait()` makes the *current stream* wait for it (same contract
    as ProcessGroupNCCL's work; the host does not block). gloo handles wrap torch's."""

    def __init__(self, event: Optional[torch.cuda.Event] = None, torch_work=None):
        self._event = event
        self._torch_work = torch_work

    def wait(self, timeout=None) -> bool:
        if self._torch_work is not None:
            self._torch_work.wait()
        elif self._event is not None:
            torch.cuda.current_stream().wait_event(self._event)
        return True

    def is_completed(self) -> bool:
        if self._torch_work is not None:
            return self._torch_work.is_completed()
        return self._event is None or self._event.query()

    def synchronize(self) -> None:
        if self._event is not None:
            self._event.synchronize()
        elif self._torch_work is not None:
            self._torch_work.wait()


class _State:
    def __init__(self):
        self.backend: Optional[str] = None
        self.rank = -1
        self.world_size = 0
        self.store: Optional[N.Store] = None
        self.groups: List[Group] = []
        self.next_gid = 0
        self.job = ""


_S = _State()


# ------------------------------------------------------------------------- lifecycle ---
def is_initialized() -> bool:
    return _S.backend is not None


def _require_init() -> None:
    if not is_initialized():
        raise ValueError("Default process group has not been initialized, please make sure to "
                         "call init_process_group.")


def init_process_group(backend: str = "gloo", rank: int = -1, world_size: int = -1,
                       init_method: Optional[str] = None, timeout=None) -> None:
    """Role of `dist.init_process_group(backend, rank=rank, world_size=size)` at main.py:94.

    backend "gloo": CPU tensors only (the reference's default, main.py:90).
    backend "b200": CUDA tensors run on the native kernels; CPU tensors still use gloo.
    Rendezvous is env:// on MASTER_ADDR / MASTER_PORT like the reference's.
    """
    if is_initialized():
        raise ValueError("trying to initialize the default process group twice!")
    backend = backend.lower()
    if backend == "nccl":
        raise ValueError("this library does not route anything through NCCL; use backend='b200' "
                         "for CUDA tensors (or 'gloo' for the CPU path)")
    if backend != "gloo" and backend not in NATIVE_BACKENDS:
        raise ValueError(f"unknown backend {backend!r}; expected 'gloo' or 'b200'")
    if rank < 0:
        rank = int(os.environ["RANK"])
    if world_size < 0:
        world_size = int(os.environ["WORLD_SIZE"])
    kwargs = {}
    if timeout is not None:
        kwargs["timeout"] = timeout
    if init_method is not None:
        kwargs["init_method"] = init_method
    tdist.init_process_group("gloo", rank=rank, world_size=world_size, **kwargs)
    _S.backend = "b200" if backend in NATIVE_BACKENDS else "gloo"
    _S.rank, _S.world_size = rank, world_size

    if _S.backend == "b200":
        # R1: rank 0 hosts the bootstrap store of the native control plane on a free port
        # and tells the others which one.
        host = os.environ.get("MASTER_ADDR", "127.0.0.1")
        tstore = tdist.distributed_c10d._get_default_store()
        if rank == 0:
            _S.store = N.Store(host, int(os.environ.get("PDCC_STORE_PORT", "0")), serve=True)
            tstore.set("pdcc/store_port", str(_S.store.port))
        else:
            port = int(tstore.get("pdcc/store_port").decode())
            _S.store = N.Store(host, port)
        _S.job = f"j{_S.store.port}"

    world = Group(_S.next_gid, range(world_size), rank, tdist.group.WORLD)
    _S.next_gid += 1
    _S.groups.append(world)
    GroupMember.WORLD = world


def destroy_process_group(group: Optional[Group] = None) -> None:
    """Tears down one group, or everything when called with no argument. The reference
    never does this (main.py:90-95); exiting without it is also safe."""
    if not is_initialized():
        return
    def _quiesce(pg) -> None:
        # A peer's kernel may still be posting its last flags into this rank's block: every
        # member finishes its device work and meets at a host barrier BEFORE anyone unmaps or
        # frees (on the legacy-IPC path the owner's cudaFree could otherwise hand the memory
        # to the next cudaMalloc while a peer still writes into it).
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        try:
            tdist.barrier(group=pg)
        except Exception:
            pass

    def _report(comm) -> None:
        # last chance to tell the user that some kernel gave up waiting for a peer
        try:
            comm.check()
        except Exception as e:  # noqa: BLE001 - teardown must go on
            warnings.warn(f"pdcc_b200: {e}")

    if group is not None and group is not GroupMember.WORLD:
        if isinstance(group, Group):
            if group.comm is not None:
                _quiesce(group.torch_pg)
                _report(group.comm)
                group.comm.destroy()
                group.comm = None
            if group in _S.groups:
                _S.groups.remove(group)
        return
    if any(g.comm is not None for g in _S.groups):
        _quiesce(None)
    for g in _S.groups:
        if g.comm is not None:
            _report(g.comm)
            g.comm.destroy()
            g.comm = None
    if _S.store is not None:
        # everyone must be done with the store before rank 0 takes the server down
        try:
            tdist.barrier()
        except Exception:
            pass
        _S.store.close()
    tdist.destroy_process_group()
    GroupMember.WORLD = None
    _S.__init__()


def get_rank(group: Optional[Group] = None) -> int:
    _require_init()
    if group is None:
        return _S.rank
    if group == GroupMember.NON_GROUP_MEMBER:
        return -1
    return group.group_rank


def get_world_size(group: Optional[Group] = None) -> int:
    _require_init()
    if group is None:
        return _S.world_size
    if group == GroupMember.NON_GROUP_MEMBER:
        return -1
    return group.size()


def get_backend(group: Optional[Group] = None) -> str:
    _require_init()
    return _S.backend


def get_group_rank(group: Group, global_rank: int) -> int:
    return group.to_group_rank(global_rank)


def get_global_rank(group: Group, group_rank: int) -> int:
    return group.ranks[group_rank]


def new_group(ranks: Optional[Sequence[int]] = None, timeout=None, backend=None):
    """Role of `dist.new_group(list(range(size)))` (main.py:11,21,31,46,63,75).

    Collective over the WORLD: every rank must call it, in the same order, members or not.
    Returns a Group, or GroupMember.NON_GROUP_MEMBER to ranks outside `ranks`."""
    _require_init()
    world = _S.world_size
    if ranks is None:
        ranks = list(range(world))
    ranks = sorted(int(r) for r in ranks)
    if len(ranks) == 0:
        raise ValueError("new_group needs at least one rank")
    if len(set(ranks)) != len(ranks):
        raise ValueError("the new group's ranks must be unique")
    if len(ranks) > world:
        raise ValueError("the new group's world size should be less or equal to the world size "
                         "set by init_process_group")
    for r in ranks:
        if r < 0 or r >= world:
            raise ValueError("The new group's rank should be within the world_size set by "
                             "init_process_group")
    kwargs = {} if timeout is None else {"timeout": timeout}
    torch_pg = tdist.new_group(ranks, backend="gloo", **kwargs)
    gid = _S.next_gid
    _S.next_gid += 1
    if _S.rank not in ranks:
        return GroupMember.NON_GROUP_MEMBER
    g = Group(gid, ranks, _S.rank, torch_pg)
    _S.groups.append(g)
    return g


# ------------------------------------------------------------------------- helpers -----
def _resolve(group) -> Optional[Group]:
    """Group to run on, or None when this rank is not a member (warn + no-op)."""
    _require_init()
    if group is None:
        return GroupMember.WORLD
    if isinstance(group, Group):
        return group if group.group_rank >= 0 else None
    if group == GroupMember.NON_GROUP_MEMBER:
        return None
    raise TypeError(f"Invalid process group specified: {group!r}")


def _warn_not_in_group(op_name: str) -> None:
    warnings.warn(f"Running {op_name} on global rank {_S.rank} which does not belong to the "
                  "given group.")


def _check_single_tensor(param, name: str) -> None:
    if not isinstance(param, torch.Tensor):
        raise TypeError(f"Invalid function argument. Expected parameter `{name}` of type "
                        f"torch.Tensor but got {type(param)} instead.")


def _check_tensor_list(param, name: str) -> None:
    if not isinstance(param, list) or not all(isinstance(p, torch.Tensor) for p in param):
        raise TypeError(f"Invalid function argument. Expected parameter `{name}` of type "
                        f"List[torch.Tensor] but got {type(param)} instead.")


def _ensure_same_dtype(*tensors) -> None:
    flat: List[torch.Tensor] = []
    for t in tensors:
        flat.extend(t if isinstance(t, (list, tuple)) else [t])
    if not flat:
        return
    first = flat[0].dtype
    for t in flat[1:]:
        if t.dtype != first:
            raise ValueError("Invalid usage of tensors with different dtypes"
                             f"Found {first} and  {t.dtype}")


def _ensure_same_shape(tensor: torch.Tensor, tensors: Sequence[torch.Tensor], name: str) -> None:
    for t in tensors:
        if t.numel() != tensor.numel():
            raise ValueError(f"All tensors in `{name}` must have the same number of elements as "
                             f"the single tensor ({t.numel()} vs {tensor.numel()})")


def _root(g: Group, global_root: Optional[int], group_root: Optional[int], name: str) -> int:
    """Global src/dst -> group rank (accepts torch>=2.6's group_src/group_dst too)."""
    if global_root is not None and group_root is not None:
        raise ValueError(f"Can't specify both group_{name} and {name}")
    if group_root is not None:
        if group_root < 0 or group_root >= g.size():
            raise ValueError(f"group_{name} {group_root} out of range for {g}")
        return group_root
    return g.to_group_rank(0 if global_root is None else global_root)


def native_comm(group: Optional[Group] = None) -> K.NativeComm:
    """The group's native communicator, created (collectively, over the store) on first use."""
    g = _resolve(group)
    if g is None:
        raise ValueError("this rank is not a member of the group")
    if g.comm is None:
        if _S.backend != "b200":
            raise RuntimeError("CUDA tensors need the native backend: call "
                               "init_process_group('b200', ...). Backend 'gloo' is the CPU "
                               "plumbing path only.")
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device is visible to this process")
        mc = os.environ.get("PDCC_MULTICAST")
        kind = {"posix_fd": N.HANDLE_POSIX_FD, "legacy_ipc": N.HANDLE_LEGACY_IPC}.get(
            os.environ.get("PDCC_HANDLE", ""), N.HANDLE_AUTO)
        g.comm = K.NativeComm(
            _S.store, f"{_S.job}g{g.gid}", g.group_rank, g.size(), torch.cuda.current_device(),
            slot_bytes=int(os.environ.get("PDCC_SLOT_BYTES", "0")),
            max_blocks=int(os.environ.get("PDCC_RINGS", "0")),
            heap_bytes=int(os.environ.get("PDCC_HEAP_BYTES", "0")),
            multicast=-1 if mc is None else int(mc), handle_kind=kind,
            device_timeout_ms=int(os.environ.get("PDCC_DEVICE_TIMEOUT_MS", "0")))
    return g.comm


def empty_symmetric(*shape, dtype: torch.dtype = torch.float32, group=None) -> torch.Tensor:
    """A CUDA tensor in the group's symmetric memory. COLLECTIVE: every member of the group
    calls it with the same shape and dtype, in the same order (like `new_group`).

    Collectives on such tensors skip every staging copy: all_reduce / reduce / broadcast /
    all_gather run as multimem instructions straight on the tensor (`csrc/sym.cuh`), which is
    what gradient buckets and other long-lived communication buffers should live in. The
    contract is NCCL's registered-buffer contract: in one call either every rank passes
    tensors of the same symmetric allocation (same offset), or none does. On a group of one,
    or without NVSwitch multicast, an ordinary tensor is returned (nothing to gain).
    No counterpart in the reference (its tensors are 4-byte CPU tensors, main.py:12)."""
    if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
        shape = tuple(shape[0])
    g = _resolve(group)
    if g is None:
        raise ValueError("this rank is not a member of the group")
    if g.size() == 1 or _S.backend != "b200":
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        return torch.empty(*shape, dtype=dtype, device=dev)
    comm = native_comm(g)
    if not comm.has_multicast:
        return torch.empty(*shape, dtype=dtype, device=torch.device("cuda", comm.device))
    return comm.empty_symmetric(shape, dtype)


def is_symmetric(tensor: torch.Tensor, group=None) -> bool:
    g = _resolve(group)
    return bool(g is not None and g.comm is not None and tensor.is_cuda and g.comm.is_symmetric(tensor))


def _finish(async_op: bool, cuda: bool, torch_work=None) -> Optional[Work]:
    if not async_op:
        return None
    if cuda:
        ev = torch.cuda.Event()
        ev.record()
        return Work(event=ev)
    return Work(torch_work=torch_work)


_TORCH_OPS = {
    ReduceOp.SUM: tdist.ReduceOp.SUM, ReduceOp.AVG: tdist.ReduceOp.AVG,
    ReduceOp.PRODUCT: tdist.ReduceOp.PRODUCT, ReduceOp.MIN: tdist.ReduceOp.MIN,
    ReduceOp.MAX: tdist.ReduceOp.MAX,
}


def _redop(op) -> ReduceOp:
    try:
        return ReduceOp(int(op))
    except (ValueError, TypeError):
        raise ValueError(f"unsupported reduce op {op!r}") from None


def _gloo_wait(work, async_op: bool):
    if async_op:
        return Work(torch_work=work)
    if work is not None:
        work.wait()
    return None


# ------------------------------------------------------------------------- collectives -
def all_reduce(tensor, op=ReduceOp.SUM, group=None, async_op: bool = False):
    """In-place reduction whose result lands on every rank, bitwise identical (main.py:23)."""
    _check_single_tensor(tensor, "tensor")
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("all_reduce")
        return None
    op = _redop(op)
    if tensor.is_cuda:
        if g.size() > 1 and tensor.numel():
            (g.comm or native_comm(g)).all_reduce(tensor, int(op))
        return _finish(async_op, True) if async_op else None
    if op == ReduceOp.AVG:  # gloo has no AVG: sum then divide
        work = tdist.all_reduce(tensor, op=tdist.ReduceOp.SUM, group=g.torch_pg, async_op=False)
        tensor.div_(g.size())
        return _finish(async_op, False, None)
    work = tdist.all_reduce(tensor, op=_TORCH_OPS[op], group=g.torch_pg, async_op=True)
    return _gloo_wait(work, async_op)


def all_reduce_avg(tensor, out=None, group=None, async_op: bool = False, scale=None,
                   nblocks: int = 0, algo=None):
    """K2 — the README's "send all gradients to all GPUs, average them" (README.md:5) as one
    kernel: sum over ranks, multiply by 1/world_size (or `scale`), optionally write fp32
    from a bf16/fp16 input. `out=None` averages in place."""
    _check_single_tensor(tensor, "tensor")
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("all_reduce_avg")
        return None
    s = (1.0 / g.size()) if scale is None else float(scale)
    if tensor.is_cuda:
        if not tensor.numel():
            return _finish(async_op, True)
        if g.size() > 1:
            # nblocks caps the CTAs of the kernel: an average that overlaps a backward pass
            # should leave most SMs to the GEMMs (parallel/data_parallel.py)
            native_comm(g).all_reduce(tensor, K.SUM, out=out, scale=s, nblocks=nblocks, algo=algo)
        elif out is not None or s != 1.0:
            K.local_scale_cast(tensor if out is None else out, tensor, s)
        return _finish(async_op, True)
    tdist.all_reduce(tensor, op=tdist.ReduceOp.SUM, group=g.torch_pg)
    if out is None:
        tensor.mul_(s)
    else:
        out.copy_(tensor.to(out.dtype) * s)
    return _finish(async_op, False)


def reduce(tensor, dst: Optional[int] = None, op=ReduceOp.SUM, group=None, async_op: bool = False,
           group_dst: Optional[int] = None):
    """Reduction whose result lands on global rank `dst` only (main.py:14). Other ranks'
    tensors are left untouched (gloo's partial sums there are an artefact)."""
    _check_single_tensor(tensor, "tensor")
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("reduce")
        return None
    op = _redop(op)
    root = _root(g, dst, group_dst, "dst")
    if tensor.is_cuda:
        if tensor.numel() and g.size() > 1:
            native_comm(g).reduce(tensor, root, int(op))
        return _finish(async_op, True)
    work = tdist.reduce(tensor, dst=g.ranks[root], op=_TORCH_OPS[op], group=g.torch_pg,
                        async_op=True)
    return _gloo_wait(work, async_op)


def broadcast(tensor, src: Optional[int] = None, group=None, async_op: bool = False,
              group_src: Optional[int] = None):
    """Copies global rank `src`'s tensor into every rank's tensor (main.py:81)."""
    _check_single_tensor(tensor, "tensor")
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("broadcast")
        return None
    root = _root(g, src, group_src, "src")
    if tensor.is_cuda:
        if tensor.numel() and g.size() > 1:
            native_comm(g).broadcast(tensor, root)
        return _finish(async_op, True)
    work = tdist.broadcast(tensor, src=g.ranks[root], group=g.torch_pg, async_op=True)
    return _gloo_wait(work, async_op)


def scatter(tensor, scatter_list=None, src: Optional[int] = None, group=None,
            async_op: bool = False, group_src: Optional[int] = None):
    """Rank i receives `scatter_list[i]` of global rank `src` (main.py:37,39)."""
    _check_single_tensor(tensor, "tensor")
    if scatter_list:
        _check_tensor_list(scatter_list, "scatter_list")
    else:
        scatter_list = []
    _ensure_same_dtype(tensor, scatter_list)
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("scatter")
        return None
    root = _root(g, src, group_src, "src")
    if g.group_rank == root:
        if not scatter_list:
            raise ValueError("Argument ``scatter_list`` must be specified on source rank.")
        if len(scatter_list) != g.size():
            raise ValueError(f"scatter_list must hold one tensor per group rank "
                             f"({len(scatter_list)} given, group size {g.size()})")
        _ensure_same_shape(tensor, scatter_list, "scatter_list")
    elif scatter_list:
        raise ValueError("Argument ``scatter_list`` must NOT be specified on non-source ranks.")
    if tensor.is_cuda:
        if tensor.numel():
            if g.size() > 1:
                native_comm(g).scatter(tensor, scatter_list, root)
            else:
                K.local_copy(tensor, scatter_list[0].contiguous())
        return _finish(async_op, True)
    work = tdist.scatter(tensor, scatter_list=scatter_list or None, src=g.ranks[root],
                         group=g.torch_pg, async_op=True)
    return _gloo_wait(work, async_op)


def gather(tensor, gather_list=None, dst: Optional[int] = None, group=None, async_op: bool = False,
           group_dst: Optional[int] = None):
    """Global rank `dst` receives every rank's tensor into `gather_list` (main.py:52,54)."""
    _check_single_tensor(tensor, "tensor")
    if gather_list:
        _check_tensor_list(gather_list, "gather_list")
    else:
        gather_list = []
    _ensure_same_dtype(tensor, gather_list)
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("gather")
        return None
    root = _root(g, dst, group_dst, "dst")
    if g.group_rank == root:
        if not gather_list:
            raise ValueError("Argument ``gather_list`` must be specified on destination rank.")
        if len(gather_list) != g.size():
            raise ValueError(f"gather_list must hold one tensor per group rank "
                             f"({len(gather_list)} given, group size {g.size()})")
        _ensure_same_shape(tensor, gather_list, "gather_list")
    elif gather_list:
        raise ValueError("Argument ``gather_list`` must NOT be specified on non-destination ranks.")
    if tensor.is_cuda:
        if tensor.numel():
            if g.size() > 1:
                native_comm(g).gather(tensor, gather_list, root)
            else:
                K.local_copy(gather_list[0], tensor.contiguous())
        return _finish(async_op, True)
    work = tdist.gather(tensor, gather_list=gather_list or None, dst=g.ranks[root],
                        group=g.torch_pg, async_op=True)
    return _gloo_wait(work, async_op)


def all_gather(tensor_list, tensor, group=None, async_op: bool = False):
    """Every rank receives every rank's tensor into `tensor_list` (main.py:68)."""
    _check_tensor_list(tensor_list, "tensor_list")
    _check_single_tensor(tensor, "tensor")
    _ensure_same_dtype(tensor_list, tensor)
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("all_gather")
        return None
    if len(tensor_list) != g.size():
        raise ValueError(f"tensor_list must hold one tensor per group rank "
                         f"({len(tensor_list)} given, group size {g.size()})")
    _ensure_same_shape(tensor, tensor_list, "tensor_list")
    if tensor.is_cuda:
        if tensor.numel():
            if g.size() > 1:
                native_comm(g).all_gather(tensor_list, tensor)
            else:
                K.local_copy(tensor_list[0], tensor.contiguous())
        return _finish(async_op, True)
    work = tdist.all_gather(tensor_list, tensor, group=g.torch_pg, async_op=True)
    return _gloo_wait(work, async_op)


def barrier(group=None, async_op: bool = False, device_ids=None):
    """Host barrier on the gloo path; with backend 'b200' and a CUDA device selected it is a
    device-side flag barrier enqueued on the current stream (no host block)."""
    g = _resolve(group)
    if g is None:
        _warn_not_in_group("barrier")
        return None
    if _S.backend == "b200" and torch.cuda.is_available() and g.size() > 1 and g.comm is not None:
        g.comm.barrier()
        return _finish(async_op, True)
    work = tdist.barrier(group=g.torch_pg, async_op=True)
    return _gloo_wait(work, async_op)
