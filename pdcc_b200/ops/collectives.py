"""Native collective ops on CUDA tensors: thin, allocation-free wrappers that turn
torch tensors into `pdcc_op_t` descriptors and enqueue ONE kernel on the current stream.

`NativeComm` is one rank's handle on a group's communicator (symmetric block, peer
mappings, NVLS multicast). `Emulator` runs N virtual ranks on one GPU in one
cooperative launch — the same kernels and flag protocol, used by the single-GPU tests.

Reference parity: these are the CUDA-side equivalents of the six torch.distributed
calls the reference makes (main.py:14,23,37/39,52/54,68,81); argument validation and
global->group rank translation live one layer up in pdcc_b200/distributed.py.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

from .. import _native as N

_DTYPES = {
    torch.float32: N.F32, torch.bfloat16: N.BF16, torch.float16: N.F16, torch.float64: N.F64,
    torch.int32: N.I32, torch.int64: N.I64, torch.uint8: N.U8, torch.int8: N.I8,
    torch.int16: N.I16, torch.bool: N.U8,
}
_REDUCIBLE = {torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int32,
              torch.int64, torch.uint8}

SUM, AVG, PRODUCT, MIN, MAX = 0, 1, 2, 3, 4
_NVTX = os.environ.get("PDCC_NVTX", "0") == "1"
_COLL_NAMES = ("all_reduce", "reduce", "broadcast", "scatter", "gather", "all_gather", "barrier")


def dtype_code(dtype: torch.dtype, reduce: bool = False) -> int:
    if reduce and dtype not in _REDUCIBLE:
        raise TypeError(f"reduction over {dtype} is not supported by the native backend")
    try:
        return _DTYPES[dtype]
    except KeyError:
        raise TypeError(f"dtype {dtype} is not supported by the native backend") from None


def _algo(algo) -> int:
    if algo is None:
        return N.ALGO_AUTO
    if isinstance(algo, str):
        return N.ALGO_NAMES[algo]
    return int(algo)


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream_ptr(stream: Optional[torch.cuda.Stream], device: int = -1) -> int:
    if stream is not None:
        return stream.cuda_stream
    if _raw_stream is not None and device >= 0:
        return _raw_stream(device)  # no Stream object on the hot path
    return torch.cuda.current_stream().cuda_stream


def _fill(op: N.Op, coll: int, count: int, dtype: int, *, out_dtype: Optional[int] = None,
          redop: int = SUM, root: int = 0, algo=None, nblocks: int = 0, scale: float = 1.0,
          in_ptr: int = 0, out_ptr: int = 0, ptrs: Sequence[int] = ()) -> N.Op:
    op.coll = coll
    op.dtype = dtype
    op.out_dtype = dtype if out_dtype is None else out_dtype
    op.redop = redop
    op.root = root
    op.algo = _algo(algo)
    op.nblocks = nblocks
    op.scale = scale
    op.count = count
    op.in_ = in_ptr
    op.out = out_ptr
    if ptrs or getattr(op, "_has_list", True):
        lst = op.list
        for j in range(N.MAX_RANKS):
            lst[j] = ptrs[j] if j < len(ptrs) else None
        op._has_list = bool(ptrs)
    return op


class _RawCudaBytes:
    """`n` bytes of device memory at `ptr`, exposed through __cuda_array_interface__ so that
    torch can alias it without copying. `keep` pins whatever owns the memory."""

    def __init__(self, ptr: int, n: int, keep):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2}
        self._keep = keep


class NativeComm:
    """One rank's communicator for one group (C type pdcc_comm_t)."""

    def __init__(self, store: Optional[N.Store], key: str, rank: int, nranks: int, device: int, *,
                 slot_bytes: int = 0, max_blocks: int = 0, heap_bytes: int = 0,
                 multicast: int = -1, handle_kind: int = N.HANDLE_AUTO, timeout_ms: int = 120000,
                 device_timeout_ms: int = 0):
        if nranks > N.MAX_RANKS:
            raise ValueError(f"native groups span at most {N.MAX_RANKS} ranks (one NVSwitch domain)")
        cfg = N.CommConfig()
        cfg.rank, cfg.nranks, cfg.device = rank, nranks, device
        cfg.handle_kind, cfg.multicast = handle_kind, multicast
        cfg.timeout_ms, cfg.device_timeout_ms = timeout_ms, device_timeout_ms
        cfg.slot_bytes, cfg.max_blocks, cfg.heap_bytes = slot_bytes, max_blocks, heap_bytes
        cfg.key = key.encode()[:63]
        self._h = C.c_void_p()
        L = N.lib()
        N.check(L.pdcc_comm_create(store.handle if store is not None else None, C.byref(cfg),
                                   C.byref(self._h)))
        info = N.CommInfo()
        N.check(L.pdcc_comm_info(self._h, C.byref(info)))
        self.rank, self.nranks, self.device = info.rank, info.nranks, info.device
        self.has_multicast = bool(info.has_multicast)
        self.handle_kind = N.HANDLE_NAMES.get(info.handle_kind, str(info.handle_kind))
        self.sm_count = info.sm_count
        self.slot_bytes, self.block_bytes = info.slot_bytes, info.block_bytes
        self.heap_base, self.heap_bytes = info.heap_base, info.heap_bytes
        self._op = N.Op()  # reused: no per-call allocation on the hot path
        self._op_ref = C.byref(self._op)
        self._dev = torch.device("cuda", self.device)
        self._collective = L.pdcc_collective
        self._launches = 0
        # A communicator's rings and flag counters are one sequence: kernels of one comm must
        # not overlap on the GPU. Calls on one stream are ordered by the stream; when the
        # caller moves to another stream the new stream is made to wait for the old one.
        self._last_stream: Optional[int] = None
        self._order_event: Optional[torch.cuda.Event] = None

    # -- plumbing ---------------------------------------------------------------------
    def _run(self, op: N.Op, stream: Optional[torch.cuda.Stream]) -> None:
        if _NVTX:  # one NVTX range per collective, off unless PDCC_NVTX=1
            torch.cuda.nvtx.range_push(f"pdcc.{_COLL_NAMES[op.coll]}[{op.count}]")
        sp = _stream_ptr(stream, self.device)
        if sp != self._last_stream:
            self._switch_stream(sp)
        rc = self._collective(self._h, self._op_ref if op is self._op else C.byref(op), sp)
        if _NVTX:
            torch.cuda.nvtx.range_pop()
        if rc != 0:
            N.check(rc)

    # -- symmetric memory ---------------------------------------------------------------
    def empty_symmetric(self, shape, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """COLLECTIVE: every rank of the group calls it with the same shape and dtype, in the
        same order. Returns a tensor in symmetric memory (same offset of the same region on
        every rank, mapped on every rank, bound to a multicast object): collectives on it
        take the zero-copy kernels (csrc/sym.cuh). Memory is carved from regions of
        PDCC_SYM_CHUNK bytes (default 256 MiB; larger requests get a region of their own) and
        lives until the communicator is destroyed."""
        if isinstance(shape, int):
            shape = (shape,)
        numel = 1
        for d in shape:
            numel *= int(d)
        if numel == 0:
            return torch.empty(*shape, dtype=dtype, device=self._dev)
        esz = torch.empty(0, dtype=dtype).element_size()
        nbytes = numel * esz
        need = (nbytes + 255) // 256 * 256
        if not hasattr(self, "_sym_base") or self._sym_left < need:
            chunk = max(need, int(os.environ.get("PDCC_SYM_CHUNK", str(256 << 20))))
            ptr = C.c_void_p()
            N.check(N.lib().pdcc_sym_alloc(self._h, chunk, C.byref(ptr)))
            self._sym_base, self._sym_left = int(ptr.value), chunk
        ptr_i = self._sym_base
        self._sym_base += need
        self._sym_left -= need
        raw = torch.as_tensor(_RawCudaBytes(ptr_i, nbytes, self), device=self._dev)
        if raw.data_ptr() != ptr_i:
            raise RuntimeError("torch copied the symmetric buffer instead of aliasing it")
        return raw.view(dtype)[:numel].view(*shape)

    def is_symmetric(self, t: torch.Tensor) -> bool:
        s, m = C.c_int(), C.c_int()
        N.check(N.lib().pdcc_sym_query(self._h, t.data_ptr(), t.numel() * t.element_size(),
                                       C.byref(s), C.byref(m)))
        return bool(s.value)

    def _switch_stream(self, sp: int) -> None:
        """Orders the next kernel (on raw stream `sp`) after everything this communicator
        has enqueued on the stream it used last. One event record + one stream wait, only
        when the stream actually changes; never a host synchronisation."""
        if self._last_stream is not None and not torch.cuda.is_current_stream_capturing():
            if self._order_event is None:
                self._order_event = torch.cuda.Event()
            old = torch.cuda.ExternalStream(self._last_stream, device=self._dev)
            new = torch.cuda.ExternalStream(sp, device=self._dev)
            self._order_event.record(old)
            new.wait_event(self._order_event)
        self._last_stream = sp

    def plan(self, coll: int, count: int, dtype: torch.dtype, *, redop: int = SUM, algo=None,
             nblocks: int = 0) -> dict:
        """What AUTO resolves to for this op (algo, CTAs, tile size): bench/debug only."""
        op = _fill(N.Op(), coll, count, dtype_code(dtype), redop=redop, algo=algo, nblocks=nblocks)
        a, b, t, n, d = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        tb, ss = C.c_uint64(), C.c_uint32()
        N.check(N.lib().pdcc_plan(self._h, C.byref(op), C.byref(a), C.byref(b), C.byref(t),
                                  C.byref(tb), C.byref(n), C.byref(d), C.byref(ss)))
        names = {v: k for k, v in N.ALGO_NAMES.items()}
        return {"algo": names.get(a.value, a.value), "nblocks": b.value, "threads": t.value,
                "tile_bytes": tb.value, "ntiles": n.value, "depth": d.value,
                "slot_stride": ss.value}

    def check(self) -> None:
        """Synchronises and raises if a device-side wait timed out in any kernel so far."""
        N.check(N.lib().pdcc_comm_check(self._h))

    def read_trace(self, cap: int = 2048) -> list:
        """Timeline CTA 0 recorded since the last read (PDCC_TRACE=1): a list of dicts
        {kind, stage, warp, tile, begin_ns, end_ns, aux}. Synchronises the device."""
        buf = (C.c_uint32 * (8 * cap))()
        n = C.c_int()
        N.check(N.lib().pdcc_comm_read_trace(self._h, buf, cap, C.byref(n)))
        out = []
        for i in range(min(n.value, cap)):
            e = buf[8 * i:8 * i + 8]
            kind = e[0] & 0xff
            out.append({"kind": "worker" if kind < 16 else "fence" if kind == 16 else "admit",
                        "stage": kind if kind < 16 else (kind - 32 if kind >= 32 else -1),
                        "warp": (e[0] >> 8) & 0xff, "tile": e[1],
                        "begin_ns": e[2] | (e[3] << 32), "end_ns": e[4] | (e[5] << 32),
                        "aux": e[6]})
        return out

    def destroy(self) -> None:
        if self._h:
            N.lib().pdcc_comm_destroy(self._h)
            self._h = C.c_void_p()

    def _own(self, t: torch.Tensor) -> None:
        if t.device != self._dev:
            raise ValueError(f"tensor on {t.device} but this group's communicator is bound to "
                             f"cuda:{self.device}")
        if not t.is_contiguous():
            raise ValueError("Tensors must be contiguous")

    # -- collectives (group ranks) ----------------------------------------------------
    def all_reduce(self, tensor: torch.Tensor, redop: int = SUM, *, out: Optional[torch.Tensor] = None,
                   scale: float = 1.0, algo=None, nblocks: int = 0, stream=None) -> None:
        """K1/K2. In place unless `out` is given; `out` may be fp32 for a 16-bit input."""
        self._own(tensor)
        dst = tensor if out is None else out
        if out is not None:
            self._own(out)
            if out.numel() != tensor.numel():
                raise ValueError("out must have as many elements as the input")
        op = _fill(self._op, N.ALL_REDUCE, tensor.numel(), dtype_code(tensor.dtype, True),
                   out_dtype=dtype_code(dst.dtype, True), redop=redop, algo=algo, nblocks=nblocks,
                   scale=scale, in_ptr=tensor.data_ptr(), out_ptr=dst.data_ptr())
        self._run(op, stream)

    def reduce(self, tensor: torch.Tensor, dst: int, redop: int = SUM, *, algo=None, nblocks: int = 0,
               stream=None) -> None:
        """K3. Only `dst` is written; every other rank's tensor is left untouched."""
        self._own(tensor)
        op = _fill(self._op, N.REDUCE, tensor.numel(), dtype_code(tensor.dtype, True), redop=redop,
                   root=dst, algo=algo, nblocks=nblocks, in_ptr=tensor.data_ptr(),
                   out_ptr=tensor.data_ptr())
        self._run(op, stream)

    def broadcast(self, tensor: torch.Tensor, src: int, *, algo=None, nblocks: int = 0,
                  stream=None) -> None:
        """K4."""
        self._own(tensor)
        nbytes = tensor.numel() * tensor.element_size()
        op = _fill(self._op, N.BROADCAST, nbytes, N.U8, root=src, algo=algo, nblocks=nblocks,
                   in_ptr=tensor.data_ptr(), out_ptr=tensor.data_ptr())
        self._run(op, stream)

    def scatter(self, tensor: torch.Tensor, scatter_list: Optional[Sequence[torch.Tensor]], src: int,
                *, algo=None, nblocks: int = 0, stream=None) -> None:
        """K5. `scatter_list` (one tensor per group rank) only on `src`."""
        self._own(tensor)
        ptrs: Sequence[int] = ()
        if self.rank == src:
            for t in scatter_list:
                self._own(t)
            ptrs = [t.data_ptr() for t in scatter_list]
        nbytes = tensor.numel() * tensor.element_size()
        op = _fill(self._op, N.SCATTER, nbytes, N.U8, root=src, algo=algo, nblocks=nblocks,
                   out_ptr=tensor.data_ptr(), ptrs=ptrs)
        self._run(op, stream)

    def gather(self, tensor: torch.Tensor, gather_list: Optional[Sequence[torch.Tensor]], dst: int,
               *, algo=None, nblocks: int = 0, stream=None) -> None:
        """K6. `gather_list` only on `dst`."""
        self._own(tensor)
        ptrs: Sequence[int] = ()
        if self.rank == dst:
            for t in gather_list:
                self._own(t)
            ptrs = [t.data_ptr() for t in gather_list]
        nbytes = tensor.numel() * tensor.element_size()
        op = _fill(self._op, N.GATHER, nbytes, N.U8, root=dst, algo=algo, nblocks=nblocks,
                   in_ptr=tensor.data_ptr(), ptrs=ptrs)
        self._run(op, stream)

    def all_gather(self, tensor_list: Sequence[torch.Tensor], tensor: torch.Tensor, *, algo=None,
                   nblocks: int = 0, stream=None) -> None:
        """K7."""
        self._own(tensor)
        for t in tensor_list:
            self._own(t)
        nbytes = tensor.numel() * tensor.element_size()
        op = _fill(self._op, N.ALL_GATHER, nbytes, N.U8, algo=algo, nblocks=nblocks,
                   in_ptr=tensor.data_ptr(), ptrs=[t.data_ptr() for t in tensor_list])
        self._run(op, stream)

    def barrier(self, stream=None) -> None:
        op = _fill(self._op, N.BARRIER, 0, N.U8)
        self._run(op, stream)


def local_copy(dst: torch.Tensor, src: torch.Tensor, stream=None) -> None:
    """K9: the world-size-1 form of scatter/gather/all_gather (a list op over one rank)."""
    if dst.data_ptr() == src.data_ptr():
        return
    nbytes = src.numel() * src.element_size()
    N.check(N.lib().pdcc_local_copy(dst.data_ptr(), src.data_ptr(), nbytes, src.device.index,
                                    _stream_ptr(stream)))


def local_scale_cast(dst: torch.Tensor, src: torch.Tensor, scale: float, stream=None) -> None:
    """K9: dst = cast(src * scale); the world-size-1 form of the fused average."""
    N.check(N.lib().pdcc_local_scale_cast(dst.data_ptr(), src.data_ptr(), src.numel(),
                                          dtype_code(src.dtype, True), dtype_code(dst.dtype, True),
                                          scale, src.device.index, _stream_ptr(stream)))


class Emulator:
    """N virtual ranks on ONE GPU, one cooperative launch per collective (P2P variants).

    Each method takes per-rank lists (index = rank), exactly what N processes would pass.
    """

    def __init__(self, nranks: int, device: int = 0, slot_bytes: int = 0, max_blocks: int = 0,
                 device_timeout_ms: int = 5000):
        self.nranks, self.device = nranks, device
        self._h = C.c_void_p()
        N.check(N.lib().pdcc_emu_create(device, nranks, slot_bytes, max_blocks, device_timeout_ms,
                                        C.byref(self._h)))
        self._ops = (N.Op * N.MAX_RANKS)()

    def _run(self, stream=None) -> None:
        N.check(N.lib().pdcc_emu_collective(self._h, self._ops, _stream_ptr(stream)))

    def check(self) -> None:
        N.check(N.lib().pdcc_emu_check(self._h))

    def destroy(self) -> None:
        if self._h:
            N.lib().pdcc_emu_destroy(self._h)
            self._h = C.c_void_p()

    def all_reduce(self, tensors, redop=SUM, *, outs=None, scale=1.0, algo=None, nblocks=0):
        for r, t in enumerate(tensors):
            o = t if outs is None else outs[r]
            _fill(self._ops[r], N.ALL_REDUCE, t.numel(), dtype_code(t.dtype, True),
                  out_dtype=dtype_code(o.dtype, True), redop=redop, algo=algo, nblocks=nblocks,
                  scale=scale, in_ptr=t.data_ptr(), out_ptr=o.data_ptr())
        self._run()

    def reduce(self, tensors, dst, redop=SUM, *, algo=None, nblocks=0):
        for r, t in enumerate(tensors):
            _fill(self._ops[r], N.REDUCE, t.numel(), dtype_code(t.dtype, True), redop=redop,
                  root=dst, algo=algo, nblocks=nblocks, in_ptr=t.data_ptr(), out_ptr=t.data_ptr())
        self._run()

    def broadcast(self, tensors, src, *, algo=None, nblocks=0):
        for r, t in enumerate(tensors):
            _fill(self._ops[r], N.BROADCAST, t.numel() * t.element_size(), N.U8, root=src,
                  algo=algo, nblocks=nblocks, in_ptr=t.data_ptr(), out_ptr=t.data_ptr())
        self._run()

    def scatter(self, outs, scatter_list, src, *, algo=None, nblocks=0):
        for r, t in enumerate(outs):
            ptrs = [x.data_ptr() for x in scatter_list] if r == src else ()
            _fill(self._ops[r], N.SCATTER, t.numel() * t.element_size(), N.U8, root=src,
                  algo=algo, nblocks=nblocks, out_ptr=t.data_ptr(), ptrs=ptrs)
        self._run()

    def gather(self, ins, gather_list, dst, *, algo=None, nblocks=0):
        for r, t in enumerate(ins):
            ptrs = [x.data_ptr() for x in gather_list] if r == dst else ()
            _fill(self._ops[r], N.GATHER, t.numel() * t.element_size(), N.U8, root=dst,
                  algo=algo, nblocks=nblocks, in_ptr=t.data_ptr(), ptrs=ptrs)
        self._run()

    def all_gather(self, out_lists, ins, *, algo=None, nblocks=0):
        for r, t in enumerate(ins):
            _fill(self._ops[r], N.ALL_GATHER, t.numel() * t.element_size(), N.U8, algo=algo,
                  nblocks=nblocks, in_ptr=t.data_ptr(), ptrs=[x.data_ptr() for x in out_lists[r]])
        self._run()

    def barrier(self):
        for r in range(self.nranks):
            _fill(self._ops[r], N.BARRIER, 0, N.U8)
        self._run()
