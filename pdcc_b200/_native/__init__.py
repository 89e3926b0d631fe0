"""ctypes binding of libpdcc.so (C ABI in include/pdcc.h).

The library is built in-tree by `build.py` (nvcc, sm_100a) and loaded from this
directory so that the GPU box — and whoever audits which `.so` files a process
mapped — sees the native code next to the package. There is no Python fallback for
the CUDA path: if the library is missing, anything that touches a CUDA tensor raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libpdcc.so"

MAX_RANKS = 8
MAX_BLOCKS = 148
RING_DEPTH = 4

# enum pdcc_coll
ALL_REDUCE, REDUCE, BROADCAST, SCATTER, GATHER, ALL_GATHER, BARRIER = range(7)
# enum pdcc_dtype
F32, BF16, F16, F64, I32, I64, U8, I8, I16 = range(9)
# enum pdcc_algo
(ALGO_AUTO, ALGO_ONESHOT_P2P, ALGO_TWOSHOT_P2P, ALGO_ONESHOT_NVLS, ALGO_TWOSHOT_NVLS, ALGO_P2P,
 ALGO_NVLS, ALGO_LL, ALGO_PULL, ALGO_SYM) = range(10)
ALGO_NAMES = {
    "auto": ALGO_AUTO, "oneshot_p2p": ALGO_ONESHOT_P2P, "twoshot_p2p": ALGO_TWOSHOT_P2P,
    "oneshot_nvls": ALGO_ONESHOT_NVLS, "twoshot_nvls": ALGO_TWOSHOT_NVLS,
    "p2p": ALGO_P2P, "nvls": ALGO_NVLS, "ll": ALGO_LL, "pull": ALGO_PULL, "sym": ALGO_SYM,
}
HANDLE_AUTO, HANDLE_POSIX_FD, HANDLE_LEGACY_IPC, HANDLE_LOCAL = range(4)
HANDLE_NAMES = {HANDLE_AUTO: "auto", HANDLE_POSIX_FD: "posix_fd", HANDLE_LEGACY_IPC: "legacy_ipc",
                HANDLE_LOCAL: "local"}

OK, EINVAL, ECUDA, ESYS, ETIMEOUT, EUNSUPPORTED, EPEER = 0, -1, -2, -3, -4, -5, -6


class PdccError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"pdcc error {code}: {msg}")
        self.code = code


class Caps(C.Structure):
    _fields_ = [
        ("device", C.c_int), ("sm_count", C.c_int), ("cc_major", C.c_int), ("cc_minor", C.c_int),
        ("vmm_supported", C.c_int), ("posix_fd_supported", C.c_int), ("fabric_supported", C.c_int),
        ("multicast_supported", C.c_int), ("cooperative_launch", C.c_int), ("device_count", C.c_int),
        ("alloc_granularity", C.c_size_t), ("mc_granularity", C.c_size_t), ("total_mem", C.c_size_t),
        ("name", C.c_char * 64),
    ]

    def as_dict(self) -> dict:
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["name"] = self.name.decode(errors="replace")
        return d


class CommConfig(C.Structure):
    _fields_ = [
        ("rank", C.c_int), ("nranks", C.c_int), ("device", C.c_int), ("handle_kind", C.c_int),
        ("multicast", C.c_int), ("timeout_ms", C.c_int), ("device_timeout_ms", C.c_int),
        ("slot_bytes", C.c_size_t), ("max_blocks", C.c_int), ("heap_bytes", C.c_size_t),
        ("key", C.c_char * 64),
    ]


class CommInfo(C.Structure):
    _fields_ = [
        ("rank", C.c_int), ("nranks", C.c_int), ("device", C.c_int), ("handle_kind", C.c_int),
        ("has_multicast", C.c_int), ("sm_count", C.c_int), ("max_blocks", C.c_int),
        ("slot_bytes", C.c_size_t), ("block_bytes", C.c_size_t), ("heap_bytes", C.c_size_t),
        ("heap_base", C.c_void_p),
    ]


class Op(C.Structure):
    _fields_ = [
        ("coll", C.c_int), ("dtype", C.c_int), ("out_dtype", C.c_int), ("redop", C.c_int),
        ("root", C.c_int), ("algo", C.c_int), ("nblocks", C.c_int), ("scale", C.c_float),
        ("count", C.c_uint64), ("in_", C.c_void_p), ("out", C.c_void_p),
        ("list", C.c_void_p * MAX_RANKS),
    ]


_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def lib() -> C.CDLL:
    """Loads (once) and returns the native library, building it if the source tree has nvcc
    and the .so is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        if os.environ.get("PDCC_NO_AUTOBUILD"):
            raise ImportError(f"{LIB_PATH} is missing; run `python -m pdcc_b200._native.build`")
        from . import build as _build

        _build.build()
    L = C.CDLL(str(LIB_PATH))
    vp, ci, cp = C.c_void_p, C.c_int, C.c_char_p
    L.pdcc_last_error.restype = cp
    L.pdcc_version.restype = cp
    L.pdcc_launch_count.restype = C.c_uint64
    L.pdcc_reload_env.restype = None
    L.pdcc_reload_env.argtypes = []
    sigs = {
        "pdcc_store_server_start": [cp, ci, C.POINTER(vp), C.POINTER(ci)],
        "pdcc_store_server_stop": [vp],
        "pdcc_store_connect": [cp, ci, ci, C.POINTER(vp)],
        "pdcc_store_close": [vp],
        "pdcc_store_set": [vp, cp, vp, ci],
        "pdcc_store_get": [vp, cp, vp, ci, C.POINTER(ci), ci],
        "pdcc_store_add": [vp, cp, C.c_longlong, C.POINTER(C.c_longlong)],
        "pdcc_store_check": [vp, cp, C.POINTER(ci)],
        "pdcc_store_delete": [vp, cp],
        "pdcc_store_barrier": [vp, cp, ci, ci],
        "pdcc_probe": [ci, C.POINTER(Caps)],
        "pdcc_comm_create": [vp, C.POINTER(CommConfig), C.POINTER(vp)],
        "pdcc_comm_destroy": [vp],
        "pdcc_comm_info": [vp, C.POINTER(CommInfo)],
        "pdcc_comm_check": [vp],
        "pdcc_comm_read_trace": [vp, C.POINTER(C.c_uint32), ci, C.POINTER(ci)],
        "pdcc_collective": [vp, C.POINTER(Op), vp],
        "pdcc_sym_alloc": [vp, C.c_size_t, C.POINTER(vp)],
        "pdcc_sym_query": [vp, vp, C.c_size_t, C.POINTER(ci), C.POINTER(ci)],
        "pdcc_plan": [vp, C.POINTER(Op), C.POINTER(ci), C.POINTER(ci), C.POINTER(ci),
                      C.POINTER(C.c_uint64), C.POINTER(ci), C.POINTER(ci), C.POINTER(C.c_uint32)],
        "pdcc_plan_dry": [ci, ci, C.c_size_t, ci, C.POINTER(Op), C.POINTER(ci), C.POINTER(ci),
                          C.POINTER(ci), C.POINTER(C.c_uint64), C.POINTER(ci), C.POINTER(ci),
                          C.POINTER(C.c_uint32)],
        "pdcc_emu_create": [ci, ci, C.c_size_t, ci, ci, C.POINTER(vp)],
        "pdcc_emu_destroy": [vp],
        "pdcc_emu_collective": [vp, C.POINTER(Op), vp],
        "pdcc_emu_check": [vp],
        "pdcc_local_copy": [vp, vp, C.c_uint64, ci, vp],
        "pdcc_local_scale_cast": [vp, vp, C.c_uint64, ci, ci, C.c_float, ci, vp],
        "pdcc_all_reduce": [vp, vp, C.c_uint64, ci, ci, vp],
        "pdcc_all_reduce_scaled": [vp, vp, vp, C.c_uint64, ci, ci, C.c_float, vp],
        "pdcc_reduce": [vp, vp, C.c_uint64, ci, ci, ci, vp],
        "pdcc_broadcast": [vp, vp, C.c_uint64, ci, ci, vp],
        "pdcc_scatter": [vp, vp, C.POINTER(vp), C.c_uint64, ci, ci, vp],
        "pdcc_gather": [vp, vp, C.POINTER(vp), C.c_uint64, ci, ci, vp],
        "pdcc_all_gather": [vp, C.POINTER(vp), vp, C.c_uint64, ci, vp],
        "pdcc_barrier": [vp, vp],
        "pdcc_emu2_all_reduce_noalloc": [ci, vp, vp, C.c_uint64, ci, ci, vp],
        "pdcc_emu2_noalloc_check": [ci],
    }
    for name, argtypes in sigs.items():
        fn = getattr(L, name)
        fn.argtypes = argtypes
        fn.restype = ci
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise PdccError(rc, lib().pdcc_last_error().decode(errors="replace"))


def plan_dry(nranks: int, coll: int, count: int, dtype: int, *, has_multicast: bool = True,
             slot_bytes: int = 0, rings: int = 0, redop: int = 0, algo: int = 0,
             nblocks: int = 0) -> dict:
    """AUTO planning without a communicator or a GPU (see pdcc_plan_dry)."""
    op = Op()
    op.coll, op.dtype, op.out_dtype, op.redop = coll, dtype, dtype, redop
    op.algo, op.nblocks, op.scale, op.count = algo, nblocks, 1.0, count
    a, b, t, n, d = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
    tb, ss = C.c_uint64(), C.c_uint32()
    check(lib().pdcc_plan_dry(nranks, int(has_multicast), slot_bytes, rings, C.byref(op),
                              C.byref(a), C.byref(b), C.byref(t), C.byref(tb), C.byref(n),
                              C.byref(d), C.byref(ss)))
    names = {v: k for k, v in ALGO_NAMES.items()}
    return {"algo": names.get(a.value, a.value), "nblocks": b.value, "threads": t.value,
            "tile_bytes": tb.value, "ntiles": n.value, "depth": d.value,
            "slot_stride": ss.value}


def reload_env() -> None:
    """Make the native planner re-read its PDCC_* knobs (it caches them at first use)."""
    lib().pdcc_reload_env()


def launch_count() -> int:
    return int(lib().pdcc_launch_count())


def probe(device: int) -> dict:
    caps = Caps()
    check(lib().pdcc_probe(device, C.byref(caps)))
    return caps.as_dict()


class Store:
    """Client of the bootstrap KV store (R1). `serve=True` also hosts it in this process."""

    def __init__(self, host: str, port: int, serve: bool = False, timeout_ms: int = 120000):
        L = lib()
        self._server = None
        self._client = None
        self.host, self.port = host, port
        if serve:
            srv = C.c_void_p()
            bound = C.c_int()
            # bind where the clients will connect (loopback for the usual 127.0.0.1), not on
            # every interface: the store is unauthenticated
            import socket

            try:
                bind = socket.gethostbyname(host)
            except OSError:
                bind = ""
            check(L.pdcc_store_server_start(bind.encode(), port, C.byref(srv), C.byref(bound)))
            self._server = srv
            self.port = bound.value
        cl = C.c_void_p()
        try:
            check(L.pdcc_store_connect(host.encode(), self.port, timeout_ms, C.byref(cl)))
        except Exception:
            self.close()
            raise
        self._client = cl

    @property
    def handle(self) -> C.c_void_p:
        return self._client

    def set(self, key: str, value: bytes) -> None:
        buf = C.create_string_buffer(value, len(value)) if value else None
        check(lib().pdcc_store_set(self._client, key.encode(), buf, len(value)))

    def get(self, key: str, timeout_ms: int = 120000, cap: int = 1 << 16) -> bytes:
        buf = C.create_string_buffer(cap)
        n = C.c_int()
        check(lib().pdcc_store_get(self._client, key.encode(), buf, cap, C.byref(n), timeout_ms))
        if n.value > cap:
            return self.get(key, timeout_ms, n.value)
        return buf.raw[: n.value]

    def add(self, key: str, delta: int) -> int:
        out = C.c_longlong()
        check(lib().pdcc_store_add(self._client, key.encode(), delta, C.byref(out)))
        return int(out.value)

    def check(self, key: str) -> bool:
        e = C.c_int()
        check(lib().pdcc_store_check(self._client, key.encode(), C.byref(e)))
        return bool(e.value)

    def delete(self, key: str) -> None:
        check(lib().pdcc_store_delete(self._client, key.encode()))

    def barrier(self, name: str, nranks: int, timeout_ms: int = 120000) -> None:
        check(lib().pdcc_store_barrier(self._client, name.encode(), nranks, timeout_ms))

    def close(self) -> None:
        L = lib()
        if self._client is not None:
            L.pdcc_store_close(self._client)
            self._client = None
        if self._server is not None:
            L.pdcc_store_server_stop(self._server)
            self._server = None

    def __del__(self):  # best effort; explicit close() is preferred
        try:
            self.close()
        except Exception:
            pass
