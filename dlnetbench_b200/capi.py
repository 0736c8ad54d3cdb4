"""ctypes binding of libdlnb.so — the C ABI declared in include/dlnb.h.

This is plumbing for the tests and bench.py; the product host side is C++
(cpp/proxy_classes.hpp: B200Communicator). There is deliberately no fallback:
if the CUDA library is missing or fails, every call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdlnb.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "dlnb.h")

F32, F16, BF16 = 0, 1, 2
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_MULTIMEM = 0, 1, 2, 3
PAT_ZERO, PAT_INT, PAT_UNIFORM = 0, 1, 2
DTYPE_SIZE = {F32: 4, F16: 2, BF16: 2}


class DlnbError(RuntimeError):
    pass


def declared_symbols(header: str = HEADER_PATH) -> list[str]:
    """Every function name include/dlnb.h declares (used by the not-gpu tests)."""
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dlnb_[a-z0-9_]+)\s*\(", text)))


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DlnbError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C dlnetbench_b200/csrc`). There is no CPU fallback."
            )
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _declare(_lib)
    return _lib


def _declare(L: C.CDLL) -> None:
    vp, sz, i, u64, cp = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_char_p
    pvp = C.POINTER(C.c_void_p)
    sig = {
        "dlnb_init": (i, [i, i, i, cp]),
        "dlnb_finalize": (i, []),
        "dlnb_rank": (i, []),
        "dlnb_world": (i, []),
        "dlnb_device": (i, []),
        "dlnb_last_error": (cp, []),
        "dlnb_version": (cp, []),
        "dlnb_arena_kind": (cp, []),
        "dlnb_host_barrier": (i, []),
        "dlnb_host_allgather": (i, [vp, sz, vp]),
        "dlnb_symm_alloc": (i, [sz, pvp]),
        "dlnb_symm_free": (i, [vp]),
        "dlnb_symm_bytes_in_use": (sz, []),
        "dlnb_comm_create": (i, [C.POINTER(i), i, i, i, pvp]),
        "dlnb_comm_destroy": (i, [vp]),
        "dlnb_comm_rank": (i, [vp]),
        "dlnb_comm_size": (i, [vp]),
        "dlnb_comm_set": (i, [vp, cp, C.c_long]),
        "dlnb_comm_get": (C.c_long, [vp, cp]),
        "dlnb_compute_spin": (i, [C.c_double]),
        "dlnb_compute_sync": (i, []),
        "dlnb_compute_mark": (i, []),
        "dlnb_slot_exposed": (i, [vp, i, C.POINTER(C.c_float)]),
        "dlnb_spin_calibrate": (i, [C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "dlnb_set": (i, [cp, C.c_long]),
        "dlnb_get": (C.c_long, [cp]),
        "dlnb_iallreduce": (i, [vp, vp, vp, sz, i]),
        "dlnb_iallgather": (i, [vp, vp, vp, sz, i]),
        "dlnb_ireduce_scatter": (i, [vp, vp, vp, sz, i]),
        "dlnb_allreduce": (i, [vp, vp, vp, sz]),
        "dlnb_allgather": (i, [vp, vp, vp, sz]),
        "dlnb_reduce_scatter": (i, [vp, vp, vp, sz]),
        "dlnb_alltoall": (i, [vp, vp, vp, sz]),
        "dlnb_ialltoall": (i, [vp, vp, vp, sz, i]),
        "dlnb_emu_alltoall": (i, [i, pvp, pvp, sz, i, i]),
        "dlnb_send": (i, [vp, vp, sz, i]),
        "dlnb_recv": (i, [vp, vp, sz, i]),
        "dlnb_isend": (i, [vp, vp, sz, i, i]),
        "dlnb_irecv": (i, [vp, vp, sz, i, i]),
        "dlnb_wait": (i, [vp, i]),
        "dlnb_waitall": (i, [vp, i]),
        "dlnb_barrier": (i, [vp]),
        "dlnb_launch_count": (u64, []),
        "dlnb_iter_begin": (i, []),
        "dlnb_iter_end": (i, [C.POINTER(C.c_float)]),
        "dlnb_foreign_begin": (i, [vp, i, pvp]),
        "dlnb_foreign_end": (i, [vp, i]),
        "dlnb_batch_begin": (i, []),
        "dlnb_batch_begin_concurrent": (i, [C.c_double]),
        "dlnb_batch_end": (i, [C.POINTER(C.c_float)]),
        "dlnb_graph_begin": (i, []),
        "dlnb_graph_end": (i, [pvp]),
        "dlnb_graph_launch": (i, [vp]),
        "dlnb_graph_kernel_nodes": (i, [vp]),
        "dlnb_graph_destroy": (i, [vp]),
        "dlnb_slot_elapsed": (i, [vp, i, C.POINTER(C.c_float)]),
        "dlnb_memcpy_h2d": (i, [vp, vp, sz]),
        "dlnb_memcpy_d2h": (i, [vp, vp, sz]),
        "dlnb_memcpy_h2d_async": (i, [vp, vp, sz]),
        "dlnb_memcpy_d2h_async": (i, [vp, vp, sz]),
        "dlnb_memcpy_h2d_after_compute": (i, [vp, vp, sz]),
        "dlnb_memcpy_d2h_after_slot": (i, [vp, i, vp, vp, sz]),
        "dlnb_host_alloc": (i, [sz, pvp]),
        "dlnb_host_free": (i, [vp]),
        "dlnb_fill_pattern": (i, [vp, sz, i, i, u64, i]),
        "dlnb_verify_sum_pattern": (i, [vp, sz, sz, i, i, u64, C.POINTER(i), i, C.POINTER(u64)]),
        "dlnb_verify_sum_pattern_tol": (i, [vp, sz, sz, i, i, u64, C.POINTER(i), i, C.c_double, C.POINTER(u64), C.POINTER(C.c_double)]),
        "dlnb_checksum": (i, [vp, sz, C.POINTER(u64)]),
        "dlnb_flush_l2": (i, []),
        "dlnb_emu_allreduce": (i, [i, pvp, pvp, sz, i, i, i]),
        "dlnb_emu_allgather": (i, [i, pvp, pvp, sz, i, i]),
        "dlnb_emu_reduce_scatter": (i, [i, pvp, pvp, sz, i, i]),
        "dlnb_emu_allreduce_noalloc": (i, [i, pvp, pvp, sz, i, i, i]),
        "dlnb_emu_sendrecv": (i, [vp, vp, sz, i, i]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args


def check(rc: int) -> None:
    if rc != 0:
        raise DlnbError(lib().dlnb_last_error().decode(errors="replace"))


# ---------------------------------------------------------------------------
# thin object layer mirroring the reference's ProxyCommunicator / Tensor names
# (cpp/proxy_classes.hpp:30-51, :381-444) so parity tests read like its call sites
# ---------------------------------------------------------------------------
def init(rank: int = 0, world: int = 1, device: int | None = None, session: str | None = None) -> None:
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", rank))
    if session is None:
        session = os.environ.get("DLNB_SESSION") or ("torchrun-" + os.environ.get("MASTER_PORT", "0"))
    check(lib().dlnb_init(rank, world, device, session.encode()))


def finalize() -> None:
    check(lib().dlnb_finalize())


class Tensor:
    """Zero-filled symmetric device buffer; `.data` is the device address, `.size` elements."""

    def __init__(self, size: int, dtype: int = F32):
        self.size = int(size)
        self.dtype = dtype
        p = C.c_void_p()
        check(lib().dlnb_symm_alloc(self.size * DTYPE_SIZE[dtype], C.byref(p)))
        self.data = p.value

    def free(self) -> None:
        """dlnb_symm_free: collective like the allocation; the block may be handed out again (cleared)."""
        if self.data:
            check(lib().dlnb_symm_free(self.data))
            self.data = None

    @property
    def nbytes(self) -> int:
        return self.size * DTYPE_SIZE[self.dtype]

    def offset(self, elems: int) -> int:
        return self.data + elems * DTYPE_SIZE[self.dtype]

    def upload(self, arr) -> None:
        import numpy as np

        a = np.ascontiguousarray(arr)
        assert a.nbytes <= self.nbytes, (a.nbytes, self.nbytes)
        check(lib().dlnb_memcpy_h2d(self.data, a.ctypes.data, a.nbytes))

    def download(self, np_dtype, count: int | None = None):
        import numpy as np

        n = self.size if count is None else count
        out = np.empty(n, dtype=np_dtype)
        assert out.nbytes <= self.nbytes
        check(lib().dlnb_memcpy_d2h(out.ctypes.data, self.data, out.nbytes))
        return out

    def fill(self, kind: int, seed: int = 0, as_rank: int = 0, count: int | None = None) -> None:
        check(lib().dlnb_fill_pattern(self.data, self.size if count is None else count, self.dtype, kind, seed, as_rank))

    def checksum(self, nbytes: int | None = None) -> int:
        out = C.c_uint64()
        check(lib().dlnb_checksum(self.data, self.nbytes if nbytes is None else nbytes, C.byref(out)))
        return out.value


class B200Communicator:
    """Python mirror of cpp/proxy_classes.hpp B200Communicator (same method names)."""

    def __init__(self, ranks=None, num_slots: int = 1, dtype: int = F32):
        L = lib()
        if ranks is None:
            ranks = list(range(L.dlnb_world()))
        arr = (C.c_int * len(ranks))(*ranks)
        h = C.c_void_p()
        check(L.dlnb_comm_create(arr, len(ranks), num_slots, dtype, C.byref(h)))
        self.h = h
        self.dtype = dtype
        self.ranks = list(ranks)

    def set(self, key: str, value: int) -> None:
        check(lib().dlnb_comm_set(self.h, key.encode(), value))

    def get(self, key: str) -> int:
        return lib().dlnb_comm_get(self.h, key.encode())

    def get_name(self) -> str:
        return "B200P2P"

    def Iallreduce(self, sendbuf, recvbuf, count, index):
        check(lib().dlnb_iallreduce(self.h, sendbuf, recvbuf, count, index))

    def Allreduce(self, sendbuf, recvbuf, count):
        check(lib().dlnb_allreduce(self.h, sendbuf, recvbuf, count))

    def Iallgather(self, sendbuf, sendcount, recvbuf, recvcount, index):
        assert sendcount == recvcount
        check(lib().dlnb_iallgather(self.h, sendbuf, recvbuf, sendcount, index))

    def Allgather(self, sendbuf, sendcount, recvbuf, recvcount):
        assert sendcount == recvcount
        check(lib().dlnb_allgather(self.h, sendbuf, recvbuf, sendcount))

    def Reduce_Scatter_block(self, sendbuf, recvbuf, recvcount):
        check(lib().dlnb_reduce_scatter(self.h, sendbuf, recvbuf, recvcount))

    def Alltoall(self, sendbuf, sendcount, recvbuf, recvcount):
        assert sendcount == recvcount
        check(lib().dlnb_alltoall(self.h, sendbuf, recvbuf, sendcount))

    def send(self, buf, count, dest):
        check(lib().dlnb_send(self.h, buf, count, dest))

    def recv(self, buf, count, source):
        check(lib().dlnb_recv(self.h, buf, count, source))

    def Isend(self, buf, count, dest, index):
        check(lib().dlnb_isend(self.h, buf, count, dest, index))

    def Irecv(self, buf, count, source, index):
        check(lib().dlnb_irecv(self.h, buf, count, source, index))

    def Wait(self, index):
        check(lib().dlnb_wait(self.h, index))

    def WaitAll(self, num_waits):
        check(lib().dlnb_waitall(self.h, num_waits))

    def Barrier(self):
        check(lib().dlnb_barrier(self.h))

    def slot_ms(self, index) -> float:
        ms = C.c_float()
        check(lib().dlnb_slot_elapsed(self.h, index, C.byref(ms)))
        return ms.value

    def finalize(self):
        check(lib().dlnb_comm_destroy(self.h))
        self.h = None


def compute_spin(usec: float) -> None:
    check(lib().dlnb_compute_spin(float(usec)))


def spin_calibrate() -> tuple[float, float]:
    a, b = C.c_double(), C.c_double()
    check(lib().dlnb_spin_calibrate(C.byref(a), C.byref(b)))
    return a.value, b.value


def iter_begin() -> None:
    check(lib().dlnb_iter_begin())


class Graph:
    """One captured proxy iteration: `with capi.Graph() as g: ...issue spins and collectives...`, then g.launch()."""

    def __init__(self):
        self.h = C.c_void_p()

    def __enter__(self):
        check(lib().dlnb_graph_begin())
        return self

    def __exit__(self, et, ev, tb):
        check(lib().dlnb_graph_end(C.byref(self.h)))
        return False

    def launch(self):
        check(lib().dlnb_graph_launch(self.h))

    def kernel_nodes(self) -> int:
        return lib().dlnb_graph_kernel_nodes(self.h)

    def destroy(self):
        check(lib().dlnb_graph_destroy(self.h))
        self.h = C.c_void_p()


def compute_sync() -> None:
    check(lib().dlnb_compute_sync())


def batch_begin() -> None:
    check(lib().dlnb_batch_begin())


def batch_begin_concurrent(gate_us: float = 0.0) -> None:
    check(lib().dlnb_batch_begin_concurrent(float(gate_us)))


def batch_end() -> float:
    ms = C.c_float()
    check(lib().dlnb_batch_end(C.byref(ms)))
    return ms.value


def set_option(key: str, value: int) -> None:
    check(lib().dlnb_set(key.encode(), value))


def iter_end() -> float:
    ms = C.c_float()
    check(lib().dlnb_iter_end(C.byref(ms)))
    return ms.value


def verify_sum_pattern(ptr: int, count: int, elem_offset: int, dtype: int, kind: int, seed: int, ranks) -> int:
    arr = (C.c_int * len(ranks))(*ranks)
    out = C.c_uint64()
    check(lib().dlnb_verify_sum_pattern(ptr, count, elem_offset, dtype, kind, seed, arr, len(ranks), C.byref(out)))
    return out.value


def verify_sum_pattern_tol(ptr: int, count: int, elem_offset: int, dtype: int, kind: int, seed: int, ranks, rel_tol: float):
    """(mismatches, worst |got-ref| / sum|x|) against the rank-order fp32 sum, tolerance rel_tol * sum|x| + half an ulp."""
    arr = (C.c_int * len(ranks))(*ranks)
    out, worst = C.c_uint64(), C.c_double()
    check(lib().dlnb_verify_sum_pattern_tol(ptr, count, elem_offset, dtype, kind, seed, arr, len(ranks), rel_tol, C.byref(out), C.byref(worst)))
    return out.value, worst.value


def _ptr_array(ptrs):
    return (C.c_void_p * len(ptrs))(*ptrs)


def emu_allreduce(send_ptrs, recv_ptrs, count, dtype, algo=ALGO_AUTO, ctas=0):
    check(lib().dlnb_emu_allreduce(len(send_ptrs), _ptr_array(send_ptrs), _ptr_array(recv_ptrs), count, dtype, algo, ctas))


def emu_allreduce_noalloc(send_ptrs, recv_ptrs, count, dtype, algo=ALGO_AUTO, ctas=0):
    """dlnb_emu_allreduce_noalloc: same result as emu_allreduce, scratch from the library image (<= 64 CTAs)."""
    check(lib().dlnb_emu_allreduce_noalloc(len(send_ptrs), _ptr_array(send_ptrs), _ptr_array(recv_ptrs), count, dtype, algo, ctas))


def emu_sendrecv(send_ptr, recv_ptr, count, dtype, ctas=0):
    """dlnb_emu_sendrecv: the Irecv / Isend / Wait kernels against one loop-back mailbox on one GPU."""
    check(lib().dlnb_emu_sendrecv(send_ptr, recv_ptr, count, dtype, ctas))


def get_option(key: str) -> int:
    return lib().dlnb_get(key.encode())


SPIN_SLEEP, SPIN_FMA, SPIN_HBM = 0, 1, 2


def emu_allgather(send_ptrs, recv_ptrs, sendcount, dtype, ctas=0):
    check(lib().dlnb_emu_allgather(len(send_ptrs), _ptr_array(send_ptrs), _ptr_array(recv_ptrs), sendcount, dtype, ctas))


def emu_alltoall(send_ptrs, recv_ptrs, sendcount, dtype, ctas=0):
    check(lib().dlnb_emu_alltoall(len(send_ptrs), _ptr_array(send_ptrs), _ptr_array(recv_ptrs), sendcount, dtype, ctas))


def emu_reduce_scatter(send_ptrs, recv_ptrs, recvcount, dtype, ctas=0):
    check(lib().dlnb_emu_reduce_scatter(len(send_ptrs), _ptr_array(send_ptrs), _ptr_array(recv_ptrs), recvcount, dtype, ctas))
