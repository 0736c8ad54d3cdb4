"""ctypes wrapper of oracle/_build/libdlnb_oracle.so (oracle/dlnb_oracle.c).

TEST INFRASTRUCTURE ONLY — the CPU restatement of the reference's algorithm. Import
it from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never from
dlnetbench_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_build", "libdlnb_oracle.so")
F32, F16, BF16 = 0, 1, 2
NP = {F32: np.float32, F16: np.float16, BF16: np.uint16}
_lib = None


def build() -> str:
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle"])
    return LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = C.CDLL(LIB)
        vp, sz, i, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64
        pvp = C.POINTER(C.c_void_p)
        L.dlnb_oracle_allreduce.argtypes = [pvp, i, vp, sz, i]
        L.dlnb_oracle_allgather.argtypes = [pvp, i, vp, sz, i]
        L.dlnb_oracle_reduce_scatter_block.argtypes = [pvp, i, i, vp, sz, i]
        L.dlnb_oracle_alltoall.argtypes = [pvp, i, i, vp, sz, i]
        L.dlnb_oracle_bucket_sizes.argtypes = [u64, i, C.POINTER(u64)]
        L.dlnb_oracle_shard_sizes.argtypes = [u64, i, i, C.POINTER(u64)]
        L.dlnb_oracle_dp_lower_bound_us.argtypes = [u64, u64, i]
        L.dlnb_oracle_dp_lower_bound_us.restype = u64
        L.dlnb_oracle_fsdp_lower_bound_us.argtypes = [u64, u64, i]
        L.dlnb_oracle_fsdp_lower_bound_us.restype = u64
        L.dlnb_oracle_hybrid2d_sizes.argtypes = [u64, u64, u64, u64, u64, u64, i, i, C.POINTER(u64)]
        L.dlnb_oracle_hybrid3d_sizes.argtypes = [u64, u64, u64, u64, u64, u64, i, i, i, C.POINTER(u64)]
        L.dlnb_oracle_hybrid3d_moe_sizes.argtypes = [u64, u64, u64, u64, u64, u64, u64, u64, i, i, i, C.POINTER(u64)]
        L.dlnb_oracle_msg_stats.argtypes = [C.POINTER(u64), i, C.c_uint, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.dlnb_oracle_parse_model_stats.argtypes = [C.c_char_p, C.POINTER(u64)]
        L.dlnb_oracle_pattern_value.argtypes = [i, u64, i, u64]
        L.dlnb_oracle_pattern_value.restype = C.c_float
        L.dlnb_oracle_fill_pattern.argtypes = [vp, sz, i, i, u64, i]
        L.dlnb_oracle_checksum.argtypes = [vp, sz]
        L.dlnb_oracle_checksum.restype = u64
        L.dlnb_oracle_half_to_float.argtypes = [C.c_uint16]
        L.dlnb_oracle_half_to_float.restype = C.c_float
        L.dlnb_oracle_float_to_half.argtypes = [C.c_float]
        L.dlnb_oracle_float_to_half.restype = C.c_uint16
        L.dlnb_oracle_bf16_to_float.argtypes = [C.c_uint16]
        L.dlnb_oracle_bf16_to_float.restype = C.c_float
        L.dlnb_oracle_float_to_bf16.argtypes = [C.c_float]
        L.dlnb_oracle_float_to_bf16.restype = C.c_uint16
        _lib = L
    return _lib


def _ptrs(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def allreduce(send, dtype):
    """send: list of equal-length contiguous arrays, one per rank -> the reduced array."""
    out = np.empty_like(send[0])
    lib().dlnb_oracle_allreduce(_ptrs(send), len(send), out.ctypes.data, send[0].size, dtype)
    return out


def allgather(send, dtype):
    out = np.empty(send[0].size * len(send), dtype=send[0].dtype)
    lib().dlnb_oracle_allgather(_ptrs(send), len(send), out.ctypes.data, send[0].size, dtype)
    return out


def reduce_scatter_block(send, me, recvcount, dtype):
    out = np.empty(recvcount, dtype=send[0].dtype)
    lib().dlnb_oracle_reduce_scatter_block(_ptrs(send), len(send), me, out.ctypes.data, recvcount, dtype)
    return out


def alltoall(send, me, count, dtype):
    out = np.empty(count * len(send), dtype=send[0].dtype)
    lib().dlnb_oracle_alltoall(_ptrs(send), len(send), me, out.ctypes.data, count, dtype)
    return out


def bucket_sizes(model_size, num_buckets):
    out = (C.c_uint64 * num_buckets)()
    lib().dlnb_oracle_bucket_sizes(model_size, num_buckets, out)
    return list(out)


def shard_sizes(model_size, num_units, sharding_factor):
    out = (C.c_uint64 * num_units)()
    lib().dlnb_oracle_shard_sizes(model_size, num_units, sharding_factor, out)
    return list(out)


def dp_lower_bound_us(fwd, bwd, n):
    return lib().dlnb_oracle_dp_lower_bound_us(fwd, bwd, n)


def fsdp_lower_bound_us(fwd, bwd, u):
    return lib().dlnb_oracle_fsdp_lower_bound_us(fwd, bwd, u)


def hybrid2d_sizes(st, num_stages, num_microbatches):
    """st: parse_model_stats() dict -> fwd/bwd us per micro-batch, pipe message and DP allreduce elements, lower bound us."""
    out = (C.c_uint64 * 5)()
    lib().dlnb_oracle_hybrid2d_sizes(st["avgForwardTime"], st["avgBackwardTime"], st["modelSize"], st["batchSize"],
                                     st["sequenceLength"], st["embeddedDim"], num_stages, num_microbatches, out)
    return dict(zip(["fwd_us", "bwd_us", "pipe_elems", "dp_elems", "lower_bound_us"], list(out)))


def hybrid3d_sizes(st, num_stages, num_microbatches, num_tensor_shards):
    out = (C.c_uint64 * 6)()
    lib().dlnb_oracle_hybrid3d_sizes(st["avgForwardTime"], st["avgBackwardTime"], st["modelSize"], st["batchSize"],
                                     st["sequenceLength"], st["embeddedDim"], num_stages, num_microbatches, num_tensor_shards, out)
    return dict(zip(["fwd_us", "bwd_us", "pipe_elems", "dp_elems", "lower_bound_us", "tp_elems"], list(out)))


def hybrid3d_moe_sizes(st, num_layers, num_stages, num_microbatches, num_expert_shards):
    out = (C.c_uint64 * 8)()
    lib().dlnb_oracle_hybrid3d_moe_sizes(st["avgForwardTime"], st["avgBackwardTime"], st["modelSize"], st["nonExpertModelSize"],
                                         st["batchSize"], st["sequenceLength"], st["embeddedDim"], num_layers, num_stages,
                                         num_microbatches, num_expert_shards, out)
    return dict(zip(["fwd_us", "bwd_us", "pipe_elems", "dp_elems", "lower_bound_us", "a2a_elems_per_peer", "non_expert_elems",
                     "a2a_per_microbatch"], list(out)))


def msg_stats(sizes, mult=1):
    arr = (C.c_uint64 * len(sizes))(*sizes)
    a, s = C.c_float(), C.c_float()
    lib().dlnb_oracle_msg_stats(arr, len(sizes), mult, C.byref(a), C.byref(s))
    return a.value, s.value


STAT_KEYS = ["forwardFlops", "backwardFlops", "modelSize", "nonExpertModelSize", "avgForwardTime",
             "avgBackwardTime", "batchSize", "ffn_avgForwardTime", "ffn_avgBackwardTime", "experts",
             "sequenceLength", "embeddedDim"]


def parse_model_stats(path):
    out = (C.c_uint64 * 12)()
    rc = lib().dlnb_oracle_parse_model_stats(str(path).encode(), out)
    if rc:
        raise RuntimeError(f"cannot parse {path} (rc={rc})")
    return dict(zip(STAT_KEYS, list(out)))


def fill_pattern(count, dtype, kind, seed, rank):
    out = np.empty(count, dtype=NP[dtype])
    lib().dlnb_oracle_fill_pattern(out.ctypes.data, count, dtype, kind, seed, rank)
    return out


def checksum(arr):
    a = np.ascontiguousarray(arr)
    return lib().dlnb_oracle_checksum(a.ctypes.data, a.nbytes)
