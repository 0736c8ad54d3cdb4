#!/usr/bin/env python3
"""One-GPU probe of the MULTI-RANK kernel bodies: n virtual ranks in one cooperative launch
(`dlnb_emu_*`), every "peer" buffer being local HBM. It exists so that the two-shot allreduce,
allgather, reduce-scatter and all-to-all bodies — which `ncu` may not wrap as a multi-process run —
can be captured by `ncu --set full -k regex:coll_kernel_emu` on a single GPU, and to show how far the
kernels themselves are from the NVLink rate when the links are taken out of the picture.

Timing is host wall clock around the (synchronous) entry point, minimum of `--reps` calls; the HBM
traffic model per launch is the sum over virtual ranks of bytes read + written by the body.

    python tools/emu_probe.py [--nranks 2 8] [--mib 64] [--ops allreduce allgather reduce_scatter alltoall]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from dlnetbench_b200 import capi  # noqa: E402


def traffic_bytes(op: str, n: int, s: int) -> int:
    """HBM bytes one launch moves (all virtual ranks): reads + writes of the kernel body."""
    if op == "allreduce":      # two-shot: every rank reads its slice from n ranks and writes it to n ranks
        return 2 * n * s
    if op == "allgather":      # read own shard (s) once, store it n times
        return n * (s + n * s)
    if op == "reduce_scatter":  # pull chunk (s) from n ranks, write it once
        return n * (n * s + s)
    if op == "alltoall":       # read n blocks (s each), write n blocks
        return n * 2 * n * s
    raise ValueError(op)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nranks", type=int, nargs="+", default=[2, 8])
    ap.add_argument("--mib", type=float, default=64.0, help="allreduce buffer / allgather+reduce-scatter+all-to-all TOTAL per rank")
    ap.add_argument("--ops", nargs="+", default=["allreduce", "allgather", "reduce_scatter", "alltoall"])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--ctas", type=int, default=0, help="CTAs per virtual rank (0 = as many as stay co-resident)")
    ap.add_argument("--odd", action="store_true", help="odd shard / block counts: rank r's chunk starts at r*count*2 bytes, the misaligned (shifted) mover runs")
    ap.add_argument("--shift-overlap", type=int, default=0, help="window overlap of the shifted mover in vectors (1, 2 or 8; 0 = library default)")
    args = ap.parse_args()

    capi.init(0, 1)
    if args.shift_overlap:
        capi.set_option("shift_overlap", args.shift_overlap)
    dtype, esz = capi.F16, 2
    total = int(args.mib * (1 << 20)) // esz          # elements of the large buffer of one rank
    rows = []
    for n in args.nranks:
        unit = total // n // 8 * 8 + (1 if args.odd else 0)  # per-peer block / shard, 16-byte aligned unless --odd
        big = [capi.Tensor(unit * n, dtype) for _ in range(n)]
        big2 = [capi.Tensor(unit * n, dtype) for _ in range(n)]
        for r, t in enumerate(big):
            t.fill(capi.PAT_INT, 7, r)
        for op in args.ops:
            ctas = args.ctas or 512
            if op == "allreduce":
                s_bytes, call = unit * n * esz, lambda: capi.emu_allreduce([t.data for t in big], [t.data for t in big2], unit * n, dtype,
                                                                           algo=capi.ALGO_TWOSHOT, ctas=ctas)
            elif op == "allgather":
                s_bytes, call = unit * esz, lambda: capi.emu_allgather([t.data for t in big], [t.data for t in big2], unit, dtype, ctas=ctas)
            elif op == "reduce_scatter":
                s_bytes, call = unit * esz, lambda: capi.emu_reduce_scatter([t.data for t in big], [t.data for t in big2], unit, dtype, ctas=ctas)
            else:
                s_bytes, call = unit * esz, lambda: capi.emu_alltoall([t.data for t in big], [t.data for t in big2], unit, dtype, ctas=ctas)
            call()  # warm-up (first use allocates the emulation flags)
            best = 1e30
            for _ in range(args.reps):
                capi.check(capi.lib().dlnb_flush_l2())
                big[0].download("uint16", 1)  # synchronous: the flush is over before the clock starts
                t0 = time.perf_counter()
                call()
                best = min(best, time.perf_counter() - t0)
            tb = traffic_bytes(op, n, s_bytes)
            rows.append({"op": op, "virtual_ranks": n, "odd": args.odd, "shift_overlap": capi.get_option("shift_overlap"), "bytes_per_rank_unit": s_bytes, "us": round(best * 1e6, 1),
                         "hbm_traffic_bytes": tb, "hbm_GBps": round(tb / best / 1e9, 1)})
            print(json.dumps(rows[-1]), flush=True)
        for t in big + big2:
            capi.check(capi.lib().dlnb_symm_free(t.data))
    capi.finalize()


if __name__ == "__main__":
    main()
