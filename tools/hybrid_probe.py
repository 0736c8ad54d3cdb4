"""Does an allreduce split between the NVSwitch (multimem) and the peer-to-peer two-shot kernel, both running at once on
the two communication lanes, beat either alone? The in-switch kernel leaves link bandwidth unused (its wire rate stalls
near 540 GB/s per direction of ~757 measured), the two-shot kernel can use the rest.

  bin/dlnb_launch -n 4 python tools/hybrid_probe.py --bytes 67108864 268435456 --fractions 1.0 0.9 0.8 0.7 0.6 0.0
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, nargs="+", default=[67108864, 268435456])
    ap.add_argument("--fractions", type=float, nargs="+", default=[1.0, 0.9, 0.8, 0.7, 0.6, 0.0])
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--mc-ctas", type=int, default=16)
    ap.add_argument("--p2p-ctas", type=int, default=64)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", os.environ.get("DLNB_RANK", 0)))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("DLNB_WORLD", 1)))
    capi.init(rank, world)
    comm = capi.B200Communicator(list(range(world)), num_slots=2 * a.iters, dtype=capi.F16)
    comm.set("mc_ctas", a.mc_ctas)
    comm.set("ctas", a.p2p_ctas)
    rows = []
    for size in a.bytes:
        count = size // 2
        nbuf = max(2, min(a.iters, (300 << 20) // size + 1))
        send = [capi.Tensor(count, capi.F16) for _ in range(nbuf)]
        recv = [capi.Tensor(count, capi.F16) for _ in range(nbuf)]
        for t in send:
            t.fill(capi.PAT_INT, 0, rank)
        for f in a.fractions:
            c_mm = int(count * f) // 4096 * 4096  # multimem part (a multiple of 8 KiB), the rest goes peer to peer
            c_pp = count - c_mm

            def issue(i):
                s, r = send[i % nbuf], recv[i % nbuf]
                if c_mm:
                    comm.set("algo", capi.ALGO_MULTIMEM)
                    comm.Iallreduce(s.data, r.data, c_mm, 2 * i)
                if c_pp:
                    comm.set("algo", capi.ALGO_TWOSHOT)
                    comm.Iallreduce(s.offset(c_mm), r.offset(c_mm), c_pp, 2 * i + 1)

            for i in range(a.iters):
                issue(i)
            comm.WaitAll(2 * a.iters)
            capi.set_option("slot_timing", 0)
            capi.compute_spin(max(3000.0, 80.0 * a.iters))
            capi.batch_begin()
            for i in range(a.iters):
                issue(i)
            ms = capi.iter_end() / a.iters
            capi.set_option("slot_timing", 1)
            comm.WaitAll(2 * a.iters)
            bad = capi.verify_sum_pattern(recv[0].data, count, 0, capi.F16, capi.PAT_INT, 0, list(range(world)))
            row = {"bytes": size, "world": world, "fraction_in_switch": f, "us": ms * 1e3, "busbw_GBps": 2 * (world - 1) / world * size / (ms * 1e-3) / 1e9,
                   "mc_ctas": a.mc_ctas, "p2p_ctas": a.p2p_ctas, "mismatches": bad}
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
        for t in send + recv:
            t.free()
    if rank == 0 and a.out:
        json.dump(rows, open(a.out, "w"), indent=1)
    comm.finalize()
    capi.finalize()


if __name__ == "__main__":
    main()
