"""Small, fast pass over every kernel body for compute-sanitizer (memcheck / racecheck / synccheck / initcheck).

  compute-sanitizer --tool memcheck python tools/sanitize_probe.py

One GPU: the multi-rank bodies run as n virtual ranks (n = 2..8, 5/6/7 included) in one cooperative launch, sizes
chosen so that heads, tails, the aligned and the shifted 128-bit movers and several overlapping runs per warp are
all walked; plus the point-to-point kernels in loop-back, the world-1 communicator path, spin kernels of every
footprint, the verifier kernels and the no-allocation entry. Results are checked (closed form) so that a clean
sanitizer log is a log of CORRECT runs. Prints one JSON line.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi  # noqa: E402


def main():
    capi.init(0, 1, device=0, session="sanitize-%d" % os.getpid())
    bad, cases = 0, 0
    counts = [1, 33, 4097, 40_001]
    for n in range(2, 9):
        ranks = list(range(n))
        for dtype in (capi.F32, capi.F16, capi.BF16):
            for count in counts:
                send = [capi.Tensor(count * n + 8, dtype) for _ in range(n)]
                recv = [capi.Tensor(count * n + 8, dtype) for _ in range(n)]
                for r, t in enumerate(send):
                    t.fill(capi.PAT_UNIFORM, 3, r)
                sp, rp = [t.data for t in send], [t.data for t in recv]
                for algo in (capi.ALGO_ONESHOT, capi.ALGO_TWOSHOT):
                    capi.emu_allreduce(sp, rp, count, dtype, algo=algo, ctas=3)
                    bad += sum(capi.verify_sum_pattern(recv[r].data, count, 0, dtype, capi.PAT_UNIFORM, 3, ranks) != 0 for r in range(n))
                    cases += n
                # misaligned views (element offset 1 / 3): shifted mover on both sides
                capi.emu_allreduce([t.offset(1) for t in send], [t.offset(3) for t in recv], count, dtype, algo=capi.ALGO_TWOSHOT, ctas=2)
                bad += sum(capi.verify_sum_pattern(recv[r].offset(3), count, 1, dtype, capi.PAT_UNIFORM, 3, ranks) != 0 for r in range(n))
                capi.emu_allgather(sp, rp, count, dtype, ctas=3)
                bad += sum(capi.verify_sum_pattern(recv[0].offset(p * count), count, 0, dtype, capi.PAT_UNIFORM, 3, [p]) != 0 for p in range(n))
                capi.emu_reduce_scatter(sp, rp, count, dtype, ctas=3)
                bad += sum(capi.verify_sum_pattern(recv[r].data, count, r * count, dtype, capi.PAT_UNIFORM, 3, ranks) != 0 for r in range(n))
                capi.emu_alltoall(sp, rp, count, dtype, ctas=3)
                bad += sum(capi.verify_sum_pattern(recv[1].offset(p * count), count, 1 * count, dtype, capi.PAT_UNIFORM, 3, [p]) != 0 for p in range(n))
                cases += 4 * n
                for t in send + recv:
                    t.free()
    # no-allocation entry, loop-back send/recv, world-1 communicator, spin footprints
    for dtype in (capi.F32, capi.F16):
        count = 40_001
        a, b = [capi.Tensor(count + 8, dtype) for _ in range(4)], [capi.Tensor(count + 8, dtype) for _ in range(4)]
        for r, t in enumerate(a):
            t.fill(capi.PAT_INT, 5, r)
        capi.emu_allreduce_noalloc([t.data for t in a], [t.data for t in b], count, dtype, algo=capi.ALGO_TWOSHOT, ctas=4)
        bad += capi.verify_sum_pattern(b[2].data, count, 0, dtype, capi.PAT_INT, 5, [0, 1, 2, 3]) != 0
        for shift in (0, 1, 3):
            capi.emu_sendrecv(a[0].offset(3 - shift), b[0].offset(shift), count, dtype, ctas=2)
            bad += capi.verify_sum_pattern(b[0].offset(shift), count, 3 - shift, dtype, capi.PAT_INT, 5, [0]) != 0
        comm = capi.B200Communicator([0], num_slots=2, dtype=dtype)
        comm.Allreduce(a[1].data, b[1].data, count)
        comm.Allgather(a[1].data, count, b[1].data, count)
        comm.Reduce_Scatter_block(a[1].data, b[1].data, count)
        bad += capi.verify_sum_pattern(b[1].data, count, 0, dtype, capi.PAT_INT, 5, [1]) != 0
        bad += capi.verify_sum_pattern_tol(b[1].data, count, 0, dtype, capi.PAT_INT, 5, [1], 0.0)[0] != 0
        comm.finalize()
        cases += 7
        for t in a + b:
            t.free()
    for mode in (capi.SPIN_SLEEP, capi.SPIN_FMA, capi.SPIN_HBM):
        capi.set_option("spin_mode", mode)
        capi.iter_begin()
        capi.compute_spin(200.0)
        capi.compute_spin(200.0)
        capi.iter_end()
    print(json.dumps({"sanitize_probe": "done", "cases": cases, "wrong": int(bad)}), flush=True)
    capi.finalize()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
