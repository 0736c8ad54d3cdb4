"""One rank of the multi-GPU parity run (started N times by bin/dlnb_launch or torchrun).

Checks, through the C ABI on real peer memory: allreduce (one-shot, two-shot, auto),
allgather, reduce-scatter for fp32/fp16/bf16 at edge counts and BASELINE-shaped counts,
against (a) the CPU oracle on the host for small sizes and (b) the closed-form on-device
verifier for large ones; sub-communicators as fsdp builds them (unit groups / replica
groups); back-to-back launches with changing grids (flag generations). Exit code 0 = pass.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi  # noqa: E402
from oracle import oracle_capi as oracle  # noqa: E402

NPDT = {0: np.float32, 1: np.float16, 2: np.uint16}


def bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a.view(np.uint16)


def main():
    rank = int(os.environ.get("DLNB_RANK", os.environ.get("RANK", 0)))
    world = int(os.environ.get("DLNB_WORLD", os.environ.get("WORLD_SIZE", 1)))
    quick = "--quick" in sys.argv
    capi.init(rank, world)
    fails = []

    def expect(ok, what):
        if not ok:
            fails.append(what)
            print(f"[rank {rank}] FAIL {what}", flush=True)

    # ---- the parity matrix bench.py also runs on the live ranks: every collective, every algorithm (in-switch
    # included), sub-groups, in place, send/recv ring, generations — verified on the device against closed forms
    from dlnetbench_b200 import selfcheck

    res = selfcheck.run(rank, world, quick=quick)
    for f in res["failed"]:
        fails.append(f)
    expect(res["failures"] == 0, f"selfcheck: {res['failures']} of {res['cases']} cases failed")
    if capi.lib().dlnb_get(b"multicast") == 1 and world > 1:
        expect(res["multimem_cases"] > 0, "selfcheck ran no in-switch case although multicast is available")
    if rank == 0:
        print(f"[selfcheck] {res}", flush=True)

    # ---- against the CPU oracle on the host (small sizes): allreduce and reduce-scatter, every dtype, every group
    groups = {"world": list(range(world))}
    if world >= 4:
        F = world // 2
        groups["unit"] = [rank // F * F + i for i in range(F)]
        groups["replica"] = [rank % F + i * F for i in range(world // F)]
    for dtype in (0, 1, 2):
        for gname, g in groups.items():
            comm = capi.B200Communicator(g, num_slots=4, dtype=dtype)
            n, me = len(g), g.index(rank)
            for count in [1, 7, 8, 15, 4095, 4097, 65536 + 17]:
                for kind in (1, 2):
                    seed = 1234 + count % 7
                    send, recv = capi.Tensor(count, dtype), capi.Tensor(count, dtype)
                    send.fill(kind, seed, rank)
                    xs = [oracle.fill_pattern(count, dtype, kind, seed, r) for r in g]
                    want = oracle.allreduce(xs, dtype)
                    for algo in (capi.ALGO_ONESHOT, capi.ALGO_TWOSHOT):
                        comm.set("algo", algo)
                        comm.Iallreduce(send.data, recv.data, count, 1)
                        comm.WaitAll(4)
                        expect(np.array_equal(bits(recv.download(NPDT[dtype])), bits(want)), f"allreduce vs oracle {gname} {dtype} {algo} {count}")
                    comm.set("algo", capi.ALGO_AUTO)
                    if count <= 4097:
                        rs_send, rs_recv = capi.Tensor(count * n, dtype), capi.Tensor(count, dtype)
                        rs_send.fill(kind, seed, rank)
                        comm.Reduce_Scatter_block(rs_send.data, rs_recv.data, count)
                        xs = [oracle.fill_pattern(count * n, dtype, kind, seed, r) for r in g]
                        want = oracle.reduce_scatter_block(xs, me, count, dtype)
                        expect(np.array_equal(bits(rs_recv.download(NPDT[dtype])), bits(want)), f"reduce_scatter vs oracle {gname} {dtype} {count}")
                        rs_send.free()
                        rs_recv.free()
                    send.free()
                    recv.free()
            comm.finalize()

    # ---- more than 64 communicators over the life of the process: ids are reused after a collective reset
    t_in, t_out = capi.Tensor(5_001, 0), capi.Tensor(5_001, 0)
    t_in.fill(2, 9, rank)
    for i in range(70):
        sub = list(range(world)) if i % 2 == 0 else ([rank // 2 * 2, rank // 2 * 2 + 1] if world % 2 == 0 else [rank])
        c = capi.B200Communicator(sub, num_slots=1, dtype=0)
        c.Allreduce(t_in.data, t_out.data, 5_001)
        if i >= 60 or i < 2:
            expect(capi.verify_sum_pattern(t_out.data, 5_001, 0, 0, 2, 9, sub) == 0, f"allreduce on communicator #{i} (recycled ids)")
        c.finalize()
    t_in.free()
    t_out.free()

    # ---- dlnb_symm_free / reuse across ranks: recycled blocks keep their symmetric offsets and come back cleared
    for dtype in (0, 1):
        count = 1_000_001
        comm = capi.B200Communicator(list(range(world)), 1, dtype)
        a, b = capi.Tensor(count, dtype), capi.Tensor(count, dtype)
        a.fill(1, 11, rank)
        comm.Allreduce(a.data, b.data, count)
        addr_a, addr_b = a.data, b.data
        a.free()
        b.free()
        c, d = capi.Tensor(count, dtype), capi.Tensor(count, dtype)
        expect((c.data, d.data) == (addr_a, addr_b), "recycled blocks land where the freed ones were")
        zero = capi.verify_sum_pattern(d.data, count, 0, dtype, 0, 0, [0])  # PAT_ZERO: every element must be 0 again
        expect(zero == 0, f"recycled block is cleared dtype={dtype}")
        c.fill(1, 12, rank)
        comm.Allreduce(c.data, d.data, count)
        expect(capi.verify_sum_pattern(d.data, count, 0, dtype, 1, 12, list(range(world))) == 0, f"allreduce on recycled blocks dtype={dtype}")
        comm.finalize()
        c.free()
        d.free()

    # ---- point-to-point (pipeline stages): ring exchange with Irecv/Isend, then a blocking chain
    # (needs the contiguous VMM windows; the cudaIpc diagnostic arena does not offer send/recv)
    if world > 1 and capi.lib().dlnb_arena_kind() == b"vmm":
        for dtype in (0, 1):
            comm = capi.B200Communicator(list(range(world)), num_slots=4, dtype=dtype)
            nxt, prv = (rank + 1) % world, (rank - 1) % world
            for count in [1, 7, 4097, 65536 + 17, 1_000_003, 8_388_608 + 3]:
                out, inn = capi.Tensor(count + 8, dtype), capi.Tensor(count + 8, dtype)
                out.fill(2, 77, rank)
                for shift in (0, 1, 3):  # shifted buffers: misaligned source and/or destination
                    inn.fill(0)
                    comm.Irecv(inn.offset(shift), count, prv, 0)
                    comm.Isend(out.offset(3 - shift), count, nxt, 1)
                    comm.WaitAll(2)
                    bad = capi.verify_sum_pattern(inn.offset(shift), count, 3 - shift, dtype, 2, 77, [prv])
                    expect(bad == 0, f"send/recv ring dtype={dtype} count={count} shift={shift}: {bad} mismatches")
            # blocking chain 0 -> 1 -> ... -> n-1 (what a pipeline's forward pass does)
            count = 2_000_001
            buf = capi.Tensor(count, dtype)
            if rank == 0:
                buf.fill(1, 5, 0)
            else:
                comm.recv(buf.data, count, rank - 1)
            if rank + 1 < world:
                comm.send(buf.data, count, rank + 1)
            expect(capi.verify_sum_pattern(buf.data, count, 0, dtype, 1, 5, [0]) == 0, f"send/recv chain dtype={dtype}")
            comm.finalize()

    # ---- a dp-shaped overlapped iteration: spin -> allreduce chains, device-timed
    comm = capi.B200Communicator(list(range(world)), num_slots=8, dtype=capi.F16)
    count = 4_000_001
    send = [capi.Tensor(count, capi.F16) for _ in range(8)]
    recv = [capi.Tensor(count, capi.F16) for _ in range(8)]
    for t in send:
        t.fill(1, 0, rank)
    capi.spin_calibrate()
    capi.iter_begin()
    capi.compute_spin(2000.0)
    for i in range(8):
        capi.compute_spin(500.0)
        comm.Iallreduce(send[i].data, recv[i].data, count, i)
    comm.WaitAll(8)
    ms = capi.iter_end()
    for t in recv:
        expect(capi.verify_sum_pattern(t.data, count, 0, capi.F16, 1, 0, list(range(world))) == 0, "dp iteration result")
    expect(6.0 <= ms < 8.0, f"dp iteration took {ms:.3f} ms for 6 ms of emulated compute")

    # ---- the same iteration captured as a CUDA graph and replayed: generations live in device memory
    for t in recv:
        t.fill(0)
    with capi.Graph() as g:
        capi.compute_spin(1000.0)
        for i in range(8):
            capi.compute_spin(250.0)
            comm.Iallreduce(send[i].data, recv[i].data, count, i)
        comm.WaitAll(8)
    for rep in range(3):
        capi.iter_begin()
        g.launch()
        gms = capi.iter_end()
        expect(3.0 <= gms < 4.0, f"graph replay {rep} took {gms:.3f} ms for 3 ms of emulated compute")
    for t in recv:
        expect(capi.verify_sum_pattern(t.data, count, 0, capi.F16, 1, 0, list(range(world))) == 0, "graph replay result")
    g.destroy()
    comm.finalize()

    # ---- an fsdp-shaped schedule captured as a graph: the blocking reduce-scatter keeps the next unit's compute
    # waiting inside the replay too (device-side edge instead of host blocking), so graph and plain times agree
    comm = capi.B200Communicator(list(range(world)), num_slots=6, dtype=capi.F16)
    shard = 8_000_000
    gather, mine = capi.Tensor(shard * world, capi.F16), capi.Tensor(shard, capi.F16)
    gather.fill(1, 0, rank)

    def fsdp_like():
        for u in range(6):
            capi.compute_spin(300.0)
            comm.Reduce_Scatter_block(gather.data, mine.data, shard)

    fsdp_like()
    capi.iter_begin()
    fsdp_like()
    plain = capi.iter_end()
    rs_ms = comm.slot_ms(-1)
    with capi.Graph() as g:
        fsdp_like()
    g.launch()
    capi.compute_sync()
    reps = []
    for _ in range(3):
        capi.iter_begin()
        g.launch()
        reps.append(capi.iter_end())
    expect(min(reps) >= 1.8 + 0.7 * 6 * rs_ms, f"captured fsdp schedule hides the reduce-scatter: {min(reps):.3f} ms, rs {rs_ms:.3f} ms")
    expect(min(reps) <= plain * 1.10 + 0.1, f"captured fsdp schedule {min(reps):.3f} ms vs plain {plain:.3f} ms")
    expect(capi.verify_sum_pattern(mine.data, shard, rank * shard, capi.F16, 1, 0, list(range(world))) == 0, "captured reduce-scatter result")
    g.destroy()
    comm.finalize()

    arena = capi.lib().dlnb_arena_kind().decode()
    mc = capi.lib().dlnb_get(b"multicast")
    capi.finalize()
    print(f"[rank {rank}/{world}] arena={arena} multicast={mc} iteration={ms:.3f} ms failures={len(fails)}", flush=True)
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
