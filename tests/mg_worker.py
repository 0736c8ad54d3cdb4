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

    small = [1, 7, 8, 15, 4095, 4097, 65536 + 17]
    big = [1_000_003, 8_388_608] if quick else [1_000_003, 33_554_432 + 5, 110_240_167]
    groups = {"world": list(range(world))}
    if world >= 4:  # fsdp with sharding_factor = world/2: unit groups and replica groups
        F = world // 2
        groups["unit"] = [rank // F * F + i for i in range(F)]
        groups["replica"] = [rank % F + i * F for i in range(world // F)]

    for dtype in (0, 1, 2):
        comms = {k: capi.B200Communicator(g, num_slots=4, dtype=dtype) for k, g in groups.items()}
        for gname, comm in comms.items():
            g = groups[gname]
            n, me = len(g), g.index(rank)
            for count in small + (big if gname == "world" else big[:1]):
                kinds = (1, 2) if count <= 65536 + 17 else (2,)
                for kind in kinds:
                    seed = 1234 + count % 7
                    # ---- allreduce, every algorithm
                    send, recv = capi.Tensor(count, dtype), capi.Tensor(count, dtype)
                    send.fill(kind, seed, rank)
                    has_mc = capi.lib().dlnb_get(b"multicast") == 1 and gname == "world" and world > 1
                    algos = [capi.ALGO_ONESHOT, capi.ALGO_TWOSHOT, capi.ALGO_AUTO] + ([capi.ALGO_MULTIMEM] if has_mc else [])
                    for algo in algos:
                        if algo == capi.ALGO_ONESHOT and count > 40_000_000:
                            continue
                        # the in-switch reduction does not promise the oracle's summation order: exact for
                        # the integer pattern, within 2(n-1)eps for random data (checked on the host below)
                        in_switch = algo == capi.ALGO_MULTIMEM or (algo == capi.ALGO_AUTO and has_mc and n >= 4 and count * capi.DTYPE_SIZE[dtype] >= comm.get("multimem_min_bytes"))
                        comm.set("algo", algo)
                        recv.fill(0)
                        comm.Iallreduce(send.data, recv.data, count, 1)
                        comm.WaitAll(4)
                        if in_switch and kind == 2:
                            if count <= 1_000_003 and dtype != 2:
                                xs = [oracle.fill_pattern(count, dtype, kind, seed, r).astype(np.float64) for r in g]
                                ref, mag = np.sum(xs, axis=0), np.sum(np.abs(xs), axis=0)
                                got = recv.download(NPDT[dtype]).astype(np.float64)
                                eps = 2.0 ** -23 if dtype == 0 else 2.0 ** -10
                                expect(bool(np.all(np.abs(got - ref) <= 2 * (n - 1) * eps * mag + 1e-30)), f"multimem allreduce tolerance dtype={dtype} count={count}")
                            # same integer-exact check as everyone else, on a second pass with the integer pattern
                            send.fill(1, 0, rank)
                            comm.Iallreduce(send.data, recv.data, count, 1)
                            comm.WaitAll(4)
                            bad = capi.verify_sum_pattern(recv.data, count, 0, dtype, 1, 0, g)
                            expect(bad == 0, f"multimem allreduce (integer pattern) dtype={dtype} algo={algo} count={count}: {bad} mismatches")
                            send.fill(kind, seed, rank)
                            continue
                        bad = capi.verify_sum_pattern(recv.data, count, 0, dtype, kind, seed, g)
                        expect(bad == 0, f"allreduce {gname} dtype={dtype} algo={algo} count={count} kind={kind}: {bad} mismatches")
                    if count <= 65536 + 17:
                        comm.set("algo", capi.ALGO_TWOSHOT)
                        comm.Iallreduce(send.data, recv.data, count, 1)
                        comm.WaitAll(4)
                        xs = [oracle.fill_pattern(count, dtype, kind, seed, r) for r in g]
                        want = oracle.allreduce(xs, dtype)
                        expect(np.array_equal(bits(recv.download(NPDT[dtype])), bits(want)), f"allreduce vs oracle {gname} {dtype} {count}")
                    comm.set("algo", capi.ALGO_AUTO)
                    # ---- allgather: every block r must be rank g[r]'s pattern
                    gathered = capi.Tensor(count * n, dtype)
                    comm.Iallgather(send.data, count, gathered.data, count, 2)
                    comm.Wait(2)
                    for r in range(n):
                        bad = capi.verify_sum_pattern(gathered.offset(r * count), count, 0, dtype, kind, seed, [g[r]])
                        expect(bad == 0, f"allgather {gname} dtype={dtype} count={count} block={r}: {bad} mismatches")
                    # ---- all-to-all: block p of the result must be block `me` of rank g[p]'s pattern
                    if count <= 1_000_003:
                        a2a_send, a2a_recv = capi.Tensor(count * n, dtype), capi.Tensor(count * n, dtype)
                        a2a_send.fill(kind, seed, rank)
                        comm.Alltoall(a2a_send.data, count, a2a_recv.data, count)
                        for r in range(n):
                            bad = capi.verify_sum_pattern(a2a_recv.offset(r * count), count, me * count, dtype, kind, seed, [g[r]])
                            expect(bad == 0, f"alltoall {gname} dtype={dtype} count={count} block={r}: {bad} mismatches")
                    # ---- reduce-scatter of the gathered buffer's shape: send n*count, receive block `me`
                    rs_send, rs_recv = capi.Tensor(count * n, dtype), capi.Tensor(count, dtype)
                    rs_send.fill(kind, seed, rank)
                    comm.Reduce_Scatter_block(rs_send.data, rs_recv.data, count)
                    bad = capi.verify_sum_pattern(rs_recv.data, count, me * count, dtype, kind, seed, g)
                    expect(bad == 0, f"reduce_scatter {gname} dtype={dtype} count={count}: {bad} mismatches")
                    if count <= 4097:
                        xs = [oracle.fill_pattern(count * n, dtype, kind, seed, r) for r in g]
                        want = oracle.reduce_scatter_block(xs, me, count, dtype)
                        expect(np.array_equal(bits(rs_recv.download(NPDT[dtype])), bits(want)), f"reduce_scatter vs oracle {gname} {dtype} {count}")
        # ---- generations: many launches, grid changing every time, results checked at the end
        comm = comms["world"]
        count = 2_000_003
        send = capi.Tensor(count, dtype)
        send.fill(2, 5, rank)
        outs = [capi.Tensor(count, dtype) for _ in range(4)]
        for it in range(40):
            comm.set("ctas", [1, 7, 32, 64, 3, 128][it % 6])
            comm.set("algo", 1 + it % (3 if capi.lib().dlnb_get(b"multicast") == 1 and world > 1 and False else 2))
            comm.Iallreduce(send.data, outs[it % 4].data, count, it % 4)
        comm.WaitAll(4)
        for o in outs:
            expect(capi.verify_sum_pattern(o.data, count, 0, dtype, 2, 5, list(range(world))) == 0, f"generations dtype={dtype}")
        comm.set("ctas", 0)
        comm.set("algo", 0)
        for c in comms.values():
            c.finalize()

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

    arena = capi.lib().dlnb_arena_kind().decode()
    mc = capi.lib().dlnb_get(b"multicast")
    capi.finalize()
    print(f"[rank {rank}/{world}] arena={arena} multicast={mc} iteration={ms:.3f} ms failures={len(fails)}", flush=True)
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
