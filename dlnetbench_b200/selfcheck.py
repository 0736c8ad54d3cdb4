"""Parity matrix on the LIVE ranks of a job, over the C ABI only (no CPU code involved).

Every rank calls run() at the same point. Each case fills seeded patterns on the device, runs one
collective of libdlnb.so across the real GPUs and verifies the result ON THE DEVICE against the
closed form of the pattern (dlnb_verify_sum_pattern: rank-order fp32 sum, one RNE rounding,
compared bit for bit; dlnb_verify_sum_pattern_tol for the in-switch reductions, whose summation
order is the NVSwitch's: integer-valued data exact, random data within 2 (n-1) eps * sum|x|).

Covered: allreduce one-shot / two-shot / in-switch / AUTO / in place, allgather and reduce-scatter
(peer and in-switch variants), all-to-all, send/recv ring, the unit / replica sub-groups fsdp builds
when sharding_factor < world (world >= 4), back-to-back launches with changing grids on both lanes.
bench.py runs this as an untimed prologue at N > 1 and reports the counts; tests/mg_worker.py runs it
and adds the comparisons against the CPU oracle.
"""
from __future__ import annotations

from . import capi

EPS = {capi.F32: 2.0 ** -23, capi.F16: 2.0 ** -10, capi.BF16: 2.0 ** -7}


class Tally:
    def __init__(self, rank):
        self.rank, self.cases, self.failures, self.multimem_cases, self.worst_rel_err = rank, 0, [], 0, 0.0

    def expect(self, ok, what):
        self.cases += 1
        if not ok:
            self.failures.append(what)
            print(f"[rank {self.rank}] PARITY FAIL {what}", flush=True)


def run(rank: int, world: int, quick: bool = True, dtypes=(capi.F32, capi.F16, capi.BF16)) -> dict:
    L = capi.lib()
    t = Tally(rank)
    has_mc = L.dlnb_get(b"multicast") == 1 and world > 1
    small = [1, 7, 4097, 65536 + 17]
    big = [1_000_003, 8_388_608] if quick else [1_000_003, 33_554_432 + 5, 110_240_167]
    groups = {"world": list(range(world))}
    if world >= 4:  # fsdp with sharding_factor = world/2: unit groups and replica groups (fsdp.cpp:258-265)
        F = world // 2
        groups["unit"] = [rank // F * F + i for i in range(F)]
        groups["replica"] = [rank % F + i * F for i in range(world // F)]

    for dtype in dtypes:
        esz = capi.DTYPE_SIZE[dtype]
        comms = {k: capi.B200Communicator(g, num_slots=4, dtype=dtype) for k, g in groups.items()}
        for gname, comm in comms.items():
            g = groups[gname]
            n, me = len(g), g.index(rank)
            whole = gname == "world"
            for count in small + (big if whole else big[:1]):
                for kind in ((capi.PAT_INT, capi.PAT_UNIFORM) if count <= 65536 + 17 else (capi.PAT_UNIFORM,)):
                    seed = 1234 + count % 7
                    send, recv = capi.Tensor(count, dtype), capi.Tensor(count, dtype)
                    send.fill(kind, seed, rank)
                    algos = [capi.ALGO_ONESHOT, capi.ALGO_TWOSHOT, capi.ALGO_AUTO] + ([capi.ALGO_MULTIMEM] if has_mc and whole else [])
                    for algo in algos:
                        if algo == capi.ALGO_ONESHOT and count > 40_000_000:
                            continue
                        in_switch = algo == capi.ALGO_MULTIMEM or (
                            algo == capi.ALGO_AUTO and has_mc and whole and n >= 4 and count * esz >= comm.get("multimem_min_bytes"))
                        comm.set("algo", algo)
                        recv.fill(capi.PAT_ZERO)
                        comm.Iallreduce(send.data, recv.data, count, 1)
                        comm.WaitAll(4)
                        if in_switch:
                            t.multimem_cases += 1
                            if kind == capi.PAT_INT:  # exact in any order
                                bad = capi.verify_sum_pattern(recv.data, count, 0, dtype, kind, seed, g)
                                t.expect(bad == 0, f"in-switch allreduce (integer) {gname} dtype={dtype} count={count}: {bad} mismatches")
                            else:
                                bad, worst = capi.verify_sum_pattern_tol(recv.data, count, 0, dtype, kind, seed, g, 2 * (n - 1) * EPS[dtype])
                                t.worst_rel_err = max(t.worst_rel_err, worst / EPS[dtype])
                                t.expect(bad == 0, f"in-switch allreduce (random, 2(n-1)eps) {gname} dtype={dtype} count={count}: {bad} outside, worst {worst:.3e}")
                        else:
                            bad = capi.verify_sum_pattern(recv.data, count, 0, dtype, kind, seed, g)
                            t.expect(bad == 0, f"allreduce {gname} dtype={dtype} algo={algo} count={count} kind={kind}: {bad} mismatches")
                    comm.set("algo", capi.ALGO_AUTO)
                    # split allreduce: part through the switch on lane 0, the rest peer to peer on lane 1, at the same time
                    if has_mc and whole and count >= 65536 + 17:
                        comm.set("algo", capi.ALGO_MULTIMEM)
                        comm.set("hybrid_min_bytes", 0)
                        for pct in (70, 30):
                            comm.set("hybrid_pct", pct)
                            recv.fill(capi.PAT_ZERO)
                            comm.Iallreduce(send.data, recv.data, count, 3)
                            comm.Wait(3)
                            if kind == capi.PAT_INT:
                                bad = capi.verify_sum_pattern(recv.data, count, 0, dtype, kind, seed, g)
                            else:
                                bad, _ = capi.verify_sum_pattern_tol(recv.data, count, 0, dtype, kind, seed, g, 2 * (n - 1) * EPS[dtype])
                            t.multimem_cases += 1
                            t.expect(bad == 0, f"split allreduce {pct}% in-switch dtype={dtype} count={count} kind={kind}: {bad} mismatches")
                        comm.set("hybrid_pct", 0)
                        comm.set("hybrid_min_bytes", 32 << 20)
                        comm.set("algo", capi.ALGO_AUTO)
                    # in place (send == recv): the runtime must keep it off the one-shot kernel
                    if kind == capi.PAT_UNIFORM and count <= 1_000_003:
                        inplace = capi.Tensor(count, dtype)
                        inplace.fill(kind, seed, rank)
                        comm.Allreduce(inplace.data, inplace.data, count)
                        if has_mc and whole and n >= 4 and count * esz >= comm.get("multimem_min_bytes"):
                            bad, _ = capi.verify_sum_pattern_tol(inplace.data, count, 0, dtype, kind, seed, g, 2 * (n - 1) * EPS[dtype])
                        else:
                            bad = capi.verify_sum_pattern(inplace.data, count, 0, dtype, kind, seed, g)
                        t.expect(bad == 0, f"allreduce in place {gname} dtype={dtype} count={count}: {bad} mismatches")
                        inplace.free()
                    # allgather (peer stores; through the switch where available): block r = rank g[r]'s pattern
                    gathered = capi.Tensor(count * n, dtype)
                    for algo in [capi.ALGO_AUTO] + ([capi.ALGO_MULTIMEM] if has_mc and whole else []):
                        comm.set("algo", algo)
                        gathered.fill(capi.PAT_ZERO)
                        comm.Iallgather(send.data, count, gathered.data, count, 2)
                        comm.Wait(2)
                        t.multimem_cases += algo == capi.ALGO_MULTIMEM
                        for r in range(n):
                            bad = capi.verify_sum_pattern(gathered.offset(r * count), count, 0, dtype, kind, seed, [g[r]])
                            t.expect(bad == 0, f"allgather {gname} dtype={dtype} algo={algo} count={count} block={r}: {bad} mismatches")
                    comm.set("algo", capi.ALGO_AUTO)
                    # all-to-all: block p of the result = block `me` of rank g[p]'s pattern
                    if count <= 1_000_003:
                        a_send, a_recv = capi.Tensor(count * n, dtype), capi.Tensor(count * n, dtype)
                        a_send.fill(kind, seed, rank)
                        comm.Alltoall(a_send.data, count, a_recv.data, count)
                        for r in range(n):
                            bad = capi.verify_sum_pattern(a_recv.offset(r * count), count, me * count, dtype, kind, seed, [g[r]])
                            t.expect(bad == 0, f"alltoall {gname} dtype={dtype} count={count} block={r}: {bad} mismatches")
                        a_send.free()
                        a_recv.free()
                    # reduce-scatter: send n*count, receive the sum of block `me`
                    rs_send, rs_recv = capi.Tensor(count * n, dtype), capi.Tensor(count, dtype)
                    rs_send.fill(kind, seed, rank)
                    for algo in [capi.ALGO_AUTO] + ([capi.ALGO_MULTIMEM] if has_mc and whole else []):
                        comm.set("algo", algo)
                        rs_recv.fill(capi.PAT_ZERO)
                        comm.Reduce_Scatter_block(rs_send.data, rs_recv.data, count)
                        through_switch = algo == capi.ALGO_MULTIMEM and (count * esz) % 16 == 0
                        t.multimem_cases += through_switch
                        if through_switch and kind == capi.PAT_UNIFORM:
                            bad, worst = capi.verify_sum_pattern_tol(rs_recv.data, count, me * count, dtype, kind, seed, g, 2 * (n - 1) * EPS[dtype])
                            t.worst_rel_err = max(t.worst_rel_err, worst / EPS[dtype])
                        else:
                            bad = capi.verify_sum_pattern(rs_recv.data, count, me * count, dtype, kind, seed, g)
                        t.expect(bad == 0, f"reduce_scatter {gname} dtype={dtype} algo={algo} count={count}: {bad} mismatches")
                    comm.set("algo", capi.ALGO_AUTO)
                    for x in (send, recv, gathered, rs_send, rs_recv):
                        x.free()
        # generations: many launches back to back, grid and algorithm changing every time, both lanes
        comm = comms["world"]
        count = 2_000_003
        send = capi.Tensor(count, dtype)
        send.fill(capi.PAT_UNIFORM, 5, rank)
        outs = [capi.Tensor(count, dtype) for _ in range(4)]
        nalg = 3 if has_mc else 2
        last_algo = [0] * 4
        for it in range(48):
            comm.set("ctas", [1, 7, 32, 64, 3, 128][it % 6])
            comm.set("mc_ctas", [4, 16, 32][it % 3])
            last_algo[it % 4] = 1 + it % nalg
            comm.set("algo", last_algo[it % 4])
            comm.Iallreduce(send.data, outs[it % 4].data, count, it % 4)
        comm.WaitAll(4)
        for o, al in zip(outs, last_algo):
            if al == capi.ALGO_MULTIMEM:
                bad, _ = capi.verify_sum_pattern_tol(o.data, count, 0, dtype, capi.PAT_UNIFORM, 5, list(range(world)), 2 * (world - 1) * EPS[dtype])
                t.multimem_cases += 1
            else:
                bad = capi.verify_sum_pattern(o.data, count, 0, dtype, capi.PAT_UNIFORM, 5, list(range(world)))
            t.expect(bad == 0, f"generations dtype={dtype} algo={al}: {bad} mismatches")
        comm.set("ctas", 0)
        comm.set("mc_ctas", 0)
        comm.set("algo", capi.ALGO_AUTO)
        for x in [send] + outs:
            x.free()
        for c in comms.values():
            c.finalize()

    # point-to-point: ring exchange with Irecv / Isend at misaligned offsets (pipeline stages, hybrid_2d.cpp:108-161)
    if world > 1 and L.dlnb_arena_kind() == b"vmm":
        for dtype in dtypes[:2]:
            comm = capi.B200Communicator(list(range(world)), num_slots=12, dtype=dtype)
            nxt, prv = (rank + 1) % world, (rank - 1) % world
            for count in [1, 7, 4097, 65536 + 17, 1_000_003]:
                out, inn = capi.Tensor(count + 8, dtype), capi.Tensor(count + 8, dtype)
                out.fill(capi.PAT_UNIFORM, 77, rank)
                for shift in (0, 1, 3):
                    inn.fill(capi.PAT_ZERO)
                    comm.Irecv(inn.offset(shift), count, prv, 0)
                    comm.Isend(out.offset(3 - shift), count, nxt, 1)
                    comm.WaitAll(2)
                    bad = capi.verify_sum_pattern(inn.offset(shift), count, 3 - shift, dtype, capi.PAT_UNIFORM, 77, [prv])
                    t.expect(bad == 0, f"send/recv ring dtype={dtype} count={count} shift={shift}: {bad} mismatches")
                out.free()
                inn.free()
            # a 9th receive from one source before any has been waited for must be refused (ring of 8 posts)
            bufs = [capi.Tensor(16, dtype) for _ in range(9)]
            for i in range(8):
                comm.Irecv(bufs[i].data, 16, prv, i)
            refused = False
            try:
                comm.Irecv(bufs[8].data, 16, prv, 8)
            except capi.DlnbError:
                refused = True
            t.expect(refused, "9th outstanding Irecv from one source is refused")
            src = capi.Tensor(16, dtype)
            src.fill(capi.PAT_INT, 3, rank)
            for i in range(8):
                comm.Isend(src.data, 16, nxt, 11)
                comm.Wait(11)
            comm.WaitAll(8)
            t.expect(capi.verify_sum_pattern(bufs[7].data, 16, 0, dtype, capi.PAT_INT, 3, [prv]) == 0, "eight posted receives all arrive")
            for b in bufs + [src]:
                b.free()
            comm.finalize()
    return {"cases": t.cases, "failures": len(t.failures), "multimem_cases": int(t.multimem_cases), "failed": t.failures[:8],
            "worst_in_switch_err_over_eps": round(t.worst_rel_err, 4), "groups": sorted(groups), "quick": quick}
