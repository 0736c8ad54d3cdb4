"""Stress of the cross-GPU flag protocol (SURVEY §7 "hard parts" #2).

N back-to-back collectives with, per launch and identical on all ranks (shared seed): random operation
(allreduce one-shot / two-shot / in-switch, allgather, reduce-scatter, all-to-all), random size (1 element
... --max-elems), random grid (1 ... 128 CTAs; in-switch 1 ... 48), alternating buffers out of a ring — and,
DIFFERENT on every rank, random host delays between launches (so ranks arrive at the barriers in every
order) and random emulated compute beside the communication. Every result is verified on the device
against the closed form of the seeded pattern (bit-exact; in-switch random data within 2(n-1)eps), every
`--verify-every` launches (1 = each launch). Both lanes are exercised: small allreduces alternate.

  bin/dlnb_launch -n 8 python tools/stress.py --launches 100000 --verify-every 1
"""
import argparse
import json
import os
import random
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi  # noqa: E402
from dlnetbench_b200.selfcheck import EPS  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches", type=int, default=100000)
    ap.add_argument("--max-elems", type=int, default=300_000)
    ap.add_argument("--verify-every", type=int, default=1)
    ap.add_argument("--seed", type=int, default=2026)
    ap.add_argument("--strict", type=int, default=0)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", os.environ.get("DLNB_RANK", 0)))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("DLNB_WORLD", 1)))
    capi.init(rank, world)
    capi.set_option("strict_barriers", a.strict)
    has_mc = capi.get_option("multicast") == 1 and world > 1
    shared = random.Random(a.seed)           # same decisions on every rank
    mine = random.Random(a.seed * 7919 + rank)  # this rank's delays
    ranks = list(range(world))
    RING = 6
    OFFMAX = 4096  # every launch reads its input at a random element offset of the pre-filled pattern: new data, new alignment
    comms, bufs = {}, {}
    for dtype in (capi.F32, capi.F16, capi.BF16):
        comms[dtype] = capi.B200Communicator(ranks, num_slots=RING, dtype=dtype)
        bufs[dtype] = [(capi.Tensor(a.max_elems * world + OFFMAX, dtype), capi.Tensor(a.max_elems * world, dtype)) for _ in range(RING)]
        for slot, (snd, _) in enumerate(bufs[dtype]):  # filled ONCE: no stream synchronisation between the launches below
            snd.fill(1 + slot % 2, 100 + slot, rank)
    ops = ["ar1", "ar2", "ag", "rs", "a2a"] + (["arm", "arm"] if has_mc else [])
    pending = []  # (slot, check closure) of launches not yet verified
    bad_total, verified, by_op = 0, 0, {}
    t_start = time.perf_counter()

    def flush():
        nonlocal bad_total, verified
        for dtype in comms:
            comms[dtype].WaitAll(RING)
        for what, fn in pending:
            bad = fn()
            verified += 1
            if bad:
                bad_total += 1
                print(f"[rank {rank}] STRESS FAIL {what}: {bad} mismatches", flush=True)
        pending.clear()

    for it in range(a.launches):
        op = shared.choice(ops)
        dtype = shared.choice((capi.F32, capi.F16, capi.BF16))
        # sizes: mostly small (latency path, both lanes), sometimes large; odd counts on purpose
        count = shared.choice((1, 3, 17, 255, 4097)) if shared.random() < 0.3 else shared.randint(1, a.max_elems)
        ctas = shared.choice((0, 1, 2, 7, 16, 33, 64, 128))
        mc_ctas = shared.choice((0, 1, 4, 16, 32, 48))
        off = shared.randint(0, OFFMAX)
        spin_us = shared.choice((0, 0, 0, 5, 40, 300))
        slot = it % RING
        comm = comms[dtype]
        send, recv = bufs[dtype][slot]
        # this rank's own pace: arrive early, late, or very late at the barriers
        r = mine.random()
        if r < 0.05:
            time.sleep(mine.random() * 2e-3)
        elif r < 0.3:
            t_end = time.perf_counter() + mine.random() * 50e-6
            while time.perf_counter() < t_end:
                pass
        if slot == 0 and it:
            flush()  # the ring of buffers is about to be reused
        kind, seed = 1 + slot % 2, 100 + slot
        if op == "arm":
            off &= ~7  # the in-switch kernel wants 16-byte aligned buffers
        src = send.offset(off)
        comm.set("ctas", ctas)
        comm.set("mc_ctas", mc_ctas)
        if spin_us:
            capi.compute_spin(spin_us)
        me = rank
        esz = capi.DTYPE_SIZE[dtype]
        if op in ("ar1", "ar2", "arm"):
            comm.set("algo", {"ar1": capi.ALGO_ONESHOT, "ar2": capi.ALGO_TWOSHOT, "arm": capi.ALGO_MULTIMEM}[op])
            comm.Iallreduce(src, recv.data, count, slot)
            if op == "arm" and kind == capi.PAT_UNIFORM:
                chk = lambda r_=recv, c=count, d=dtype, k=kind, s=seed, o=off: capi.verify_sum_pattern_tol(r_.data, c, o, d, k, s, ranks, 2 * (world - 1) * EPS[d])[0]
            else:
                chk = lambda r_=recv, c=count, d=dtype, k=kind, s=seed, o=off: capi.verify_sum_pattern(r_.data, c, o, d, k, s, ranks)
        elif op == "ag":
            comm.set("algo", capi.ALGO_AUTO)
            comm.Iallgather(src, count, recv.data, count, slot)
            chk = lambda r_=recv, c=count, d=dtype, k=kind, s=seed, o=off: sum(
                capi.verify_sum_pattern(r_.offset(p * c), c, o, d, k, s, [p]) for p in range(world))
        elif op == "rs":
            comm.set("algo", capi.ALGO_AUTO)
            capi.check(capi.lib().dlnb_ireduce_scatter(comm.h, src, recv.data, count, slot))
            chk = lambda r_=recv, c=count, d=dtype, k=kind, s=seed, o=off: capi.verify_sum_pattern(r_.data, c, o + me * c, d, k, s, ranks)
        else:
            comm.set("algo", capi.ALGO_AUTO)
            capi.check(capi.lib().dlnb_ialltoall(comm.h, src, recv.data, count, slot))
            chk = lambda r_=recv, c=count, d=dtype, k=kind, s=seed, o=off: sum(
                capi.verify_sum_pattern(r_.offset(p * c), c, o + me * c, d, k, s, [p]) for p in range(world))
        by_op[op] = by_op.get(op, 0) + 1
        if a.verify_every <= 1 or it % a.verify_every == 0:
            pending.append((f"launch {it} op={op} dtype={dtype} count={count} ctas={ctas}/{mc_ctas} kind={kind}", chk))
    flush()
    secs = time.perf_counter() - t_start
    # every rank must have seen zero failures
    all_bad = (capi.C.c_long * world)()
    mine_bad = capi.C.c_long(bad_total)
    capi.check(capi.lib().dlnb_host_allgather(capi.C.byref(mine_bad), 8, all_bad))
    out = {"stress": "flag protocol", "world": world, "launches": a.launches, "verified": verified, "failures_per_rank": list(all_bad),
           "by_op": by_op, "seconds": round(secs, 1), "strict_barriers": a.strict, "multicast": int(has_mc),
           "lanes": capi.get_option("lanes"), "max_elems": a.max_elems}
    if rank == 0:
        print(json.dumps(out), flush=True)
        if a.out:
            json.dump(out, open(a.out, "w"), indent=1)
    for c in comms.values():
        c.finalize()
    capi.finalize()
    sys.exit(1 if any(all_bad) else 0)


if __name__ == "__main__":
    main()
