"""Collective micro-sweep through the C ABI (device-timed, slot events).

Single process: world=1 (HBM copy path). Under `torchrun`/dlnb_launch (RANK, WORLD_SIZE
set): the real multi-GPU kernels; rank 0 prints. busBW follows the nccl-tests
convention: allreduce 2(n-1)/n*S/t, allgather/reduce-scatter (n-1)/n*S_tot/t;
at n=1: 2*S/t of HBM traffic.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="allreduce", choices=["allreduce", "allgather", "reduce_scatter", "sendrecv"])
    ap.add_argument("--dtype", default="f16", choices=["f32", "f16", "bf16"])
    ap.add_argument("--min-bytes", type=int, default=1 << 10)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--algo", type=int, default=0)
    ap.add_argument("--threads", type=int, default=512)
    ap.add_argument("--out", default="")
    ap.add_argument("--mc-ctas", default="0", help="grid(s) of the in-switch kernels (comma list, 0 = default by group size)")
    ap.add_argument("--counts", default="", help="explicit element counts per rank message (comma list) instead of the size ladder; "
                                                  "for allgather / reduce_scatter: the SHARD count (e.g. 110240167, the odd llama3_70b shard)")
    ap.add_argument("--nvlink", action="store_true", help="NVML link counters around each batch: measured bytes per launch per direction")
    ap.add_argument("--strict", type=int, default=0, help="1: st.release / ld.acquire on every flag barrier")
    ap.add_argument("--per-slot", dest="b2b", action="store_false", help="busBW from per-launch slot events instead of the back-to-back batch")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", os.environ.get("DLNB_RANK", 0)))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("DLNB_WORLD", 1)))
    if a.op == "sendrecv":
        a.iters = min(a.iters, 8)  # at most 8 receives from one source may be outstanding
    capi.init(rank, world)
    dt = {"f32": capi.F32, "f16": capi.F16, "bf16": capi.BF16}[a.dtype]
    esz = capi.DTYPE_SIZE[dt]
    comm = capi.B200Communicator(list(range(world)), num_slots=2 * a.iters, dtype=dt)
    comm.set("algo", a.algo)
    comm.set("threads", a.threads)
    capi.set_option("strict_barriers", a.strict)
    links = None
    if a.nvlink:
        from dlnetbench_b200 import nvlink

        links = nvlink.LinkCounters(int(os.environ.get("LOCAL_RANK", os.environ.get("DLNB_LOCAL_RANK", rank))))
        if not links.ok:
            if rank == 0:
                print(json.dumps({"nvlink_counters_unavailable": links.why}), flush=True)
            links = None
    # ring of buffers larger than L2 so no iteration finds its input cached
    nxt, prv = (rank + 1) % world, (rank - 1) % world

    def issue(s, r, sc, rc, i):
        if a.op == "allreduce":
            comm.Iallreduce(s.data, r.data, sc, i)
        elif a.op == "allgather":
            comm.Iallgather(s.data, sc, r.data, sc, i)
        elif a.op == "reduce_scatter":
            capi.check(capi.lib().dlnb_ireduce_scatter(comm.h, s.data, r.data, rc, i))
        else:  # ring exchange: every rank receives from its left and sends to its right
            comm.Irecv(r.data, rc, prv, i)  # a slot holds ONE pending operation: the send gets a slot of its own
            comm.Isend(s.data, sc, nxt, a.iters + i)

    rows = []
    ladder = []
    size = a.min_bytes
    while size <= a.max_bytes and not a.counts:
        ladder.append((size, None))
        size *= 4
    for c in [int(x) for x in a.counts.split(",") if x]:
        ladder.append((c * esz * (1 if a.op in ("allreduce", "sendrecv") else world), c))
    for size, shard in ladder:
        count = size // esz
        if a.op in ("allreduce", "sendrecv"):
            sc, rc = count, count
        elif a.op == "allgather":
            sc = shard if shard else count // world
            rc = sc * world
        else:
            rc = shard if shard else count // world
            sc = rc * world
        nbuf = max(2, min(a.iters, (300 << 20) // max(size, 1) + 1))
        send = [capi.Tensor(max(sc, 1), dt) for _ in range(nbuf)]
        recv = [capi.Tensor(max(rc, 1), dt) for _ in range(nbuf)]
        for t in send:
            t.fill(capi.PAT_INT, 0, rank)
        for ctas, mc_ctas in [(int(c), int(m)) for c in a.ctas.split(",") for m in a.mc_ctas.split(",")]:
            comm.set("ctas", ctas)
            comm.set("mc_ctas", mc_ctas)
            for rep in range(2):  # first pass = warm-up
                for i in range(a.iters):
                    issue(send[i % nbuf], recv[i % nbuf], sc, rc, i)
                comm.WaitAll(2 * a.iters)
            ms = sorted(comm.slot_ms(i) for i in range(a.iters))
            # device-only batch time: a spin blocks the communication stream while the host queues
            # the whole batch (no per-slot timing events), one event pair around the batch
            capi.set_option("slot_timing", 0)
            before = links.settle() if links else None
            capi.compute_spin(max(3000.0, 40.0 * a.iters))
            capi.batch_begin()
            for i in range(a.iters):
                issue(send[i % nbuf], recv[i % nbuf], sc, rc, i)
            b2b = capi.iter_end() / a.iters
            comm.WaitAll(2 * a.iters)  # posted receives count as outstanding until they have been waited for
            wire = None
            if links:
                from dlnetbench_b200 import nvlink

                wire = nvlink.per_launch(before, links.settle(), a.iters)
            capi.set_option("slot_timing", 1)
            med = b2b if a.b2b else ms[len(ms) // 2]
            if world == 1:
                bus = 2 * sc * esz / (med * 1e-3) / 1e9
            elif a.op == "sendrecv":
                bus = sc * esz / (med * 1e-3) / 1e9  # bytes per direction per GPU
            elif a.op == "allreduce":
                bus = 2 * (world - 1) / world * sc * esz / (med * 1e-3) / 1e9
            else:
                bus = (world - 1) / world * max(sc, rc) * esz / (med * 1e-3) / 1e9
            row = dict(op=a.op, dtype=a.dtype, world=world, bytes=sc * esz if a.op != "reduce_scatter" else sc * esz,
                       ctas=ctas, mc_ctas=mc_ctas, algo=a.algo, us=med * 1e3, us_slot_med=ms[len(ms) // 2] * 1e3, us_b2b=b2b * 1e3, busbw_GBps=bus)
            if wire:
                row.update({k: round(v) for k, v in wire.items()})
                row["nvlink_tx_GBps"] = wire["data_tx_bytes_per_launch"] / (b2b * 1e-3) / 1e9
                row["nvlink_rx_GBps"] = wire["data_rx_bytes_per_launch"] / (b2b * 1e-3) / 1e9
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
        for t in send + recv:
            t.free()
    if rank == 0 and a.out:
        json.dump(rows, open(a.out, "w"), indent=1)
    capi.finalize()


if __name__ == "__main__":
    main()
