"""Round-2 GPU parity and behaviour tests (one B200; through the C ABI).

Group sizes 5, 6, 7 (every body), the rewritten shifted mover on misaligned chunks with many
CTAs, in-place allreduce, the point-to-point kernels in loop-back, communicator-id reuse,
the footprint of the emulated compute (co-residency with a collective, no deadlock when it does
not fit), and the captured fsdp schedule against the call-by-call one.
Bar: bit-exact against the CPU oracle (rank-order fp32 sum, one RNE rounding).
"""
import os
import subprocess
import time

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

NPDT = {0: np.float32, 1: np.float16, 2: np.uint16}
COUNTS = [1, 7, 33, 4097, 65536 + 17, 1_000_003]


def bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a.view(np.uint16)


def make_dev(dl, arrays, dtype):
    ts = []
    for a in arrays:
        t = dl.Tensor(a.size, dtype)
        t.upload(a)
        ts.append(t)
    return ts


def free_all(*lists):
    for ts in lists:
        for t in ts:
            t.free()


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("n", [5, 6, 7])
def test_emu_group_sizes_5_6_7_every_body(dl, oracle, n, dtype):
    """Group sizes the round-1 suites never executed: allreduce (both algorithms), allgather,
    reduce-scatter, all-to-all — integer and random patterns, odd counts (shifted mover)."""
    for count in COUNTS:
        for kind in (1, 2):
            xs = [oracle.fill_pattern(count, dtype, kind, 77, r) for r in range(n)]
            send = make_dev(dl, xs, dtype)
            recv = [dl.Tensor(count, dtype) for _ in range(n)]
            want = oracle.allreduce(xs, dtype)
            for algo in (1, 2):
                for t in recv:
                    t.fill(0)
                dl.emu_allreduce([t.data for t in send], [t.data for t in recv], count, dtype, algo=algo, ctas=5)
                for r in range(n):
                    assert np.array_equal(bits(recv[r].download(NPDT[dtype])), bits(want)), ("allreduce", n, dtype, algo, kind, count, r)
            free_all(recv)
            if kind == 2:
                gathered = [dl.Tensor(count * n, dtype) for _ in range(n)]
                dl.emu_allgather([t.data for t in send], [t.data for t in gathered], count, dtype, ctas=5)
                wantg = oracle.allgather(xs, dtype)
                for r in range(n):
                    assert np.array_equal(bits(gathered[r].download(NPDT[dtype])), bits(wantg)), ("allgather", n, dtype, count, r)
                free_all(gathered)
            free_all(send)
            # reduce-scatter and all-to-all on n * count elements per rank
            xs = [oracle.fill_pattern(count * n, dtype, kind, 78, r) for r in range(n)]
            send = make_dev(dl, xs, dtype)
            recv = [dl.Tensor(count, dtype) for _ in range(n)]
            dl.emu_reduce_scatter([t.data for t in send], [t.data for t in recv], count, dtype, ctas=5)
            for r in range(n):
                want = oracle.reduce_scatter_block(xs, r, count, dtype)
                assert np.array_equal(bits(recv[r].download(NPDT[dtype])), bits(want)), ("reduce_scatter", n, dtype, kind, count, r)
            free_all(recv)
            if kind == 2 and count <= 65536 + 17:
                out = [dl.Tensor(count * n, dtype) for _ in range(n)]
                dl.emu_alltoall([t.data for t in send], [t.data for t in out], count, dtype, ctas=5)
                for r in range(n):
                    want = oracle.alltoall(xs, r, count, dtype)
                    assert np.array_equal(bits(out[r].download(NPDT[dtype])), bits(want)), ("alltoall", n, dtype, count, r)
                free_all(out)
            free_all(send)


@pytest.mark.parametrize("dtype", [0, 1, 2])
@pytest.mark.parametrize("n", [2, 3, 4, 8])
def test_shifted_mover_many_ctas(dl, oracle, n, dtype):
    """Misaligned chunks (odd recvcount puts rank r's chunk at r*count*sizeof(T)) through the 128-bit
    shifted path with enough CTAs and vectors that warps take several overlapping runs each."""
    cap = 148 // n
    for count, ctas in ((3_000_001, min(16, cap)), (999_999, min(7, cap)), (40_001, 3), (515, 2)):
        xs = [oracle.fill_pattern(count * n, dtype, 2, 5, r) for r in range(n)]
        send = make_dev(dl, xs, dtype)
        recv = [dl.Tensor(count, dtype) for _ in range(n)]
        dl.emu_reduce_scatter([t.data for t in send], [t.data for t in recv], count, dtype, ctas=ctas)
        for r in range(n):
            want = oracle.reduce_scatter_block(xs, r, count, dtype)
            assert np.array_equal(bits(recv[r].download(NPDT[dtype])), bits(want)), (n, dtype, count, r)
        # allgather of the same odd count: misaligned destinations
        small = [oracle.fill_pattern(count, dtype, 2, 6, r) for r in range(n)]
        s2 = make_dev(dl, small, dtype)
        g = [dl.Tensor(count * n, dtype) for _ in range(n)]
        dl.emu_allgather([t.data for t in s2], [t.data for t in g], count, dtype, ctas=ctas)
        wantg = oracle.allgather(small, dtype)
        for r in (0, n - 1):
            assert np.array_equal(bits(g[r].download(NPDT[dtype])), bits(wantg)), (n, dtype, count, r)
        free_all(send, recv, s2, g)


@pytest.mark.parametrize("n", [2, 4, 8])
def test_two_shot_allreduce_in_place(dl, oracle, n):
    """send == recv (MPI_IN_PLACE / ncclAllReduce in place): the two-shot kernel reads and writes every
    element through its one owner, so the result is the same sum (the runtime sends overlapping
    buffers to it; the one-shot kernel is not in-place safe)."""
    dtype = 1
    for count in (4097, 1_000_003):
        xs = [oracle.fill_pattern(count, dtype, 2, 9, r) for r in range(n)]
        buf = make_dev(dl, xs, dtype)
        dl.emu_allreduce([t.data for t in buf], [t.data for t in buf], count, dtype, algo=dl.ALGO_TWOSHOT, ctas=6)
        want = oracle.allreduce(xs, dtype)
        for r in range(n):
            assert np.array_equal(bits(buf[r].download(np.float16)), bits(want)), (n, count, r)
        free_all(buf)
    # world 1 through the communicator: in place is a copy onto itself
    comm = dl.B200Communicator([0], 1, dtype)
    x = oracle.fill_pattern(100_001, dtype, 2, 3, 0)
    t, = make_dev(dl, [x], dtype)
    comm.Allreduce(t.data, t.data, 100_001)
    assert np.array_equal(bits(t.download(np.float16)), bits(x))
    comm.finalize()
    t.free()


@pytest.mark.parametrize("dtype", [0, 1])
def test_send_recv_loopback(dl, oracle, dtype):
    """The Irecv / Isend / Wait kernels (rendezvous post, put with the shifted mover, last-CTA "done")
    on one GPU against one loop-back mailbox — counts and misalignments as the multi-GPU ring test."""
    for count in (1, 7, 4097, 65536 + 17, 1_000_003, 8_388_608 + 3):
        x = oracle.fill_pattern(count + 8, dtype, 2, 77, 0)
        src, = make_dev(dl, [x], dtype)
        dst = dl.Tensor(count + 8, dtype)
        for shift in (0, 1, 3):
            dst.fill(0)
            dl.emu_sendrecv(src.offset(3 - shift), dst.offset(shift), count, dtype, ctas=0 if count > 4097 else 2)
            got = dst.download(NPDT[dtype])
            assert np.array_equal(bits(got[shift:shift + count]), bits(x[3 - shift:3 - shift + count])), (dtype, count, shift)
            assert not got[:shift].view(np.uint8).any() and not got[shift + count:].view(np.uint8).any(), "wrote outside the message"
        free_all([src, dst])


def test_communicator_ids_are_recycled(dl, oracle):
    """More than 64 communicators over a process's life: a destroyed communicator's flag rows are
    handed out again (reset), and collectives on the recycled one work."""
    x = oracle.fill_pattern(10_001, 0, 2, 1, 0)
    t, = make_dev(dl, [x], 0)
    u = dl.Tensor(10_001, 0)
    ids = set()
    for i in range(80):
        comm = dl.B200Communicator([0], num_slots=2, dtype=0)
        ids.add(comm.get("id"))
        u.fill(0)
        comm.Iallreduce(t.data, u.data, 10_001, i % 2)
        comm.WaitAll(2)
        assert np.array_equal(bits(u.download(np.float32)), bits(x))
        comm.finalize()
    assert len(ids) <= 64
    free_all([t, u])


def test_emulated_compute_footprints(dl):
    """Every footprint of the emulated compute lasts what was asked (device-timed chain), and the default
    one (one 256-thread FMA-busy CTA per SM) leaves room for a collective CTA on the same SM."""
    L = dl.lib()
    sms = dl.get_option("sms")
    assert dl.get_option("spin_ctas") == sms and dl.get_option("spin_threads") == 256 and dl.get_option("spin_mode") == dl.SPIN_FMA
    assert dl.get_option("spin_coresident") == 1 and 0 < dl.get_option("spin_regs") <= 64
    try:
        for mode in (dl.SPIN_SLEEP, dl.SPIN_FMA, dl.SPIN_HBM):
            for ctas, threads in ((1, 32), (0, 256), (2 * sms, 128)):
                dl.set_option("spin_mode", mode)
                dl.set_option("spin_ctas", ctas)
                dl.set_option("spin_threads", threads)
                dl.iter_begin()
                dl.compute_spin(50.0)  # start of the chain
                dl.iter_end()
                dl.iter_begin()
                for _ in range(20):
                    dl.compute_spin(500.0)
                got = dl.iter_end() * 1000 / 20
                assert abs(got - 500.0) <= 5.0, (mode, ctas, threads, got)
    finally:
        dl.set_option("spin_mode", dl.SPIN_FMA)
        dl.set_option("spin_ctas", 0)
        dl.set_option("spin_threads", 256)


def test_collective_shares_the_sms_with_the_emulated_compute(dl, oracle):
    """SM contention is real, and it cannot deadlock. With the default footprint a 128-CTA collective
    (8 virtual ranks x 16 CTAs, all of which must be resident at once: they wait on one another) runs WHILE
    a long spin holds a CTA on every SM — it finishes long before the spin does. With a footprint that leaves no
    room (1024 threads per SM) the collective has to wait for SMs, and still completes: spin CTAs never wait on
    anything, so they always make way."""
    n, dtype, count = 8, 1, 4_000_001
    xs = [oracle.fill_pattern(count, dtype, 2, 4, r) for r in range(n)]
    send = make_dev(dl, xs, dtype)
    recv = [dl.Tensor(count, dtype) for _ in range(n)]
    want = oracle.allreduce(xs, dtype)
    dl.emu_allreduce([t.data for t in send], [t.data for t in recv], count, dtype, algo=2, ctas=16)  # warm-up
    try:
        for mode in (dl.SPIN_FMA, dl.SPIN_HBM):
            dl.set_option("spin_mode", mode)
            dl.compute_sync()
            for t in recv:
                t.fill(0)
            t0 = time.perf_counter()
            dl.compute_spin(200_000.0)  # 200 ms of "compute" on every SM
            dl.emu_allreduce([t.data for t in send], [t.data for t in recv], count, dtype, algo=2, ctas=16)
            dt_ms = (time.perf_counter() - t0) * 1e3
            assert dt_ms < 100.0, f"collective did not overlap the spin (mode {mode}): {dt_ms:.1f} ms"
            for r in (0, n - 1):
                assert np.array_equal(bits(recv[r].download(np.float16)), bits(want))
            dl.compute_sync()
        # no room: 1024 spin threads per SM
        dl.set_option("spin_mode", dl.SPIN_FMA)
        dl.set_option("spin_threads", 1024)
        assert dl.get_option("spin_coresident") == 0
        for t in recv:
            t.fill(0)
        t0 = time.perf_counter()
        dl.compute_spin(50_000.0)
        dl.compute_spin(50_000.0)
        dl.emu_allreduce([t.data for t in send], [t.data for t in recv], count, dtype, algo=2, ctas=16)
        dt_ms = (time.perf_counter() - t0) * 1e3
        assert dt_ms < 5_000.0, dt_ms
        for r in (0, n - 1):
            assert np.array_equal(bits(recv[r].download(np.float16)), bits(want))
        dl.compute_sync()
    finally:
        dl.set_option("spin_mode", dl.SPIN_FMA)
        dl.set_option("spin_threads", 256)
        dl.set_option("spin_ctas", 0)
    free_all(send, recv)


def test_strict_barriers_give_the_same_bits(dl, oracle):
    n, dtype, count = 4, 0, 300_001
    xs = [oracle.fill_pattern(count, dtype, 2, 8, r) for r in range(n)]
    send = make_dev(dl, xs, dtype)
    recv = [dl.Tensor(count, dtype) for _ in range(n)]
    want = oracle.allreduce(xs, dtype)
    try:
        dl.set_option("strict_barriers", 1)
        for algo in (1, 2):
            dl.emu_allreduce([t.data for t in send], [t.data for t in recv], count, dtype, algo=algo, ctas=8)
            assert np.array_equal(bits(recv[1].download(np.float32)), bits(want))
    finally:
        dl.set_option("strict_barriers", 0)
    free_all(send, recv)


def test_graph_wait_orders_compute_behind_the_collective(dl):
    """Inside a capture Wait / the blocking collectives cannot block the host; they make the compute stream
    wait for the collective instead. A captured fsdp-like schedule — spin, blocking reduce-scatter, spin, ... —
    must therefore expose every reduce-scatter, exactly as the call-by-call schedule does."""
    dtype, count, units = dl.F16, 100_000_000, 6  # world-1 reduce-scatter = 200 MB read + 200 MB written: ~70 us each
    comm = dl.B200Communicator([0], num_slots=units, dtype=dtype)
    big, shard = dl.Tensor(count, dtype), dl.Tensor(count, dtype)

    def schedule():
        for u in range(units):
            dl.compute_spin(100.0)
            comm.Reduce_Scatter_block(big.data, shard.data, count)

    schedule()
    dl.iter_begin()
    schedule()
    plain_ms = dl.iter_end()
    rs_ms = comm.slot_ms(-1)
    with dl.Graph() as g:
        schedule()
    g.launch()
    dl.compute_sync()
    times = []
    for _ in range(3):
        dl.iter_begin()
        g.launch()
        times.append(dl.iter_end())
    graph_ms = min(times)
    g.destroy()
    comm.finalize()
    free_all([big, shard])
    compute_ms = units * 0.1
    assert rs_ms > 0.04, rs_ms
    assert graph_ms >= compute_ms + 0.8 * units * rs_ms, (graph_ms, compute_ms, rs_ms)  # nothing hidden
    assert graph_ms <= plain_ms * 1.05 + 0.05, (graph_ms, plain_ms)


def test_fsdp_binary_graph_and_plain_agree(dl):
    """`fsdp -g` must report the same schedule as plain fsdp when the reduce-scatter is not small
    against the compute (synthetic comm-heavy table, one GPU, sharding factor 1)."""
    from dlnetbench_b200.parser import parse_sections

    exe = os.path.join(ROOT, "bin", "b200", "cpp", "data_parallel", "fsdp")
    assert os.path.exists(exe), "bin/b200/cpp/data_parallel/fsdp missing: run __graft_entry__.build()"
    out = {}
    for flag in ([], ["-g"]):
        res = subprocess.run([exe, "commheavy_1_bfloat16", "8", "1", "tests/data", "-t", "fp16", "-w", "2", "-r", "5", *flag],
                             capture_output=True, text=True, timeout=600, cwd=ROOT)
        assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
        out[bool(flag)] = parse_sections(res.stdout)["fsdp"]["json_data"]
    plain, graph = out[False]["device_iter_ms_median"], out[True]["device_iter_ms_median"]
    lb_ms = out[False]["lower_bound_us"] / 1e3
    rs_ms = out[False]["reduce_scatter_last_ms"]
    assert rs_ms > 0.03, rs_ms
    assert graph >= lb_ms + 0.8 * 8 * rs_ms, (graph, lb_ms, rs_ms)
    assert abs(graph - plain) <= 0.12 * plain, (graph, plain)


@pytest.mark.parametrize("n", [2, 5, 8])
def test_kernels_write_only_inside_their_buffers(dl, oracle, n):
    """compute-sanitizer is closed on this pool, so the bounds check is ours: every receive buffer is a window inside one
    big allocation whose every other byte holds a canary pattern; after each collective (odd counts, odd element offsets:
    head / tail peeling, aligned and shifted 128-bit movers, several CTAs) the canaries around the windows are intact and
    the windows hold the oracle's result."""
    dtype, guard = 1, 4099
    for count, off in ((1, 1), (33, 3), (4097, 5), (300_001, 7)):
        span = count * n + 2 * guard + 16
        arena = [dl.Tensor(span, dtype) for _ in range(n)]      # receive side: canary everywhere, window in the middle
        for r, t in enumerate(arena):
            t.fill(2, 999, r)
        lo = guard + off                                          # window [lo, lo + len) in elements
        xs = [oracle.fill_pattern(count * n, dtype, 2, 31, r) for r in range(n)]
        send = make_dev(dl, xs, dtype)

        def canaries_intact(length):
            for r, t in enumerate(arena):
                assert dl.verify_sum_pattern(t.data, lo, 0, dtype, 2, 999, [r]) == 0, ("below the window", n, count, r)
                hi = lo + length
                assert dl.verify_sum_pattern(t.offset(hi), span - hi, hi, dtype, 2, 999, [r]) == 0, ("above the window", n, count, r)

        # allreduce (both algorithms) into a window of `count`
        want = oracle.allreduce([x[:count] for x in xs], dtype)
        for algo in (1, 2):
            dl.emu_allreduce([t.data for t in send], [t.offset(lo) for t in arena], count, dtype, algo=algo, ctas=3)
            canaries_intact(count)
            got = arena[n - 1].download(np.float16)[lo:lo + count]
            assert np.array_equal(bits(got), bits(want)), (n, count, algo)
        # reduce-scatter into a window of `count`
        dl.emu_reduce_scatter([t.data for t in send], [t.offset(lo) for t in arena], count, dtype, ctas=3)
        canaries_intact(count)
        # allgather / all-to-all into a window of n * count
        dl.emu_allgather([t.data for t in send], [t.offset(lo) for t in arena], count, dtype, ctas=3)
        canaries_intact(count * n)
        dl.emu_alltoall([t.data for t in send], [t.offset(lo) for t in arena], count, dtype, ctas=3)
        canaries_intact(count * n)
        got = arena[0].download(np.float16)[lo:lo + count * n]
        assert np.array_equal(bits(got), bits(oracle.alltoall(xs, 0, count, dtype)))
        free_all(arena, send)
