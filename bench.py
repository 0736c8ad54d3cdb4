#!/usr/bin/env python
"""bench.py — the reference's headline benchmark on B200: dp proxy iteration time.

    python bench.py --gpus N --steps K --warmup W            # this repo (libdlnb.so)
    python bench.py --impl reference --gpus N ...            # the reference's mpi_cpu path on host cores
    torchrun --nproc-per-node N bench.py --gpus N ...        # N > 1: one rank per GPU

Workload (BASELINE.json configs[1]): dp proxy, llama3_8b_64_bfloat16 stats, fp16 gradient
buckets (default 240 buckets of 66.9 MB = 16.06 GB of gradients per GPU per step). One step =
one proxy iteration: emulated forward, then per bucket an emulated backward slice followed
by that bucket's Allreduce, then WaitAll (reference: cpp/data_parallel/dp.cpp:87-106).
Metric: iteration time in ms, device-timed (CUDA events), max over ranks; lower is better.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "dp_proxy_iter_time_ms"


def parse_stats(path):
    """model_stats/<name>.txt, positional (reference: cpp/utils.hpp:200-269)."""
    vals = []
    with open(path) as f:
        for line in f:
            vals.append(line.split(":", 1)[1].strip())
    return {"modelSize": int(vals[2]), "fwd_us": int(float(vals[4])), "bwd_us": int(float(vals[5])), "batch": int(vals[6])}


def bucket_sizes(model_size, n):
    return [model_size // n + (1 if i < model_size % n else 0) for i in range(n)]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "500", "-i", str(self.device)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower() == "active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own dp.cpp (oracle/_ref, built from /root/reference
# against oracle/compat) on the box's host cores, on a bounded 1/scale sample of the workload
# ---------------------------------------------------------------------------------------------
def run_reference_dp(model, buckets, ranks, warmup, steps, scale):
    ref = os.path.join(ROOT, "oracle", "_ref")
    dp, mpirun = os.path.join(ref, "dp"), os.path.join(ref, "dlnb_mpirun")
    if not (os.path.exists(dp) and os.path.exists(mpirun)):
        return None
    src = open(os.path.join(ROOT, "model_stats", model + ".txt")).read().splitlines()
    st = parse_stats(os.path.join(ROOT, "model_stats", model + ".txt"))
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "model_stats"))
        lines = list(src)
        lines[2] = f"Model_Size:{st['modelSize'] // scale}"
        lines[4] = f"Average_Forward_Time (us):{st['fwd_us'] / scale:.2f}"
        lines[5] = f"Average_Backward_Time (us):{st['bwd_us'] / scale:.2f}"
        open(os.path.join(d, "model_stats", "sample.txt"), "w").write("\n".join(lines) + "\n")
        cmd = [mpirun, "-n", str(ranks), dp, "sample", str(buckets), ".", "-w", str(warmup), "-r", str(steps)]
        t0 = time.time()
        res = subprocess.run(cmd, cwd=d, capture_output=True, text=True)
        wall = time.time() - t0
    if res.returncode != 0:
        return None
    runtimes = []
    for line in res.stdout.splitlines():
        if line.startswith("#CCUTILS_RANK_JSON dp "):
            runtimes.append(json.loads(line.split(" ", 3)[3])["runtimes"])
    per_iter = [max(r[i] for r in runtimes) for i in range(len(runtimes[0]))]  # max over ranks
    ms = statistics.median(per_iter) * 1e3
    return {"sample_ms": ms, "value_ms": ms * scale, "wall_s": wall, "scale": scale, "ranks": ranks, "extrapolated": scale > 1,
            "sample": (f"reference dp.cpp (mpi_cpu, fp32 MPI_SUM over oracle/compat's shared-memory MPI), {buckets} buckets, {ranks} rank(s), "
                       + ("the WHOLE workload (scale 1: measured, not extrapolated)" if scale == 1 else
                          f"1/{scale} of the workload (Model_Size and fwd/bwd us divided by {scale}: what fits half of the host RAM); "
                          f"value = median iteration x {scale} (EXTRAPOLATED)"))}


def run_reference_c1():
    """BASELINE config C1 exactly as specified: `dp gpt2_l_16_bfloat16 10 . -r 5`, fp32, mpi_cpu, 2 ranks on host cores."""
    ref = os.path.join(ROOT, "oracle", "_ref")
    dp, mpirun = os.path.join(ref, "dp"), os.path.join(ref, "dlnb_mpirun")
    if not (os.path.exists(dp) and os.path.exists(mpirun)):
        return None
    res = subprocess.run([mpirun, "-n", "2", dp, "gpt2_l_16_bfloat16", "10", ".", "-r", "5"], cwd=ROOT, capture_output=True, text=True)
    if res.returncode != 0:
        return None
    runtimes = [json.loads(l.split(" ", 3)[3])["runtimes"] for l in res.stdout.splitlines() if l.startswith("#CCUTILS_RANK_JSON dp ")]
    if not runtimes:
        return None
    per_iter = [max(r[i] for r in runtimes) for i in range(len(runtimes[0]))]
    st = parse_stats(os.path.join(ROOT, "model_stats", "gpt2_l_16_bfloat16.txt"))
    lb = (st["fwd_us"] + 10 * (st["bwd_us"] // 10)) / 1e3
    return {"command": "dp gpt2_l_16_bfloat16 10 . -r 5 (2 ranks, fp32, mpi_cpu)", "iter_ms_median": statistics.median(per_iter) * 1e3,
            "lower_bound_ms": lb, "ratio_to_lower_bound": statistics.median(per_iter) * 1e3 / lb, "ranks": 2}


class Plumbing:
    """Cross-rank plumbing of the bench (torch.distributed when there is more than one rank):
    barrier + device synchronize on both sides of the timed region, and max over ranks."""

    def __init__(self, dist, device):
        self.dist, self.device = dist, device

    def _sync(self):
        if self.device == "cuda":
            import torch

            torch.cuda.synchronize()

    def barrier(self):
        self._sync()
        if self.dist is not None:
            self.dist.barrier()
        self._sync()

    def max_over_ranks(self, x):
        if self.dist is None:
            return x
        import torch

        t = torch.tensor([x], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def host_resources():
    """Cores and available RAM of the box this runs on (the CPU baseline is sized from them)."""
    avail = None
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                avail = int(line.split()[1]) * 1024
    except OSError:
        pass
    return {"cores": os.cpu_count() or 1, "mem_available_bytes": avail}


def pick_scale(model_size, ranks, res, lower_bound_ms=None, iterations=None, time_budget_s=600.0):
    """Largest sample (smallest divisor) of the workload (a) whose fp32 buffers (2 x 4 bytes x Model_Size per rank, dp.cpp:227-232)
    fit half of the box's available RAM and (b) whose run is expected to end within the time budget: `iterations` iterations of
    lower bound x a communication factor seen on this pool's hosts (1.1 / 1.6 / 2.2 / 3.0 at 1 / 2 / 4 / 8 ranks).
    scale 1 = the whole workload, nothing extrapolated; otherwise the caller reports "extrapolated": true and the scale."""
    budget = 0.5 * res["mem_available_bytes"] if res["mem_available_bytes"] else 16e9
    scale = 1
    while ranks * 2 * 4 * model_size / scale > budget:
        scale *= 2
    if lower_bound_ms and iterations:
        factor = {1: 1.1, 2: 1.6, 4: 2.2}.get(ranks, 3.0)
        while iterations * lower_bound_ms * 1e-3 * factor / scale > time_budget_s:
            scale *= 2
    return scale


def run_proxy(rank, world, local_rank, prog, args, env_extra=None, timeout=900):
    """One rank of a proxy binary (bin/b200/cpp/data_parallel/<prog>) as a child process of this bench rank: the N
    bench ranks start the N ranks of the binary, which find each other through libdlnb's own rendezvous.
    Returns (global JSON of the section, per-rank JSONs seen on this rank's stdout) — rank 0 prints the globals."""
    exe = os.path.join(ROOT, "bin", "b200", "cpp", "data_parallel", prog)
    if not os.path.exists(exe):
        return {"unavailable": f"{exe} not built"}
    env = dict(os.environ)
    env.update({"DLNB_RANK": str(rank), "DLNB_WORLD": str(world), "DLNB_LOCAL_RANK": str(local_rank),
                "DLNB_SESSION": f"bench-{prog}-{os.environ.get('MASTER_PORT', '0')}-{'-'.join(args[:3])}"})
    env.update(env_extra or {})
    try:
        res = subprocess.run([exe, *args], cwd=ROOT, capture_output=True, text=True, timeout=timeout, env=env)
    except subprocess.TimeoutExpired:
        return {"failed": f"{prog} {' '.join(args)} timed out after {timeout} s"}
    if res.returncode != 0:
        return {"failed": f"{prog} {' '.join(args)} rc={res.returncode}: {(res.stderr or res.stdout)[-300:]}"}
    g = None
    for line in res.stdout.splitlines():
        if line.startswith("#CCUTILS_GLOBAL_JSON "):
            g = json.loads(line.split(" ", 2)[2])
    return g if g is not None else {}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama3_8b_64_bfloat16")
    ap.add_argument("--buckets", type=int, default=240)
    ap.add_argument("--dtype", default="fp16", choices=["fp32", "fp16", "bf16"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--spin-ctas", type=int, default=0, help="CTAs of the emulated compute, 0 = one per SM")
    ap.add_argument("--spin-threads", type=int, default=256)
    ap.add_argument("--spin-mode", default="fma", choices=["sleep", "fma", "hbm"])
    ap.add_argument("--no-sub", action="store_true", help="skip the fsdp / dp_loop / fsdp_loop sub-results")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity matrix prologue (never do this for a reported run)")
    ap.add_argument("--no-contention", action="store_true")
    ap.add_argument("--ctas", type=int, default=0)
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    st = parse_stats(os.path.join(ROOT, "model_stats", a.model + ".txt"))
    sizes = bucket_sizes(st["modelSize"], a.buckets)
    bwd_per_bucket = st["bwd_us"] // a.buckets
    lower_bound_ms = (st["fwd_us"] + a.buckets * bwd_per_bucket) / 1e3
    esz = {"fp32": 4, "fp16": 2, "bf16": 2}[a.dtype]
    config = {"workload": f"dp proxy, {a.model} stats, {a.dtype} gradient buckets", "num_buckets": a.buckets,
              "bucket_bytes": sizes[0] * esz, "grad_bytes_per_gpu": st["modelSize"] * esz, "world_size": world,
              "fwd_us": st["fwd_us"], "bwd_us_per_bucket": bwd_per_bucket, "lower_bound_ms": lower_bound_ms,
              "l2": "inputs larger than L2: every step reads %.2f GB of distinct buckets" % (st["modelSize"] * esz / 1e9)}

    # ------------------------------------------------------------------ reference arm
    if a.impl == "reference":
        if rank != 0:
            return
        res = host_resources()
        scale = pick_scale(st["modelSize"], a.gpus, res, lower_bound_ms, a.warmup + a.steps)
        r = run_reference_dp(a.model, a.buckets, a.gpus, a.warmup, a.steps, scale)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/dp missing or failed (build with make -C oracle ref where /root/reference is mounted)"}))
            return
        config.update({"extrapolated": r["extrapolated"], "scale": scale})
        line = {"metric": METRIC, "value": r["value_ms"], "unit": "ms", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": r["value_ms"], "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic (zero-filled buffers, as the reference allocates them)", "config": config,
                "impl": "reference",
                "cpu_baseline": {"value": r["value_ms"], "unit": "ms", "cores": r["ranks"], "kind": "reference", "sample": r["sample"],
                                 "sample_ms": r["sample_ms"], "extrapolated": r["extrapolated"], "scale": scale, "wall_s": r["wall_s"],
                                 "host": res},
                "c1_exact": run_reference_c1(),
                "e2e": {"value": r["value_ms"], "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return


    # ------------------------------------------------------------------ this repo
    import torch  # plumbing only: barrier / max over ranks / synchronize

    from dlnetbench_b200 import capi, nvlink, selfcheck

    dist = None
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    capi.init(rank, world, device=local_rank, session="bench-" + os.environ.get("MASTER_PORT", str(os.getpid())))
    L = capi.lib()
    dt = {"fp32": capi.F32, "fp16": capi.F16, "bf16": capi.BF16}[a.dtype]
    modes = {"sleep": capi.SPIN_SLEEP, "fma": capi.SPIN_FMA, "hbm": capi.SPIN_HBM}

    def set_footprint(ctas, threads, mode):
        capi.set_option("spin_ctas", ctas)
        capi.set_option("spin_threads", threads)
        capi.set_option("spin_mode", modes[mode])

    set_footprint(a.spin_ctas, a.spin_threads, a.spin_mode)
    plumb = Plumbing(dist, "cuda")
    barrier, max_over_ranks = plumb.barrier, plumb.max_over_ranks

    def over_ranks(x):
        """min / median / max over ranks of a per-rank number."""
        if dist is None:
            return {"min": x, "median": x, "max": x}
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = x
        dist.all_reduce(t)
        v = sorted(t.tolist())
        return {"min": v[0], "median": statistics.median(v), "max": v[-1]}

    # ---- parity on the live ranks, untimed: every collective and algorithm of the library across the real GPUs
    parity = None
    if not a.no_parity:
        t0 = time.perf_counter()
        parity = selfcheck.run(rank, world, quick=True)
        fails = max_over_ranks(float(parity["failures"]))
        parity["failures_max_over_ranks"] = int(fails)
        parity["seconds"] = round(time.perf_counter() - t0, 1)
        if fails:
            raise SystemExit(f"[rank {rank}] parity matrix failed: {parity['failed']}")

    comm = capi.B200Communicator(list(range(world)), num_slots=a.buckets, dtype=dt)
    if a.ctas:
        comm.set("ctas", a.ctas)
    grad = [capi.Tensor(s, dt) for s in sizes]
    summed = [capi.Tensor(s, dt) for s in sizes]
    for g in grad:
        g.fill(capi.PAT_INT, 0, rank)  # integer-valued gradients: exact sums, checked below
    spin_overhead_us, spin_slope = capi.spin_calibrate()

    def step(e2e=False, host_in=None, host_out=None):
        """run_data_parallel (dp.cpp:87-106) over the C ABI."""
        capi.compute_spin(st["fwd_us"])
        for i in range(a.buckets):
            capi.compute_spin(bwd_per_bucket)
            if e2e:  # this bucket's gradients arrive from pinned host memory when its backward slice has ended
                capi.check(L.dlnb_memcpy_h2d_after_compute(grad[i].data, host_in[i % len(host_in)], sizes[i] * esz))
            comm.Iallreduce(grad[i].data, summed[i].data, sizes[i], i)
            if e2e:  # and the reduced bucket goes back to the host as soon as it is complete
                capi.check(L.dlnb_memcpy_d2h_after_slot(comm.h, i, host_out[i % len(host_out)], summed[i].data, sizes[i] * esz))
        comm.WaitAll(a.buckets)

    def verify_all_buckets(what):
        bad = sum(capi.verify_sum_pattern(summed[i].data, sizes[i], 0, dt, capi.PAT_INT, 0, list(range(world))) for i in range(a.buckets))
        if bad:
            raise SystemExit(f"[rank {rank}] {what}: allreduce result wrong in {bad} elements")
        return a.buckets

    for _ in range(a.warmup):
        step()
    buckets_verified = verify_all_buckets("after warm-up")  # every summed bucket, closed form on the device
    for t in summed:
        t.fill(capi.PAT_ZERO)  # the timed steps must produce the sums again

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    launches0 = L.dlnb_launch_count()
    capi.iter_begin()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    total_ms = capi.iter_end()
    host_ms = (time.perf_counter() - t0) * 1e3
    launches = L.dlnb_launch_count() - launches0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = max_over_ranks(total_ms / a.steps)
    # per-bucket device time of the allreduce inside the LAST timed step (slot events), read after the timed region
    in_step = [comm.slot_ms(i) for i in range(a.buckets)]
    in_step_stats = over_ranks(statistics.median(in_step))
    in_step_mean = over_ranks(sum(in_step) / len(in_step))
    k8 = min(8, a.buckets)
    in_step_first, in_step_last = over_ranks(sum(in_step[:k8]) / k8), over_ranks(sum(in_step[-k8:]) / k8)
    buckets_verified += verify_all_buckets("after the timed steps")

    # ---- the dominant kernel alone: every bucket's allreduce back to back (16 GB of distinct inputs, nothing
    # cached), ONE CUDA-event pair around the batch, started on all ranks together by a device-side barrier.
    # NVLink byte counters (NVML) are read on both sides: measured bytes per launch per direction.
    def isolated_batch():
        capi.set_option("slot_timing", 0)  # no per-launch timing events between the kernels of the batch
        capi.batch_begin()
        for i in range(a.buckets):
            comm.Iallreduce(grad[i].data, summed[i].data, sizes[i], i)
        comm.WaitAll(a.buckets)
        ms = capi.iter_end() / a.buckets
        capi.set_option("slot_timing", 1)
        return ms

    barrier()
    isolated_batch()  # warm
    links = nvlink.LinkCounters(local_rank) if world > 1 else None
    barrier()
    before = links.settle() if links and links.ok else None
    barrier()
    kern_avg_ms = max_over_ranks(isolated_batch())
    barrier()
    traffic = None
    if before is not None:
        after = links.settle()
        traffic = nvlink.per_launch(before, after, a.buckets)
    elif links is not None:
        traffic = {"unavailable": links.why}

    # ---- the same batch while emulated compute of each footprint runs beside it (SM contention, measured)
    contention = None
    if not a.no_contention:
        contention = {}
        sms = capi.get_option("sms")
        for name, (c, th, m) in {"sleep_1warp": (1, 32, "sleep"), "sleep_1cta_per_sm": (0, 256, "sleep"), "fma_1cta_per_sm": (0, 256, "fma"),
                                  "hbm_1cta_per_sm": (0, 256, "hbm"), "fma_1024thr_per_sm_no_room": (0, 1024, "fma")}.items():
            set_footprint(c, th, m)
            barrier()
            spin_us = a.buckets * kern_avg_ms * 1000 * 6 + 20000
            t_spin = time.perf_counter()
            capi.compute_spin(spin_us)  # compute that outlasts the batch, NOT waited for
            capi.set_option("slot_timing", 0)
            capi.batch_begin_concurrent(a.buckets * 25.0)
            for i in range(a.buckets):
                comm.Iallreduce(grad[i].data, summed[i].data, sizes[i], i)
            comm.WaitAll(a.buckets)
            ms_local = capi.batch_end() / a.buckets
            t_batch = (time.perf_counter() - t_spin) * 1e3  # (before any torch collective: NCCL's own kernel would queue for SMs too)
            capi.set_option("slot_timing", 1)
            capi.compute_sync()
            ms = max_over_ranks(ms_local)
            contention[name] = {"avg_launch_ms": ms, "slowdown": ms / kern_avg_ms, "coresident": capi.get_option("spin_coresident"),
                                "spin_ctas": capi.get_option("spin_ctas"), "spin_threads": th, "sms": sms,
                                "spin_ms": spin_us / 1e3, "batch_done_after_ms": max_over_ranks(t_batch),
                                "ran_beside_the_compute": bool(max_over_ranks(t_batch) < 0.8 * spin_us / 1e3)}
        set_footprint(a.spin_ctas, a.spin_threads, a.spin_mode)

    # ---- end to end: host buffers in, host result out, copies inside the timed region
    e2e = None
    if not a.no_e2e:
        def pinned(n, nbytes):
            out = []
            for _ in range(n):
                p = capi.C.c_void_p()
                capi.check(L.dlnb_host_alloc(nbytes, capi.C.byref(p)))
                out.append(p.value)
            return out

        host_in, host_out = pinned(2, max(sizes) * esz), pinned(2, max(sizes) * esz)
        for h in host_in:  # the rank's gradient pattern, staged once (device -> host, untimed)
            capi.check(L.dlnb_memcpy_d2h(h, grad[0].data, sizes[0] * esz))
        step(e2e=True, host_in=host_in, host_out=host_out)  # warm-up
        barrier()
        capi.iter_begin()
        for _ in range(a.steps):
            step(e2e=True, host_in=host_in, host_out=host_out)
        e2e_ms = capi.iter_end() / a.steps
        barrier()
        # the last reduced bucket as the host received it: exact sums of the integer pattern
        import numpy as np

        npdt = {"fp32": np.float32, "fp16": np.float16, "bf16": np.uint16}[a.dtype]
        last = (a.buckets - 1) % 2
        got = np.ctypeslib.as_array((capi.C.c_uint8 * (sizes[-1] * esz)).from_address(host_out[last])).view(npdt)[:4096]
        dev = summed[-1].download(npdt, 4096)
        if not np.array_equal(got.view(np.uint8), dev.view(np.uint8)):
            raise SystemExit(f"[rank {rank}] end-to-end: the host copy of the last bucket differs from the device result")
        e2e = {"value": max_over_ranks(e2e_ms), "unit": "ms", "h2d_bytes_per_step": st["modelSize"] * esz,
               "d2h_bytes_per_step": st["modelSize"] * esz,
               "how": "per bucket: cudaMemcpyAsync from pinned host memory once the bucket's backward slice has ended -> Iallreduce -> the whole "
                      "reduced bucket back to pinned host memory; the iteration ends when the last copy has landed"}
        for h in host_in + host_out:
            L.dlnb_host_free(h)

    # ---- the other half of the metric, through the drop-in binaries themselves (each bench rank starts one rank of the
    # binary): C3 fsdp llama3_8b, C4 dp_loop vit_h (1000 buckets call by call, 4000 as a replayed graph), C5 fsdp_loop
    # llama3_70b on 8 GPUs. Bounded iteration counts; the bench's own buffers are released first.
    sub = {}
    if not a.no_sub and a.model == "llama3_8b_64_bfloat16":
        for t in grad + summed:
            t.free()
        grad, summed = [], []
        barrier()

        def keep(g, keys):
            return g if ("failed" in g or "unavailable" in g) else {k: g[k] for k in keys if k in g}

        common = ["world_size", "device_iter_ms_median", "lower_bound_us", "ratio_to_lower_bound", "dtype", "backend", "spin_mode", "spin_ctas",
                  "spin_threads", "cuda_graph"]
        if world >= 2:
            g = run_proxy(rank, world, local_rank, "fsdp", ["llama3_8b_64_bfloat16", "32", str(world), ".", "-t", "fp16", "-w", "1", "-r", "2"])
            sub["fsdp"] = keep(g, common + ["num_units", "sharding_factor", "allgather_msg_size_bytes", "reducescatter_msg_size_bytes",
                                            "reduce_scatter_last_ms", "reduce_scatter_busbw_GBps", "allgather_first_ms", "allgather_busbw_GBps"])
            if "reduce_scatter_busbw_GBps" in sub["fsdp"]:
                sub["fsdp"]["reduce_scatter_frac_of_900"] = sub["fsdp"]["reduce_scatter_busbw_GBps"] / 900.0
            barrier()
        dpl = {}
        for nb, extra in ((1000, []), (4000, ["-g"])):
            g = run_proxy(rank, world, local_rank, "dp_loop", ["vit_h_128_bfloat16", str(nb), ".", "-t", "fp32", "-w", "3", *extra],
                          env_extra={"DLNB_LOOP_ITERS": "60"})
            dpl[f"buckets_{nb}"] = keep(g, common + ["num_buckets", "msg_size_avg_bytes", "allreduce_busbw_GBps_avg", "allreduce_slowest_ms"])
            barrier()
        sub["dp_loop"] = dpl
        if world == 8:
            g = run_proxy(rank, world, local_rank, "fsdp_loop", ["llama3_70b_16_bfloat16", "80", "8", ".", "-t", "fp16", "-w", "1"],
                          env_extra={"DLNB_LOOP_ITERS": os.environ.get("DLNB_BENCH_C5_ITERS", "2")}, timeout=1200)
            sub["fsdp_loop"] = keep(g, common + ["num_units", "sharding_factor", "allgather_msg_size_bytes", "reducescatter_msg_size_bytes",
                                                 "reduce_scatter_last_ms", "reduce_scatter_busbw_GBps"])
            barrier()

    # ---- roofline of the dominant kernel (the bucket allreduce), timed live with CUDA events
    S = sizes[0] * esz
    if world == 1:
        hbm_peak, which = load_peaks()
        algo_bytes = 2 * S  # read + write of local HBM
        ach = algo_bytes / (kern_avg_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "peak_source": which,
                "kernel": "coll_kernel<T,1,OP_AR1> (world 1: vectorised HBM copy)", "algorithmic_bytes_per_launch": algo_bytes,
                "avg_launch_ms": kern_avg_ms, "traffic": None}
    else:
        algo_bytes = 2 * (world - 1) / world * S  # per GPU per direction, whatever the algorithm moves
        ach = algo_bytes / (kern_avg_ms * 1e-3) / 1e9
        in_switch = L.dlnb_get(b"multicast") == 1 and world >= 4 and S >= comm.get("multimem_min_bytes")
        kname = (f"coll_kernel<T,{world},OP_ARM> (in-switch allreduce: multimem.ld_reduce + multimem.st over the NVSwitch multicast mapping)"
                 if in_switch else f"coll_kernel<T,{world},OP_AR2> (fused two-shot allreduce over peer HBM)")
        wire = (1 + 1 / world) * S if in_switch else algo_bytes
        roof = {"bound": "nvlink", "achieved": ach, "peak": 900.0, "unit": "GB/s", "frac": ach / 900.0,
                "peak_source": "nominal NVLink 5 per direction; achieved is bus bandwidth 2(n-1)/n*S/t, which in-switch reduction may push "
                               "past the link rate because it moves fewer bytes (expected_wire_bytes_per_launch)",
                "kernel": kname, "algorithmic_bytes_per_launch": algo_bytes, "expected_wire_bytes_per_launch": wire,
                "avg_launch_ms": kern_avg_ms,
                # measured NVLink payload per launch per direction on rank 0 (NVML link counters around the isolated batch)
                "traffic": (traffic or {}).get("data_tx_bytes_per_launch"), "nvlink": traffic}
        if traffic and "data_tx_bytes_per_launch" in traffic:
            roof["nvlink_tx_GBps"] = traffic["data_tx_bytes_per_launch"] / (kern_avg_ms * 1e-3) / 1e9
            roof["nvlink_rx_GBps"] = traffic["data_rx_bytes_per_launch"] / (kern_avg_ms * 1e-3) / 1e9
    roof["avg_launch_ms_inside_step"] = {"median_per_rank": in_step_stats, "mean_per_rank": in_step_mean,
                                         "first_8_buckets": in_step_first, "last_8_buckets": in_step_last,
                                         "note": "slot events of the last timed step; min / median / max over ranks. The minimum is the kernel's own time beside "
                                                 "the emulated compute; the maximum adds the wait for the rank whose %globaltimer runs slowest (the ranks start "
                                                 "together, then drift apart by tens of ppm: the spread grows from the first to the last buckets)"}
    if world == 1:
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            roof["traffic"] = json.load(open(tp)).get(f"allreduce_w{world}_{a.dtype}_{S}")

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        res = host_resources()
        scale = pick_scale(st["modelSize"], 1, res)
        r = run_reference_dp(a.model, a.buckets, 1, 1, 2, scale)
        if r:
            cpu = {"value": r["value_ms"], "unit": "ms", "cores": 1, "kind": "reference", "sample": r["sample"], "sample_ms": r["sample_ms"],
                   "extrapolated": r["extrapolated"], "scale": scale, "host": res, "c1_exact": run_reference_c1()}

    if rank == 0:
        line = {"metric": METRIC, "value": ms_per_step, "unit": "ms", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": a.dtype.replace("fp", "f"),
                "data": "synthetic (integer-valued gradients, exact sums verified on device)", "config": config,
                "ratio_to_lower_bound": ms_per_step / lower_bound_ms,
                "allreduce_busbw_GBps": (algo_bytes / (kern_avg_ms * 1e-3) / 1e9) if world > 1 else None,
                "host_ms_per_step": host_ms / a.steps,
                "spin": {"overhead_us": spin_overhead_us, "slope": spin_slope, "ctas": capi.get_option("spin_ctas"), "threads": a.spin_threads,
                         "mode": a.spin_mode, "regs_per_thread": capi.get_option("spin_regs"), "sms": capi.get_option("sms"),
                         "coresident_with_collective_cta": capi.get_option("spin_coresident"),
                         "footprint": "one CTA of the emulated compute on every SM, FMA-busy; a collective CTA (512 threads x <= 96 registers) fits beside it"},
                "aligned_start": capi.get_option("align_iterations"),
                "arena": L.dlnb_arena_kind().decode(), "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
                "parity": dict(parity or {}, buckets_verified=buckets_verified,
                               note="fp32 goldens come from the reference's MPICommunicator; fp16/bf16 are pinned by the CPU restatement (tests/)"),
                "roofline": roof, "contention": contention, "cpu_baseline": cpu}
        line.update(sub)
        print(json.dumps(line), flush=True)
    comm.finalize()
    capi.finalize()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
