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
    return {"sample_ms": ms, "value_ms": ms * scale, "wall_s": wall, "scale": scale, "ranks": ranks,
            "sample": f"reference dp.cpp (mpi_cpu, fp32 MPI_SUM) on 1/{scale} of the workload: Model_Size and fwd/bwd us "
                      f"divided by {scale}, {buckets} buckets, {ranks} rank(s); value = median iteration x {scale}"}


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


def pick_scale(model_size, ranks, budget_bytes=16e9):
    scale = 1
    while ranks * 2 * 4 * model_size / scale > budget_bytes:
        scale *= 2
    return scale


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
    ap.add_argument("--spin-ctas", type=int, default=1)
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
        scale = pick_scale(st["modelSize"], a.gpus)
        r = run_reference_dp(a.model, a.buckets, a.gpus, a.warmup, a.steps, scale)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/dp missing or failed (build with make -C oracle ref where /root/reference is mounted)"}))
            return
        line = {"metric": METRIC, "value": r["value_ms"], "unit": "ms", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": r["value_ms"], "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic (zero-filled buffers, as the reference allocates them)", "config": config,
                "impl": "reference",
                "cpu_baseline": {"value": r["value_ms"], "unit": "ms", "cores": r["ranks"], "kind": "reference", "sample": r["sample"],
                                 "sample_ms": r["sample_ms"]},
                "e2e": {"value": r["value_ms"], "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    # ------------------------------------------------------------------ this repo
    import torch  # plumbing only: barrier / max over ranks / synchronize

    from dlnetbench_b200 import capi

    dist = None
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    capi.init(rank, world, device=local_rank, session="bench-" + os.environ.get("MASTER_PORT", str(os.getpid())))
    L = capi.lib()
    dt = {"fp32": capi.F32, "fp16": capi.F16, "bf16": capi.BF16}[a.dtype]
    capi.check(L.dlnb_set(b"spin_ctas", a.spin_ctas))
    comm = capi.B200Communicator(list(range(world)), num_slots=a.buckets, dtype=dt)
    if a.ctas:
        comm.set("ctas", a.ctas)
    grad = [capi.Tensor(s, dt) for s in sizes]
    summed = [capi.Tensor(s, dt) for s in sizes]
    for g in grad:
        g.fill(capi.PAT_INT, 0, rank)  # integer-valued gradients: exact sums, checked below
    spin_overhead_us, spin_slope = capi.spin_calibrate()

    plumb = Plumbing(dist, "cuda")
    barrier, max_over_ranks = plumb.barrier, plumb.max_over_ranks

    def step(e2e=False, host=None, probe=None):
        """run_data_parallel (dp.cpp:87-106) over the C ABI."""
        capi.compute_spin(st["fwd_us"])
        for i in range(a.buckets):
            capi.compute_spin(bwd_per_bucket)
            if e2e:  # this bucket's gradients arrive from pinned host memory
                capi.check(L.dlnb_memcpy_h2d_async(grad[i].data, host[i % len(host)], sizes[i] * esz))
            comm.Iallreduce(grad[i].data, summed[i].data, sizes[i], i)
            if e2e:  # and a slice of the reduced bucket goes back to the host
                capi.check(L.dlnb_memcpy_d2h_async(probe + i * 4096, summed[i].data, min(4096, sizes[i] * esz)))
        comm.WaitAll(a.buckets)

    for _ in range(a.warmup):
        step()
    # parity gate inside the bench: integer-valued gradients must sum exactly (closed form on device)
    bad = capi.verify_sum_pattern(summed[0].data, sizes[0], 0, dt, capi.PAT_INT, 0, list(range(world)))
    bad += capi.verify_sum_pattern(summed[-1].data, sizes[-1], 0, dt, capi.PAT_INT, 0, list(range(world)))
    if bad:
        raise SystemExit(f"[rank {rank}] allreduce result wrong in {bad} elements")

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    launches0 = L.dlnb_launch_count()
    capi.iter_begin()
    kernel_ms = []
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
        kernel_ms.extend(comm.slot_ms(i) for i in range(a.buckets))
    total_ms = capi.iter_end()
    host_ms = (time.perf_counter() - t0) * 1e3
    launches = L.dlnb_launch_count() - launches0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = max_over_ranks(total_ms / a.steps)
    kern_in_step_ms = max_over_ranks(sum(kernel_ms) / len(kernel_ms))
    # the dominant kernel alone: every bucket's allreduce back to back (16 GB of distinct inputs,
    # nothing cached), ONE CUDA-event pair around the batch -> average launch duration. The
    # per-slot events inside a step (kern_in_step_ms) also contain the stream-wait-to-launch gap.
    barrier()
    capi.set_option("slot_timing", 0)  # no per-launch timing events between the kernels of the batch
    capi.iter_begin()
    for i in range(a.buckets):
        comm.Iallreduce(grad[i].data, summed[i].data, sizes[i], i)
    comm.WaitAll(a.buckets)
    kern_avg_ms = max_over_ranks(capi.iter_end() / a.buckets)
    capi.set_option("slot_timing", 1)

    # ---- end to end: host buffers in, host result out, copies inside the timed region
    e2e = None
    if not a.no_e2e:
        host = []
        for _ in range(2):
            p = capi.C.c_void_p()
            capi.check(L.dlnb_host_alloc(max(sizes) * esz, capi.C.byref(p)))
            host.append(p.value)
        pr = capi.C.c_void_p()
        capi.check(L.dlnb_host_alloc(a.buckets * 4096, capi.C.byref(pr)))
        # fill the pinned staging buffers with rank's gradient pattern (device -> host once, untimed)
        for h in host:
            capi.check(L.dlnb_memcpy_d2h(h, grad[0].data, sizes[0] * esz))
        step(e2e=True, host=host, probe=pr.value)  # warm-up
        barrier()
        capi.iter_begin()
        for _ in range(a.steps):
            step(e2e=True, host=host, probe=pr.value)
        e2e_ms = capi.iter_end() / a.steps
        barrier()
        e2e = {"value": max_over_ranks(e2e_ms), "unit": "ms", "h2d_bytes_per_step": st["modelSize"] * esz,
               "d2h_bytes_per_step": sum(min(4096, s * esz) for s in sizes),
               "how": "per bucket: cudaMemcpyAsync from pinned host memory -> Iallreduce -> 4 KiB of the result back to the host"}

    # ---- roofline of the dominant kernel (the bucket allreduce), timed live with CUDA events
    S = sizes[0] * esz
    if world == 1:
        hbm_peak, which = load_peaks()
        algo_bytes = 2 * S  # read + write of local HBM
        ach = algo_bytes / (kern_avg_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "peak_source": which,
                "kernel": "coll_kernel<T,1,OP_AR1> (world 1: vectorised HBM copy)", "algorithmic_bytes_per_launch": algo_bytes,
                "avg_launch_ms": kern_avg_ms, "avg_launch_ms_inside_step": kern_in_step_ms, "traffic": None}
    else:
        algo_bytes = 2 * (world - 1) / world * S  # per GPU per direction, whatever the algorithm moves
        ach = algo_bytes / (kern_avg_ms * 1e-3) / 1e9
        in_switch = L.dlnb_get(b"multicast") == 1 and world >= 4 and S >= comm.get("multimem_min_bytes")
        kname = (f"coll_kernel<T,{world},OP_ARM> (in-switch allreduce: multimem.ld_reduce + multimem.st over the NVSwitch multicast mapping)"
                 if in_switch else f"coll_kernel<T,{world},OP_AR2> (fused two-shot allreduce over peer HBM)")
        roof = {"bound": "nvlink", "achieved": ach, "peak": 900.0, "unit": "GB/s", "frac": ach / 900.0,
                "peak_source": "nominal NVLink 5 per direction (measured peer copy on this pool: 770 GB/s); achieved is bus bandwidth "
                               "2(n-1)/n*S/t, which in-switch reduction may push past the link rate because it moves fewer bytes",
                "kernel": kname, "algorithmic_bytes_per_launch": algo_bytes, "avg_launch_ms": kern_avg_ms,
                "avg_launch_ms_inside_step": kern_in_step_ms, "traffic": None}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        roof["traffic"] = json.load(open(tp)).get(f"allreduce_w{world}_{a.dtype}_{S}")

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        scale = pick_scale(st["modelSize"], 1)
        r = run_reference_dp(a.model, a.buckets, 1, 1, 3, scale)
        if r:
            cpu = {"value": r["value_ms"], "unit": "ms", "cores": 1, "kind": "reference", "sample": r["sample"], "sample_ms": r["sample_ms"]}

    if rank == 0:
        line = {"metric": METRIC, "value": ms_per_step, "unit": "ms", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": a.dtype.replace("fp", "f"),
                "data": "synthetic (integer-valued gradients, exact sums verified on device)", "config": config,
                "ratio_to_lower_bound": ms_per_step / lower_bound_ms,
                "allreduce_busbw_GBps": (algo_bytes / (kern_avg_ms * 1e-3) / 1e9) if world > 1 else None,
                "host_ms_per_step": host_ms / a.steps, "spin": {"overhead_us": spin_overhead_us, "slope": spin_slope, "ctas": a.spin_ctas},
                "arena": L.dlnb_arena_kind().decode(), "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
                "roofline": roof, "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    comm.finalize()
    capi.finalize()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
