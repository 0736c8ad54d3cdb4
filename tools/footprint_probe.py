"""What each footprint of the emulated compute does to a collective beside it, on one GPU.

For every footprint: (1) how long a 10 ms spin really lasts, (2) how long 8 virtual ranks' two-shot allreduce of
64 MiB takes while a 300 ms spin runs, against the same call alone. Prints JSON lines."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi  # noqa: E402

capi.init(0, 1, device=0, session="fp-%d" % os.getpid())
n, count = 8, 4 * 1024 * 1024
send = [capi.Tensor(count, capi.F16) for _ in range(n)]
recv = [capi.Tensor(count, capi.F16) for _ in range(n)]
for r, t in enumerate(send):
    t.fill(capi.PAT_INT, 0, r)
sp, rp = [t.data for t in send], [t.data for t in recv]


def alone():
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        capi.emu_allreduce(sp, rp, count, capi.F16, algo=2, ctas=16)
        best = min(best, time.perf_counter() - t0)
    return best * 1e3


base = alone()
print(json.dumps({"alone_ms": base}), flush=True)
for name, ctas, threads, mode in [("sleep 1 warp", 1, 32, 0), ("sleep 1 CTA/SM x256", 0, 256, 0), ("fma 1 CTA/SM x256", 0, 256, 1), ("fma 1 CTA/SM x512", 0, 512, 1),
                                  ("fma 1 CTA/SM x1024", 0, 1024, 1), ("fma 2 CTA/SM x256", 296, 256, 1), ("hbm 1 CTA/SM x256", 0, 256, 2), ("hbm 1 CTA/SM x1024", 0, 1024, 2)]:
    capi.set_option("spin_ctas", ctas)
    capi.set_option("spin_threads", threads)
    capi.set_option("spin_mode", mode)
    capi.iter_begin()
    capi.compute_spin(10000.0)
    spin_ms = capi.iter_end()
    capi.compute_sync()
    t0 = time.perf_counter()
    capi.compute_spin(300000.0)
    time.sleep(0.005)
    t1 = time.perf_counter()
    capi.emu_allreduce(sp, rp, count, capi.F16, algo=2, ctas=16)
    beside = (time.perf_counter() - t1) * 1e3
    capi.compute_sync()
    total = (time.perf_counter() - t0) * 1e3
    ok = capi.verify_sum_pattern(recv[3].data, count, 0, capi.F16, capi.PAT_INT, 0, list(range(n))) == 0
    print(json.dumps({"footprint": name, "coresident": capi.get_option("spin_coresident"), "spin_10ms_lasted_ms": round(spin_ms, 3),
                      "allreduce_beside_ms": round(beside, 3), "slowdown": round(beside / base, 2), "spin_300ms_total_ms": round(total, 1), "result_ok": ok}), flush=True)
capi.finalize()
