"""Parser for the proxies' stdout sections (the output-format contract).

Returns, per section id, the shapes the reference's plots/parser.py reads from ccutils'
MPIOutputParser: `json_data` (the global object) and `rank_outputs` (rank -> per-rank
object, what it finds under mpi_all_prints["ccutils_rank_json"])."""
from __future__ import annotations

import json


def parse_sections(text: str) -> dict:
    out: dict = {}
    for line in text.splitlines():
        if line.startswith("#CCUTILS_SECTION_BEGIN "):
            _, sid, title = line.split(" ", 2)
            out.setdefault(sid, {"title": title.strip('"'), "json_data": {}, "rank_outputs": {}})
        elif line.startswith("#CCUTILS_GLOBAL_JSON "):
            _, sid, js = line.split(" ", 2)
            out.setdefault(sid, {"title": "", "json_data": {}, "rank_outputs": {}})["json_data"] = json.loads(js)
        elif line.startswith("#CCUTILS_RANK_JSON "):
            _, sid, rank, js = line.split(" ", 3)
            out.setdefault(sid, {"title": "", "json_data": {}, "rank_outputs": {}})["rank_outputs"][int(rank)] = json.loads(js)
    return out


# ---------------------------------------------------------------------------------------------
# Tables for the analysis flow. The reference turns a section into pandas frames in
# plots/parser.py (extract_dp_metrics_df :139-196, extract_fsdp_metrics_df :19-100) with the job
# variables of its batch system; here the same columns come from the section alone, `job_vars`
# (optional) overriding like there. Extra columns of this build (device-timed iteration) are
# appended when the section carries them.
# ---------------------------------------------------------------------------------------------
def dp_metrics_df(section: dict, job_vars: dict | None = None, network: str = "b200-nvswitch"):
    """One row per rank and run: runtime (s, host clock), barrier_time (s, exposed communication tail)."""
    import pandas as pd

    jv = job_vars or {}
    g = section["json_data"]
    base = {
        "network": network,
        "world_size": int(jv.get("world_size", g.get("world_size"))),
        "model_name": jv.get("models", g.get("model_name")),
        "local_batch_size": g.get("local_batch_size"),
        "num_buckets": jv.get("num_buckets", g.get("num_buckets")),
        "protocol": jv.get("protocol"), "channels": jv.get("channels"), "threads": jv.get("threads"),
        "algorithm": jv.get("algorithm"),
        "fwd_rt_whole_model": g.get("fwd_rt_whole_model"),
        "bwd_rt_per_bucket": g.get("bwd_rt_per_bucket"),
        "msg_size_avg_bytes": g.get("msg_size_avg_bytes"),
        "msg_size_std_bytes": g.get("msg_size_std_bytes"),
    }
    rows = []
    for rank, r in sorted(section["rank_outputs"].items()):
        runtimes, barriers = r["runtimes"], r["barrier_time"]
        energy = r.get("energy_consumed", [None] * len(runtimes))
        device = r.get("device_runtimes_ms", [None] * len(runtimes))
        for i, (rt, bt) in enumerate(zip(runtimes, barriers)):
            rows.append({**base, "rank": rank, "hostname": r.get("hostname", "unknown"), "run": i, "runtime": rt, "barrier_time": bt,
                         "energy_consumed": energy[i] if i < len(energy) else None,
                         "device_runtime_ms": device[i] if i < len(device) else None})
    return pd.DataFrame(rows)


def fsdp_metrics_df(section: dict, job_vars: dict | None = None, network: str = "b200-nvswitch"):
    """(runtime_df, comm_df): one row per rank and run / per rank, run and unit."""
    import pandas as pd

    jv = job_vars or {}
    g = section["json_data"]
    keys = ("sharding_factor", "num_replicas", "model_size_bytes", "local_batch_size", "num_units", "fwd_time_per_unit_us",
            "bwd_time_per_unit_us", "allgather_msg_size_bytes", "reducescatter_msg_size_bytes")
    base = {"network": network, "world_size": int(jv.get("world_size", g.get("world_size"))),
            "model_name": jv.get("models_fsdp", g.get("model_name")), **{k: g.get(k) for k in keys}}
    runtime_rows, comm_rows = [], []
    for rank, r in sorted(section["rank_outputs"].items()):
        runtimes = r.get("runtime", [])
        n_runs = len(runtimes)
        if not n_runs:
            continue
        barriers, ag = r.get("barrier", []), r.get("allgather", [])
        wf, wb, rs = r.get("allgather_wait_fwd", []), r.get("allgather_wait_bwd", []), r.get("reduce_scatter", [])
        # b200 addition: the part of each prefetch allgather that was still running when the compute hiding it had ended
        # (device-timed, ms) — the reference's meaning of the two wait timers (its host sleeps through the compute first)
        xf, xb = r.get("allgather_exposed_wait_fwd_ms", []), r.get("allgather_exposed_wait_bwd_ms", [])
        device = r.get("device_runtimes_ms", [])
        host = r.get("hostname", "unknown")
        units = base["num_units"] or max(len(rs) // n_runs, 1)
        for i in range(n_runs):
            runtime_rows.append({**base, "rank": rank, "hostname": host, "run": i, "runtime": runtimes[i],
                                 "allgather": ag[i] if i < len(ag) else 0, "barrier": barriers[i] if i < len(barriers) else 0,
                                 "device_runtime_ms": device[i] if i < len(device) else None})
            # Per run the proxies time (fsdp.cpp:84-141): units-1 forward waits — the k-th is for the prefetch of unit
            # k+1 —, units-1 backward waits — the k-th is for unit units-2-k —, and `units` reduce-scatters, the k-th
            # for unit units-1-k. Rows are per UNIT; a unit without that wait (unit 0's gather is the blocking
            # "allgather" timer, the last unit has no backward prefetch) gets 0.
            for u in range(units):
                kf, kb, ks = i * (units - 1) + (u - 1), i * (units - 1) + (units - 2 - u), i * units + (units - 1 - u)
                comm_rows.append({**base, "rank": rank, "hostname": host, "run": i, "unit_idx": u,
                                  "allgather_wait_fwd": wf[kf] if u >= 1 and kf < len(wf) else 0,
                                  "allgather_wait_bwd": wb[kb] if u <= units - 2 and kb < len(wb) else 0,
                                  "allgather_exposed_wait_fwd_ms": xf[kf] if u >= 1 and kf < len(xf) else 0,
                                  "allgather_exposed_wait_bwd_ms": xb[kb] if u <= units - 2 and kb < len(xb) else 0,
                                  "reduce_scatter": rs[ks] if ks < len(rs) else 0})
    return pd.DataFrame(runtime_rows), pd.DataFrame(comm_rows)


def summarize(section_id: str, section: dict) -> dict:
    """Headline numbers of one section: iteration time (max over ranks, median over runs; device-timed when the build
    records it, host-timed otherwise), the exposed-communication share and, when present, the ratio to the lower bound."""
    import statistics

    g, ranks = section["json_data"], section["rank_outputs"]
    rt_key = "runtimes" if section_id == "dp" or any("runtimes" in r for r in ranks.values()) else "runtime"
    per_rank = [r[rt_key] for r in ranks.values() if rt_key in r]
    out = {"section": section_id, "model_name": g.get("model_name"), "world_size": g.get("world_size"), "backend": g.get("backend"),
           "runs": len(per_rank[0]) if per_rank else 0}
    if per_rank and per_rank[0]:
        n = min(len(v) for v in per_rank)
        out["host_iter_ms_median"] = statistics.median(max(v[i] for v in per_rank) for i in range(n)) * 1e3
    dev = [r["device_runtimes_ms"] for r in ranks.values() if r.get("device_runtimes_ms")]
    if dev:
        n = min(len(v) for v in dev)
        out["device_iter_ms_median"] = statistics.median(max(v[i] for v in dev) for i in range(n))
    for k in ("lower_bound_us", "ratio_to_lower_bound", "allreduce_busbw_GBps_avg", "reduce_scatter_busbw_GBps"):
        if k in g:
            out[k] = g[k]
    # hybrid proxies: blocking communication per category (pp / dp / tp / ep), seconds per run summed, slowest rank
    runs = out["runs"]
    for k in sorted({k for r in ranks.values() for k in r if k.endswith("_comm_time")}):
        tot = [sum(r[k]) / runs for r in ranks.values() if isinstance(r.get(k), list) and r[k] and runs]
        if tot:
            out[k.replace("_time", "") + "_ms_per_iter_max_rank"] = max(tot) * 1e3
    tail_key = "barrier_time" if rt_key == "runtimes" and any("barrier_time" in r for r in ranks.values()) else "barrier"
    tails = [r[tail_key] for r in ranks.values() if r.get(tail_key)]
    if tails and "host_iter_ms_median" in out and out["host_iter_ms_median"] > 0:
        n = min(len(v) for v in tails)
        out["exposed_comm_tail_ms_median"] = statistics.median(max(v[i] for v in tails) for i in range(n)) * 1e3
    return out
