"""CPU suite, part 2: the boundary and the host logic — no GPU, no compute calls.

 * libdlnb.so loads here and exports every symbol include/dlnb.h declares;
 * without a GPU the product path fails loudly (no CPU fallback anywhere);
 * the rendezvous (abstract unix socket hub + SCM_RIGHTS fd exchange) works across processes;
 * the launcher, the drivers' CLI and exit codes behave like the reference's;
 * the reference's own dp/fsdp (oracle/_ref) run multi-rank on host cores and their output
   parses with dlnetbench_b200.parser — that is also the multi-process (world_size 2) CPU path
   of bench.py's reference arm.
"""
import ctypes
import json
import os
import re
import subprocess
import sys

import pytest

from conftest import GPU_COUNT, ROOT

BIN = os.path.join(ROOT, "bin", "b200", "cpp", "data_parallel")
REF = os.path.join(ROOT, "oracle", "_ref")


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g

    g.build()


def test_library_loads_and_exports_every_declared_symbol():
    from dlnetbench_b200 import capi

    names = capi.declared_symbols()
    assert len(names) >= 45 and "dlnb_iallreduce" in names and "dlnb_emu_allreduce" in names
    lib = ctypes.CDLL(capi.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # and nothing is exported beyond the header (the ABI is exactly the header)
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (dlnb_[a-z0-9_]+)", out))
    assert exported == set(names), exported ^ set(names)


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """The boundary is a C ABI: the header compiles as C99 (no C++ or torch types) and a C program links against it."""
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                           os.path.join(ROOT, "include", "dlnb.h")])
    src = tmp_path / "c_client.c"
    src.write_text('#include <dlnb.h>\n#include <stdio.h>\n'
                   'int main(void) { printf("%s world=%d\\n", dlnb_version(), dlnb_world());\n'
                   '  void *p = 0; return dlnb_symm_alloc(16, &p) != 0 ? 0 : 1; /* must fail: no dlnb_init */ }\n')
    exe = tmp_path / "c_client"
    subprocess.check_call(["gcc", "-std=c99", "-I" + os.path.join(ROOT, "include"), "-o", str(exe), str(src),
                           "-L" + os.path.join(ROOT, "dlnetbench_b200"), "-ldlnb", "-Wl,-rpath," + os.path.join(ROOT, "dlnetbench_b200")])
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0 and "dlnb" in res.stdout and "world=1" in res.stdout, res.stdout + res.stderr


def test_header_cites_the_reference_for_every_replaced_call():
    text = open(os.path.join(ROOT, "include", "dlnb.h")).read()
    for needle in ("proxy_classes.hpp:32", "proxy_classes.hpp:34-35", "proxy_classes.hpp:38", "proxy_classes.hpp:41-43",
                   "dp.cpp:93,98", "utils.hpp:74-117", "proxy_classes.hpp:401-405"):
        assert needle in text, needle
    assert "torch" not in text.lower()


@pytest.mark.skipif(GPU_COUNT > 0, reason="checks the no-GPU failure mode")
def test_no_gpu_means_loud_failure_not_fallback():
    from dlnetbench_b200 import capi

    with pytest.raises(capi.DlnbError) as e:
        capi.init(0, 1, 0, "nogpu")
    assert "CUDA" in str(e.value) or "device" in str(e.value)
    res = subprocess.run([os.path.join(BIN, "dp"), "gpt2_l_16_bfloat16", "10", "."], cwd=ROOT, capture_output=True, text=True)
    assert res.returncode == 1 and "[rank 0]" in res.stderr and "dlnb_init" in res.stderr
    assert "#CCUTILS_SECTION" not in res.stdout


def test_product_never_touches_the_oracle():
    """Nothing the product ships may import, include, link or execute anything under oracle/."""
    bad = re.compile(r"import\s+oracle|from\s+oracle|oracle/|oracle_capi|dlnb_oracle|libdlnb_oracle")
    for top in ("dlnetbench_b200", "cpp", "include", "tools"):
        for base, _, files in os.walk(os.path.join(ROOT, top)):
            if os.sep + "build" in base:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".inc", ".h", ".mk")) or f.startswith("Makefile"):
                    text = open(os.path.join(base, f), errors="ignore").read()
                    assert not bad.search(text), os.path.join(base, f)


def test_rendezvous_allgather_and_fd_exchange(tmp_path):
    exe = tmp_path / "bootstrap_test"
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "tests", "cpp", "bootstrap_test.cpp"),
                           os.path.join(ROOT, "dlnetbench_b200", "csrc", "bootstrap.cpp")])
    for n in (2, 8):
        res = subprocess.run([str(exe), str(n)], capture_output=True, text=True, timeout=60)
        assert res.returncode == 0 and "OK" in res.stdout, res.stdout + res.stderr


def test_arena_layout_placement_logic(tmp_path):
    """ChunkLayout (dlnb_symm_alloc / dlnb_symm_free placement): no overlap, deterministic, coalescing, recycled flag."""
    exe = tmp_path / "arena_layout_test"
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-Wall", "-o", str(exe), os.path.join(ROOT, "tests", "cpp", "arena_layout_test.cpp")])
    res = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert res.returncode == 0 and "OK" in res.stdout, res.stdout + res.stderr


def test_launcher_sets_rank_env_and_propagates_failure():
    launch = os.path.join(ROOT, "bin", "dlnb_launch")
    res = subprocess.run([launch, "-n", "3", "sh", "-c", "echo r=$DLNB_RANK w=$DLNB_WORLD s=$DLNB_SESSION"], capture_output=True, text=True)
    lines = sorted(res.stdout.split())
    assert res.returncode == 0
    assert [x for x in lines if x.startswith("r=")] == ["r=0", "r=1", "r=2"]
    assert len({x for x in lines if x.startswith("s=")}) == 1
    res = subprocess.run([launch, "-n", "2", "sh", "-c", "test $DLNB_RANK = 0 || exit 7"], capture_output=True, text=True)
    assert res.returncode == 7
    assert subprocess.run([launch, "-n", "9", "true"], capture_output=True).returncode == 2


@pytest.mark.parametrize("prog,nreq", [("dp", 3), ("fsdp", 4), ("dp_loop", 3), ("fsdp_loop", 4)])
def test_driver_cli_matches_reference_conventions(prog, nreq):
    exe = os.path.join(BIN, prog)
    res = subprocess.run([exe, "-h"], capture_output=True, text=True, cwd=ROOT)
    assert res.returncode in (1, 255)  # help -> 1 (dp.cpp:132-135); fsdp with argc<5 -> -1 (fsdp.cpp:198-202)
    assert "Usage" in res.stdout
    args = ["gpt2_l_16_bfloat16", "4", "2", "."][: nreq - 1] if prog.startswith("fsdp") else ["gpt2_l_16_bfloat16", "4"]
    res = subprocess.run([exe, *args], capture_output=True, text=True, cwd=ROOT)  # one positional short
    assert res.returncode in (1, 255) and "Usage" in res.stdout
    full = ["nope_model", "4", "1", "."] if prog.startswith("fsdp") else ["nope_model", "4", "."]
    res = subprocess.run([exe, *full], capture_output=True, text=True, cwd=ROOT)
    assert res.returncode == 255 and "model stats file does not exist" in res.stderr  # return -1 (dp.cpp:145-148)


def test_hybrid_2d_cli_and_argument_checks():
    exe = os.path.join(ROOT, "bin", "b200", "cpp", "hybrid_parallel", "hybrid_2d")
    res = subprocess.run([exe, "-h"], capture_output=True, text=True, cwd=ROOT)
    assert res.returncode == 1 and "num_microbatches" in res.stdout
    res = subprocess.run([exe, "nope_model", "2", "4", "."], capture_output=True, text=True, cwd=ROOT)
    assert res.returncode == 255 and "model stats file does not exist" in res.stderr
    # 36 layers over 5 stages / batch 16 over 3 micro-batches: refused like the reference's asserts (hybrid_2d.cpp:263-267)
    for args in (["gpt2_l_16_bfloat16", "5", "4", "."], ["gpt2_l_16_bfloat16", "2", "3", "."]):
        res = subprocess.run([exe, *args], capture_output=True, text=True, cwd=ROOT, env={**os.environ, "DLNB_WORLD": "10", "DLNB_RANK": "0"})
        assert res.returncode == 1 and "must be multiples" in res.stderr, res.stderr


def test_hybrid_3d_cli_and_argument_checks():
    exe = os.path.join(ROOT, "bin", "b200", "cpp", "hybrid_parallel", "hybrid_3d")
    res = subprocess.run([exe, "-h"], capture_output=True, text=True, cwd=ROOT)
    assert res.returncode == 1 and "num_tensor_shards" in res.stdout
    res = subprocess.run([exe, "gpt2_l_16_bfloat16", "2", "4", "."], capture_output=True, text=True, cwd=ROOT)  # one short
    assert res.returncode == 1 and "Usage" in res.stdout
    res = subprocess.run([exe, "gpt2_l_16_bfloat16", "2", "4", "3", "."], capture_output=True, text=True, cwd=ROOT,
                         env={**os.environ, "DLNB_WORLD": "8", "DLNB_RANK": "0"})  # 8 ranks over 2 x 3
    assert res.returncode == 1 and "stages x tensor shards" in res.stderr


def test_hybrid_3d_moe_cli_and_argument_checks():
    exe = os.path.join(ROOT, "bin", "b200", "cpp", "hybrid_parallel", "hybrid_3d_moe")
    res = subprocess.run([exe, "-h"], capture_output=True, text=True, cwd=ROOT)
    assert res.returncode == 1 and "num_expert_shards" in res.stdout
    env = {**os.environ, "DLNB_WORLD": "8", "DLNB_RANK": "0"}
    # 4 experts over 3 shards / 8 ranks over 2 x 3: refused like the reference's checks (hybrid_3d_moe.cpp:296-311)
    res = subprocess.run([exe, "moe_tiny_8_bfloat16", "2", "4", "3", "tests/data"], capture_output=True, text=True, cwd=ROOT, env=env)
    assert res.returncode == 1 and "expert shards" in res.stderr, res.stderr
    # a dense table has Experts:1, which two shards do not divide
    res = subprocess.run([exe, "gpt2_l_16_bfloat16", "2", "4", "2", "."], capture_output=True, text=True, cwd=ROOT, env=env)
    assert res.returncode == 1 and "the experts (1) of the expert shards (2)" in res.stderr, res.stderr
    # the one MoE table upstream ships: 12.1e9 gradient elements per rank at this layout. The reference's int count
    # would wrap; here the sizing passes (64-bit allreduce entry) and the run stops only because this box has no GPU
    res = subprocess.run([exe, "mixtral_8x7b_16_bfloat16", "2", "4", "2", "."], capture_output=True, text=True, cwd=ROOT, env=env)
    if GPU_COUNT == 0:
        assert res.returncode != 0 and "dlnb_init" in res.stderr and "int count" not in res.stderr, res.stderr
    # a stage-to-stage message beyond int is still refused (it goes through send / recv's int)
    res = subprocess.run([exe, "mixtral_8x7b_128_bfloat16", "2", "1", "2", "."], capture_output=True, text=True, cwd=ROOT, env=env)
    assert res.returncode == 1 and "does not fit the interface's int count" in res.stderr, res.stderr


def test_count_layers_reads_the_model_json(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text('#include "utils.hpp"\nint main(int c,char**v){printf("%u\\n",count_layers(v[1]));}\n')
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-std=c++17", "-I" + os.path.join(ROOT, "cpp"), "-I" + os.path.join(ROOT, "include"), "-o", str(exe), str(src),
                           "-L" + os.path.join(ROOT, "dlnetbench_b200"), "-ldlnb", "-Wl,-rpath," + os.path.join(ROOT, "dlnetbench_b200")])
    got = {m: int(subprocess.check_output([str(exe), os.path.join(ROOT, "models", m + ".json")])) for m in ("gpt2_l", "vit_b", "llama3_70b", "vit_h")}
    assert got == {"gpt2_l": 36, "vit_b": 12, "llama3_70b": 80, "vit_h": 32}


def test_fsdp_rejects_sharding_factor_not_dividing_world():
    res = subprocess.run([os.path.join(BIN, "fsdp"), "gpt2_l_16_bfloat16", "4", "3", "."], capture_output=True, text=True, cwd=ROOT,
                         env={**os.environ, "DLNB_WORLD": "4", "DLNB_RANK": "0"})
    assert res.returncode == 1 and "multiple of the sharding factor" in res.stderr


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "dp")), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_reference_proxies_run_multirank_on_cpu_and_parse():
    from dlnetbench_b200.parser import parse_sections

    mpirun = os.path.join(REF, "dlnb_mpirun")
    res = subprocess.run([mpirun, "-n", "2", os.path.join(REF, "dp"), "vit_b_16_bfloat16", "7", ".", "-w", "0", "-r", "2"],
                         capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert res.returncode == 0, res.stderr
    sec = parse_sections(res.stdout)["dp"]
    assert sec["title"] == "Data Parallelism" and sec["json_data"]["backend"] == "MPI" and sec["json_data"]["world_size"] == 2
    assert sorted(sec["rank_outputs"]) == [0, 1] and len(sec["rank_outputs"][1]["runtimes"]) == 2
    res = subprocess.run([mpirun, "-n", "4", os.path.join(REF, "fsdp"), "vit_b_16_bfloat16", "3", "2", ".", "-w", "0", "-r", "1"],
                         capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert res.returncode == 0, res.stderr
    sec = parse_sections(res.stdout)["fsdp"]
    assert sec["json_data"]["num_replicas"] == 2 and len(sec["rank_outputs"][3]["reduce_scatter"]) == 3
    assert len(sec["rank_outputs"][0]["barrier"]) == 1


def test_metric_tables_from_sections():
    """The frames the reference's plots read (plots/parser.py:19-196), built from a section alone. The per-unit rows
    follow the order in which run_fsdp records its timers (fsdp.cpp:84-141)."""
    from dlnetbench_b200.parser import dp_metrics_df, fsdp_metrics_df, parse_sections, summarize

    text = "\n".join([
        '#CCUTILS_SECTION_BEGIN dp "Data Parallelism"',
        '#CCUTILS_GLOBAL_JSON dp {"model_name":"m","num_buckets":3,"local_batch_size":16,"world_size":2,"fwd_rt_whole_model":100,'
        '"bwd_rt_per_bucket":10.0,"msg_size_avg_bytes":4096.0,"msg_size_std_bytes":0.0,"backend":"B200P2P","lower_bound_us":130,'
        '"ratio_to_lower_bound":1.5}',
        '#CCUTILS_RANK_JSON dp 0 {"runtimes":[0.2,0.1],"barrier_time":[0.05,0.01],"hostname":"h","device_runtimes_ms":[190.0,95.0]}',
        '#CCUTILS_RANK_JSON dp 1 {"runtimes":[0.1,0.3],"barrier_time":[0.02,0.04],"hostname":"h","device_runtimes_ms":[100.0,290.0]}',
        '#CCUTILS_SECTION_END dp',
        '#CCUTILS_SECTION_BEGIN fsdp "Fully Sharded Data Parallelism"',
        '#CCUTILS_GLOBAL_JSON fsdp {"model_name":"m","world_size":2,"num_units":3,"sharding_factor":2,"num_replicas":1,'
        '"local_batch_size":8,"backend":"MPI"}',
        # 2 runs, 3 units: 2 forward waits, 2 backward waits and 3 reduce-scatters per run, in recording order
        '#CCUTILS_RANK_JSON fsdp 0 {"runtime":[1.0,2.0],"allgather":[0.1,0.2],"allgather_wait_fwd":[11,12,21,22],'
        '"allgather_wait_bwd":[13,14,23,24],"reduce_scatter":[15,16,17,25,26,27],"barrier":[],"hostname":"h","rank":0}',
        '#CCUTILS_SECTION_END fsdp'])
    secs = parse_sections(text)
    df = dp_metrics_df(secs["dp"])
    assert len(df) == 4 and list(df[df["rank"] == 1]["runtime"]) == [0.1, 0.3]
    assert {"network", "world_size", "model_name", "num_buckets", "fwd_rt_whole_model", "bwd_rt_per_bucket", "msg_size_avg_bytes",
            "rank", "runtime", "barrier_time", "energy_consumed"} <= set(df.columns)
    assert dp_metrics_df(secs["dp"], job_vars={"num_buckets": 99})["num_buckets"].iloc[0] == 99
    s = summarize("dp", secs["dp"])
    assert s["host_iter_ms_median"] == pytest.approx(250.0) and s["device_iter_ms_median"] == pytest.approx(240.0)  # max over ranks, median over runs
    assert s["exposed_comm_tail_ms_median"] == pytest.approx(45.0) and s["ratio_to_lower_bound"] == 1.5
    rt, comm = fsdp_metrics_df(secs["fsdp"])
    assert len(rt) == 2 and list(rt["allgather"]) == [0.1, 0.2] and len(comm) == 6
    run1 = comm[comm["run"] == 1].set_index("unit_idx")
    # forward wait k is for unit k+1; backward wait k for unit U-2-k; reduce-scatter k for unit U-1-k
    assert list(run1["allgather_wait_fwd"]) == [0, 21, 22]
    assert list(run1["allgather_wait_bwd"]) == [24, 23, 0]
    assert list(run1["reduce_scatter"]) == [27, 26, 25]
    assert summarize("fsdp", secs["fsdp"])["host_iter_ms_median"] == pytest.approx(1500.0)


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "dp")), reason="oracle/_ref not built")
def test_bench_reference_arm_two_ranks_on_cpu():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                          "--model", "vit_b_16_bfloat16", "--buckets", "5"], capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert res.returncode == 0, res.stderr
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["unit"] == "ms" and line["higher_is_better"] is False
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] == 2
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["value"] > 0


def test_bucket_split_in_bench_matches_oracle(oracle):
    import bench

    for m, n in ((8_030_261_248, 240), (632_404_480, 1000), (774_030_080, 10), (17, 5)):
        assert bench.bucket_sizes(m, n) == oracle.bucket_sizes(m, n)


def _gloo_rank(rank, world, port, q):
    import torch.distributed as dist

    import bench

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    p = bench.Plumbing(dist, "cpu")
    p.barrier()
    q.put((rank, p.max_over_ranks(10.0 + rank), p.max_over_ranks(5.0 - rank)))
    dist.destroy_process_group()


def test_bench_plumbing_world_size_2_gloo():
    """The N>1 plumbing of bench.py (barrier, max over ranks) on CPU with gloo, world_size 2."""
    import socket

    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert got == [(0, 11.0, 5.0), (1, 11.0, 5.0)]


def test_link_counters_degrade_gracefully_without_nvml():
    """No NVML / nvidia-smi (this container) or counters not exposed (the pool's boxes): ok is False and `why` says so — bench.py
    then reports the reason in roofline.nvlink instead of a byte count."""
    from dlnetbench_b200 import nvlink

    lc = nvlink.LinkCounters(0)
    if not lc.ok:
        assert lc.why and ("nvml" in lc.why.lower() or "nvidia" in lc.why.lower())
    d = nvlink.per_launch({"data_tx": 1000, "data_rx": 0}, {"data_tx": 21000, "data_rx": 4000}, 20)
    assert d == {"data_tx_bytes_per_launch": 1000.0, "data_rx_bytes_per_launch": 200.0}


def test_cpu_baseline_scale_follows_the_host_ram():
    """bench.py sizes the reference's CPU run from the box: the whole workload when its fp32 buffers (2 x 4 x Model_Size bytes per
    rank, dp.cpp:227-232) fit half of the available RAM, else the largest power-of-two fraction that does — and says which."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    m = 8_030_261_248  # llama3_8b
    assert bench.pick_scale(m, 1, {"cores": 16, "mem_available_bytes": 206e9}) == 1      # 64 GB of 103 GB
    assert bench.pick_scale(m, 2, {"cores": 16, "mem_available_bytes": 206e9}) == 2      # 128 GB does not fit 103 GB
    assert bench.pick_scale(m, 8, {"cores": 128, "mem_available_bytes": 2e12}) == 1      # 514 GB of 1 TB
    assert bench.pick_scale(m, 8, {"cores": 16, "mem_available_bytes": None}) == 64       # unknown RAM: 16 GB budget, 514 GB / 64 = 8 GB
    res = bench.host_resources()
    assert res["cores"] >= 1 and (res["mem_available_bytes"] is None or res["mem_available_bytes"] > 0)


def test_selfcheck_and_session_tools_import_without_a_gpu():
    import ast

    for rel in ("dlnetbench_b200/selfcheck.py", "tools/stress.py", "tools/hybrid_probe.py", "tools/footprint_probe.py", "tools/sanitize_probe.py"):
        ast.parse(open(os.path.join(ROOT, rel)).read())
    from dlnetbench_b200 import selfcheck

    assert selfcheck.EPS[1] == 2.0 ** -10


def test_reference_arm_scale_respects_the_time_budget():
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench_mod2", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    m, lb, big = 8_030_261_248, 11258.759, {"cores": 128, "mem_available_bytes": 4e12}
    assert bench.pick_scale(m, 1, big, lb, 25) == 1   # 25 x 11.26 s x 1.1 = 310 s
    assert bench.pick_scale(m, 2, big, lb, 25) == 1   # 450 s
    assert bench.pick_scale(m, 8, big, lb, 25) == 2   # 845 s -> half the workload, reported as extrapolated
    assert bench.pick_scale(m, 8, big, lb, 4) == 1
