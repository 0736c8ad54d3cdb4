"""Real multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise).

One process per GPU, started by bin/dlnb_launch; each rank checks its results against the
oracle / the closed-form verifier (tests/mg_worker.py)."""
import os
import subprocess
import sys

import pytest

from conftest import GPU_COUNT, ROOT

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _launch(n, extra=(), env=None):
    launcher = os.path.join(ROOT, "bin", "dlnb_launch")
    assert os.path.exists(launcher), "bin/dlnb_launch missing: run __graft_entry__.build()"
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([launcher, "-n", str(n), sys.executable, os.path.join(ROOT, "tests", "mg_worker.py"), *extra],
                          capture_output=True, text=True, timeout=900, env=e)


@pytest.mark.parametrize("n", [2, 4, 8])
def test_collectives_on_n_gpus(n):
    if GPU_COUNT < n:
        pytest.skip(f"needs {n} GPUs, have {GPU_COUNT}")
    res = _launch(n, ["--quick"] if n > 2 else [])
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "arena=vmm" in res.stdout


def test_ipc_arena_fallback_on_2_gpus():
    if GPU_COUNT < 2:
        pytest.skip("needs 2 GPUs")
    res = _launch(2, ["--quick"], env={"DLNB_ARENA": "ipc"})
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "arena=ipc" in res.stdout


@pytest.mark.parametrize("prog,args,section", [
    ("data_parallel/dp", ["gpt2_l_16_bfloat16", "10", ".", "-t", "fp32", "-r", "3"], "dp"),
    ("data_parallel/dp", ["gpt2_l_16_bfloat16", "100", ".", "-t", "fp32", "-r", "3", "-g"], "dp"),
    ("data_parallel/fsdp", ["gpt2_l_16_bfloat16", "36", "2", ".", "-t", "fp16", "-r", "2"], "fsdp"),
    ("hybrid_parallel/hybrid_2d", ["gpt2_l_16_bfloat16", "2", "4", ".", "-t", "fp16", "-r", "3"], "dp_pp"),
    # one stage, two tensor shards: compute + the four tensor-parallel allreduces per micro-batch
    ("hybrid_parallel/hybrid_3d", ["gpt2_l_16_bfloat16", "1", "4", "2", ".", "-t", "fp16", "-r", "3"], "dp_pp_tp")])
def test_proxy_binaries_on_2_gpus(prog, args, section):
    if GPU_COUNT < 2:
        pytest.skip("needs 2 GPUs")
    from dlnetbench_b200.parser import parse_sections

    res = subprocess.run([os.path.join(ROOT, "bin", "dlnb_launch"), "-n", "2",
                          os.path.join(ROOT, "bin", "b200", "cpp", prog), *args],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    sec = parse_sections(res.stdout)[section]
    g = sec["json_data"]
    assert g["backend"] == "B200P2P" and g["world_size"] == 2 and sorted(sec["rank_outputs"]) == [0, 1]
    # emulated compute dominates these configurations: the device-timed iteration stays near its lower bound
    assert 1.0 <= g["ratio_to_lower_bound"] < 1.25, g["ratio_to_lower_bound"]


def test_hybrid_3d_moe_on_2_gpus_two_expert_shards():
    """One stage, two expert shards on the synthetic MoE table: 2 * layers all-to-alls per micro-batch and direction."""
    if GPU_COUNT < 2:
        pytest.skip("needs 2 GPUs")
    from dlnetbench_b200.parser import parse_sections

    res = subprocess.run([os.path.join(ROOT, "bin", "dlnb_launch"), "-n", "2",
                          os.path.join(ROOT, "bin", "b200", "cpp", "hybrid_parallel", "hybrid_3d_moe"),
                          "moe_tiny_8_bfloat16", "1", "2", "2", "tests/data", "-t", "fp16", "-r", "3"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    sec = parse_sections(res.stdout)["dp_pp_ep"]
    g, ro = sec["json_data"], sec["rank_outputs"]
    assert g["backend"] == "B200P2P" and g["num_experts"] == 4 and g["dp_size"] == 1
    assert g["ep_alltoall_size_bytes"] == 4 * 256 * 2 * 512 // 2 * 2 and g["dp_allreduce_size_bytes"] == (8_000_000 + 16_000_000) * 2
    assert [(ro[r]["stage_id"], ro[r]["ep_id"], ro[r]["dp_id"]) for r in range(2)] == [(0, 0, 0), (0, 1, 0)]
    assert len(ro[0]["ep_comm_time"]) == 3 * 2 * 2 * (2 * 4) and len(ro[0]["dp_ep_comm_time"]) == 3
    # 32 blocking all-to-alls of 1 MiB per peer against 12 ms of emulated compute
    assert 1.0 <= g["ratio_to_lower_bound"] < 1.5, g["ratio_to_lower_bound"]


def test_hybrid_3d_on_4_gpus_two_stages_two_shards():
    if GPU_COUNT < 4:
        pytest.skip("needs 4 GPUs")
    from dlnetbench_b200.parser import parse_sections

    res = subprocess.run([os.path.join(ROOT, "bin", "dlnb_launch"), "-n", "4",
                          os.path.join(ROOT, "bin", "b200", "cpp", "hybrid_parallel", "hybrid_3d"),
                          "gpt2_l_16_bfloat16", "2", "4", "2", ".", "-t", "fp16", "-r", "3"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    sec = parse_sections(res.stdout)["dp_pp_tp"]
    ro = sec["rank_outputs"]
    assert [(ro[r]["stage_id"], ro[r]["tp_id"], ro[r]["dp_id"]) for r in range(4)] == [(0, 0, 0), (0, 1, 0), (1, 0, 0), (1, 1, 0)]
    assert len(ro[0]["tp_comm_time"]) == 3 * 4 * 4 and sec["json_data"]["dp_size"] == 1


def test_hybrid_2d_on_4_gpus_two_pipelines():
    if GPU_COUNT < 4:
        pytest.skip("needs 4 GPUs")
    from dlnetbench_b200.parser import parse_sections

    res = subprocess.run([os.path.join(ROOT, "bin", "dlnb_launch"), "-n", "4",
                          os.path.join(ROOT, "bin", "b200", "cpp", "hybrid_parallel", "hybrid_2d"),
                          "gpt2_l_16_bfloat16", "2", "4", ".", "-t", "fp16", "-r", "3"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    sec = parse_sections(res.stdout)["dp_pp"]
    assert sec["json_data"]["dp_size"] == 2 and [sec["rank_outputs"][r]["stage_id"] for r in range(4)] == [0, 1, 0, 1]
    assert len(sec["rank_outputs"][0]["pp_comm_time"]) == 3 * 2 * 4  # runs x directions x micro-batches


def test_nccl_comparison_build_on_2_gpus():
    """The `nccl` comparison build (same drivers, collectives through NCCL as in the reference's CCLCommunicator): runs, reports
    backend NCCL, and stays near the lower bound on a compute-dominated configuration."""
    if GPU_COUNT < 2:
        pytest.skip("needs 2 GPUs")
    exe = os.path.join(ROOT, "bin", "nccl", "cpp", "data_parallel", "dp")
    if not os.path.exists(exe):
        pytest.skip("nccl build not present (no nccl.h at build time)")
    from dlnetbench_b200.parser import parse_sections

    res = subprocess.run([os.path.join(ROOT, "bin", "dlnb_launch"), "-n", "2", exe, "gpt2_l_16_bfloat16", "10", ".", "-t", "fp16", "-r", "3"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    g = parse_sections(res.stdout)["dp"]["json_data"]
    assert g["backend"] == "NCCL" and g["world_size"] == 2
    assert 1.0 <= g["ratio_to_lower_bound"] < 1.5, g["ratio_to_lower_bound"]
