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


@pytest.mark.parametrize("prog,args", [("dp", ["gpt2_l_16_bfloat16", "10", ".", "-t", "fp32", "-r", "3"]),
                                       ("fsdp", ["gpt2_l_16_bfloat16", "36", "2", ".", "-t", "fp16", "-r", "2"])])
def test_proxy_binaries_on_2_gpus(prog, args):
    if GPU_COUNT < 2:
        pytest.skip("needs 2 GPUs")
    res = subprocess.run([os.path.join(ROOT, "bin", "dlnb_launch"), "-n", "2",
                          os.path.join(ROOT, "bin", "b200", "cpp", "data_parallel", prog), *args],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert f"#CCUTILS_SECTION_END {prog}" in res.stdout
