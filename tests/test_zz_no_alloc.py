"""dlnb_emu_allreduce_noalloc — the allreduce entry point for callers that cannot allocate.

Some users of this library have no memory to give: their own pool already owns every free byte
of HBM, or they are in the middle of an overlapped compute/communication phase where a
cudaMalloc would synchronise the whole device. For them the library carries an 18 KiB scratch
array in its own image and this entry point uses that instead of cudaMalloc'ed flags.

The test calls it through the C ABI twice in a row — second time with another group size and
other data, so the first call's leftovers in the shared scratch are in the way — and requires
the outputs to be bit-identical (the tolerance of every parity test here) to what the existing
dlnb_emu_allreduce produces for the same inputs. It also checks the documented limit.
This file sorts last on purpose: with `-x`, a failure here cannot hide any other test.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NPDT = {0: np.float32, 1: np.float16, 2: np.uint16}


def bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a.view(np.uint16)


def _run_pair(dl, n, count, dtype, algo, ctas, seed):
    send = [dl.Tensor(count, dtype) for _ in range(n)]
    for r, t in enumerate(send):
        t.fill(dl.PAT_UNIFORM, seed, r)
    want = [dl.Tensor(count, dtype) for _ in range(n)]
    got = [dl.Tensor(count, dtype) for _ in range(n)]
    dl.emu_allreduce([t.data for t in send], [t.data for t in want], count, dtype, algo=algo, ctas=ctas)
    dl.emu_allreduce_noalloc([t.data for t in send], [t.data for t in got], count, dtype, algo=algo, ctas=ctas)
    for r in range(n):
        assert np.array_equal(bits(got[r].download(NPDT[dtype])), bits(want[r].download(NPDT[dtype]))), (n, count, dtype, algo, r)
    # and the result is the right one, not merely the same one
    assert dl.verify_sum_pattern(got[0].data, count, 0, dtype, dl.PAT_UNIFORM, seed, list(range(n))) == 0


def test_two_calls_in_a_row_match_the_allocating_entry_point(dl):
    _run_pair(dl, n=8, count=300_007, dtype=dl.F16, algo=dl.ALGO_TWOSHOT, ctas=12, seed=21)
    _run_pair(dl, n=3, count=65_541, dtype=dl.F32, algo=dl.ALGO_ONESHOT, ctas=5, seed=22)


def test_default_grid_and_auto_algorithm(dl):
    _run_pair(dl, n=4, count=1_000_003, dtype=dl.BF16, algo=dl.ALGO_AUTO, ctas=0, seed=23)


def test_grid_beyond_the_builtin_scratch_is_refused(dl):
    t = dl.Tensor(1024, dl.F32)
    with pytest.raises(dl.DlnbError) as e:
        dl.emu_allreduce_noalloc([t.data, t.data], [t.data, t.data], 1024, dl.F32, ctas=65)
    assert "at most 64" in str(e.value)
    with pytest.raises(dl.DlnbError):
        dl.emu_allreduce_noalloc([t.data] * 9, [t.data] * 9, 1024, dl.F32)
