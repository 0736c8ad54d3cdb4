"""Size-independent properties of the oracle's collectives (CPU only; hypothesis-generated shapes).

These are the identities the GPU parity tests lean on at sizes where no golden vector exists:
  * reduce_scatter_block of rank r == the r-th block of allreduce (same fixed summation order),
  * allgather of the reduce-scatter results == allreduce,
  * alltoall is a block transpose: applying it twice returns every rank's input,
  * integer-valued inputs sum exactly, whatever the rank order,
  * the sizing rules partition the model exactly.
"""
import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

NP = {0: np.float32, 1: np.float16, 2: np.uint16}


def inputs(oracle, n, count, dtype, kind, seed):
    return [oracle.fill_pattern(count, dtype, kind, seed, r) for r in range(n)]


def bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a.view(np.uint16)


@settings(max_examples=40, deadline=None)
@given(n=st.integers(1, 8), block=st.integers(1, 300), dtype=st.sampled_from([0, 1, 2]), seed=st.integers(0, 2**31))
def test_reduce_scatter_then_allgather_is_allreduce(oracle, n, block, dtype, seed):
    xs = inputs(oracle, n, n * block, dtype, 2, seed)
    full = oracle.allreduce(xs, dtype)
    parts = [oracle.reduce_scatter_block(xs, r, block, dtype) for r in range(n)]
    for r in range(n):
        assert np.array_equal(bits(parts[r]), bits(full[r * block:(r + 1) * block]))
    assert np.array_equal(bits(oracle.allgather(parts, dtype)), bits(full))


@settings(max_examples=40, deadline=None)
@given(n=st.integers(1, 8), block=st.integers(1, 200), dtype=st.sampled_from([0, 1]), seed=st.integers(0, 2**31))
def test_alltoall_is_a_block_transpose(oracle, n, block, dtype, seed):
    xs = inputs(oracle, n, n * block, dtype, 2, seed)
    once = [oracle.alltoall(xs, r, block, dtype) for r in range(n)]
    for r in range(n):
        for p in range(n):  # block p of rank r's result is block r of rank p's input
            assert np.array_equal(bits(once[r][p * block:(p + 1) * block]), bits(xs[p][r * block:(r + 1) * block]))
    twice = [oracle.alltoall(once, r, block, dtype) for r in range(n)]
    for r in range(n):
        assert np.array_equal(bits(twice[r]), bits(xs[r]))


@settings(max_examples=30, deadline=None)
@given(n=st.integers(2, 8), count=st.integers(1, 2000), dtype=st.sampled_from([0, 1, 2]), seed=st.integers(0, 2**31), data=st.data())
def test_integer_valued_inputs_sum_exactly_in_any_rank_order(oracle, n, count, dtype, seed, data):
    xs = inputs(oracle, n, count, dtype, 1, seed)
    perm = data.draw(st.permutations(list(range(n))))
    a, b = oracle.allreduce(xs, dtype), oracle.allreduce([xs[i] for i in perm], dtype)
    assert np.array_equal(bits(a), bits(b))
    if dtype != 2:  # against plain integer arithmetic
        assert np.array_equal(a.astype(np.int64), sum(x.astype(np.int64) for x in xs))


@settings(max_examples=60, deadline=None)
@given(model=st.integers(1, 10**11), buckets=st.integers(1, 5000))
def test_bucket_split_partitions_the_model(oracle, model, buckets):
    sizes = oracle.bucket_sizes(model, buckets)
    assert sum(sizes) == model and max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


@settings(max_examples=60, deadline=None)
@given(model=st.integers(1, 10**11), units=st.integers(1, 200), factor=st.integers(1, 8))
def test_shards_cover_every_unit(oracle, model, units, factor):
    shards = oracle.shard_sizes(model, units, factor)
    unit_sizes = oracle.bucket_sizes(model, units)
    assert len(shards) == units
    for s, u in zip(shards, unit_sizes):  # ceil(unit / F): F shards cover the unit with less than F elements of padding
        assert s * factor >= u and s * factor - u < factor
