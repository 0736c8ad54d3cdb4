"""CPU suite, part 1: the oracle against the golden vectors and the reference's own outputs.

tests/golden/ref_collectives_f32.npz and ref_proxy_outputs.json were produced by the REFERENCE
(its MPICommunicator / dp.cpp / fsdp.cpp compiled unmodified against oracle/compat; generator:
tests/golden/make_golden.py). Nothing here reads /root/reference."""
import json
import os

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "ref_collectives_f32.npz"))


def test_golden_file_covers_every_collective(gold):
    ops = {k.split("_w")[0] for k in gold.files}
    assert ops == {"iallreduce", "allreduce", "iallgather", "reduce_scatter", "alltoall"}
    assert len(gold.files) == 5 * 2 * (2 * 11 + 3 * 9)  # worlds x patterns x (sum ops x counts + block ops x counts)


def _case(key):
    op, w, k, c = key.rsplit("_", 3)
    return op, int(w[1:]), int(k[1:]), int(c[1:])


def test_oracle_reproduces_reference_collectives_bit_for_bit(oracle, gold):
    for key in gold.files:
        op, world, kind, count = _case(key)
        want = gold[key]
        if op in ("allreduce", "iallreduce"):
            xs = [oracle.fill_pattern(count, 0, kind, 1234, r) for r in range(world)]
            got = oracle.allreduce(xs, 0)
        elif op == "iallgather":
            xs = [oracle.fill_pattern(count, 0, kind, 1234, r) for r in range(world)]
            got = oracle.allgather(xs, 0)
        elif op == "alltoall":
            xs = [oracle.fill_pattern(count * world, 0, kind, 1234, r) for r in range(world)]
            got = np.concatenate([oracle.alltoall(xs, r, count, 0) for r in range(world)])
        else:
            xs = [oracle.fill_pattern(count * world, 0, kind, 1234, r) for r in range(world)]
            got = np.concatenate([oracle.reduce_scatter_block(xs, r, count, 0) for r in range(world)])
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), key


def test_integer_pattern_sums_are_exact_in_fp16(oracle):
    n, count = 8, 10007
    xs16 = [oracle.fill_pattern(count, 1, 1, 0, r) for r in range(n)]
    xs32 = [x.astype(np.float32) for x in xs16]
    assert max(np.abs(x).max() for x in xs32) <= 8
    exact = np.sum(np.stack(xs32).astype(np.int64), axis=0)
    got = oracle.allreduce(xs16, 1).astype(np.float32)
    assert np.array_equal(got, exact.astype(np.float32))
    # any association gives the same bits
    rev = oracle.allreduce(xs16[::-1], 1)
    assert np.array_equal(rev.view(np.uint16), oracle.allreduce(xs16, 1).view(np.uint16))


def test_random_sums_within_north_star_tolerance(oracle):
    """2*(n-1)*eps relative error bound of the north star, checked against float64."""
    for dtype, eps, npdt in ((0, 2.0**-23, np.float32), (1, 2.0**-10, np.float16)):
        for n in (2, 4, 8):
            xs = [oracle.fill_pattern(50021, dtype, 2, 9, r) for r in range(n)]
            got = oracle.allreduce(xs, dtype).astype(np.float64)
            ref = np.sum(np.stack([x.astype(np.float64) for x in xs]), axis=0)
            scale = np.sum(np.stack([np.abs(x.astype(np.float64)) for x in xs]), axis=0)
            assert np.all(np.abs(got - ref) <= 2 * (n - 1) * eps * scale + 1e-30)


def test_half_and_bf16_conversions_match_numpy(oracle):
    L = oracle.lib()
    allh = np.arange(65536, dtype=np.uint16)
    finite = np.isfinite(allh.view(np.float16))
    ref = allh.view(np.float16).astype(np.float32)
    got = np.array([L.dlnb_oracle_half_to_float(int(h)) for h in allh[finite]], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), ref[finite].view(np.uint32))
    rng = np.random.default_rng(0)
    f = np.concatenate([rng.standard_normal(3000).astype(np.float32) * s for s in (1e-8, 1e-5, 1e-3, 1, 100, 7e4)]
                       + [np.array([65519.9, 65520, 65504, 2**-25, 2**-24, 2**-25 * 1.0000001, 0, -0.0, np.inf, -np.inf], np.float32)])
    hv = ref[finite]
    hv = np.sort(hv)
    mid = ((hv[:-1].astype(np.float64) + hv[1:].astype(np.float64)) / 2).astype(np.float32)[::7]
    f = np.concatenate([f, mid])
    with np.errstate(over="ignore"):
        want = f.astype(np.float16).view(np.uint16)
    got = np.array([L.dlnb_oracle_float_to_half(float(x)) for x in f], dtype=np.uint16)
    assert np.array_equal(got, want)
    # bf16: round-to-nearest-even of the top 16 bits
    u = f[np.isfinite(f)].view(np.uint32)
    want_bf = ((u.astype(np.uint64) + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    got_bf = np.array([L.dlnb_oracle_float_to_bf16(float(x)) for x in f[np.isfinite(f)]], dtype=np.uint16)
    assert np.array_equal(got_bf, want_bf)


def test_sizing_rules_match_reference_outputs(oracle):
    """Bucket/shard sizes and per-unit times as the reference's own binaries printed them."""
    gold = json.load(open(os.path.join(GOLD, "ref_proxy_outputs.json")))
    assert {g["prog"] for g in gold} == {"dp", "fsdp", "hybrid_2d", "hybrid_3d", "hybrid_3d_moe"}
    for g in gold:
        if g["prog"].startswith("hybrid"):
            continue  # checked in test_hybrid2d_sizing_matches_reference_outputs
        G = g["global"]
        st = oracle.parse_model_stats(os.path.join(ROOT, "model_stats", g["args"][0] + ".txt"))
        if g["prog"] == "dp":
            n = int(g["args"][1])
            sizes = oracle.bucket_sizes(st["modelSize"], n)
            assert sum(sizes) == st["modelSize"] == G["total_model_size_params"]
            assert G["fwd_rt_whole_model"] == st["avgForwardTime"]
            assert G["bwd_rt_per_bucket"] == st["avgBackwardTime"] // n
            avg, std = oracle.msg_stats(sizes)
            assert G["msg_size_avg_bytes"] == pytest.approx(avg * 4, rel=1e-6)  # the reference's _FLOAT is 4 bytes
            assert G["msg_size_std_bytes"] == pytest.approx(std * 4, rel=1e-4, abs=1e-3)
            assert set(g["rank_keys"]) == {"runtimes", "barrier_time", "hostname"}
        else:
            u, f = int(g["args"][1]), int(g["args"][2])
            shards = oracle.shard_sizes(st["modelSize"], u, f)
            assert G["allgather_msg_size_bytes"] == shards[0] * f * 4
            assert G["reducescatter_msg_size_bytes"] == shards[0] * 4
            assert G["num_replicas"] == g["world"] // f
            assert G["fwd_time_per_unit_us"] == pytest.approx(st["avgForwardTime"] / u, rel=1e-6)
            assert G["bwd_time_per_unit_us"] == pytest.approx(st["avgBackwardTime"] / u, rel=1e-6)
            rk = g["rank_keys"]
            runs = rk["runtime"]
            assert rk["allgather"] == runs and rk["reduce_scatter"] == runs * u
            assert rk["allgather_wait_fwd"] == runs * (u - 1) == rk["allgather_wait_bwd"]
            assert rk["barrier"] == (runs if G["num_replicas"] > 1 else 0)
            assert ("allreduce_msg_size_bytes" in G) == (G["num_replicas"] > 1)


def test_hybrid2d_sizing_matches_reference_outputs(oracle):
    """Per-micro-batch times, pipeline message and DP allreduce sizes as the reference's hybrid_2d printed them."""
    gold = [g for g in json.load(open(os.path.join(GOLD, "ref_proxy_outputs.json"))) if g["prog"] == "hybrid_2d"]
    assert len(gold) == 2
    for g in gold:
        G = g["global"]
        stages, mbs = int(g["args"][1]), int(g["args"][2])
        st = oracle.parse_model_stats(os.path.join(ROOT, "model_stats", g["args"][0] + ".txt"))
        z = oracle.hybrid2d_sizes(st, stages, mbs)
        assert G["fwd_rt_per_microbatch"] == z["fwd_us"] and G["bwd_rt_per_microbatch"] == z["bwd_us"]
        assert G["pipe_msg_size_bytes"] == z["pipe_elems"] * 4 and G["dp_allreduce_size_bytes"] == z["dp_elems"] * 4
        assert G["dp_size"] == g["world"] // stages and G["samples_per_microbatch"] == st["batchSize"] // mbs
        assert G["global_batch_size"] == g["world"] * st["batchSize"] // stages
        runs = g["rank_keys"]["runtimes"]
        # first / last stage time one transfer per micro-batch and direction; a middle stage reports the merged pairs
        assert g["rank_keys"]["pp_comm_time"] == runs * 2 * mbs and g["rank_keys"]["dp_comm_time"] == runs
        if stages > 2:
            assert g["rank1_keys"]["pp_comm_time"] in (2 * mbs, runs * 2 * mbs)
        assert z["lower_bound_us"] == (mbs + stages - 1) * (z["fwd_us"] + z["bwd_us"])


def test_hybrid3d_sizing_matches_reference_outputs(oracle):
    gold = [g for g in json.load(open(os.path.join(GOLD, "ref_proxy_outputs.json"))) if g["prog"] == "hybrid_3d"]
    assert len(gold) == 1
    for g in gold:
        G = g["global"]
        stages, mbs, tp = int(g["args"][1]), int(g["args"][2]), int(g["args"][3])
        st = oracle.parse_model_stats(os.path.join(ROOT, "model_stats", g["args"][0] + ".txt"))
        z = oracle.hybrid3d_sizes(st, stages, mbs, tp)
        assert G["fwd_rt_per_microbatch"] == z["fwd_us"] and G["bwd_rt_per_microbatch"] == z["bwd_us"]
        assert G["pipe_msg_size_bytes"] == z["pipe_elems"] * 4 and G["tp_allreduce_size_bytes"] == z["tp_elems"] * 4
        assert G["dp_allreduce_size_bytes"] == z["dp_elems"] * 4 and G["dp_size"] == g["world"] // (stages * tp)
        assert G["global_batch_size"] == G["dp_size"] * st["batchSize"]
        runs = g["rank_keys"]["runtimes"]
        assert g["rank_keys"]["tp_comm_time"] == runs * 4 * mbs and g["rank_keys"]["dp_comm_time"] == runs
        assert set(g["rank_keys"]) >= {"stage_id", "tp_id", "dp_id", "pp_comm_time"}


def test_hybrid3d_moe_sizing_matches_reference_outputs(oracle):
    """DP + PP + EP sizing against what the reference binary printed for the synthetic MoE table (tests/data)."""
    gold = [g for g in json.load(open(os.path.join(GOLD, "ref_proxy_outputs.json"))) if g["prog"] == "hybrid_3d_moe"]
    assert len(gold) == 2
    for g in gold:
        G = g["global"]
        stages, mbs, ep = int(g["args"][1]), int(g["args"][2]), int(g["args"][3])
        st = oracle.parse_model_stats(os.path.join(ROOT, g["args"][4], "model_stats", g["args"][0] + ".txt"))
        layers = json.load(open(os.path.join(ROOT, g["args"][4], "models", "moe_tiny.json")))["num_decoder_blocks"]
        z = oracle.hybrid3d_moe_sizes(st, layers, stages, mbs, ep)
        assert G["fwd_rt_per_microbatch"] == z["fwd_us"] and G["bwd_rt_per_microbatch"] == z["bwd_us"]
        assert G["pipe_msg_size_bytes"] == z["pipe_elems"] * 4
        assert G["ep_alltoall_size_bytes"] == z["a2a_elems_per_peer"] * 4
        assert G["dp_allreduce_size_bytes"] == z["dp_elems"] * 4 and G["dp_size"] == g["world"] // (stages * ep)
        assert G["num_experts"] == st["experts"] and G["global_batch_size"] == G["dp_size"] * st["batchSize"]
        runs = g["rank_keys"]["runtimes"]
        assert g["rank_keys"]["ep_comm_time"] == runs * 2 * mbs * z["a2a_per_microbatch"]
        assert g["rank_keys"]["dp_ep_comm_time"] == runs and g["rank_keys"]["dp_comm_time"] == runs
        assert set(g["rank_keys"]) >= {"stage_id", "ep_id", "dp_id", "pp_comm_time"}
        assert z["non_expert_elems"] <= z["dp_elems"]


def test_baseline_sizes_and_lower_bounds(oracle):
    """The numbers BASELINE.md / SURVEY §8a quote for C1..C5."""
    st = {m: oracle.parse_model_stats(os.path.join(ROOT, "model_stats", m + ".txt"))
          for m in ("gpt2_l_16_bfloat16", "llama3_8b_64_bfloat16", "vit_h_128_bfloat16", "llama3_70b_16_bfloat16")}
    assert st["llama3_8b_64_bfloat16"]["modelSize"] == 8_030_261_248
    assert (st["llama3_8b_64_bfloat16"]["avgForwardTime"], st["llama3_8b_64_bfloat16"]["avgBackwardTime"]) == (3_752_999, 7_505_999)
    assert oracle.bucket_sizes(774_030_080, 10) == [77_403_008] * 10
    b = oracle.bucket_sizes(8_030_261_248, 10)
    assert b[:8] == [803_026_125] * 8 and b[8:] == [803_026_124] * 2
    v = oracle.bucket_sizes(632_404_480, 1000)
    assert v.count(632_405) == 480 and v.count(632_404) == 520
    assert oracle.shard_sizes(8_030_261_248, 32, 8)[0] == 31_368_208
    assert oracle.shard_sizes(8_030_261_248, 32, 2)[0] == 125_472_832
    s70 = oracle.shard_sizes(70_553_706_496, 80, 8)
    assert s70[0] == 110_240_167 and len(s70) == 80
    lb = lambda m, n: oracle.dp_lower_bound_us(st[m]["avgForwardTime"], st[m]["avgBackwardTime"], n)  # noqa: E731
    assert lb("gpt2_l_16_bfloat16", 10) == 35_042
    assert lb("vit_h_128_bfloat16", 1000) == 57_012
    assert lb("llama3_8b_64_bfloat16", 240) == 11_258_759
    assert oracle.fsdp_lower_bound_us(3_752_999, 7_505_999, 32) == 11_141_695
    assert oracle.fsdp_lower_bound_us(8_131_499, 16_262_998, 80) == 24_292_757


def test_every_model_stats_table_parses(oracle):
    d = os.path.join(ROOT, "model_stats")
    files = sorted(f for f in os.listdir(d) if f.endswith(".txt"))
    assert len(files) == 72
    for f in files:
        st = oracle.parse_model_stats(os.path.join(d, f))
        assert st["modelSize"] > 0 and st["avgBackwardTime"] >= st["avgForwardTime"] > 0 and st["batchSize"] > 0


def test_checksum_detects_permutation(oracle):
    x = oracle.fill_pattern(1001, 0, 2, 1, 0)
    y = x.copy()
    y[[3, 4]] = y[[4, 3]]
    assert oracle.checksum(x) != oracle.checksum(y)
    assert oracle.checksum(x) == oracle.checksum(x.copy())


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_collectives")),
                    reason="oracle/_ref not built (needs /root/reference at build time)")
def test_restatement_against_the_reference_live_on_random_shapes(oracle, tmp_path):
    """Beyond the committed vectors: the reference's MPICommunicator (oracle/_ref/ref_collectives, built from the reference's
    own proxy_classes.hpp) run HERE on shapes drawn at test time — group sizes 1..8, counts up to 200k incl. a seed-dependent
    odd one — against the C restatement, bit for bit. Uses only oracle/_ref; nothing reads /root/reference."""
    import subprocess

    ref = os.path.join(ROOT, "oracle", "_ref")
    rng = np.random.default_rng(20261107)
    ops = ["allreduce", "iallreduce", "allgather", "iallgather", "reduce_scatter", "alltoall"]
    for case in range(18):
        op = ops[case % len(ops)]
        world = int(rng.integers(1, 9))
        count = int(rng.choice([1, 3, 1000, 4099, 65537, int(rng.integers(2, 200_000))]))
        if op in ("reduce_scatter", "alltoall", "allgather", "iallgather"):
            count = min(count, 30_000)
        kind, seed = int(rng.integers(1, 3)), int(rng.integers(0, 2**31))
        prefix = str(tmp_path / f"c{case}")
        subprocess.check_call([os.path.join(ref, "dlnb_mpirun"), "-n", str(world), os.path.join(ref, "ref_collectives"),
                               op, str(count), str(kind), str(seed), prefix], timeout=120)
        outs = [np.fromfile(f"{prefix}.r{r}.bin", dtype=np.float32) for r in range(world)]
        per_rank = count if op in ("allreduce", "iallreduce", "allgather", "iallgather") else count * world
        xs = [oracle.fill_pattern(per_rank, 0, kind, seed, r) for r in range(world)]
        for r in range(world):
            if op in ("allreduce", "iallreduce"):
                want = oracle.allreduce(xs, 0)
            elif op in ("allgather", "iallgather"):
                want = oracle.allgather(xs, 0)
            elif op == "reduce_scatter":
                want = oracle.reduce_scatter_block(xs, r, count, 0)
            else:
                want = oracle.alltoall(xs, r, count, 0)
            assert np.array_equal(outs[r].view(np.uint32), want.view(np.uint32)), (op, world, count, kind, seed, r)
