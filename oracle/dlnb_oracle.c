/*
 * dlnb_oracle.c — CPU restatement of DLNetBench's data-parallel hot path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and only as the checker or as the timed CPU baseline.
 *
 * PARITY STATUS: "parity unpinned" by the reference's own tests — the reference
 * ships no tests, golden vectors or known answers for this path, its
 * buffers are all zeros and never inspected (proxy_classes.hpp:397,405). The
 * arithmetic it invokes lives in the site's MPI / NCCL library (not vendored, no
 * version pinned). What pins this file instead:
 *   (1) the MPI-standard definition of MPI_Allreduce / MPI_Allgather /
 *       MPI_Reduce_scatter_block with MPI_SUM on MPI_FLOAT, which is what the
 *       reference's MPICommunicator calls (proxy_classes.hpp:66-101);
 *   (2) oracle/_ref: the reference's own dp.cpp / fsdp.cpp / proxy_classes.hpp
 *       compiled UNMODIFIED against oracle/compat (a shared-memory mini-MPI and a
 *       ccutils subset) — tests/golden/ holds its outputs, and tests check this
 *       restatement against them;
 *   (3) closed forms: integer-valued inputs sum exactly in fp16/fp32 in any order.
 *
 * Reduction order is fixed: rank 0, 1, ..., n-1, accumulated in fp32
 * (((x0 + x1) + x2) + ...). For 16-bit types the inputs are widened to fp32,
 * summed in fp32 and rounded once (RNE); the reference's MPI path itself is
 * fp32-only (README.md:33, proxy_classes.hpp:67).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- 16-bit float conversions (bit-exact, RNE), no compiler support needed ---- */
static inline float u32_as_f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f_as_u32(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

float dlnb_oracle_half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ffu;
    if (exp == 0) {
        if (man == 0) return u32_as_f(sign);
        /* subnormal: normalise */
        int e = -1;
        do { man <<= 1; e++; } while (!(man & 0x400u));
        man &= 0x3ffu;
        return u32_as_f(sign | (uint32_t)(127 - 15 - e) << 23 | man << 13);
    }
    if (exp == 31) return u32_as_f(sign | 0x7f800000u | man << 13);
    return u32_as_f(sign | (exp + 112) << 23 | man << 13);
}

uint16_t dlnb_oracle_float_to_half(float f) {
    const uint32_t x = f_as_u32(f);
    const uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    const uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u) return (uint16_t)(sign | 0x7c00u | (ax > 0x7f800000u ? 0x200u | ((ax >> 13) & 0x3ffu) : 0));
    if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u); /* rounds to inf (>= 65520) */
    if (ax < 0x33000001u) return sign;                        /* rounds to zero (<= 2^-25) */
    int32_t e = (int32_t)(ax >> 23) - 127;
    uint32_t m = (ax & 0x7fffffu) | 0x800000u;
    int shift;
    uint32_t base;
    if (e < -14) { shift = 13 + (-14 - e); base = 0; }  /* subnormal half */
    else { shift = 13; base = (uint32_t)(e + 15) << 10; m &= 0x7fffffu; }
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (q & 1u))) q++;
    return (uint16_t)(sign | (base + q)); /* carry into the exponent is correct by construction */
}

float dlnb_oracle_bf16_to_float(uint16_t h) { return u32_as_f((uint32_t)h << 16); }
uint16_t dlnb_oracle_float_to_bf16(float f) {
    uint32_t x = f_as_u32(f);
    if ((x & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((x >> 16) | 0x40u);
    x += 0x7fffu + ((x >> 16) & 1u);
    return (uint16_t)(x >> 16);
}

enum { ORACLE_F32 = 0, ORACLE_F16 = 1, ORACLE_BF16 = 2 };
static size_t esize(int dt) { return dt == ORACLE_F32 ? 4 : 2; }

static inline float load_elem(const void *p, size_t i, int dt) {
    if (dt == ORACLE_F32) return ((const float *)p)[i];
    if (dt == ORACLE_F16) return dlnb_oracle_half_to_float(((const uint16_t *)p)[i]);
    return dlnb_oracle_bf16_to_float(((const uint16_t *)p)[i]);
}
static inline void store_elem(void *p, size_t i, int dt, float v) {
    if (dt == ORACLE_F32) ((float *)p)[i] = v;
    else if (dt == ORACLE_F16) ((uint16_t *)p)[i] = dlnb_oracle_float_to_half(v);
    else ((uint16_t *)p)[i] = dlnb_oracle_float_to_bf16(v);
}

/* ---- collectives --------------------------------------------------------------
 * MPICommunicator::Iallreduce / Allreduce — MPI_SUM over MPI_FLOAT
 * (cpp/proxy_classes.hpp:66-72). send[r] is rank r's buffer; every rank receives
 * the same `recv`. */
void dlnb_oracle_allreduce(const void *const *send, int nranks, void *recv, size_t count, int dtype) {
    for (size_t i = 0; i < count; ++i) {
        if (nranks == 1) { /* bit copy */
            memcpy((char *)recv + i * esize(dtype), (const char *)send[0] + i * esize(dtype), esize(dtype));
            continue;
        }
        float acc = load_elem(send[0], i, dtype);
        for (int r = 1; r < nranks; ++r) acc += load_elem(send[r], i, dtype);
        store_elem(recv, i, dtype, acc);
    }
}

/* MPICommunicator::Allgather / Iallgather (cpp/proxy_classes.hpp:91-97): recv holds
 * nranks blocks of sendcount elements in rank order; identical on every rank. */
void dlnb_oracle_allgather(const void *const *send, int nranks, void *recv, size_t sendcount, int dtype) {
    const size_t b = sendcount * esize(dtype);
    for (int r = 0; r < nranks; ++r) memcpy((char *)recv + (size_t)r * b, send[r], b);
}

/* MPICommunicator::Reduce_Scatter_block (cpp/proxy_classes.hpp:99-101): send[r] holds
 * nranks*recvcount elements; rank `me` receives the sum over ranks of block `me`. */
void dlnb_oracle_reduce_scatter_block(const void *const *send, int nranks, int me, void *recv, size_t recvcount,
                                      int dtype) {
    const size_t off = (size_t)me * recvcount;
    for (size_t i = 0; i < recvcount; ++i) {
        if (nranks == 1) {
            memcpy((char *)recv + i * esize(dtype), (const char *)send[0] + (off + i) * esize(dtype), esize(dtype));
            continue;
        }
        float acc = load_elem(send[0], off + i, dtype);
        for (int r = 1; r < nranks; ++r) acc += load_elem(send[r], off + i, dtype);
        store_elem(recv, i, dtype, acc);
    }
}

/* MPICommunicator::Alltoall (cpp/proxy_classes.hpp:74-77): send[r] holds nranks blocks of `count`
 * elements; rank `me` receives block `me` of every rank's send buffer, in rank order. */
void dlnb_oracle_alltoall(const void *const *send, int nranks, int me, void *recv, size_t count, int dtype) {
    const size_t b = count * esize(dtype);
    for (int r = 0; r < nranks; ++r) memcpy((char *)recv + (size_t)r * b, (const char *)send[r] + (size_t)me * b, b);
}

/* ---- sizing rules -----------------------------------------------------------------
 * dp.cpp:159-164 — bucket i gets floor(M/N) + (i < M mod N). */
void dlnb_oracle_bucket_sizes(uint64_t model_size, int num_buckets, uint64_t *out) {
    const uint64_t base = model_size / (uint64_t)num_buckets, rem = model_size % (uint64_t)num_buckets;
    for (int i = 0; i < num_buckets; ++i) out[i] = base + ((uint64_t)i < rem ? 1 : 0);
}
/* fsdp.cpp:244-255 — unit sizes as buckets; shard = ceil(unit / sharding_factor). */
void dlnb_oracle_shard_sizes(uint64_t model_size, int num_units, int sharding_factor, uint64_t *out) {
    const uint64_t base = model_size / (uint64_t)num_units, rem = model_size % (uint64_t)num_units;
    for (int u = 0; u < num_units; ++u) {
        const uint64_t unit = base + ((uint64_t)u < rem ? 1 : 0);
        out[u] = unit / (uint64_t)sharding_factor + (unit % (uint64_t)sharding_factor ? 1 : 0);
    }
}
/* dp.cpp:93,98,155 — pure emulated compute per iteration, microseconds:
 * fwd + N * floor(bwd / N), with fwd/bwd truncated to integers (utils.hpp:228,232). */
uint64_t dlnb_oracle_dp_lower_bound_us(uint64_t fwd_us, uint64_t bwd_us, int num_buckets) {
    return fwd_us + (uint64_t)num_buckets * (bwd_us / (uint64_t)num_buckets);
}
/* fsdp.cpp:95-108,111-145,356-357 — (U-1) forward sleeps and U backward sleeps of
 * (float)t/U microseconds each, truncated by usleep's integer argument. */
uint64_t dlnb_oracle_fsdp_lower_bound_us(uint64_t fwd_us, uint64_t bwd_us, int num_units) {
    const float f = (float)fwd_us / (float)num_units, b = (float)bwd_us / (float)num_units;
    return (uint64_t)(num_units - 1) * (uint64_t)f + (uint64_t)num_units * (uint64_t)b;
}
/* hybrid_2d.cpp:232-250 — GPipe proxy sizing: out[0] = fwd us per micro-batch, out[1] = bwd us per
 * micro-batch (whole-model time / stages / micro-batches, integer divisions), out[2] = elements of
 * one stage-to-stage message (seq_len * embed_dim * batch/micro-batches), out[3] = elements of the
 * per-stage DP allreduce (model size / stages), out[4] = pure-compute critical path of one GPipe
 * iteration in us: (M + S - 1) * (fwd + bwd). */
void dlnb_oracle_hybrid2d_sizes(uint64_t fwd_us, uint64_t bwd_us, uint64_t model_size, uint64_t batch, uint64_t seq_len,
                                uint64_t embed_dim, int num_stages, int num_microbatches, uint64_t *out) {
    out[0] = fwd_us / (uint64_t)num_stages / (uint64_t)num_microbatches;
    out[1] = bwd_us / (uint64_t)num_stages / (uint64_t)num_microbatches;
    out[2] = (uint64_t)((unsigned)seq_len * (unsigned)embed_dim * (batch / (uint64_t)num_microbatches));
    out[3] = model_size / (uint64_t)num_stages;
    out[4] = (uint64_t)(num_microbatches + num_stages - 1) * (out[0] + out[1]);
}

/* hybrid_3d.cpp:308-323 — the same with T tensor shards: times divided by (micro-batches * T),
 * out[2] pipe message as above, out[3] = DP allreduce elements = model / (stages * T),
 * out[4] = compute critical path, out[5] = TP allreduce elements = pipe message / T. */
void dlnb_oracle_hybrid3d_sizes(uint64_t fwd_us, uint64_t bwd_us, uint64_t model_size, uint64_t batch, uint64_t seq_len,
                                uint64_t embed_dim, int num_stages, int num_microbatches, int num_tensor_shards, uint64_t *out) {
    const uint64_t div = (uint64_t)num_microbatches * (uint64_t)num_tensor_shards;
    out[0] = fwd_us / (uint64_t)num_stages / div;
    out[1] = bwd_us / (uint64_t)num_stages / div;
    out[2] = (uint64_t)((unsigned)seq_len * (unsigned)embed_dim * (batch / (uint64_t)num_microbatches));
    out[3] = model_size / ((uint64_t)num_stages * (uint64_t)num_tensor_shards);
    out[4] = (uint64_t)(num_microbatches + num_stages - 1) * (out[0] + out[1]);
    out[5] = out[2] / (uint64_t)num_tensor_shards;
}

/* hybrid_3d_moe.cpp:330-352 — DP + PP + expert parallel (E expert shards, top_k = 2 routing): out[0], out[1] =
 * fwd / bwd us per micro-batch (time / stages / micro-batches), out[2] = pipe message elements
 * (seq_len * samples per micro-batch * embed_dim, 64-bit), out[3] = DP allreduce elements = non-expert share
 * + expert share, out[4] = compute critical path, out[5] = all-to-all elements PER PEER = tokens per
 * micro-batch * top_k * embed_dim / E, out[6] = non-expert elements per stage (the expert group's allreduce),
 * out[7] = all-to-alls per micro-batch and direction = 2 * layers / stages. */
void dlnb_oracle_hybrid3d_moe_sizes(uint64_t fwd_us, uint64_t bwd_us, uint64_t model_size, uint64_t non_expert_size,
                                    uint64_t batch, uint64_t seq_len, uint64_t embed_dim, uint64_t num_layers, int num_stages,
                                    int num_microbatches, int num_expert_shards, uint64_t *out) {
    const uint64_t S = (uint64_t)num_stages, M = (uint64_t)num_microbatches, E = (uint64_t)num_expert_shards;
    const uint64_t samples = batch / M, top_k = 2;
    out[0] = fwd_us / S / M;
    out[1] = bwd_us / S / M;
    out[2] = seq_len * samples * embed_dim;
    out[6] = non_expert_size / S;
    out[3] = out[6] + (model_size - non_expert_size) / S / E;
    out[4] = (M + S - 1) * (out[0] + out[1]);
    out[5] = samples * seq_len * top_k * embed_dim / E;
    out[7] = 2 * (num_layers / S);
}

/* utils.hpp:145-159 — mean and population standard deviation, float arithmetic. */
void dlnb_oracle_msg_stats(const uint64_t *sizes, int n, unsigned mult, float *avg_out, float *std_out) {
    float avg = 0.0f;
    for (int i = 0; i < n; ++i) avg += sizes[i] * mult;
    avg /= n;
    float sd = 0.0f;
    for (int i = 0; i < n; ++i) { float d = sizes[i] * mult - avg; sd += d * d; }
    *avg_out = avg;
    *std_out = sqrtf(sd / n);
}

/* utils.hpp:167-179,200-269 — positional parse of model_stats/<name>.txt.
 * out[12] = forwardFlops, backwardFlops, modelSize, nonExpertModelSize, avgForwardTime,
 * avgBackwardTime, batchSize, ffn_avgForwardTime, ffn_avgBackwardTime, experts,
 * sequenceLength, embeddedDim. Returns 0 on success. */
int dlnb_oracle_parse_model_stats(const char *path, uint64_t *out) {
    FILE *f = fopen(path, "r");
    if (!f) return 1;
    char line[512];
    for (int k = 0; k < 12; ++k) {
        if (!fgets(line, sizeof(line), f)) { fclose(f); return 2; }
        const char *c = strchr(line, ':');
        if (!c) { fclose(f); return 3; }
        c++;
        while (*c == ' ' || *c == '\t') c++;
        if (k == 4 || k == 5) out[k] = (uint64_t)strtod(c, NULL); /* stod then truncation (utils.hpp:228,232) */
        else out[k] = strtoull(c, NULL, 10);                       /* stoull: integer prefix */
    }
    fclose(f);
    return 0;
}

/* ---- buffer patterns (SURVEY §8c pins), restated from the kernel-side definition ---- */
static inline uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
float dlnb_oracle_pattern_value(int kind, uint64_t seed, int r, uint64_t i) {
    if (kind == 1) {
        const uint64_t h = i * 2654435761ull + (uint64_t)r * 40503ull;
        return (float)((int)(h % 17ull) - 8);
    } else if (kind == 2) {
        uint32_t h = mix32((uint32_t)i ^ mix32((uint32_t)(i >> 32) + 0x9e3779b9u * (uint32_t)(r + 1) + (uint32_t)seed));
        return (float)((int)(h >> 8) - (1 << 23)) * (1.0f / 8388608.0f);
    }
    return 0.0f;
}
void dlnb_oracle_fill_pattern(void *p, size_t count, int dtype, int kind, uint64_t seed, int r) {
    for (size_t i = 0; i < count; ++i) store_elem(p, i, dtype, dlnb_oracle_pattern_value(kind, seed, r, i));
}
uint64_t dlnb_oracle_checksum(const void *p, size_t bytes) {
    const size_t nwords = bytes / 4;
    uint64_t s = 0;
    for (size_t i = 0; i < nwords; ++i) s += (uint64_t)mix32(((const uint32_t *)p)[i] + (uint32_t)i * 0x9e3779b9u);
    if (bytes % 4) {
        uint32_t tail = 0;
        memcpy(&tail, (const char *)p + nwords * 4, bytes % 4);
        s += (uint64_t)mix32(tail + (uint32_t)nwords * 0x9e3779b9u);
    }
    return s;
}
