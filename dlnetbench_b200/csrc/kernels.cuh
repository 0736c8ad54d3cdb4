// kernels.cuh — sm_100a device code of the DLNetBench data-parallel hot path.
//
// Kernels here replace the library calls behind the reference's communicator
// virtuals (cpp/proxy_classes.hpp:66-101 MPI, :149-209 NCCL):
//   allreduce one-shot / fused two-shot / in-switch (multimem)  <- MPI_Iallreduce / ncclAllReduce (SUM)
//   allgather (push)                                            <- MPI_(I)allgather / ncclAllGather
//   reduce-scatter (pull)                                       <- MPI_Reduce_scatter_block / ncclReduceScatter
//   all-to-all (push)                                           <- MPI_Alltoall / grouped ncclSend + ncclRecv
// (spin kernel, send/recv and the test helpers are in aux_kernels.cuh)
//
// All of them address the peers' buffers directly (VMM-mapped peer HBM over
// NVLink 5 / NVSwitch): 128-bit coalesced loads/stores, fp32 accumulation in fixed
// rank order 0..n-1 (bit-identical on every rank and to the CPU oracle), and a
// flag barrier per CTA index with system-scope release/acquire.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dlnb {

constexpr int MAXR = 8;            // ranks of one NVSwitch box
constexpr int MAX_FLAG_CTAS = 512; // flag rows per communicator
constexpr int FLAG_ROW = 8;        // uint32 flags per row (one per source rank)

// Per-communicator control block, local to each (virtual) rank: one barrier generation per
// CTA row (monotonic, wraps). Row b advances by 2 in every launch whose grid contains CTA b;
// all ranks of a communicator launch the same grid for the same call, so row b holds the same
// value on every rank. Kept in device memory (not a kernel argument) so a captured launch
// replays correctly.
struct CommCtl {
    uint32_t epoch[MAX_FLAG_CTAS];
};

// What one rank sees of a collective: its own and its peers' buffers and flags.
struct RankView {
    int rank;                  // my index in the group
    int nranks;
    const void *send[MAXR];    // send[p]: rank p's send buffer (my mapping of it)
    void *recv[MAXR];          // recv[p]: rank p's recv buffer
    uint32_t *flags[MAXR];     // flags[p]: rank p's flag rows for this communicator
    const void *mc_send;       // multicast address of the send buffer (all ranks' copies), or null
    void *mc_recv;             // multicast address of the recv buffer, or null
    CommCtl *ctl;              // my control block
    uint32_t *host_error;      // mapped pinned word, set by the watchdog (may be null)
    unsigned long long timeout_ns;
    int strict;                // 1: every barrier uses release/acquire flag traffic (see cta_barrier)
    int shift_overlap;         // shifted mover: runs overlap by 1, 2 or 8 vectors (16 B / one 32 B sector / one 128 B line)
};

struct EmuViews {
    RankView v[MAXR];
};

// ----------------------------------------------------------------------------
// memory access flavours
// ----------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Streaming 128-bit load of data that may live in peer HBM and may have been
// written by a peer before the barrier we just passed: system-scope relaxed, so it
// is never served from a stale L1 line.
__device__ __forceinline__ uint4 ld_v4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
// Read-only streaming load (buffer not written by anyone during the kernel).
__device__ __forceinline__ uint4 ld_v4_ro(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_v4(uint4 *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// ----------------------------------------------------------------------------
// element arithmetic: widen to fp32, add in rank order, round once (RNE)
// ----------------------------------------------------------------------------
template <typename T> struct Elem;
template <> struct Elem<float> {
    static constexpr int PER_VEC = 4;
    static constexpr float HALF_ULP = 5.9604645e-8f; // 2^-24: one rounding to T, relative
    static __device__ __forceinline__ float to_f(float x) { return x; }
    static __device__ __forceinline__ float from_f(float x) { return x; }
};
template <> struct Elem<__half> {
    static constexpr int PER_VEC = 8;
    static constexpr float HALF_ULP = 4.8828125e-4f; // 2^-11
    static __device__ __forceinline__ float to_f(__half x) { return __half2float(x); }
    static __device__ __forceinline__ __half from_f(float x) { return __float2half_rn(x); }
};
template <> struct Elem<__nv_bfloat16> {
    static constexpr int PER_VEC = 8;
    static constexpr float HALF_ULP = 3.90625e-3f; // 2^-8
    static __device__ __forceinline__ float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }
    static __device__ __forceinline__ __nv_bfloat16 from_f(float x) { return __float2bfloat16_rn(x); }
};

template <typename T> struct Acc;
template <> struct Acc<float> {
    float a[4];
    __device__ __forceinline__ void set(const uint4 &v) {
        a[0] = __uint_as_float(v.x); a[1] = __uint_as_float(v.y);
        a[2] = __uint_as_float(v.z); a[3] = __uint_as_float(v.w);
    }
    __device__ __forceinline__ void add(const uint4 &v) {
        a[0] += __uint_as_float(v.x); a[1] += __uint_as_float(v.y);
        a[2] += __uint_as_float(v.z); a[3] += __uint_as_float(v.w);
    }
    __device__ __forceinline__ uint4 pack() const {
        return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]),
                          __float_as_uint(a[3]));
    }
};
template <> struct Acc<__half> {
    float a[8];
    static __device__ __forceinline__ float2 cvt(uint32_t w) {
        return __half22float2(*reinterpret_cast<const __half2 *>(&w));
    }
    __device__ __forceinline__ void set(const uint4 &v) {
        float2 f;
        f = cvt(v.x); a[0] = f.x; a[1] = f.y;
        f = cvt(v.y); a[2] = f.x; a[3] = f.y;
        f = cvt(v.z); a[4] = f.x; a[5] = f.y;
        f = cvt(v.w); a[6] = f.x; a[7] = f.y;
    }
    __device__ __forceinline__ void add(const uint4 &v) {
        float2 f;
        f = cvt(v.x); a[0] += f.x; a[1] += f.y;
        f = cvt(v.y); a[2] += f.x; a[3] += f.y;
        f = cvt(v.z); a[4] += f.x; a[5] += f.y;
        f = cvt(v.w); a[6] += f.x; a[7] += f.y;
    }
    static __device__ __forceinline__ uint32_t pk(float x, float y) {
        __half2 h = __floats2half2_rn(x, y);
        return *reinterpret_cast<uint32_t *>(&h);
    }
    __device__ __forceinline__ uint4 pack() const {
        return make_uint4(pk(a[0], a[1]), pk(a[2], a[3]), pk(a[4], a[5]), pk(a[6], a[7]));
    }
};
template <> struct Acc<__nv_bfloat16> {
    float a[8];
    // bf16 -> fp32 is a 16-bit shift
    __device__ __forceinline__ void set(const uint4 &v) {
        a[0] = __uint_as_float(v.x << 16); a[1] = __uint_as_float(v.x & 0xffff0000u);
        a[2] = __uint_as_float(v.y << 16); a[3] = __uint_as_float(v.y & 0xffff0000u);
        a[4] = __uint_as_float(v.z << 16); a[5] = __uint_as_float(v.z & 0xffff0000u);
        a[6] = __uint_as_float(v.w << 16); a[7] = __uint_as_float(v.w & 0xffff0000u);
    }
    __device__ __forceinline__ void add(const uint4 &v) {
        a[0] += __uint_as_float(v.x << 16); a[1] += __uint_as_float(v.x & 0xffff0000u);
        a[2] += __uint_as_float(v.y << 16); a[3] += __uint_as_float(v.y & 0xffff0000u);
        a[4] += __uint_as_float(v.z << 16); a[5] += __uint_as_float(v.z & 0xffff0000u);
        a[6] += __uint_as_float(v.w << 16); a[7] += __uint_as_float(v.w & 0xffff0000u);
    }
    static __device__ __forceinline__ uint32_t pk(float x, float y) {
        __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
        return *reinterpret_cast<uint32_t *>(&h);
    }
    __device__ __forceinline__ uint4 pack() const {
        return make_uint4(pk(a[0], a[1]), pk(a[2], a[3]), pk(a[4], a[5]), pk(a[6], a[7]));
    }
};

// ----------------------------------------------------------------------------
// cross-GPU barrier between the CTAs of equal index on every rank.
//   `gen` is the generation to reach. Flags are monotonic (never reset) and compared with
//   wrap-safe >=, so a rank that runs ahead by one barrier is fine.
//   After a __syncthreads, thread p (< n) writes flag[peer p][cta][me] = gen and polls
//   flag[me][cta][p] >= gen.
//   PUBLISH = true  : this CTA wrote into peer memory before the barrier. The flag is stored
//                     with release and read with acquire at system scope, so those writes are
//                     visible to whoever passes the barrier (costs a wait for the write acks).
//   PUBLISH = false : only reads happened (or nothing yet): relaxed flag traffic, no fence.
//                     What peers read after it was produced by kernels that had completed
//                     before the peer's kernel could write its flag.
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cta_epoch(const RankView &v, int cta) {
    return v.nranks > 1 ? *(volatile uint32_t *)&v.ctl->epoch[cta] : 0u;
}

template <bool PUBLISH>
__device__ __forceinline__ void cta_barrier(const RankView &v, uint32_t gen, int cta) {
    if (v.nranks <= 1) return;
    __syncthreads();
    if (threadIdx.x < (unsigned)v.nranks) {
        const int p = threadIdx.x;
        uint32_t *theirs = v.flags[p] + (size_t)cta * FLAG_ROW + v.rank;
        const uint32_t *mine = v.flags[v.rank] + (size_t)cta * FLAG_ROW + p;
        const bool fenced = PUBLISH || v.strict;
        if (fenced) st_release_sys(theirs, gen);
        else st_relaxed_sys(theirs, gen);
        unsigned long long t0 = 0;
        unsigned polls = 0;
        while ((int32_t)((fenced ? ld_acquire_sys(mine) : ld_relaxed_sys(mine)) - gen) < 0) {
            if ((++polls & 0x3ff) == 0) {
                unsigned long long now = globaltimer_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > v.timeout_ns) {
                    if (v.host_error) *(volatile uint32_t *)v.host_error = 1;
                    __threadfence_system();
                    break;
                }
            }
        }
    }
    __syncthreads();
}

// Row `cta` has passed both barriers of this launch.
__device__ __forceinline__ void cta_finish(const RankView &v, uint32_t new_epoch, int cta) {
    if (v.nranks > 1 && threadIdx.x == 0) v.ctl->epoch[cta] = new_epoch;
}

// ----------------------------------------------------------------------------
// The data mover: out[i] = sum_{p<nsrc} src[p][i] (fp32, p ascending), stored to
// every dst[d]. All pointers 16-byte aligned; `nvec` 16-byte vectors. Each thread
// issues U*NSRC independent 128-bit loads before the first use.
// ----------------------------------------------------------------------------
template <typename T, int NSRC, int NDST, int U, bool RO>
__device__ __forceinline__ void reduce_store_vec(const uint4 *const *src, uint4 *const *dst,
                                                 size_t nvec, size_t tid, size_t nthr) {
    const uint4 *s[NSRC];
    uint4 *d[NDST];
#pragma unroll
    for (int p = 0; p < NSRC; ++p) s[p] = src[p];
#pragma unroll
    for (int q = 0; q < NDST; ++q) d[q] = dst[q];

    // Tiled: a CTA streams contiguous tiles of U*blockDim.x vectors (tile index strides by the
    // grid), a warp touches 512 contiguous bytes per load instruction. Tiles are laid out from the
    // 128-byte line below the buffers of the side that crosses NVLink (the sources of a pull, the
    // destinations of a push), so a warp's 512 bytes are 4 whole lines even when the buffer itself is
    // only 16-byte aligned (bucket slices and shards rarely start on a line): `a0` vectors of the first
    // tile lie before the data and are skipped (unsigned wrap makes i >= nvec for them).
    const size_t bdim = blockDim.x, tile = (size_t)U * bdim;
    const size_t cta = tid / bdim, nctas = nthr / bdim;
    const size_t a0 = (size_t)((((NSRC == 1) ? (uintptr_t)d[0] : (uintptr_t)s[0]) >> 4) & 7u);
    const size_t nvirt = nvec + a0;
    for (size_t t = cta; t * tile < nvirt; t += nctas) {
        const size_t base = t * tile + threadIdx.x - a0;
        uint4 val[U][NSRC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < nvec) {
#pragma unroll
                for (int p = 0; p < NSRC; ++p) val[u][p] = RO ? ld_v4_ro(s[p] + i) : ld_v4(s[p] + i);
            }
        }
        // Programmatic dependent launch (world-1 copies only, see coll_kernel): once this CTA's LAST tile is
        // loaded, the next kernel of the stream may start filling the machine while this one drains. A no-op
        // for kernels that were not launched with the attribute.
        if (RO && (t + nctas) * tile >= nvirt) asm volatile("griddepcontrol.launch_dependents;");
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < nvec) {
                uint4 out;
                if (NSRC == 1) {
                    out = val[u][0];
                } else {
                    Acc<T> acc;
                    acc.set(val[u][0]);
#pragma unroll
                    for (int p = 1; p < NSRC; ++p) acc.add(val[u][p]);
                    out = acc.pack();
                }
#pragma unroll
                for (int q = 0; q < NDST; ++q) st_v4(d[q] + i, out);
            }
        }
    }
}

// Element-wise version for heads, tails and buffers with non-uniform alignment.
template <typename T>
__device__ __forceinline__ void reduce_store_elem(const T *const *src, int nsrc, T *const *dst,
                                                  int ndst, size_t n, size_t tid, size_t nthr) {
    for (size_t i = tid; i < n; i += nthr) {
        T out;
        if (nsrc == 1) {
            out = *(const volatile T *)(src[0] + i);
        } else {
            float acc = Elem<T>::to_f(*(const volatile T *)(src[0] + i));
            for (int p = 1; p < nsrc; ++p) acc += Elem<T>::to_f(*(const volatile T *)(src[p] + i));
            out = Elem<T>::from_f(acc);
        }
        for (int q = 0; q < ndst; ++q) dst[q][i] = out;
    }
}

template <int NR> struct Unroll {
    // U*NR independent 128-bit loads in flight per thread. The multi-rank kernels are held to 96
    // registers (see coll_kernel) so that a compute CTA can share the SM with a collective CTA:
    // at most 12 data vectors (48 registers) in flight next to the peer pointers.
    static constexpr int U = NR == 1 ? 4 : NR == 2 ? 6 : NR == 3 ? 4 : NR == 4 ? 3 : 2;
    // allgather: one local source, NR destinations (pointers, not data, fill the registers)
    static constexpr int U_AG = NR <= 4 ? 4 : 2;
    // shifted path: same loads in flight as the aligned path plus the shuffled neighbour
    static constexpr int U_SH = NR == 1 ? 4 : NR == 2 ? 4 : NR <= 6 ? 2 : 1;
    // ... of the bodies that also hold NR destination pointers (fused two-shot allreduce)
    static constexpr int U_SH_PUSH = NR <= 6 ? U_SH : 1;
    static constexpr int U_PUSH = NR <= 6 ? U : 1; // 7-8 ranks: 8 loads in flight x 65,536 threads = 8 MB per GPU, three times the link's bandwidth-delay product
};

// ----------------------------------------------------------------------------
// Shifted data mover for chunks whose sources are NOT 16-byte aligned relative to their
// (aligned) destinations — odd shard sizes put rank r's chunk at r*count*sizeof(T), e.g. the
// 110,240,167-element fp16 shards of llama3_70b. All sources share the misalignment `m`
// (0 < m < 16, a multiple of sizeof(T)); `src[p]` are the ALIGNED addresses just below the
// data. Output vector j = bytes [m, m+16) of the pair (A[j], A[j+1]) where A[j] is the
// rank-order fp32 sum of the sources' aligned vectors (the sum is element-wise, so summing
// in the aligned frame and shifting afterwards gives the same bits). A thread gets A[j+1]
// from its right-hand lane by shuffle; lane 31 loads it. Loads A[0..nvec], stores nvec vectors.
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint4 shift_pair(const uint4 &a, const uint4 &b, unsigned m) {
    const unsigned mw = m >> 2, sh = (m & 3u) * 8u;
    uint32_t w0, w1, w2, w3, w4;
    switch (mw) {
    case 0: w0 = a.x; w1 = a.y; w2 = a.z; w3 = a.w; w4 = b.x; break;
    case 1: w0 = a.y; w1 = a.z; w2 = a.w; w3 = b.x; w4 = b.y; break;
    case 2: w0 = a.z; w1 = a.w; w2 = b.x; w3 = b.y; w4 = b.z; break;
    default: w0 = a.w; w1 = b.x; w2 = b.y; w3 = b.z; w4 = b.w; break;
    }
    return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                      __funnelshift_r(w3, w4, sh));
}

template <typename T, int NSRC, int NDST, int U>
__device__ __forceinline__ void reduce_store_vec_shifted(const uint4 *const *src, uint4 *const *dst, size_t nvec,
                                                         unsigned m, size_t tid, size_t nthr, unsigned overlap) {
    const uint4 *s[NSRC];
    uint4 *d[NDST];
#pragma unroll
    for (int p = 0; p < NSRC; ++p) s[p] = src[p];
#pragma unroll
    for (int q = 0; q < NDST; ++q) d[q] = dst[q];
    // out[j] = shift(A[j], A[j+1]) over the aligned input vectors A[0..nvec]. A warp owns a WINDOW of `rows` rows of 32
    // consecutive vectors and walks it U rows at a time (U*NSRC loads in flight per thread); A[j+1] comes from the right-hand
    // lane by shuffle, for lane 31 from lane 0 of the next row — across steps the last row of a step is CARRIED (one summed
    // vector per thread) and finished when the next step's first row is there. Consecutive windows overlap by `overlap`
    // vectors (the last outputs of a window need the first vector of the next): nothing is reloaded by a divergent lane, no
    // raw data is held across steps, and with overlap 8 every window starts on a 128-byte line of the side that crosses
    // NVLink (a 512-byte warp access then touches 4 lines, not 5). The redundant loads are overlap / (32*rows): 1.6 % for
    // the 16-row windows of a large message, whatever U is (a window of only 32*U vectors, as in the first version of this
    // path, cost 25 % at 7-8 ranks where U = 1). Windows shrink for small messages so that all warps get work.
    const unsigned lane = threadIdx.x & 31u;
    const long long wpc = blockDim.x >> 5, warp = threadIdx.x >> 5;
    const long long cta = (long long)(tid / blockDim.x), nctas = (long long)(nthr / blockDim.x);
    const long long nv = (long long)nvec;
    long long rows = (nv / 32 + nctas * wpc - 1) / (nctas * wpc); // rows per warp if every warp took one window
    rows = (min(max(rows, 1ll), 16ll) + U - 1) / U * U;           // 1..16, a multiple of U
    const long long run = 32 * rows - (long long)overlap;         // outputs per window
    // ... of the side that crosses NVLink: the sources of a pull, the destinations of a push (one local source)
    const uintptr_t ref = NSRC == 1 ? (uintptr_t)d[0] : (uintptr_t)s[0];
    const long long a0 = overlap > 1 ? (long long)((ref >> 4) & (overlap - 1)) : 0;
    for (long long r = cta * wpc + warp; r * run - a0 < nv; r += nctas * wpc) { // warp-uniform trip count
        const long long base = r * run - a0;
        uint4 carry = make_uint4(0, 0, 0, 0);
        for (long long row0 = 0; row0 < rows && base + row0 * 32 <= nv; row0 += U) { // warp-uniform
            uint4 val[U][NSRC];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long long i = min(max(base + (row0 + u) * 32 + (long long)lane, 0ll), nv); // A[nvec] is always readable
#pragma unroll
                for (int p = 0; p < NSRC; ++p) val[u][p] = ld_v4(s[p] + i);
            }
            uint4 cur[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (NSRC == 1) {
                    cur[u] = val[u][0];
                } else {
                    Acc<T> a;
                    a.set(val[u][0]);
#pragma unroll
                    for (int p = 1; p < NSRC; ++p) a.add(val[u][p]);
                    cur[u] = a.pack();
                }
            }
            // rows row0-1 (carried), row0 .. row0+U-2: each is complete now that its successor row is summed
#pragma unroll
            for (int u = -1; u < U - 1; ++u) {
                const uint4 mine = u < 0 ? carry : cur[u < 0 ? 0 : u];
                const uint4 succ = cur[u + 1];
                // lane l takes `mine` of lane l+1; lane 31 takes `succ` of lane 0
                const uint4 give = lane == 0 ? succ : mine;
                uint4 nxt;
                nxt.x = __shfl_sync(0xffffffffu, give.x, (lane + 1) & 31);
                nxt.y = __shfl_sync(0xffffffffu, give.y, (lane + 1) & 31);
                nxt.z = __shfl_sync(0xffffffffu, give.z, (lane + 1) & 31);
                nxt.w = __shfl_sync(0xffffffffu, give.w, (lane + 1) & 31);
                const long long slot = (row0 + u) * 32 + (long long)lane, j = base + slot;
                if (row0 + u >= 0 && slot < run && j >= 0 && j < nv) {
                    const uint4 out = shift_pair(mine, nxt, m);
#pragma unroll
                    for (int q = 0; q < NDST; ++q) st_v4(d[q] + j, out);
                }
            }
            carry = cur[U - 1];
            // the window's (or the data's) last row has no successor row: its lanes 0..30 finish among themselves
            const bool last_step = row0 + U >= rows || base + (row0 + U) * 32 > nv;
            if (last_step) {
                uint4 nxt;
                nxt.x = __shfl_sync(0xffffffffu, carry.x, (lane + 1) & 31);
                nxt.y = __shfl_sync(0xffffffffu, carry.y, (lane + 1) & 31);
                nxt.z = __shfl_sync(0xffffffffu, carry.z, (lane + 1) & 31);
                nxt.w = __shfl_sync(0xffffffffu, carry.w, (lane + 1) & 31);
                const long long slot = (row0 + U - 1) * 32 + (long long)lane, j = base + slot;
                if (lane < 31u && slot < run && j >= 0 && j < nv) {
                    const uint4 out = shift_pair(carry, nxt, m);
#pragma unroll
                    for (int q = 0; q < NDST; ++q) st_v4(d[q] + j, out);
                }
            }
        }
    }
}

// ----------------------------------------------------------------------------
// out[i] = sum_p src[p][i] for i in [0, count), stored to every dst[q] — any alignment.
// Sources must share one alignment class and destinations another (true for every
// collective here: all ranks use the same offsets into equally aligned buffers); otherwise
// the element-wise path does the whole range. Head elements are peeled until the
// destinations are 16-byte aligned, the bulk goes through the aligned or the shifted
// 128-bit path, the last few elements element-wise.
// ----------------------------------------------------------------------------
template <typename T, int NSRC, int NDST, int U, int USH>
__device__ __forceinline__ void move_any(const T *const *src, T *const *dst, size_t count, int cta, int nctas, unsigned overlap = 8) {
    const size_t tid = (size_t)cta * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)nctas * blockDim.x;
    constexpr int PV = Elem<T>::PER_VEC;
    const unsigned ms = (unsigned)((uintptr_t)src[0] & 15u), md = (unsigned)((uintptr_t)dst[0] & 15u);
    bool uniform = true;
#pragma unroll
    for (int p = 1; p < NSRC; ++p) uniform = uniform && ((uintptr_t)src[p] & 15u) == ms;
#pragma unroll
    for (int q = 1; q < NDST; ++q) uniform = uniform && ((uintptr_t)dst[q] & 15u) == md;
    if (!uniform) {
        reduce_store_elem<T>(src, NSRC, dst, NDST, count, tid, nthr);
        return;
    }
    const size_t head = min((size_t)(((16u - md) & 15u) / sizeof(T)), count);
    const unsigned m = (unsigned)((ms + head * sizeof(T)) & 15u);
    const size_t rest = count - head;
    size_t nvec = rest / PV;
    if (m != 0 && nvec > 0) nvec -= 1; // the shifted path reads one aligned vector more than it writes
    const size_t tail0 = head + nvec * PV;
    if (nvec > 0) {
        const uint4 *vs[NSRC];
        uint4 *vd[NDST];
#pragma unroll
        for (int p = 0; p < NSRC; ++p) vs[p] = (const uint4 *)((const char *)(src[p] + head) - m);
#pragma unroll
        for (int q = 0; q < NDST; ++q) vd[q] = (uint4 *)(dst[q] + head);
        if (m == 0) reduce_store_vec<T, NSRC, NDST, U, NSRC == 1 && NDST == 1>(vs, vd, nvec, tid, nthr);
        else reduce_store_vec_shifted<T, NSRC, NDST, USH>(vs, vd, nvec, m, tid, nthr, overlap == 1 || overlap == 2 ? overlap : 8u);
    }
    // head and tail: a handful of elements, one CTA
    if (cta == 0) {
        if (head) reduce_store_elem<T>(src, NSRC, dst, NDST, head, threadIdx.x, blockDim.x);
        if (tail0 < count) {
            const T *ts[NSRC];
            T *td[NDST];
#pragma unroll
            for (int p = 0; p < NSRC; ++p) ts[p] = src[p] + tail0;
#pragma unroll
            for (int q = 0; q < NDST; ++q) td[q] = dst[q] + tail0;
            reduce_store_elem<T>(ts, NSRC, td, NDST, count - tail0, threadIdx.x, blockDim.x);
        }
    }
}

// ----------------------------------------------------------------------------
// Allreduce, one-shot: every rank reads all n send buffers and writes its own recv.
// Latency path: one barrier before (peers' inputs final), one after (peers done
// reading mine). Wire: (n-1)*S per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allreduce_oneshot_body(const RankView &v, size_t count, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const T *src[NR];
#pragma unroll
    for (int p = 0; p < NR; ++p) src[p] = (const T *)v.send[p];
    T *dst[1] = {(T *)v.recv[v.rank]};
    move_any<T, NR, 1, Unroll<NR>::U, Unroll<NR>::U_SH>(src, dst, count, cta, nctas, v.shift_overlap);

    cta_barrier<false>(v, e + 2, cta); // only reads happened
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Allreduce, two-shot fused in one pass: rank r owns slice r (slice boundaries are whole
// 16-byte vectors of the buffer), pulls that slice from every rank, sums, and pushes the
// result into every rank's recv. Wire: 2(n-1)/n*S per direction, reads and writes overlap.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allreduce_twoshot_body(const RankView &v, size_t count, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    constexpr int PV = Elem<T>::PER_VEC;
    // slices in units of PV elements; the last rank also takes the sub-vector remainder
    const size_t nvec = count / PV;
    const size_t per = ((nvec + NR - 1) / NR + 7) / 8 * 8; // slices of whole 128-byte lines
    const size_t e0 = min(per * (size_t)v.rank, nvec) * PV;
    const size_t e1 = v.rank == NR - 1 ? count : min(per * (size_t)(v.rank + 1), nvec) * PV;
    const T *src[NR];
    T *dst[NR];
#pragma unroll
    for (int p = 0; p < NR; ++p) {
        src[p] = (const T *)v.send[p] + e0;
        dst[p] = (T *)v.recv[p] + e0;
    }
    move_any<T, NR, NR, Unroll<NR>::U_PUSH, Unroll<NR>::U_SH_PUSH>(src, dst, e1 - e0, cta, nctas, v.shift_overlap);

    cta_barrier<true>(v, e + 2, cta); // publishes the pushed slices
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Allgather, push: read my shard once from local HBM, store it into slot `rank`
// of every rank's recv (n-1 of them over NVLink). Wire: (n-1)*shard per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allgather_body(const RankView &v, size_t sendcount, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const size_t off = sendcount * (size_t)v.rank; // elements into every recv
    const T *src[1] = {(const T *)v.send[v.rank]};
    T *dst[NR];
#pragma unroll
    for (int p = 0; p < NR; ++p) dst[p] = (T *)v.recv[p] + off;
    move_any<T, 1, NR, Unroll<NR>::U_AG, Unroll<NR>::U_SH>(src, dst, sendcount, cta, nctas, v.shift_overlap);

    cta_barrier<true>(v, e + 2, cta); // publishes the pushed shard
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Reduce-scatter, pull: rank r reads chunk r (recvcount elements at offset
// r*recvcount) of every rank's send, sums in rank order, writes its own recv.
// Wire: (n-1)*recvcount per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void reduce_scatter_body(const RankView &v, size_t recvcount, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const size_t off = recvcount * (size_t)v.rank;
    const T *src[NR];
#pragma unroll
    for (int p = 0; p < NR; ++p) src[p] = (const T *)v.send[p] + off;
    T *dst[1] = {(T *)v.recv[v.rank]};
    move_any<T, NR, 1, Unroll<NR>::U, Unroll<NR>::U_SH>(src, dst, recvcount, cta, nctas, v.shift_overlap);

    cta_barrier<false>(v, e + 2, cta); // only reads happened
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// All-to-all, push: block p of my send buffer (sendcount elements at p*sendcount) goes into
// slot `rank` of rank p's recv — ProxyCommunicator::Alltoall (cpp/proxy_classes.hpp:74-77 MPI_Alltoall,
// :160-182 grouped ncclSend/ncclRecv), the expert-parallel token exchange of hybrid_3d_moe.cpp:161-165.
// Peers are visited starting from my right-hand neighbour so the n ranks do not all hit the same
// destination at once. Wire: (n-1)*sendcount per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void alltoall_body(const RankView &v, size_t sendcount, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    for (int k = 0; k < NR; ++k) {
        const int p = (v.rank + 1 + k) % NR;
        const T *src[1] = {(const T *)v.send[v.rank] + (size_t)p * sendcount};
        T *dst[1] = {(T *)v.recv[p] + (size_t)v.rank * sendcount};
        move_any<T, 1, 1, 4, 4>(src, dst, sendcount, cta, nctas, v.shift_overlap);
    }

    cta_barrier<true>(v, e + 2, cta); // publishes the pushed blocks
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Allreduce through NVSwitch multicast ("NVLS"): rank r owns slice r as in the two-shot
// kernel, but ONE multimem.ld_reduce fetches the slice from all ranks and has the switch add
// them (fp32 accumulation), and ONE multimem.st writes the result into every rank's recv.
// Wire per GPU per direction: ~(1 + 1/n)*S instead of 2(n-1)/n*S — the only way past the
// per-direction link rate. The summation order inside the switch is not specified: results
// are bit-exact for integer-valued data and within 2(n-1)eps otherwise, and identical on all
// ranks (one owner computes each slice).
// ----------------------------------------------------------------------------
template <typename T> struct Multimem;
template <> struct Multimem<float> {
    static __device__ __forceinline__ uint4 ld_reduce(const uint4 *p) {
        uint4 r;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(uint4 *p, const uint4 &v) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                     "r"(v.z), "r"(v.w) : "memory");
    }
};
template <> struct Multimem<__half> {
    static __device__ __forceinline__ uint4 ld_reduce(const uint4 *p) {
        uint4 r;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(uint4 *p, const uint4 &v) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                     "r"(v.z), "r"(v.w) : "memory");
    }
};
template <> struct Multimem<__nv_bfloat16> {
    static __device__ __forceinline__ uint4 ld_reduce(const uint4 *p) {
        uint4 r;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(uint4 *p, const uint4 &v) {
        asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                     "r"(v.z), "r"(v.w) : "memory");
    }
};

template <typename T, int NR>
__device__ __forceinline__ void allreduce_multimem_body(const RankView &v, size_t count, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    constexpr int PV = Elem<T>::PER_VEC;
    constexpr int U = 8;
    const size_t nvec = count / PV;
    const size_t per = ((nvec + NR - 1) / NR + 7) / 8 * 8; // slices of whole 128-byte lines
    const size_t v0 = min(per * (size_t)v.rank, nvec);
    const size_t n = min(v0 + per, nvec) - v0;
    const uint4 *src = (const uint4 *)v.mc_send + v0;
    uint4 *dst = (uint4 *)v.mc_recv + v0;
    const size_t bdim = blockDim.x, tile = (size_t)U * bdim;
    for (size_t t = cta; t * tile < n; t += nctas) {
        const size_t base = t * tile + threadIdx.x;
        uint4 val[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < n) val[u] = Multimem<T>::ld_reduce(src + i);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < n) Multimem<T>::st(dst + i, val[u]);
        }
    }
    // sub-vector tail (< 16 bytes): the last rank reduces it over the unicast mappings and pushes
    // it into every recv (one reader-then-writer per element, like the slices: in-place safe)
    const size_t done = nvec * PV;
    if (done < count && cta == 0 && v.rank == NR - 1) {
        const T *ts[NR];
        T *td[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) {
            ts[p] = (const T *)v.send[p] + done;
            td[p] = (T *)v.recv[p] + done;
        }
        reduce_store_elem<T>(ts, NR, td, NR, count - done, threadIdx.x, blockDim.x);
    }

    cta_barrier<true>(v, e + 2, cta); // publishes the multicast stores
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Reduce-scatter through the switch: rank r fetches ONLY its chunk with multimem.ld_reduce (the
// switch pulls the chunk from all n ranks and adds) and stores it locally. Receive side: one
// chunk instead of n-1; send side: n chunks leave every GPU. Needs 16-byte aligned chunks.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void reduce_scatter_multimem_body(const RankView &v, size_t recvcount, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    constexpr int PV = Elem<T>::PER_VEC;
    constexpr int U = 8;
    const size_t off = recvcount * (size_t)v.rank;
    const size_t n = recvcount / PV;
    const uint4 *src = (const uint4 *)((const T *)v.mc_send + off);
    uint4 *dst = (uint4 *)v.recv[v.rank];
    const size_t bdim = blockDim.x, tile = (size_t)U * bdim;
    for (size_t t = cta; t * tile < n; t += nctas) {
        const size_t base = t * tile + threadIdx.x;
        uint4 val[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < n) val[u] = Multimem<T>::ld_reduce(src + i);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < n) st_v4(dst + i, val[u]);
        }
    }
    const size_t done = n * PV;
    if (done < recvcount && cta == 0) {
        const T *ts[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) ts[p] = (const T *)v.send[p] + off + done;
        T *td[1] = {(T *)v.recv[v.rank] + done};
        reduce_store_elem<T>(ts, NR, td, 1, recvcount - done, threadIdx.x, blockDim.x);
    }

    cta_barrier<false>(v, e + 2, cta); // only reads of peer memory happened
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Allgather through the switch: read my shard once, ONE multimem.st per vector lands it in slot
// `rank` of every rank's recv (the switch replicates). Send side: one shard instead of n-1.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allgather_multimem_body(const RankView &v, size_t sendcount, int cta, int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    constexpr int PV = Elem<T>::PER_VEC;
    constexpr int U = 8;
    const size_t off = sendcount * (size_t)v.rank;
    const size_t n = sendcount / PV;
    const uint4 *src = (const uint4 *)v.send[v.rank];
    uint4 *dst = (uint4 *)((T *)v.mc_recv + off);
    const size_t bdim = blockDim.x, tile = (size_t)U * bdim;
    for (size_t t = cta; t * tile < n; t += nctas) {
        const size_t base = t * tile + threadIdx.x;
        uint4 val[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < n) val[u] = ld_v4(src + i);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < n) Multimem<T>::st(dst + i, val[u]);
        }
    }
    const size_t done = n * PV;
    if (done < sendcount && cta == 0) {
        const T *ts[1] = {(const T *)v.send[v.rank] + done};
        T *td[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) td[p] = (T *)v.recv[p] + off + done;
        reduce_store_elem<T>(ts, 1, td, NR, sendcount - done, threadIdx.x, blockDim.x);
    }

    cta_barrier<true>(v, e + 2, cta); // publishes the multicast stores
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// __global__ wrappers. Real mode: one rank per GPU, gridDim.y == 1. Emulation
// mode: gridDim.y virtual ranks on one GPU in one cooperative launch.
// ----------------------------------------------------------------------------
enum Op { OP_AR1 = 0, OP_AR2 = 1, OP_AG = 2, OP_RS = 3, OP_ARM = 4, OP_A2A = 5, OP_RSM = 6, OP_AGM = 7 };

template <typename T, int NR, int OP>
__device__ __forceinline__ void run_op(const RankView &v, size_t count, int cta, int nctas) {
    if (OP == OP_AR1) allreduce_oneshot_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_AR2) allreduce_twoshot_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_AG) allgather_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_ARM) allreduce_multimem_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_A2A) alltoall_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_RSM) reduce_scatter_multimem_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_AGM) allgather_multimem_body<T, NR>(v, count, cta, nctas);
    else reduce_scatter_body<T, NR>(v, count, cta, nctas);
}

// The rank's view travels as a __grid_constant__ kernel parameter (constant bank: no dependent
// global load on the latency path). Real mode: one view, gridDim.y == 1. Emulation mode: one
// view per virtual rank, gridDim.y == nranks, cooperative launch.
// NR == 1 keeps <= 64 registers so two CTAs share an SM (the world-1 HBM copy wants the
// occupancy). NR > 1 is held to 96 registers: a 512-thread collective CTA then takes 48 K of the
// SM's 64 K registers and leaves 16 K (and 1536 thread slots) to a CTA of the emulated compute,
// so compute and communication share SMs instead of queueing for them (DESIGN: SM contention).
#define DLNB_COLL_REGS(NR) ((NR) == 1 ? 64 : 96)

template <typename T, int NR, int OP>
__global__ void __maxnreg__(DLNB_COLL_REGS(NR))
coll_kernel(const __grid_constant__ RankView v, size_t count) {
    // World 1: consecutive collectives of a stream are launched with programmatic stream serialization, so this
    // kernel may be resident before its predecessor has finished. Nothing is read or written before the
    // predecessor's memory operations are complete and visible (a later collective may read what an earlier one
    // wrote, e.g. fsdp's gather buffer). Groups > 1 are never launched that way: a kernel that waits on its
    // peers must not hold SMs its predecessor's peers are waiting for.
    if (NR == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
    run_op<T, NR, OP>(v, count, blockIdx.x, gridDim.x);
}

template <typename T, int NR, int OP>
__global__ void __maxnreg__(DLNB_COLL_REGS(NR))
coll_kernel_emu(const __grid_constant__ EmuViews ev, size_t count) {
    run_op<T, NR, OP>(ev.v[blockIdx.y], count, blockIdx.x, gridDim.x);
}

} // namespace dlnb
