// kernels.cuh — sm_100a device code of the DLNetBench data-parallel hot path.
//
// Kernels here replace the library calls behind the reference's communicator
// virtuals (cpp/proxy_classes.hpp:66-101 MPI, :149-209 NCCL):
//   allreduce_oneshot / allreduce_twoshot  <- MPI_Iallreduce / ncclAllReduce (SUM)
//   allgather_push                         <- MPI_(I)allgather / ncclAllGather
//   reduce_scatter_pull                    <- MPI_Reduce_scatter_block / ncclReduceScatter
//   spin_kernel                            <- usleep() compute emulation (dp.cpp:93,98)
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
    CommCtl *ctl;              // my control block
    uint32_t *host_error;      // mapped pinned word, set by the watchdog (may be null)
    unsigned long long timeout_ns;
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
    static __device__ __forceinline__ float to_f(float x) { return x; }
    static __device__ __forceinline__ float from_f(float x) { return x; }
};
template <> struct Elem<__half> {
    static constexpr int PER_VEC = 8;
    static __device__ __forceinline__ float to_f(__half x) { return __half2float(x); }
    static __device__ __forceinline__ __half from_f(float x) { return __float2half_rn(x); }
};
template <> struct Elem<__nv_bfloat16> {
    static constexpr int PER_VEC = 8;
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
        if (PUBLISH) st_release_sys(theirs, gen);
        else st_relaxed_sys(theirs, gen);
        unsigned long long t0 = 0;
        unsigned polls = 0;
        while ((int32_t)((PUBLISH ? ld_acquire_sys(mine) : ld_relaxed_sys(mine)) - gen) < 0) {
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
    // grid), a warp touches 512 contiguous bytes per load instruction.
    const size_t bdim = blockDim.x, tile = (size_t)U * bdim;
    const size_t cta = tid / bdim, nctas = nthr / bdim;
    for (size_t t = cta; t * tile < nvec; t += nctas) {
        const size_t base = t * tile + threadIdx.x;
        uint4 val[U][NSRC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * bdim;
            if (i < nvec) {
#pragma unroll
                for (int p = 0; p < NSRC; ++p) val[u][p] = RO ? ld_v4_ro(s[p] + i) : ld_v4(s[p] + i);
            }
        }
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

// Element-wise version for heads, tails and (for now) misaligned chunks.
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
    // U*NR independent 128-bit loads in flight per thread, at most 16 (64 data registers)
    static constexpr int U = NR == 1 ? 4 : NR == 2 ? 8 : NR <= 4 ? 4 : 2;
    // allgather: one local source, NR destinations (pointers, not data, fill the registers)
    static constexpr int U_AG = NR <= 2 ? 4 : NR <= 4 ? 4 : 2;
};

__device__ __forceinline__ bool all_aligned16(const RankView &v, size_t send_off_bytes,
                                              size_t recv_off_bytes) {
    bool ok = true;
    for (int p = 0; p < v.nranks; ++p) {
        ok = ok && (((uintptr_t)v.send[p] + send_off_bytes) & 15) == 0;
        ok = ok && (((uintptr_t)v.recv[p] + recv_off_bytes) & 15) == 0;
    }
    return ok;
}

// ----------------------------------------------------------------------------
// Allreduce, one-shot: every rank reads all n send buffers and writes its own recv.
// Latency path: one barrier before (peers' inputs final), one after (peers done
// reading mine). Wire: (n-1)*S per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allreduce_oneshot_body(const RankView &v, size_t count, int cta,
                                                       int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const size_t tid = (size_t)cta * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)nctas * blockDim.x;
    T *mine = (T *)v.recv[v.rank];
    if (all_aligned16(v, 0, 0)) {
        constexpr int PV = Elem<T>::PER_VEC;
        const size_t nvec = count / PV;
        const uint4 *src[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) src[p] = (const uint4 *)v.send[p];
        uint4 *dst[1] = {(uint4 *)mine};
        reduce_store_vec<T, NR, 1, Unroll<NR>::U, NR == 1>(src, dst, nvec, tid, nthr);
        const size_t done = nvec * PV;
        if (done < count && cta == 0) {
            const T *ts[MAXR];
            for (int p = 0; p < NR; ++p) ts[p] = (const T *)v.send[p] + done;
            T *td[1] = {mine + done};
            reduce_store_elem<T>(ts, NR, td, 1, count - done, threadIdx.x, blockDim.x);
        }
    } else {
        const T *ts[MAXR];
        for (int p = 0; p < NR; ++p) ts[p] = (const T *)v.send[p];
        T *td[1] = {mine};
        reduce_store_elem<T>(ts, NR, td, 1, count, tid, nthr);
    }

    cta_barrier<false>(v, e + 2, cta); // only reads happened
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Allreduce, two-shot fused in one pass: rank r owns slice r (16-byte aligned
// slice boundaries), pulls that slice from every rank, sums, and pushes the result
// into every rank's recv. Wire: 2(n-1)/n*S per direction, reads and writes overlap.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allreduce_twoshot_body(const RankView &v, size_t count, int cta,
                                                       int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const size_t tid = (size_t)cta * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)nctas * blockDim.x;
    constexpr int PV = Elem<T>::PER_VEC;
    if (all_aligned16(v, 0, 0)) {
        const size_t nvec = count / PV;
        const size_t per = (nvec + NR - 1) / NR;
        const size_t v0 = min(per * (size_t)v.rank, nvec);
        const size_t v1 = min(v0 + per, nvec);
        const uint4 *src[NR];
        uint4 *dst[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) {
            src[p] = (const uint4 *)v.send[p] + v0;
            dst[p] = (uint4 *)v.recv[p] + v0;
        }
        reduce_store_vec<T, NR, NR, Unroll<NR>::U, NR == 1>(src, dst, v1 - v0, tid, nthr);
        // sub-vector tail (< 16 bytes): every rank reduces it for itself
        const size_t done = nvec * PV;
        if (done < count && cta == 0) {
            const T *ts[MAXR];
            for (int p = 0; p < NR; ++p) ts[p] = (const T *)v.send[p] + done;
            T *td[1] = {(T *)v.recv[v.rank] + done};
            reduce_store_elem<T>(ts, NR, td, 1, count - done, threadIdx.x, blockDim.x);
        }
    } else {
        const size_t per = (count + NR - 1) / NR;
        const size_t e0 = min(per * (size_t)v.rank, count);
        const size_t e1 = min(e0 + per, count);
        const T *ts[MAXR];
        T *td[MAXR];
        for (int p = 0; p < NR; ++p) {
            ts[p] = (const T *)v.send[p] + e0;
            td[p] = (T *)v.recv[p] + e0;
        }
        reduce_store_elem<T>(ts, NR, td, NR, e1 - e0, tid, nthr);
    }

    cta_barrier<true>(v, e + 2, cta); // publishes the pushed slices
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Allgather, push: read my shard once from local HBM, store it into slot `rank`
// of every rank's recv (n-1 of them over NVLink). Wire: (n-1)*shard per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void allgather_body(const RankView &v, size_t sendcount, int cta,
                                               int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const size_t tid = (size_t)cta * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)nctas * blockDim.x;
    constexpr int PV = Elem<T>::PER_VEC;
    const size_t off = sendcount * (size_t)v.rank; // elements into every recv
    const T *s = (const T *)v.send[v.rank];
    bool aligned = (((uintptr_t)s) & 15) == 0;
    for (int p = 0; p < NR; ++p) aligned = aligned && ((((uintptr_t)v.recv[p]) + off * sizeof(T)) & 15) == 0;
    if (aligned) {
        const size_t nvec = sendcount / PV;
        const uint4 *src[1] = {(const uint4 *)s};
        uint4 *dst[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) dst[p] = (uint4 *)((T *)v.recv[p] + off);
        reduce_store_vec<T, 1, NR, Unroll<NR>::U_AG, true>(src, dst, nvec, tid, nthr);
        const size_t done = nvec * PV;
        if (done < sendcount && cta == 0) {
            const T *ts[1] = {s + done};
            T *td[MAXR];
            for (int p = 0; p < NR; ++p) td[p] = (T *)v.recv[p] + off + done;
            reduce_store_elem<T>(ts, 1, td, NR, sendcount - done, threadIdx.x, blockDim.x);
        }
    } else {
        const T *ts[1] = {s};
        T *td[MAXR];
        for (int p = 0; p < NR; ++p) td[p] = (T *)v.recv[p] + off;
        reduce_store_elem<T>(ts, 1, td, NR, sendcount, tid, nthr);
    }

    cta_barrier<true>(v, e + 2, cta); // publishes the pushed shard
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// Reduce-scatter, pull: rank r reads chunk r (recvcount elements at offset
// r*recvcount) of every rank's send, sums in rank order, writes its own recv.
// Wire: (n-1)*recvcount per direction.
// ----------------------------------------------------------------------------
template <typename T, int NR>
__device__ __forceinline__ void reduce_scatter_body(const RankView &v, size_t recvcount, int cta,
                                                    int nctas) {
    const uint32_t e = cta_epoch(v, cta);
    cta_barrier<false>(v, e + 1, cta);

    const size_t tid = (size_t)cta * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)nctas * blockDim.x;
    constexpr int PV = Elem<T>::PER_VEC;
    const size_t off = recvcount * (size_t)v.rank;
    T *mine = (T *)v.recv[v.rank];
    bool aligned = (((uintptr_t)mine) & 15) == 0;
    for (int p = 0; p < NR; ++p) aligned = aligned && ((((uintptr_t)v.send[p]) + off * sizeof(T)) & 15) == 0;
    if (aligned) {
        const size_t nvec = recvcount / PV;
        const uint4 *src[NR];
#pragma unroll
        for (int p = 0; p < NR; ++p) src[p] = (const uint4 *)((const T *)v.send[p] + off);
        uint4 *dst[1] = {(uint4 *)mine};
        reduce_store_vec<T, NR, 1, Unroll<NR>::U, NR == 1>(src, dst, nvec, tid, nthr);
        const size_t done = nvec * PV;
        if (done < recvcount && cta == 0) {
            const T *ts[MAXR];
            for (int p = 0; p < NR; ++p) ts[p] = (const T *)v.send[p] + off + done;
            T *td[1] = {mine + done};
            reduce_store_elem<T>(ts, NR, td, 1, recvcount - done, threadIdx.x, blockDim.x);
        }
    } else {
        const T *ts[MAXR];
        for (int p = 0; p < NR; ++p) ts[p] = (const T *)v.send[p] + off;
        T *td[1] = {mine};
        reduce_store_elem<T>(ts, NR, td, 1, recvcount, tid, nthr);
    }

    cta_barrier<false>(v, e + 2, cta); // only reads happened
    cta_finish(v, e + 2, cta);
}

// ----------------------------------------------------------------------------
// __global__ wrappers. Real mode: one rank per GPU, gridDim.y == 1. Emulation
// mode: gridDim.y virtual ranks on one GPU in one cooperative launch.
// ----------------------------------------------------------------------------
enum Op { OP_AR1 = 0, OP_AR2 = 1, OP_AG = 2, OP_RS = 3 };

template <typename T, int NR, int OP>
__device__ __forceinline__ void run_op(const RankView &v, size_t count, int cta, int nctas) {
    if (OP == OP_AR1) allreduce_oneshot_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_AR2) allreduce_twoshot_body<T, NR>(v, count, cta, nctas);
    else if (OP == OP_AG) allgather_body<T, NR>(v, count, cta, nctas);
    else reduce_scatter_body<T, NR>(v, count, cta, nctas);
}

// The rank's view travels as a __grid_constant__ kernel parameter (constant bank: no dependent
// global load on the latency path). Real mode: one view, gridDim.y == 1. Emulation mode: one
// view per virtual rank, gridDim.y == nranks, cooperative launch.
// NR == 1 keeps <= 64 registers so two CTAs share an SM (the world-1 HBM copy wants the
// occupancy), larger groups hold U*NR vectors in flight.
template <typename T, int NR, int OP>
__global__ void __launch_bounds__(512, (NR == 1 ? 2 : 1))
coll_kernel(const __grid_constant__ RankView v, size_t count) {
    run_op<T, NR, OP>(v, count, blockIdx.x, gridDim.x);
}

template <typename T, int NR, int OP>
__global__ void __launch_bounds__(512, (NR == 1 ? 2 : 1))
coll_kernel_emu(const __grid_constant__ EmuViews ev, size_t count) {
    run_op<T, NR, OP>(ev.v[blockIdx.y], count, blockIdx.x, gridDim.x);
}

} // namespace dlnb
