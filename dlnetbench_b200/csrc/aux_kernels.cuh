// aux_kernels.cuh — compute-emulation spin kernel and test/bench helper kernels.
// Included by runtime.cu only.
#pragma once
#include "kernels.cuh"

namespace dlnb {

// ----------------------------------------------------------------------------
// Emulated compute: occupy the GPU until a %globaltimer deadline `ns` nanoseconds away.
// Replaces usleep(fwd) / usleep(bwd) of the reference (dp.cpp:93,98; fsdp.cpp:103,120,145).
// ----------------------------------------------------------------------------
// Consecutive spins form one seamless stretch of "compute": a CTA that starts within `slack_ns`
// of the previous spin's NOMINAL end counts from that end, so launch gaps and per-kernel timer
// granularity do not accumulate over a chain (a dp iteration is up to 4001 spins). A CTA that
// starts later than that (first spin of a chain, or its SM was not free) spins its full `ns`
// from its own start: compute that could not get an SM is late, as it would be for real.
// stamps[parity] receives this spin's nominal end; stamps[parity ^ 1] holds the previous one's
// (two words, so a late CTA never reads a stamp written by its own kernel).
//
// What the CTAs do until the deadline is the FOOTPRINT of the emulated compute (dlnb_set
// "spin_mode"): what a collective running beside it has to share the SM with.
//   SPIN_SLEEP  __nanosleep back-off: warps are resident but issue nothing (no contention)
//   SPIN_FMA    four dependent FMA chains per thread: issue slots and the FP32 pipe stay busy
//   SPIN_HBM    streams a buffer larger than L2 with 128-bit loads: HBM and L2 stay busy
// Dynamic shared memory (`spin_smem`) is ballast only: it limits what else fits on the SM.
enum SpinMode { SPIN_SLEEP = 0, SPIN_FMA = 1, SPIN_HBM = 2 };

__global__ void __launch_bounds__(1024)
spin_kernel(unsigned long long ns, unsigned long long *stamps, unsigned parity, unsigned long long slack_ns, int mode,
            const uint4 *stream, size_t stream_vecs, float *sink) {
    const unsigned long long t0 = globaltimer_ns();
    const unsigned long long prev = stamps[parity ^ 1u];
    const unsigned long long start = (t0 >= prev && t0 - prev <= slack_ns) ? prev : t0;
    const unsigned long long deadline = start + ns;
    if (mode == SPIN_FMA) {
        float a = 1.0f + threadIdx.x * 1e-6f, b = 0.5f, c = 0.25f, d = 0.125f;
        const float k = 0.999f, z = 1e-3f;
        while (globaltimer_ns() < deadline) {
#pragma unroll
            for (int i = 0; i < 64; ++i) { // 256 FMAs between two looks at the clock
                a = fmaf(a, k, z);
                b = fmaf(b, k, z);
                c = fmaf(c, k, z);
                d = fmaf(d, k, z);
            }
        }
        if (a + b + c + d == -1.0f) *sink = a; // never true: keeps the chains alive
    } else if (mode == SPIN_HBM && stream_vecs) {
        const size_t nthr = (size_t)gridDim.x * blockDim.x;
        size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        unsigned acc = 0;
        while (globaltimer_ns() < deadline) {
#pragma unroll
            for (int r = 0; r < 4; ++r) { // 16 x 128-bit loads per thread between two looks at the clock
                uint4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    v[u] = ld_v4_ro(stream + i);
                    i += nthr;
                    if (i >= stream_vecs) i -= stream_vecs;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
            }
        }
        if (acc == 0x9e3779b9u) *sink = 1.0f; // practically never: keeps the loads alive
    } else {
        for (;;) {
            const unsigned long long now = globaltimer_ns();
            if (now >= deadline) break;
            const unsigned long long left = deadline - now;
            // back off while far from the deadline so the SM's issue slots stay free
            if (left > 4000) __nanosleep(left > 64000 ? 16000 : (unsigned)(left >> 2));
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) stamps[parity] = deadline;
}

// One flag barrier between all ranks on whatever stream it is launched on: dlnb_iter_begin uses
// it so that every rank's emulated compute starts within a flag round trip of the others (the
// host barrier before it leaves ranks ~1 ms apart, and a spin chain preserves that skew for the
// whole iteration: every collective then waits for the slowest rank to arrive).
__global__ void rank_sync_kernel(const __grid_constant__ RankView v) {
    const uint32_t e = cta_epoch(v, 0);
    cta_barrier<false>(v, e + 1, 0);
    cta_finish(v, e + 1, 0);
}

// ----------------------------------------------------------------------------
// Test / bench helpers
// ----------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
// value of element i of rank r's buffer, as fp32 before rounding to the dtype
__host__ __device__ __forceinline__ float pattern_value(int kind, uint64_t seed, int r, uint64_t i) {
    if (kind == 1) {
        const uint64_t h = i * 2654435761ull + (uint64_t)r * 40503ull;
        return (float)((int)(h % 17ull) - 8);
    } else if (kind == 2) {
        uint32_t h = mix32((uint32_t)i ^ mix32((uint32_t)(i >> 32) + 0x9e3779b9u * (uint32_t)(r + 1) +
                                               (uint32_t)seed));
        return (float)((int)(h >> 8) - (1 << 23)) * (1.0f / 8388608.0f);
    }
    return 0.0f;
}

template <typename T>
__global__ void fill_kernel(T *p, size_t n, int kind, uint64_t seed, int r) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = Elem<T>::from_f(pattern_value(kind, seed, r, i));
}

struct RankList { int r[MAXR]; int n; };

template <typename T>
__global__ void verify_sum_kernel(const T *p, size_t n, size_t elem_off, int kind, uint64_t seed,
                                  RankList rl, unsigned long long *mismatches) {
    unsigned long long bad = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        // each rank's element is first rounded to T (that is what the buffer held)
        float acc = Elem<T>::to_f(Elem<T>::from_f(pattern_value(kind, seed, rl.r[0], elem_off + i)));
        for (int k = 1; k < rl.n; ++k)
            acc += Elem<T>::to_f(Elem<T>::from_f(pattern_value(kind, seed, rl.r[k], elem_off + i)));
        const T want = rl.n == 1 ? Elem<T>::from_f(pattern_value(kind, seed, rl.r[0], elem_off + i))
                                 : Elem<T>::from_f(acc);
        const T got = p[i];
        bool same;
        if (sizeof(T) == 4) same = *(const uint32_t *)&want == *(const uint32_t *)&got;
        else same = *(const uint16_t *)&want == *(const uint16_t *)&got;
        bad += same ? 0 : 1;
    }
    for (int o = 16; o > 0; o >>= 1) bad += __shfl_down_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(mismatches, bad);
}

// Tolerance form of the verifier, for reductions whose summation order is not ours (the NVSwitch's):
// |got - ref| <= rel_tol * sum_r |x_r|, ref = rank-order fp32 sum (SURVEY 8c pin 2: rel_tol = 2(n-1)eps).
// worst[0] receives the largest |got - ref| / sum|x| seen, as float bits (non-negative floats order like uints).
template <typename T>
__global__ void verify_sum_tol_kernel(const T *p, size_t n, size_t elem_off, int kind, uint64_t seed, RankList rl,
                                      float rel_tol, unsigned long long *mismatches, unsigned *worst) {
    unsigned long long bad = 0;
    float w = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float acc = 0.f, mag = 0.f;
        for (int k = 0; k < rl.n; ++k) {
            const float x = Elem<T>::to_f(Elem<T>::from_f(pattern_value(kind, seed, rl.r[k], elem_off + i)));
            acc += x;
            mag += fabsf(x);
        }
        const float got = Elem<T>::to_f(p[i]);
        const float err = fabsf(got - acc);
        // one rounding of the result to T is part of the contract as well: half an ulp of T at |ref|
        const float ulp_half = Elem<T>::HALF_ULP * fabsf(acc);
        const float allow = rel_tol * mag + ulp_half;
        if (!(err <= allow)) bad++;
        if (mag > 0.f) w = fmaxf(w, err / mag);
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_down_sync(0xffffffffu, bad, o);
        w = fmaxf(w, __shfl_down_sync(0xffffffffu, w, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (bad) atomicAdd(mismatches, bad);
        atomicMax(worst, __float_as_uint(w));
    }
}

__global__ void checksum_kernel(const uint32_t *p, size_t nwords, unsigned long long *out) {
    unsigned long long s = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x)
        s += (unsigned long long)mix32(p[i] + (uint32_t)i * 0x9e3779b9u);
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, s);
}

__global__ void flush_kernel(uint4 *p, size_t nvec, uint32_t tag) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_uint4(tag, tag, tag, tag);
}

} // namespace dlnb

// ----------------------------------------------------------------------------
// Point-to-point put with a rendezvous — ProxyCommunicator::send / recv (reference:
// cpp/proxy_classes.hpp:103-117 MPI_Send/MPI_Recv, :211-227 ncclSend/ncclRecv), the
// stage-to-stage traffic of the pipeline proxies (hybrid_2d.cpp:108-161).
//   recv_post / recv_wait  post the arena offset of the receive buffer into the SENDER's mailbox,
//                and (separately) wait for the sender's "done" flag;
//   send_kernel  waits for that post, stores its data straight into the receiver's buffer
//                (window base of the receiver + posted offset; 128-bit, shifted if needed),
//                and the last CTA to finish publishes "done" with a system-scope release.
// The k-th send from a to b matches the k-th recv on b from a (sequence numbers are counted on
// the host: point-to-point calls are not captured into graphs).
// ----------------------------------------------------------------------------
namespace dlnb {

constexpr int P2P_RING = 8; // outstanding receives per (source, destination) pair

struct alignas(128) P2PBox {
    unsigned long long post[P2P_RING]; // written by the peer as RECEIVER: (arena offset << 16) | (seq & 0xffff), slot seq % 8
    uint32_t done;                     // written by the peer as SENDER: seq of the last delivered message
};
struct P2PState {
    uint32_t send_ctas_done[MAXR];
};

__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// A receive is two tiny kernels. The POST runs in order on the communication stream (whatever
// used the buffer before is finished) and returns at once. The WAIT is only enqueued when the
// host asks for completion (Wait / WaitAll / the blocking recv): a kernel that spins on a flag
// another GPU sets must never sit in front of work that GPU is waiting for — streams may share a
// hardware queue, and a posted receive followed by a send (ring exchange) would deadlock.
__global__ void recv_post_kernel(P2PBox *box_at_sender, uint32_t seq, unsigned long long arena_offset) {
    if (threadIdx.x != 0) return;
    st_relaxed_sys_u64(&box_at_sender->post[seq % P2P_RING], (arena_offset << 16) | (unsigned long long)(seq & 0xffffu));
}
__global__ void recv_wait_kernel(const P2PBox *my_box, uint32_t seq, unsigned long long timeout_ns, uint32_t *host_error) {
    if (threadIdx.x != 0) return;
    unsigned long long t0 = 0;
    unsigned polls = 0;
    while ((int32_t)(ld_acquire_sys(&my_box->done) - seq) < 0) {
        if ((++polls & 0x3ff) == 0) {
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > timeout_ns) {
                if (host_error) *(volatile uint32_t *)host_error = 1;
                __threadfence_system();
                break;
            }
        }
    }
}

__global__ void __launch_bounds__(512, 1)
send_kernel(const P2PBox *my_box, P2PBox *box_at_receiver, char *receiver_window, const void *src, size_t bytes,
            P2PState *st, int dst, uint32_t seq, unsigned long long timeout_ns, uint32_t *host_error) {
    __shared__ unsigned long long s_off;
    __shared__ int s_abort;
    if (threadIdx.x == 0) {
        s_abort = 0;
        unsigned long long post, t0 = 0;
        unsigned polls = 0;
        while (((post = ld_relaxed_sys_u64(&my_box->post[seq % P2P_RING])) & 0xffffull) != (unsigned long long)(seq & 0xffffu)) {
            if ((++polls & 0x3ff) == 0) {
                const unsigned long long now = globaltimer_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > timeout_ns) {
                    if (host_error) *(volatile uint32_t *)host_error = 1;
                    __threadfence_system();
                    s_abort = 1; // the post never came: the word holds some other receive's offset
                    break;
                }
            }
        }
        s_off = post >> 16;
    }
    __syncthreads();
    if (s_abort) return; // no copy, no "done": the host reports the time-out
    // payload as 2-byte elements (every dtype here has an even size): pure copy, any alignment
    const __half *s[1] = {(const __half *)src};
    __half *d[1] = {(__half *)(receiver_window + s_off)};
    move_any<__half, 1, 1, 4, 4>(s, d, bytes / 2, blockIdx.x, gridDim.x);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system(); // this CTA's stores are out before it is counted
        if (atomicAdd(&st->send_ctas_done[dst], 1u) == gridDim.x - 1) {
            st->send_ctas_done[dst] = 0;
            st_release_sys(&box_at_receiver->done, seq);
        }
    }
}

} // namespace dlnb
