// Exploration only (not product, not built by build()): how low can the small-message allreduce floor go on
// NVSwitch with flags carried INSIDE the data ("LL"-style: every 8-byte store is {4 bytes of payload, 4 bytes of
// generation}, the receiver polls the payload slot itself), against the library's protocol for small messages
// (flag barrier, pull from every peer, flag barrier = four one-way NVLink trips)? Background: DESIGN.md §11 item 1 —
// the one-shot allreduce costs 8-10 us on 2 GPUs and ~15 us on 8 for <= 256 KiB, and a push variant with ONE
// release/acquire barrier measured slower. LL needs no barrier and no fence: one one-way trip.
//
// Single process driving n GPUs with peer access (cudaMalloc + cudaDeviceEnablePeerAccess), one stream per GPU;
// kernels on DIFFERENT GPUs wait on one another (never two waiting kernels on one GPU).
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ll_probe ll_allreduce_probe.cu && ./ll_probe [ngpus]
//
// Reuse of the staging slots without a barrier: two parities. A peer writes generation g+2 into the slots of parity
// g&1 only after it finished g+1, which needed my g+1 contribution, which I sent after I had consumed generation g.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s (line %d)\n", #x, cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
#define MAXR 8

struct Peers {
    uint2 *ll[MAXR];        // rank p's staging area: [parity 2][source rank n][count] of {payload, generation}
    const float *send[MAXR];
    uint32_t *flags[MAXR];  // rank p's barrier flags: [n] (baseline protocol only)
};

__device__ __forceinline__ void st_v2(uint2 *p, uint32_t a, uint32_t b) {
    asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ uint2 ld_v2(const uint2 *p) {
    uint2 r;
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_flag(uint32_t *p, uint32_t v) { asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_flag(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_sys_f32(const float *p) {
    float v;
    asm volatile("ld.relaxed.sys.global.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// LL one-shot: push my payload (with the generation) into everybody's staging area, then sum what arrived, rank order.
__global__ void ll_allreduce(Peers P, int me, int n, float *recv, size_t count, uint32_t gen, unsigned long long timeout_ns, int *error) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    const size_t par = (size_t)(gen & 1u) * n;
    for (size_t i = tid; i < count; i += nthr) {
        const uint32_t bits = __float_as_uint(P.send[me][i]);
        for (int k = 1; k < n; ++k) {
            const int p = (me + k) % n;
            st_v2(P.ll[p] + (par + me) * count + i, bits, gen);
        }
    }
    unsigned long long t0 = 0;
    for (size_t i = tid; i < count; i += nthr) {
        float acc = 0.f;
        for (int s = 0; s < n; ++s) {
            float v;
            if (s == me) {
                v = P.send[me][i];
            } else {
                const uint2 *slot = P.ll[me] + (par + s) * count + i;
                uint2 w = ld_v2(slot);
                while (w.y != gen) {
                    unsigned long long now;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                    if (!t0) t0 = now;
                    if (now - t0 > timeout_ns) { *error = 1; return; }
                    w = ld_v2(slot);
                }
                v = __uint_as_float(w.x);
            }
            acc = s == 0 ? v : acc + v;
        }
        recv[i] = acc;
    }
}

// Baseline, the library's small-message protocol in miniature: barrier, pull from every peer, barrier.
__device__ __forceinline__ void flag_barrier(const Peers &P, int me, int n, uint32_t gen, unsigned long long timeout_ns, int *error) {
    __syncthreads();
    if ((int)threadIdx.x < n) {
        const int p = threadIdx.x;
        st_flag(P.flags[p] + (size_t)blockIdx.x * MAXR + me, gen);
        unsigned long long t0 = 0;
        while ((int32_t)(ld_flag(P.flags[me] + (size_t)blockIdx.x * MAXR + p) - gen) < 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (!t0) t0 = now;
            if (now - t0 > timeout_ns) { *error = 1; break; }
        }
    }
    __syncthreads();
}
__global__ void pull_allreduce(Peers P, int me, int n, float *recv, size_t count, uint32_t gen, unsigned long long timeout_ns, int *error) {
    flag_barrier(P, me, n, 2 * gen - 1, timeout_ns, error);
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < count; i += nthr) {
        float acc = ld_sys_f32(P.send[0] + i);
        for (int s = 1; s < n; ++s) acc += ld_sys_f32(P.send[s] + i);
        recv[i] = acc;
    }
    flag_barrier(P, me, n, 2 * gen, timeout_ns, error);
}

int main(int argc, char **argv) {
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    int n = argc > 1 ? atoi(argv[1]) : ndev;
    if (n < 2 || n > ndev || n > MAXR) { printf("need 2..%d GPUs (have %d)\n", MAXR, ndev); return 1; }
    const size_t max_count = 64 * 1024; // 256 KiB of fp32
    std::vector<float *> send(n), recv(n);
    std::vector<uint2 *> ll(n);
    std::vector<uint32_t *> flags(n);
    std::vector<int *> err(n);
    std::vector<cudaStream_t> st(n);
    std::vector<cudaEvent_t> e0(n), e1(n);
    for (int d = 0; d < n; ++d) {
        CK(cudaSetDevice(d));
        for (int p = 0; p < n; ++p)
            if (p != d) CK(cudaDeviceEnablePeerAccess(p, 0));
        CK(cudaMalloc(&send[d], max_count * 4));
        CK(cudaMalloc(&recv[d], max_count * 4));
        CK(cudaMalloc(&ll[d], 2 * (size_t)n * max_count * sizeof(uint2)));
        CK(cudaMemset(ll[d], 0, 2 * (size_t)n * max_count * sizeof(uint2)));
        CK(cudaMalloc(&flags[d], 64 * MAXR * 4));
        CK(cudaMemset(flags[d], 0, 64 * MAXR * 4));
        CK(cudaMalloc(&err[d], 4));
        CK(cudaMemset(err[d], 0, 4));
        std::vector<float> h(max_count);
        for (size_t i = 0; i < max_count; ++i) h[i] = (float)((int)((i * 2654435761ull + (uint64_t)d * 40503ull) % 17ull) - 8);
        CK(cudaMemcpy(send[d], h.data(), max_count * 4, cudaMemcpyHostToDevice));
        CK(cudaStreamCreateWithFlags(&st[d], cudaStreamNonBlocking));
        CK(cudaEventCreate(&e0[d]));
        CK(cudaEventCreate(&e1[d]));
    }
    Peers P;
    for (int d = 0; d < n; ++d) { P.ll[d] = ll[d]; P.send[d] = send[d]; P.flags[d] = flags[d]; }
    const unsigned long long timeout = 5ull * 1000000000ull;
    uint32_t gen_ll = 0, gen_pull = 0;
    const size_t counts[] = {4, 256, 1024, 4096, 16384, 65536};
    for (size_t count : counts) {
        // the LL slots are addressed with this call's count: start every size from clean staging areas
        for (int d = 0; d < n; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); CK(cudaMemset(ll[d], 0, 2 * (size_t)n * max_count * sizeof(uint2))); CK(cudaDeviceSynchronize()); }
        for (int variant = 0; variant < 2; ++variant) {
            const int iters = 200;
            const int grid = (int)((count + 255) / 256 < 8 ? (count + 255) / 256 : 8);
            float worst = 0;
            for (int rep = 0; rep < 2; ++rep) {
                for (int d = 0; d < n; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(e0[d], st[d])); }
                for (int it = 0; it < iters; ++it) {
                    const uint32_t g = variant == 0 ? ++gen_ll : ++gen_pull;
                    for (int d = 0; d < n; ++d) {
                        CK(cudaSetDevice(d));
                        if (variant == 0) ll_allreduce<<<grid, 256, 0, st[d]>>>(P, d, n, recv[d], count, g, timeout, err[d]);
                        else pull_allreduce<<<grid, 256, 0, st[d]>>>(P, d, n, recv[d], count, g, timeout, err[d]);
                    }
                }
                worst = 0;
                for (int d = 0; d < n; ++d) {
                    CK(cudaSetDevice(d));
                    CK(cudaEventRecord(e1[d], st[d]));
                    CK(cudaEventSynchronize(e1[d]));
                    float ms;
                    CK(cudaEventElapsedTime(&ms, e0[d], e1[d]));
                    worst = ms > worst ? ms : worst;
                }
            }
            // verify on rank n-1
            std::vector<float> h(count);
            CK(cudaSetDevice(n - 1));
            CK(cudaMemcpy(h.data(), recv[n - 1], count * 4, cudaMemcpyDeviceToHost));
            size_t bad = 0;
            for (size_t i = 0; i < count; ++i) {
                float want = 0;
                for (int d = 0; d < n; ++d) want += (float)((int)((i * 2654435761ull + (uint64_t)d * 40503ull) % 17ull) - 8);
                bad += h[i] != want;
            }
            int herr = 0;
            CK(cudaMemcpy(&herr, err[n - 1], 4, cudaMemcpyDeviceToHost));
            printf("n=%d bytes=%7zu %-5s %7.2f us/call (back to back, max over GPUs) mismatches=%zu timeout=%d\n", n, count * 4,
                   variant == 0 ? "LL" : "pull", worst * 1e3 / iters, bad, herr);
        }
    }
    return 0;
}
