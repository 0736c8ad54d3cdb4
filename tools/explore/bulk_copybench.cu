// Exploration (not product): does a shared-memory-staged bulk-async copy (cp.async.bulk global -> smem -> global
// behind an mbarrier ring, one issuing thread per CTA) beat the register copy of coll_kernel<T,1,*> for the
// copy-only bodies (world-1 allreduce = bucket copy, allgather, all-to-all, send)? north_star names "shared-memory
// staging"; round 1 declined it by argument. This measures it: 66.9 MB (the llama3_8b bucket), 256 MiB, 1 GiB,
// every buffer distinct (ring of 3 GiB, nothing cached), events around a batch.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o bin/bulk_copybench tools/explore/bulk_copybench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, unsigned n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, unsigned parity) {
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem, const void *g, unsigned bytes, uint64_t *b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem)), "l"(g), "r"(bytes),
                 "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *g, const void *smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(smem)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }

// One thread per CTA drives a ring of STAGES chunks: load i+STAGES is issued as soon as store i-1 has read its stage.
template <int STAGES>
__global__ void __launch_bounds__(32) bulk_copy(const char *__restrict__ src, char *__restrict__ dst, size_t bytes, unsigned chunk) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *bar = (uint64_t *)smem; // STAGES barriers, then the stages
    unsigned char *stage = smem + 128;
    if (threadIdx.x != 0) return;
    for (int s = 0; s < STAGES; ++s) mbar_init(bar + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const size_t nchunks = (bytes + chunk - 1) / chunk;
    // chunks blockIdx.x, blockIdx.x + gridDim.x, ...
    const size_t mine = nchunks > blockIdx.x ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto off = [&](size_t k) { return (blockIdx.x + k * gridDim.x) * (size_t)chunk; };
    auto len = [&](size_t k) { size_t o = off(k); return (unsigned)(bytes - o < chunk ? bytes - o : chunk); };
    for (size_t k = 0; k < mine && k < STAGES; ++k) {
        mbar_expect(bar + k, len(k));
        bulk_g2s(stage + k * chunk, src + off(k), len(k), bar + k);
    }
    for (size_t k = 0; k < mine; ++k) {
        const int s = (int)(k % STAGES);
        mbar_wait(bar + s, (unsigned)((k / STAGES) & 1));
        bulk_s2g(dst + off(k), stage + (size_t)s * chunk, len(k));
        if (k >= 1 && k - 1 + STAGES < mine) { // refill the stage store k-1 used, once that store has read it
            bulk_wait_read<1>();
            const size_t r = k - 1 + STAGES;
            const int rs = (int)(r % STAGES);
            mbar_expect(bar + rs, len(r));
            bulk_g2s(stage + (size_t)rs * chunk, src + off(r), len(r), bar + rs);
        }
    }
    bulk_wait_read<0>();
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// the product's register copy (coll_kernel<T,1,*>): 512 threads, 4 x 128-bit loads in flight, tiles strided by the grid
__global__ void __launch_bounds__(512, 2) reg_copy(const uint4 *__restrict__ s, uint4 *__restrict__ d, size_t nvec) {
    const size_t tile = 4 * (size_t)blockDim.x;
    for (size_t t = blockIdx.x; t * tile < nvec; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const size_t i = base + (size_t)u * blockDim.x;
            if (i < nvec) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(s + i));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const size_t i = base + (size_t)u * blockDim.x;
            if (i < nvec) asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(d + i), "r"(v[u].x), "r"(v[u].y), "r"(v[u].z), "r"(v[u].w) : "memory");
        }
    }
}

template <int STAGES> static void bench_bulk(const char *a, char *b, size_t S, size_t ring, int sms, unsigned chunk, int per_sm, cudaEvent_t e0, cudaEvent_t e1) {
    const size_t smem = 128 + (size_t)STAGES * chunk;
    if (smem * per_sm > 227 * 1024) return;
    CK(cudaFuncSetAttribute(bulk_copy<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int nbuf = (int)(ring / S), iters = nbuf < 4 ? 8 : (nbuf > 40 ? 40 : nbuf);
    float ms = 0;
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0));
        for (int i = 0; i < iters; ++i) bulk_copy<STAGES><<<sms * per_sm, 32, smem>>>(a + (size_t)(i % nbuf) * S, b + (size_t)(i % nbuf) * S, S, chunk);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
    }
    CK(cudaGetLastError());
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("{\"S\": %zu, \"kernel\": \"bulk_async\", \"stages\": %d, \"chunk_kib\": %u, \"ctas_per_sm\": %d, \"GBps\": %.1f, \"us\": %.2f}\n", S, STAGES, chunk >> 10,
           per_sm, 2.0 * S * iters / (ms * 1e-3) / 1e9, ms * 1e3 / iters);
    fflush(stdout);
}

int main() {
    int sms;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const size_t sizes[] = {66918844 / 16 * 16, 268435456, 1073741824};
    const size_t ring = (size_t)3 << 30;
    char *a, *b;
    CK(cudaMalloc(&a, ring));
    CK(cudaMalloc(&b, ring));
    CK(cudaMemset(a, 1, ring));
    CK(cudaMemset(b, 2, ring));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (size_t S : sizes) {
        const int nbuf = (int)(ring / S), iters = nbuf < 4 ? 8 : (nbuf > 40 ? 40 : nbuf);
        float ms = 0;
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaEventRecord(e0));
            for (int i = 0; i < iters; ++i) CK(cudaMemcpyAsync(b + (size_t)(i % nbuf) * S, a + (size_t)(i % nbuf) * S, S, cudaMemcpyDeviceToDevice));
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("{\"S\": %zu, \"kernel\": \"cudaMemcpyAsync D2D\", \"GBps\": %.1f, \"us\": %.2f}\n", S, 2.0 * S * iters / (ms * 1e-3) / 1e9, ms * 1e3 / iters);
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaEventRecord(e0));
            for (int i = 0; i < iters; ++i) reg_copy<<<sms * 4, 512>>>((const uint4 *)(a + (size_t)(i % nbuf) * S), (uint4 *)(b + (size_t)(i % nbuf) * S), S / 16);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("{\"S\": %zu, \"kernel\": \"register copy (product)\", \"GBps\": %.1f, \"us\": %.2f}\n", S, 2.0 * S * iters / (ms * 1e-3) / 1e9, ms * 1e3 / iters);
        for (unsigned chunk : {8192u, 16384u, 32768u, 65536u}) {
            for (int per_sm : {1, 2, 4}) {
                bench_bulk<2>(a, b, S, ring, sms, chunk, per_sm, e0, e1);
                bench_bulk<3>(a, b, S, ring, sms, chunk, per_sm, e0, e1);
                bench_bulk<4>(a, b, S, ring, sms, chunk, per_sm, e0, e1);
                bench_bulk<6>(a, b, S, ring, sms, chunk, per_sm, e0, e1);
            }
        }
    }
    // correctness of the staged copy on an odd size
    const size_t S = 66918844 / 16 * 16;
    CK(cudaMemset(b, 0, S));
    const size_t smem = 128 + 4 * 32768;
    CK(cudaFuncSetAttribute(bulk_copy<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    bulk_copy<4><<<sms, 32, smem>>>(a, b, S, 32768);
    CK(cudaDeviceSynchronize());
    unsigned char *h = (unsigned char *)malloc(S);
    CK(cudaMemcpy(h, b, S, cudaMemcpyDeviceToHost));
    size_t bad = 0;
    for (size_t i = 0; i < S; ++i) bad += h[i] != 1;
    printf("{\"check\": \"bulk copy of %zu bytes\", \"wrong_bytes\": %zu}\n", S, bad);
    return bad != 0;
}
