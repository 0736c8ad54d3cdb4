// Exploration only (not product, not built by build()): does programmatic dependent launch (PDL) hide the
// per-launch fill/drain of back-to-back bucket copies on B200?  Background: DESIGN.md §11 item 4 — the world-1
// bucket "allreduce" (a 66.9 MB HBM copy, 592 CTAs x 512 threads, 2 CTAs/SM) runs at 83 % of the copy peak in a
// back-to-back batch against 95 % at 1 GiB; the difference is per-launch, not steady state.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pdl_copybench pdl_copybench.cu && ./pdl_copybench
//
// Variants (same copy loop as the library's NR = 1 path: tiles of 4 x 512 128-bit vectors, ld.global.nc / st no-allocate):
//   plain      ordinary stream-ordered launches
//   pdl_wait   launched with programmaticStreamSerialization; griddepcontrol.launch_dependents first thing,
//              griddepcontrol.wait before the first load (what the library would have to do: a later collective
//              may read what an earlier one wrote)
//   pdl_late   as pdl_wait, but launch_dependents only after the CTA's last tile is loaded
//   pdl_nowait no wait at all (upper bound: only valid when consecutive launches are independent)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint4 ld_nc(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_na(uint4 *p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

enum { PLAIN = 0, PDL_WAIT = 1, PDL_LATE = 2, PDL_NOWAIT = 3 };

template <int MODE> __global__ void __launch_bounds__(512, 2) copyk(const uint4 *__restrict__ s, uint4 *__restrict__ d, size_t nvec) {
    constexpr int U = 4;
    if (MODE == PDL_WAIT || MODE == PDL_NOWAIT) asm volatile("griddepcontrol.launch_dependents;");
    if (MODE == PDL_WAIT || MODE == PDL_LATE) asm volatile("griddepcontrol.wait;" ::: "memory");
    const size_t tile = (size_t)U * blockDim.x;
    for (size_t t = blockIdx.x; t * tile < nvec; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * blockDim.x;
            if (i < nvec) v[u] = ld_nc(s + i);
        }
        if (MODE == PDL_LATE && (t + gridDim.x) * tile >= nvec) asm volatile("griddepcontrol.launch_dependents;");
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = base + (size_t)u * blockDim.x;
            if (i < nvec) st_na(d + i, v[u]);
        }
    }
}

template <int MODE> static void launch(int grid, cudaStream_t st, const uint4 *s, uint4 *d, size_t nvec) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(512);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = MODE == PLAIN ? 0 : 1;
    CK(cudaLaunchKernelEx(&cfg, copyk<MODE>, s, d, nvec));
}

int main() {
    int sms;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const size_t ring = (size_t)4 << 30;
    char *a, *b;
    CK(cudaMalloc(&a, ring));
    CK(cudaMalloc(&b, ring));
    CK(cudaMemset(a, 1, ring));
    CK(cudaMemset(b, 2, ring));
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const size_t sizes[] = {1u << 20, 4u << 20, 16u << 20, 66918844 / 16 * 16, 256u << 20};
    const char *names[] = {"plain", "pdl_wait", "pdl_late", "pdl_nowait"};
    void (*fns[])(int, cudaStream_t, const uint4 *, uint4 *, size_t) = {launch<PLAIN>, launch<PDL_WAIT>, launch<PDL_LATE>, launch<PDL_NOWAIT>};
    for (size_t S : sizes) {
        const int nbuf = (int)(ring / S);
        const int iters = nbuf > 240 ? 240 : nbuf; // distinct buffers: nothing is served from L2
        const size_t nvec = S / 16;
        const size_t tiles = (nvec + 2047) / 2048;
        const int grid = (int)(tiles < (size_t)sms * 4 ? tiles : (size_t)sms * 4);
        for (int m = 0; m < 4; ++m) {
            float best = 1e30f;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaEventRecord(e0, st));
                for (int i = 0; i < iters; ++i)
                    fns[m](grid, st, (const uint4 *)(a + (size_t)i * S), (uint4 *)(b + (size_t)i * S), nvec);
                CK(cudaEventRecord(e1, st));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                best = ms < best ? ms : best;
            }
            printf("S=%10zu grid=%4d %-10s %8.2f us/launch %8.1f GB/s\n", S, grid, names[m], best * 1e3 / iters, 2.0 * S * iters / (best * 1e-3) / 1e9);
        }
    }
    // spot check of the last variant's result
    unsigned char h = 0;
    CK(cudaMemcpy(&h, b + 12345, 1, cudaMemcpyDeviceToHost));
    printf("check: %s\n", h == 1 ? "ok" : "WRONG");
    return 0;
}
