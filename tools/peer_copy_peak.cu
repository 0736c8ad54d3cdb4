// peer_copy_peak.cu — what one NVLink direction of this box actually carries (the measured link peak the
// collective kernels are scored against, next to the nominal 900 GB/s). Single process, peer access between
// the first two (or --gpus N) devices. Prints one JSON line per measurement.
//
//   copy engine   cudaMemcpyPeerAsync 0 -> 1            (one direction busy)
//   copy engine   0 -> 1 and 1 -> 0 at once             (both directions busy: what every collective does)
//   SM kernel     128-bit stores into the peer (push)   one direction / both
//   SM kernel     128-bit loads from the peer (pull)    one direction / both
//   all-to-all    every GPU pushes to every other one   (n GPUs: the NVSwitch fabric under full load)
//
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o bin/peer_copy_peak tools/peer_copy_peak.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            fprintf(stderr, "%s failed: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

__global__ void __launch_bounds__(512) copy_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t nvec) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
        dst[i] = a;
        dst[i + stride] = b;
        dst[i + 2 * stride] = c;
        dst[i + 3 * stride] = d;
    }
    for (; i < nvec; i += stride) dst[i] = src[i];
}

// bulk-async staged copy: one thread per CTA drives a ring of STAGES chunks, cp.async.bulk global -> smem (mbarrier
// complete_tx) and smem -> global (bulk_group); either side may be peer memory. Same code as tools/explore/bulk_copybench.cu.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, unsigned parity) {
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem, const void *g, unsigned bytes, uint64_t *b) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem)), "l"(g), "r"(bytes),
                 "r"(smem_u32(b)) : "memory");
}
template <int STAGES>
__global__ void __launch_bounds__(32) bulk_copy(const char *__restrict__ src, char *__restrict__ dst, size_t bytes, unsigned chunk) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *bar = (uint64_t *)smem;
    unsigned char *stage = smem + 128;
    if (threadIdx.x != 0) return;
    for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + s)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const size_t nchunks = (bytes + chunk - 1) / chunk;
    const size_t mine = nchunks > blockIdx.x ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto off = [&](size_t k) { return (blockIdx.x + k * gridDim.x) * (size_t)chunk; };
    auto len = [&](size_t k) { size_t o = off(k); return (unsigned)(bytes - o < chunk ? bytes - o : chunk); };
    for (size_t k = 0; k < mine && k < STAGES; ++k) bulk_g2s(stage + k * chunk, src + off(k), len(k), bar + k);
    for (size_t k = 0; k < mine; ++k) {
        const int s = (int)(k % STAGES);
        mbar_wait(bar + s, (unsigned)((k / STAGES) & 1));
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + off(k)), "r"(smem_u32(stage + (size_t)s * chunk)), "r"(len(k)) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (k >= 1 && k - 1 + STAGES < mine) {
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            const size_t r = k - 1 + STAGES;
            bulk_g2s(stage + (size_t)(r % STAGES) * chunk, src + off(r), len(r), bar + (r % STAGES));
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

struct Dev {
    int id;
    char *a, *b; // two buffers on this device
    cudaStream_t s;
    cudaEvent_t e0, e1;
};

int main(int argc, char **argv) {
    int want = 2;
    size_t bytes = (size_t)256 << 20;
    int reps = 10, ctas = 128;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--gpus") && i + 1 < argc) want = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--mib") && i + 1 < argc) bytes = (size_t)atol(argv[++i]) << 20;
        else if (!strcmp(argv[i], "--reps") && i + 1 < argc) reps = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--ctas") && i + 1 < argc) ctas = atoi(argv[++i]);
    }
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) {
        printf("{\"peer_copy_peak\": \"unavailable\", \"why\": \"%d GPU(s) visible, need 2\"}\n", ndev);
        return 0;
    }
    const int n = want < ndev ? want : ndev;
    std::vector<Dev> d((size_t)n);
    for (int i = 0; i < n; ++i) {
        d[i].id = i;
        CK(cudaSetDevice(i));
        for (int j = 0; j < n; ++j)
            if (j != i) {
                int can = 0;
                CK(cudaDeviceCanAccessPeer(&can, i, j));
                if (!can) {
                    printf("{\"peer_copy_peak\": \"unavailable\", \"why\": \"no peer access %d -> %d\"}\n", i, j);
                    return 0;
                }
                cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
                cudaGetLastError();
            }
        CK(cudaMalloc((void **)&d[i].a, bytes));
        CK(cudaMalloc((void **)&d[i].b, bytes));
        CK(cudaMemset(d[i].a, 1, bytes));
        CK(cudaMemset(d[i].b, 2, bytes));
        CK(cudaStreamCreateWithFlags(&d[i].s, cudaStreamNonBlocking));
        CK(cudaEventCreate(&d[i].e0));
        CK(cudaEventCreate(&d[i].e1));
    }
    const size_t nvec = bytes / 16;

    const unsigned chunk = 32768;
    const size_t bulk_smem = 128 + 4 * (size_t)chunk;
    for (int i = 0; i < n; ++i) {
        CK(cudaSetDevice(i));
        CK(cudaFuncSetAttribute(bulk_copy<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bulk_smem));
    }
    // mode: 0 copy engine, 1 kernel push (runs on the source), 2 kernel pull (runs on the destination),
    //       3 / 4 the same two with the bulk-async staged copy (32 KiB chunks, 4 stages, one CTA per SM)
    // pairs: list of (src, dst) transfers active at once; the time is the slowest stream's
    auto run = [&](const char *name, int mode, const std::vector<std::pair<int, int>> &pairs) {
        double best = 1e30;
        for (int rep = 0; rep < reps + 2; ++rep) {
            for (int i = 0; i < n; ++i) { CK(cudaSetDevice(i)); CK(cudaDeviceSynchronize()); }
            std::vector<int> used;
            for (auto &p : pairs) {
                const int runner = (mode == 2 || mode == 4) ? p.second : p.first;
                CK(cudaSetDevice(runner));
                bool first = true;
                for (int u : used) first = first && u != runner;
                if (first) { CK(cudaEventRecord(d[runner].e0, d[runner].s)); used.push_back(runner); }
                if (mode == 0) CK(cudaMemcpyPeerAsync(d[p.second].b, p.second, d[p.first].a, p.first, bytes, d[runner].s));
                else if (mode >= 3) bulk_copy<4><<<148, 32, bulk_smem, d[runner].s>>>(d[p.first].a, d[p.second].b, bytes, chunk);
                else copy_kernel<<<ctas, 512, 0, d[runner].s>>>((const uint4 *)d[p.first].a, (uint4 *)d[p.second].b, nvec);
            }
            double worst = 0;
            for (int u : used) {
                CK(cudaSetDevice(u));
                CK(cudaEventRecord(d[u].e1, d[u].s));
            }
            for (int u : used) {
                CK(cudaSetDevice(u));
                CK(cudaEventSynchronize(d[u].e1));
                float ms = 0;
                CK(cudaEventElapsedTime(&ms, d[u].e0, d[u].e1));
                if (ms > worst) worst = ms;
            }
            if (rep >= 2 && worst < best) best = worst;
        }
        // bytes one GPU sends in one direction during the measurement
        size_t out0 = 0;
        for (auto &p : pairs)
            if (p.first == pairs[0].first) out0 += bytes;
        printf("{\"measure\": \"%s\", \"gpus\": %d, \"mib_per_transfer\": %zu, \"transfers\": %zu, \"ms\": %.4f, \"GBps_per_direction_per_gpu\": %.1f}\n",
               name, n, bytes >> 20, pairs.size(), best, (double)out0 / (best * 1e-3) / 1e9);
        fflush(stdout);
    };

    run("copy_engine_one_direction", 0, {{0, 1}});
    run("copy_engine_both_directions", 0, {{0, 1}, {1, 0}});
    run("kernel_push_one_direction", 1, {{0, 1}});
    run("kernel_push_both_directions", 1, {{0, 1}, {1, 0}});
    run("kernel_pull_one_direction", 2, {{0, 1}});
    run("kernel_pull_both_directions", 2, {{0, 1}, {1, 0}});
    run("bulk_async_push_one_direction", 3, {{0, 1}});
    run("bulk_async_push_both_directions", 3, {{0, 1}, {1, 0}});
    run("bulk_async_pull_one_direction", 4, {{0, 1}});
    run("bulk_async_pull_both_directions", 4, {{0, 1}, {1, 0}});
    if (n > 2) {
        std::vector<std::pair<int, int>> all;
        for (int k = 1; k < n; ++k)
            for (int i = 0; i < n; ++i) all.push_back({i, (i + k) % n});
        run("kernel_push_all_to_all", 1, all);
        run("kernel_pull_all_to_all", 2, all);
    }
    return 0;
}
