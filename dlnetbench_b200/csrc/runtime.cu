// runtime.cu — host side of libdlnb.so: the C ABI declared in include/dlnb.h.
//
// One process per GPU. Owns: the rendezvous (bootstrap.cpp), the symmetric arena
// (cuMem VMM chunks exported as POSIX fds and mapped by every peer, plus their NVSwitch
// multicast view), the streams (emulated compute, two communication lanes, copies),
// per-slot events, communicators (rank lists + flag rows + control blocks), the
// point-to-point mailboxes, CUDA-graph capture of whole iterations and every kernel launch.
//
// Reference counterparts: cpp/proxy_classes.hpp (CCLCommunicator :135-253, Tensor
// :381-444), cpp/utils.hpp:74-117 (set_local_device), cpp/data_parallel/dp.cpp:166-232.
#include "../../include/dlnb.h"

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "arena_layout.hpp"
#include "aux_kernels.cuh"
#include "bootstrap.hpp"
#include "launch.hpp"

using namespace dlnb;

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
static std::string g_err;
static int fail(const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}
#define CU_TRY(expr)                                                                               \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---------------------------------------------------------------------------
// driver API entry points, resolved through the runtime so that libdlnb.so has no
// link-time dependency on libcuda (it must load on a box without a driver).
// ---------------------------------------------------------------------------
struct Drv {
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags);
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t);
    CUresult (*GetErrorString)(CUresult, const char **);
    // NVSwitch multicast (optional: absent or failing => the P2P kernels are used for everything)
    CUresult (*DeviceGet)(CUdevice *, int);
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice);
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *);
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags);
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
    bool ok = false, mc_ok = false;
};
static Drv drv;

static int load_driver() {
    if (drv.ok) return 0;
#define DLNB_SYM(field, name)                                                                      \
    do {                                                                                           \
        void *p = nullptr;                                                                         \
        cudaDriverEntryPointQueryResult q;                                                         \
        cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);                  \
        if (e != cudaSuccess || !p) return fail("driver symbol %s not available", name);           \
        drv.field = (decltype(drv.field))p;                                                        \
    } while (0)
    DLNB_SYM(MemGetAllocationGranularity, "cuMemGetAllocationGranularity");
    DLNB_SYM(MemAddressReserve, "cuMemAddressReserve");
    DLNB_SYM(MemAddressFree, "cuMemAddressFree");
    DLNB_SYM(MemCreate, "cuMemCreate");
    DLNB_SYM(MemRelease, "cuMemRelease");
    DLNB_SYM(MemExportToShareableHandle, "cuMemExportToShareableHandle");
    DLNB_SYM(MemImportFromShareableHandle, "cuMemImportFromShareableHandle");
    DLNB_SYM(MemMap, "cuMemMap");
    DLNB_SYM(MemUnmap, "cuMemUnmap");
    DLNB_SYM(MemSetAccess, "cuMemSetAccess");
    DLNB_SYM(GetErrorString, "cuGetErrorString");
#undef DLNB_SYM
    {
        struct { void **slot; const char *name; } opt[] = {
            {(void **)&drv.DeviceGet, "cuDeviceGet"},
            {(void **)&drv.DeviceGetAttribute, "cuDeviceGetAttribute"},
            {(void **)&drv.MulticastCreate, "cuMulticastCreate"},
            {(void **)&drv.MulticastAddDevice, "cuMulticastAddDevice"},
            {(void **)&drv.MulticastBindMem, "cuMulticastBindMem"},
            {(void **)&drv.MulticastGetGranularity, "cuMulticastGetGranularity"},
            {(void **)&drv.MulticastUnbind, "cuMulticastUnbind"},
        };
        drv.mc_ok = true;
        for (auto &o : opt) {
            void *p = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(o.name, &p, cudaEnableDefault, &q) != cudaSuccess || !p) drv.mc_ok = false;
            *o.slot = p;
        }
    }
    drv.ok = true;
    return 0;
}
static const char *drv_err(CUresult r) {
    const char *s = nullptr;
    if (drv.GetErrorString) drv.GetErrorString(r, &s);
    return s ? s : "unknown driver error";
}
#define DRV_TRY(expr)                                                                              \
    do {                                                                                           \
        CUresult r_ = (expr);                                                                      \
        if (r_ != CUDA_SUCCESS)                                                                    \
            return fail("%s failed: %s (%s:%d)", #expr, drv_err(r_), __FILE__, __LINE__);          \
    } while (0)

// ---------------------------------------------------------------------------
// state
// ---------------------------------------------------------------------------
enum ArenaKind { ARENA_LOCAL = 0, ARENA_VMM = 1, ARENA_IPC = 2 };

struct Chunk {
    char *local = nullptr;
    size_t size = 0;
    ChunkLayout layout; // which bytes of the chunk are handed out (identical on every rank)
    char *peer[MAXR] = {};                   // my mapping of world rank p's copy
    CUmemGenericAllocationHandle handle = 0; // VMM only
    std::vector<CUmemGenericAllocationHandle> imported;
    char *mc = nullptr;                         // multicast mapping of this chunk on all ranks
    CUmemGenericAllocationHandle mc_handle = 0;
};

struct dlnb_comm {
    int id = 0, n = 1, me = 0, dtype = DLNB_F32, num_slots = 1;
    int ranks[MAXR] = {};
    std::vector<cudaEvent_t> ev_start, ev_end; // timing pair, recorded when slot_timing is on
    std::vector<cudaEvent_t> ev_done;          // completion (no timestamp: cheaper), always recorded
    std::vector<char> slot_timed;
    int next_lane = 0;
    // a posted receive whose completion wait has not been enqueued yet: source world rank and message number
    std::vector<int> recv_src;
    std::vector<uint32_t> recv_seq;
    std::vector<char> slot_used;
    long ctas = 0, threads = 512, algo = DLNB_ALGO_AUTO, twoshot_min_bytes = -1 /* auto: 2 MiB / n */, timeout_ms = 30000,
         multimem_min_bytes = 512 << 10, mc_ctas = 0 /* grid of the in-switch kernels; 0 = by group size */,
         hybrid_pct = 0 /* share (percent) of a large allreduce that goes through the switch while the rest goes peer to peer */,
         hybrid_min_bytes = 32 << 20;
};

static constexpr int MAX_COMMS = 64;
static constexpr int SYS_FLAG_SET = MAX_COMMS * 2; // one more set of flag rows / generations: the rank_sync barrier
static constexpr size_t FLAG_BYTES_PER_COMM = (size_t)MAX_FLAG_CTAS * FLAG_ROW * sizeof(uint32_t);

struct State {
    bool inited = false;
    int rank = 0, world = 1, device = 0, sms = 0;
    ArenaKind kind = ARENA_LOCAL;
    Bootstrap boot;
    cudaStream_t s_compute = nullptr, s_comm = nullptr, s_comm2 = nullptr, s_copy = nullptr, s_out = nullptr;
    cudaEvent_t ev_gate = nullptr, ev_out = nullptr, ev_mark = nullptr;
    bool mark_set = false;
    bool out_dirty = false; // device -> host copies enqueued on s_out that the end of the iteration must wait for
    // Two communication lanes: consecutive small allreduces alternate between s_comm (lane 0) and
    // s_comm2 (lane 1) so their barrier latencies overlap; everything else runs on lane 0 alone.
    long lanes = 2;
    cudaEvent_t ev_lane1 = nullptr, ev_excl = nullptr;
    uint32_t p2p_sent[MAXR] = {}, p2p_posted[MAXR] = {}; // messages sent to / receives posted for each world rank
    uint32_t p2p_waited[MAXR] = {};                      // receives whose completion wait has been enqueued
    bool lane1_dirty = false, excl_pending = false, comm2_in_capture = false;
    cudaEvent_t ev_join = nullptr, ev_copy = nullptr, ev_iter0 = nullptr, ev_iter1 = nullptr, ev_tmp0 = nullptr,
                ev_tmp1 = nullptr;
    bool compute_dirty = false, copy_dirty = false;
    bool detach_compute = false; // between dlnb_batch_begin_concurrent and dlnb_batch_end: collectives do not wait for compute
    bool capturing = false, comm_in_capture = false; // between dlnb_graph_begin and dlnb_graph_end
    // arena
    std::vector<Chunk> chunks;
    struct Alloc { int chunk; size_t off, size; };
    std::map<const void *, Alloc> allocs; // live dlnb_symm_alloc blocks by local address
    size_t gran = 2u << 20, min_chunk = (size_t)1 << 30 /* a 134 MB bucket wastes 9 % of a 1 GiB chunk, 48 % of a 256 MiB one */, va_size = 0, va_used = 0;
    CUdeviceptr window[MAXR] = {};
    size_t bytes_in_use = 0;
    // NVSwitch multicast view of the arena (in-switch reduction for large allreduces)
    bool mc_enabled = false;
    size_t mc_gran = 0, mc_va_used = 0;
    CUdeviceptr mc_window = 0;
    CUdevice cu_dev = 0;
    // flags / control
    char *flag_area = nullptr; // symmetric: MAX_COMMS * FLAG_BYTES_PER_COMM
    CommCtl *ctl = nullptr;    // symmetric allocation too (only used locally)
    uint32_t *host_error = nullptr, *host_error_dev = nullptr;
    int next_comm = 0;
    std::vector<int> free_comm_ids; // ids of destroyed communicators (same order on every rank), reused after a reset
    std::vector<dlnb_comm *> comms;
    long strict_barriers = 0, align_iterations = 1, shift_overlap = 8, pdl = 1;
    // point-to-point mailboxes (symmetric, one per peer world rank) and sequence state (local)
    P2PBox *p2p_boxes = nullptr;
    P2PState *p2p_state = nullptr;
    // emulation scratch (single-GPU n-rank mode)
    uint32_t *emu_flags = nullptr;
    CommCtl *emu_ctl = nullptr;
    int emu_last_n = 0;
    // spin
    // footprint of the emulated compute: one 256-thread CTA on every SM, FMA-busy (spin_ctas 0 = one per SM)
    long spin_ctas = 0, spin_threads = 256, spin_mode = SPIN_FMA, spin_smem = 0, slot_timing = 1, spin_slack_ns = 8000;
    uint4 *spin_stream = nullptr; // SPIN_HBM: buffer larger than L2 that the spin kernel reads
    size_t spin_stream_vecs = 0;
    float *spin_sink = nullptr;
    int spin_regs = 0;
    unsigned long long *spin_stamps = nullptr;
    unsigned spin_seq = 0;
    double spin_overhead_us = 0.0, spin_slope = 1.0;
    // misc
    uint64_t launches = 0;
    void *flush_buf = nullptr;
    size_t flush_bytes = 0;
    unsigned long long *d_scalar = nullptr; // 8-byte device scratch for verify/checksum
    unsigned *d_worst = nullptr;            // 4-byte device scratch of the tolerance verifier
};
static State G;

#define NEED_INIT()                                                                                \
    do {                                                                                           \
        if (!G.inited) return fail("dlnb_init has not been called");                               \
    } while (0)

static size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static size_t dtype_size(int dt) { return dt == DLNB_F32 ? 4 : 2; }

// ---------------------------------------------------------------------------
// arena
// ---------------------------------------------------------------------------
// true iff `mine` is true on every rank (ranks must take the same decision about optional features)
static bool all_ranks_agree(bool mine) {
    std::vector<char> all((size_t)G.world);
    char m = mine ? 1 : 0;
    if (!G.boot.allgather(&m, 1, all.data()).empty()) return false;
    for (char c : all)
        if (!c) return false;
    return true;
}

// Multicast view of a freshly mapped chunk: rank 0 creates the multicast object, everybody adds
// its device and binds its own physical chunk at offset 0, then maps the object. Any failure on
// any rank switches multicast off everywhere (the P2P kernels need nothing of this).
static void chunk_add_multicast(Chunk &c) {
    if (!G.mc_enabled) return;
    CUmulticastObjectProp mp = {};
    mp.numDevices = (unsigned)G.world;
    mp.size = c.size;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle mc = 0;
    int fd = -1, got = -1;
    bool ok = true;
    if (G.rank == 0) {
        ok = drv.MulticastCreate(&mc, &mp) == CUDA_SUCCESS &&
             drv.MemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
    }
    if (!all_ranks_agree(ok)) { if (mc) drv.MemRelease(mc); if (fd >= 0) ::close(fd); G.mc_enabled = false; return; }
    if (!G.boot.bcast_fd(fd, got).empty()) ok = false;
    if (fd >= 0) ::close(fd);
    if (ok && G.rank != 0)
        ok = drv.MemImportFromShareableHandle(&mc, (void *)(uintptr_t)got, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
    if (got >= 0) ::close(got);
    if (ok) ok = drv.MulticastAddDevice(mc, G.cu_dev) == CUDA_SUCCESS;
    if (!all_ranks_agree(ok)) { if (mc) drv.MemRelease(mc); G.mc_enabled = false; return; } // all devices added before any bind
    ok = drv.MulticastBindMem(mc, 0, c.handle, 0, c.size, 0) == CUDA_SUCCESS;
    CUdeviceptr va = G.mc_window + G.mc_va_used;
    if (ok) ok = drv.MemMap(va, c.size, 0, mc, 0) == CUDA_SUCCESS;
    if (ok) {
        CUmemAccessDesc acc = {};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = G.device;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        ok = drv.MemSetAccess(va, c.size, &acc, 1) == CUDA_SUCCESS;
    }
    if (!all_ranks_agree(ok)) { G.mc_enabled = false; return; } // leaks this one object; the process goes on with P2P
    c.mc = (char *)va;
    c.mc_handle = mc;
    G.mc_va_used += c.size;
}

static int arena_grow(size_t need) {
    Chunk c;
    const size_t want = std::max(need, G.min_chunk);
    if (G.kind == ARENA_LOCAL) {
        c.size = round_up(want, 2u << 20);
        void *p = nullptr;
        CU_TRY(cudaMalloc(&p, c.size));
        c.local = (char *)p;
        c.peer[0] = c.local;
        for (int r = 0; r < MAXR; ++r) c.peer[r] = (r == G.rank) ? c.local : nullptr;
    } else if (G.kind == ARENA_VMM) {
        c.size = round_up(want, G.gran);
        if (G.va_used + c.size > G.va_size)
            return fail("symmetric arena exhausted its %zu GiB of reserved address space", G.va_size >> 30);
        CUmemAllocationProp prop = {};
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = G.device;
        prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        DRV_TRY(drv.MemCreate(&c.handle, c.size, &prop, 0));
        int fd = -1;
        DRV_TRY(drv.MemExportToShareableHandle(&fd, c.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
        // consistency: every rank must be growing by the same amount
        std::vector<size_t> sizes((size_t)G.world);
        std::string e = G.boot.allgather(&c.size, sizeof(size_t), sizes.data());
        if (!e.empty()) return fail("%s", e.c_str());
        for (int r = 0; r < G.world; ++r)
            if (sizes[r] != c.size)
                return fail("dlnb_symm_alloc is collective: rank %d grows by %zu bytes, rank %d by %zu", G.rank,
                            c.size, r, sizes[r]);
        std::vector<int> fds;
        e = G.boot.exchange_fds(fd, fds);
        if (!e.empty()) return fail("%s", e.c_str());
        ::close(fd);
        CUmemAccessDesc acc = {};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = G.device;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        c.imported.assign((size_t)G.world, 0);
        for (int r = 0; r < G.world; ++r) {
            CUdeviceptr va = G.window[r] + G.va_used;
            CUmemGenericAllocationHandle h = c.handle;
            if (r != G.rank) {
                DRV_TRY(drv.MemImportFromShareableHandle(&h, (void *)(uintptr_t)fds[r],
                                                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
                c.imported[r] = h;
            }
            DRV_TRY(drv.MemMap(va, c.size, 0, h, 0));
            DRV_TRY(drv.MemSetAccess(va, c.size, &acc, 1));
            c.peer[r] = (char *)va;
            ::close(fds[r]);
        }
        c.local = c.peer[G.rank];
        G.va_used += c.size;
        chunk_add_multicast(c);
    } else { // ARENA_IPC (diagnostic fallback: cudaMalloc + cudaIpc handles)
        c.size = round_up(want, 2u << 20);
        void *p = nullptr;
        CU_TRY(cudaMalloc(&p, c.size));
        c.local = (char *)p;
        cudaIpcMemHandle_t mine;
        CU_TRY(cudaIpcGetMemHandle(&mine, p));
        std::vector<cudaIpcMemHandle_t> all((size_t)G.world);
        std::string e = G.boot.allgather(&mine, sizeof(mine), all.data());
        if (!e.empty()) return fail("%s", e.c_str());
        for (int r = 0; r < G.world; ++r) {
            if (r == G.rank) {
                c.peer[r] = c.local;
            } else {
                void *q = nullptr;
                CU_TRY(cudaIpcOpenMemHandle(&q, all[r], cudaIpcMemLazyEnablePeerAccess));
                c.peer[r] = (char *)q;
            }
        }
    }
    CU_TRY(cudaMemset(c.local, 0, c.size));
    CU_TRY(cudaDeviceSynchronize());
    if (G.boot.active()) {
        std::string e = G.boot.barrier(); // peers' mappings exist before anyone uses them
        if (!e.empty()) return fail("%s", e.c_str());
    }
    c.layout = ChunkLayout(c.size);
    G.chunks.push_back(std::move(c));
    return 0;
}

static int arena_sync_all_ranks() {
    CU_TRY(cudaDeviceSynchronize());
    if (G.boot.active()) {
        std::string e = G.boot.barrier();
        if (!e.empty()) return fail("%s", e.c_str());
    }
    return 0;
}

static int arena_alloc(size_t bytes, void **out) {
    const size_t need = round_up(std::max<size_t>(bytes, 1), 1024);
    int ci = -1;
    size_t off = 0;
    bool recycled = false;
    for (size_t i = 0; i < G.chunks.size() && ci < 0; ++i)
        if (G.chunks[i].layout.take(need, &off, &recycled)) ci = (int)i;
    if (ci < 0) {
        if (arena_grow(need)) return 1;
        ci = (int)G.chunks.size() - 1;
        if (!G.chunks[(size_t)ci].layout.take(need, &off, &recycled)) return fail("symmetric arena: a fresh chunk cannot hold %zu bytes", need);
    }
    Chunk &c = G.chunks[(size_t)ci];
    if (recycled) {
        // The block has held data. Peers may still be reading or writing the old contents through
        // their mappings, and a peer's kernel may push into the new block as soon as that peer's
        // dlnb_symm_alloc returns: everybody idle -> clear -> everybody cleared.
        if (arena_sync_all_ranks()) return 1;
        CU_TRY(cudaMemset(c.local + off, 0, need));
        if (arena_sync_all_ranks()) return 1;
    }
    *out = c.local + off;
    G.allocs[*out] = State::Alloc{ci, off, need};
    G.bytes_in_use += need;
    return 0;
}

static int arena_free(void *p) {
    auto it = G.allocs.find(p);
    if (it == G.allocs.end()) return fail("dlnb_symm_free: %p is not a live dlnb_symm_alloc block", p);
    G.chunks[(size_t)it->second.chunk].layout.give_back(it->second.off, it->second.size);
    G.bytes_in_use -= it->second.size;
    G.allocs.erase(it);
    return 0;
}

// my mapping of world rank `wr`'s copy of the symmetric allocation holding `p`
static char *translate(const void *p, int wr) {
    if (wr == G.rank) return (char *)p;
    const char *q = (const char *)p;
    for (const Chunk &c : G.chunks)
        if (q >= c.local && q < c.local + c.size) return c.peer[wr] ? c.peer[wr] + (q - c.local) : nullptr;
    return nullptr;
}

// multicast address of the symmetric allocation holding `p` (null if it has none)
static char *translate_mc(const void *p) {
    const char *q = (const char *)p;
    for (const Chunk &c : G.chunks)
        if (q >= c.local && q < c.local + c.size) return c.mc ? c.mc + (q - c.local) : nullptr;
    return nullptr;
}

// ---------------------------------------------------------------------------
// init / finalize
// ---------------------------------------------------------------------------
extern "C" const char *dlnb_last_error(void) { return g_err.c_str(); }
extern "C" const char *dlnb_version(void) { return "dlnb 0.1 (sm_100a)"; }
extern "C" int dlnb_rank(void) { return G.rank; }
extern "C" int dlnb_world(void) { return G.world; }
extern "C" int dlnb_device(void) { return G.device; }
extern "C" const char *dlnb_arena_kind(void) {
    return G.kind == ARENA_VMM ? "vmm" : G.kind == ARENA_IPC ? "ipc" : "local";
}
extern "C" uint64_t dlnb_launch_count(void) { return G.launches; }
extern "C" size_t dlnb_symm_bytes_in_use(void) { return G.bytes_in_use; }

static int ensure_spin_stream();
extern "C" int dlnb_init(int rank, int world, int device, const char *session) {
    if (G.inited) return fail("dlnb_init called twice");
    if (world < 1 || world > MAXR) return fail("world size %d not in 1..%d (one NVSwitch box)", world, MAXR);
    if (rank < 0 || rank >= world) return fail("rank %d out of range for world %d", rank, world);
    int ndev = 0;
    CU_TRY(cudaGetDeviceCount(&ndev));
    if (ndev <= 0) return fail("no CUDA device visible: libdlnb has no CPU fallback");
    if (device < 0 || device >= ndev) return fail("device %d out of range (%d visible)", device, ndev);
    CU_TRY(cudaSetDevice(device));
    CU_TRY(cudaFree(0));
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10 && !getenv("DLNB_ALLOW_ANY_ARCH"))
        return fail("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major,
                    prop.minor);
    G.rank = rank;
    G.world = world;
    G.device = device;
    G.sms = prop.multiProcessorCount;

    if (world > 1) {
        const char *k = getenv("DLNB_ARENA");
        G.kind = (k && !strcmp(k, "ipc")) ? ARENA_IPC : ARENA_VMM;
        std::string e = G.boot.connect(rank, world, session ? session : "default", 120);
        if (!e.empty()) return fail("%s", e.c_str());
    } else {
        G.kind = ARENA_LOCAL;
    }
    if (const char *m = getenv("DLNB_CHUNK_MB")) G.min_chunk = (size_t)atol(m) << 20;
    if (const char *m = getenv("DLNB_SPIN_SLACK_NS")) G.spin_slack_ns = std::max(0L, atol(m));

    if (G.kind == ARENA_VMM) {
        if (load_driver()) return 1;
        CUmemAllocationProp ap = {};
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = device;
        ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        DRV_TRY(drv.MemGetAllocationGranularity(&G.gran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
        size_t gb = 256;
        if (const char *v = getenv("DLNB_ARENA_VA_GB")) gb = (size_t)atol(v);
        G.va_size = gb << 30;
        for (int r = 0; r < world; ++r) DRV_TRY(drv.MemAddressReserve(&G.window[r], G.va_size, G.gran, 0, 0));
        // optional: NVSwitch multicast view of the arena
        const char *mcenv = getenv("DLNB_MULTICAST");
        bool want = drv.mc_ok && !(mcenv && !strcmp(mcenv, "0"));
        if (want) {
            int sup = 0;
            want = drv.DeviceGet(&G.cu_dev, device) == CUDA_SUCCESS &&
                   drv.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, G.cu_dev) == CUDA_SUCCESS && sup;
        }
        if (want) {
            CUmulticastObjectProp mp = {};
            mp.numDevices = (unsigned)world;
            mp.size = G.min_chunk;
            mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
            size_t g = 0;
            want = drv.MulticastGetGranularity(&g, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g > 0 && g <= (1u << 30);
            if (want) G.mc_gran = g;
        }
        if (want) want = drv.MemAddressReserve(&G.mc_window, G.va_size, std::max(G.gran, G.mc_gran), 0, 0) == CUDA_SUCCESS;
        G.mc_enabled = all_ranks_agree(want);
        if (G.mc_enabled) {
            G.gran = std::max(G.gran, G.mc_gran); // chunk sizes must suit both the allocation and the multicast object
            G.min_chunk = std::max(G.min_chunk, G.gran);
        }
        if (getenv("DLNB_DEBUG") && rank == 0)
            fprintf(stderr, "[dlnb] multicast %s (granularity %zu MiB)\n", G.mc_enabled ? "on" : "off", G.mc_gran >> 20);
    }

    int lo = 0, hi = 0;
    CU_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi)); // lo = least priority (numerically greatest)
    CU_TRY(cudaStreamCreateWithPriority(&G.s_compute, cudaStreamNonBlocking, lo));
    CU_TRY(cudaStreamCreateWithPriority(&G.s_comm, cudaStreamNonBlocking, hi));
    CU_TRY(cudaStreamCreateWithPriority(&G.s_comm2, cudaStreamNonBlocking, hi));
    CU_TRY(cudaEventCreateWithFlags(&G.ev_lane1, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&G.ev_excl, cudaEventDisableTiming));
    if (const char *v = getenv("DLNB_LANES")) G.lanes = atol(v) >= 2 ? 2 : 1;
    CU_TRY(cudaStreamCreateWithPriority(&G.s_copy, cudaStreamNonBlocking, lo));
    CU_TRY(cudaStreamCreateWithPriority(&G.s_out, cudaStreamNonBlocking, lo));
    CU_TRY(cudaEventCreateWithFlags(&G.ev_gate, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&G.ev_out, cudaEventDisableTiming));
    CU_TRY(cudaEventCreate(&G.ev_mark));
    CU_TRY(cudaEventCreateWithFlags(&G.ev_join, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&G.ev_copy, cudaEventDisableTiming));
    CU_TRY(cudaEventCreate(&G.ev_iter0));
    CU_TRY(cudaEventCreate(&G.ev_iter1));
    CU_TRY(cudaEventCreate(&G.ev_tmp0));
    CU_TRY(cudaEventCreate(&G.ev_tmp1));
    CU_TRY(cudaHostAlloc((void **)&G.host_error, sizeof(uint32_t), cudaHostAllocMapped));
    *G.host_error = 0;
    CU_TRY(cudaHostGetDevicePointer((void **)&G.host_error_dev, G.host_error, 0));
    CU_TRY(cudaMalloc((void **)&G.d_scalar, sizeof(unsigned long long)));
    CU_TRY(cudaMalloc((void **)&G.spin_stamps, 4 * sizeof(unsigned long long))); // [0,1] compute chain, [2,3] gate spins
    CU_TRY(cudaMemset(G.spin_stamps, 0, 4 * sizeof(unsigned long long)));
    CU_TRY(cudaMalloc((void **)&G.spin_sink, sizeof(float)));
    {
        cudaFuncAttributes fa;
        CU_TRY(cudaFuncGetAttributes(&fa, spin_kernel));
        G.spin_regs = fa.numRegs;
    }
    if (const char *v = getenv("DLNB_SPIN_MODE")) G.spin_mode = std::min(2L, std::max(0L, atol(v)));
    if (G.spin_mode == SPIN_HBM && ensure_spin_stream()) return 1;
    if (const char *v = getenv("DLNB_SPIN_CTAS")) G.spin_ctas = std::max(0L, atol(v));
    if (const char *v = getenv("DLNB_SPIN_THREADS")) G.spin_threads = std::min(1024L, std::max(32L, atol(v)));
    if (const char *v = getenv("DLNB_STRICT_BARRIERS")) G.strict_barriers = atol(v) != 0;
    if (const char *v = getenv("DLNB_ALIGN")) G.align_iterations = atol(v) != 0;
    if (const char *v = getenv("DLNB_PDL")) G.pdl = atol(v) != 0;
    if (const char *v = getenv("DLNB_SHIFT_OVERLAP")) { const long o = atol(v); if (o == 1 || o == 2 || o == 8) G.shift_overlap = o; }

    G.inited = true;
    // first symmetric allocations: flag rows and control blocks of all communicators
    void *p = nullptr;
    if (arena_alloc((size_t)(MAX_COMMS * 2 + 1) * FLAG_BYTES_PER_COMM, &p)) { G.inited = false; return 1; } // x2: one set per lane; +1: rank_sync
    G.flag_area = (char *)p;
    if (arena_alloc((size_t)(MAX_COMMS * 2 + 1) * sizeof(CommCtl), &p)) { G.inited = false; return 1; }
    G.ctl = (CommCtl *)p;
    if (arena_alloc((size_t)MAXR * sizeof(P2PBox), &p)) { G.inited = false; return 1; }
    G.p2p_boxes = (P2PBox *)p;
    CU_TRY(cudaMalloc((void **)&G.p2p_state, sizeof(P2PState)));
    CU_TRY(cudaMemset(G.p2p_state, 0, sizeof(P2PState)));
    return 0;
}

extern "C" int dlnb_finalize(void) {
    if (!G.inited) return 0;
    cudaDeviceSynchronize();
    if (G.boot.active()) G.boot.barrier(); // nobody unmaps while a peer may still touch it
    for (dlnb_comm *c : G.comms) {
        if (!c) continue;
        for (auto e : c->ev_start) cudaEventDestroy(e);
        for (auto e : c->ev_end) cudaEventDestroy(e);
        for (auto e : c->ev_done) cudaEventDestroy(e);
        delete c;
    }
    G.comms.clear();
    for (Chunk &c : G.chunks) {
        if (G.kind == ARENA_VMM) {
            if (c.mc) {
                drv.MemUnmap((CUdeviceptr)c.mc, c.size);
                drv.MulticastUnbind(c.mc_handle, G.cu_dev, 0, c.size);
                drv.MemRelease(c.mc_handle);
            }
            for (int r = 0; r < G.world; ++r)
                if (c.peer[r]) drv.MemUnmap((CUdeviceptr)c.peer[r], c.size);
            for (auto h : c.imported)
                if (h) drv.MemRelease(h);
            drv.MemRelease(c.handle);
        } else if (G.kind == ARENA_IPC) {
            for (int r = 0; r < G.world; ++r)
                if (r != G.rank && c.peer[r]) cudaIpcCloseMemHandle(c.peer[r]);
            if (G.boot.active()) G.boot.barrier();
            cudaFree(c.local);
        } else {
            cudaFree(c.local);
        }
    }
    G.chunks.clear();
    G.allocs.clear();
    if (G.kind == ARENA_VMM)
        for (int r = 0; r < G.world; ++r)
            if (G.window[r]) drv.MemAddressFree(G.window[r], G.va_size);
    if (G.kind == ARENA_VMM && G.mc_window) drv.MemAddressFree(G.mc_window, G.va_size);
    if (G.emu_flags) cudaFree(G.emu_flags);
    if (G.emu_ctl) cudaFree(G.emu_ctl);
    if (G.flush_buf) cudaFree(G.flush_buf);
    if (G.d_scalar) cudaFree(G.d_scalar);
    if (G.d_worst) cudaFree(G.d_worst);
    if (G.p2p_state) cudaFree(G.p2p_state);
    if (G.spin_stamps) cudaFree(G.spin_stamps);
    if (G.spin_sink) cudaFree(G.spin_sink);
    if (G.spin_stream) cudaFree(G.spin_stream);
    if (G.host_error) cudaFreeHost(G.host_error);
    cudaStreamDestroy(G.s_compute);
    cudaStreamDestroy(G.s_comm);
    cudaStreamDestroy(G.s_comm2);
    cudaEventDestroy(G.ev_lane1);
    cudaEventDestroy(G.ev_excl);
    cudaStreamDestroy(G.s_copy);
    cudaStreamDestroy(G.s_out);
    cudaEventDestroy(G.ev_gate);
    cudaEventDestroy(G.ev_out);
    cudaEventDestroy(G.ev_mark);
    cudaEventDestroy(G.ev_join);
    cudaEventDestroy(G.ev_copy);
    cudaEventDestroy(G.ev_iter0);
    cudaEventDestroy(G.ev_iter1);
    cudaEventDestroy(G.ev_tmp0);
    cudaEventDestroy(G.ev_tmp1);
    G.boot.close_all();
    G = State();
    return 0;
}

extern "C" int dlnb_host_barrier(void) {
    NEED_INIT();
    std::string e = G.boot.barrier();
    return e.empty() ? 0 : fail("%s", e.c_str());
}
extern "C" int dlnb_host_allgather(const void *in, size_t bytes, void *out) {
    NEED_INIT();
    std::string e = G.boot.allgather(in, bytes, out);
    return e.empty() ? 0 : fail("%s", e.c_str());
}

extern "C" int dlnb_symm_alloc(size_t bytes, void **ptr) {
    NEED_INIT();
    if (!ptr) return fail("dlnb_symm_alloc: null out pointer");
    return arena_alloc(bytes, ptr);
}
extern "C" int dlnb_symm_free(void *ptr) {
    NEED_INIT();
    if (!ptr) return 0;
    return arena_free(ptr);
}

// SPIN_HBM streams a buffer larger than L2: allocated on first use (1 GiB: eight times the L2)
static int ensure_spin_stream() {
    if (G.spin_stream) return 0;
    G.spin_stream_vecs = ((size_t)1 << 30) / sizeof(uint4);
    CU_TRY(cudaMalloc((void **)&G.spin_stream, G.spin_stream_vecs * sizeof(uint4)));
    CU_TRY(cudaMemset(G.spin_stream, 0x11, G.spin_stream_vecs * sizeof(uint4)));
    return 0;
}

extern "C" int dlnb_set(const char *key, long value) {
    if (!key) return fail("dlnb_set: null key");
    if (!strcmp(key, "spin_ctas")) G.spin_ctas = std::max(0L, value); // 0 = one CTA per SM
    else if (!strcmp(key, "spin_threads")) G.spin_threads = std::min(1024L, std::max(1L, value));
    else if (!strcmp(key, "spin_mode")) {
        if (value < SPIN_SLEEP || value > SPIN_HBM) return fail("dlnb_set: spin_mode %ld not in 0..2", value);
        if (value == SPIN_HBM && !G.spin_stream) {
            NEED_INIT();
            if (G.capturing) return fail("dlnb_set: spin_mode cannot change inside a graph capture");
            if (ensure_spin_stream()) return 1;
        }
        G.spin_mode = value;
    } else if (!strcmp(key, "spin_smem")) {
        if (value < 0 || value > 227 * 1024) return fail("dlnb_set: spin_smem %ld not in 0..232448", value);
        if (value > 48 * 1024) CU_TRY(cudaFuncSetAttribute(spin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)value));
        G.spin_smem = value;
    } else if (!strcmp(key, "strict_barriers")) G.strict_barriers = value != 0;
    else if (!strcmp(key, "align_iterations")) G.align_iterations = value != 0;
    else if (!strcmp(key, "pdl")) G.pdl = value != 0;
    else if (!strcmp(key, "shift_overlap")) {
        if (value != 1 && value != 2 && value != 8) return fail("dlnb_set: shift_overlap must be 1, 2 or 8 (vectors), got %ld", value);
        G.shift_overlap = value;
    }
    else if (!strcmp(key, "slot_timing")) G.slot_timing = value;
    else if (!strcmp(key, "spin_slack_ns")) G.spin_slack_ns = std::max(0L, value);
    else if (!strcmp(key, "lanes")) G.lanes = value >= 2 ? 2 : 1;
    else return fail("dlnb_set: unknown key %s", key);
    return 0;
}
extern "C" long dlnb_get(const char *key) {
    if (!key) return -1;
    if (!strcmp(key, "spin_ctas")) return G.spin_ctas > 0 ? G.spin_ctas : G.sms;
    if (!strcmp(key, "spin_threads")) return G.spin_threads;
    if (!strcmp(key, "spin_mode")) return G.spin_mode;
    if (!strcmp(key, "spin_smem")) return G.spin_smem;
    if (!strcmp(key, "spin_regs")) return G.spin_regs;
    if (!strcmp(key, "strict_barriers")) return G.strict_barriers;
    if (!strcmp(key, "align_iterations")) return G.align_iterations;
    if (!strcmp(key, "shift_overlap")) return G.shift_overlap;
    if (!strcmp(key, "pdl")) return G.pdl;
    // 1 when a CTA of the emulated compute and a 512-thread collective CTA (96 registers) fit on one SM together
    if (!strcmp(key, "spin_coresident")) {
        const long per_sm = ((G.spin_ctas > 0 ? G.spin_ctas : G.sms) + G.sms - 1) / G.sms;
        const long regs = per_sm * G.spin_threads * ((G.spin_regs + 7) / 8 * 8);
        return regs <= 65536 - 512 * 96 && per_sm * G.spin_threads <= 2048 - 512 && per_sm * G.spin_smem <= 227 * 1024;
    }
    if (!strcmp(key, "slot_timing")) return G.slot_timing;
    if (!strcmp(key, "spin_slack_ns")) return G.spin_slack_ns;
    if (!strcmp(key, "sms")) return G.sms;
    if (!strcmp(key, "multicast")) return G.mc_enabled ? 1 : 0;
    if (!strcmp(key, "lanes")) return G.lanes;
    return -1;
}

// ---------------------------------------------------------------------------
// communicators
// ---------------------------------------------------------------------------
extern "C" int dlnb_comm_create(const int *ranks, int nranks, int num_slots, int dtype, dlnb_comm_t *out) {
    NEED_INIT();
    if (!ranks || !out) return fail("dlnb_comm_create: null argument");
    if (nranks < 1 || nranks > MAXR) return fail("dlnb_comm_create: group size %d not in 1..%d", nranks, MAXR);
    if (dtype < DLNB_F32 || dtype > DLNB_BF16) return fail("dlnb_comm_create: unknown dtype %d", dtype);
    if (G.next_comm >= MAX_COMMS && G.free_comm_ids.empty())
        return fail("dlnb_comm_create: more than %d live communicators", MAX_COMMS);
    if (G.capturing) return fail("dlnb_comm_create inside a graph capture");
    dlnb_comm *c = new dlnb_comm;
    c->me = -1;
    for (int i = 0; i < nranks; ++i) {
        if (ranks[i] < 0 || ranks[i] >= G.world) {
            delete c;
            return fail("dlnb_comm_create: rank %d outside world %d", ranks[i], G.world);
        }
        c->ranks[i] = ranks[i];
        if (ranks[i] == G.rank) c->me = i;
    }
    if (c->me < 0) {
        delete c;
        return fail("dlnb_comm_create: calling rank %d is not in the group", G.rank);
    }
    if (G.next_comm < MAX_COMMS) {
        c->id = G.next_comm++;
    } else {
        // Reuse the flag rows of a destroyed communicator (creation and destruction are collective over
        // the world in the same order, so every rank picks the same id). The rows and their generations
        // carry the old group's history — and the new group may be another subset of ranks — so they
        // restart from zero on every rank: everybody idle -> clear -> everybody cleared.
        c->id = G.free_comm_ids.front();
        G.free_comm_ids.erase(G.free_comm_ids.begin());
        if (arena_sync_all_ranks()) { delete c; return 1; }
        CU_TRY(cudaMemset(G.flag_area + (size_t)c->id * 2 * FLAG_BYTES_PER_COMM, 0, 2 * FLAG_BYTES_PER_COMM));
        CU_TRY(cudaMemset(G.ctl + c->id * 2, 0, 2 * sizeof(CommCtl)));
        if (arena_sync_all_ranks()) { delete c; return 1; }
    }
    c->n = nranks;
    c->dtype = dtype;
    c->num_slots = std::max(1, num_slots);
    // one extra, internal slot (index num_slots) for the blocking calls
    c->ev_start.resize(c->num_slots + 1);
    c->ev_end.resize(c->num_slots + 1);
    c->ev_done.resize(c->num_slots + 1);
    c->slot_timed.assign(c->num_slots + 1, 0);
    c->recv_src.assign(c->num_slots + 1, -1);
    c->recv_seq.assign(c->num_slots + 1, 0);
    c->slot_used.assign(c->num_slots + 1, 0);
    for (int i = 0; i <= c->num_slots; ++i) {
        CU_TRY(cudaEventCreate(&c->ev_start[i]));
        CU_TRY(cudaEventCreate(&c->ev_end[i]));
        CU_TRY(cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming));
    }
    if (const char *v = getenv("DLNB_CTAS")) c->ctas = atol(v);
    if (const char *v = getenv("DLNB_THREADS")) {
        const long t = atol(v);
        if (t >= 32 && t <= 512 && t % 32 == 0) c->threads = t;
    }
    if (const char *v = getenv("DLNB_ALGO")) c->algo = atol(v);
    if (const char *v = getenv("DLNB_TWOSHOT_MIN_BYTES")) c->twoshot_min_bytes = atol(v);
    if (const char *v = getenv("DLNB_MULTIMEM_MIN_BYTES")) c->multimem_min_bytes = atol(v);
    if (const char *v = getenv("DLNB_MC_CTAS")) c->mc_ctas = atol(v);
    if (const char *v = getenv("DLNB_HYBRID_PCT")) c->hybrid_pct = std::min(100L, std::max(0L, atol(v)));
    G.comms.push_back(c);
    *out = c;
    return 0;
}
extern "C" int dlnb_comm_destroy(dlnb_comm_t c) {
    if (!c) return 0;
    bool live = false;
    for (auto &slot : G.comms)
        if (slot == c) { slot = nullptr; live = true; }
    if (!live) return fail("dlnb_comm_destroy: not a live communicator");
    G.free_comm_ids.push_back(c->id);
    for (auto e : c->ev_start) cudaEventDestroy(e);
    for (auto e : c->ev_end) cudaEventDestroy(e);
    for (auto e : c->ev_done) cudaEventDestroy(e);
    delete c;
    return 0;
}
extern "C" int dlnb_comm_rank(dlnb_comm_t c) { return c ? c->me : -1; }
extern "C" int dlnb_comm_size(dlnb_comm_t c) { return c ? c->n : -1; }
extern "C" int dlnb_comm_set(dlnb_comm_t c, const char *key, long value) {
    if (!c || !key) return fail("dlnb_comm_set: null argument");
    if (!strcmp(key, "ctas")) c->ctas = value;
    else if (!strcmp(key, "threads")) {
        // the tile layout and the flag barrier assume whole warps and the kernels' launch bound
        if (value < 32 || value > 512 || value % 32) return fail("dlnb_comm_set: threads must be a multiple of 32 in 32..512, got %ld", value);
        c->threads = value;
    }
    else if (!strcmp(key, "algo")) c->algo = value;
    else if (!strcmp(key, "twoshot_min_bytes")) c->twoshot_min_bytes = value;
    else if (!strcmp(key, "multimem_min_bytes")) c->multimem_min_bytes = value;
    else if (!strcmp(key, "mc_ctas")) c->mc_ctas = value;
    else if (!strcmp(key, "hybrid_pct")) {
        if (value < 0 || value > 100) return fail("dlnb_comm_set: hybrid_pct %ld not in 0..100", value);
        c->hybrid_pct = value;
    } else if (!strcmp(key, "hybrid_min_bytes")) c->hybrid_min_bytes = value;
    else if (!strcmp(key, "timeout_ms")) c->timeout_ms = value;
    else return fail("dlnb_comm_set: unknown key %s", key);
    return 0;
}
extern "C" long dlnb_comm_get(dlnb_comm_t c, const char *key) {
    if (!c || !key) return -1;
    if (!strcmp(key, "ctas")) return c->ctas;
    if (!strcmp(key, "threads")) return c->threads;
    if (!strcmp(key, "algo")) return c->algo;
    if (!strcmp(key, "twoshot_min_bytes")) return c->twoshot_min_bytes;
    if (!strcmp(key, "multimem_min_bytes")) return c->multimem_min_bytes;
    if (!strcmp(key, "mc_ctas")) return c->mc_ctas;
    if (!strcmp(key, "hybrid_pct")) return c->hybrid_pct;
    if (!strcmp(key, "hybrid_min_bytes")) return c->hybrid_min_bytes;
    if (!strcmp(key, "timeout_ms")) return c->timeout_ms;
    if (!strcmp(key, "id")) return c->id;
    return -1;
}

// ---------------------------------------------------------------------------
// emulated compute
// ---------------------------------------------------------------------------
static int launch_spin(double ns) {
    const unsigned ctas = (unsigned)(G.spin_ctas > 0 ? G.spin_ctas : G.sms);
    spin_kernel<<<ctas, (unsigned)G.spin_threads, (size_t)G.spin_smem, G.s_compute>>>(
        (unsigned long long)(ns + 0.5), G.spin_stamps, G.spin_seq++ & 1u, (unsigned long long)G.spin_slack_ns,
        (int)G.spin_mode, G.spin_stream, G.spin_stream_vecs, G.spin_sink);
    CU_TRY(cudaGetLastError());
    G.launches++;
    G.compute_dirty = true;
    return 0;
}
extern "C" int dlnb_compute_spin(double usec) {
    NEED_INIT();
    if (!(usec > 0)) return 0;
    // only the clock-rate correction is applied: the start-up latency of the first spin of a
    // chain is genuine emulation overhead and stays in the measurement
    return launch_spin(usec * 1000.0 / G.spin_slope);
}
extern "C" int dlnb_compute_sync(void) {
    NEED_INIT();
    if (G.capturing) return 0;
    CU_TRY(cudaStreamSynchronize(G.s_compute));
    return 0;
}
// "The compute enqueued so far ends HERE": a timestamp on the compute stream, against which dlnb_slot_exposed()
// measures how much of a collective was still running after that compute had finished — the part of a Wait() that
// the reference's host, which sleeps through the compute first (usleep), would actually have waited.
extern "C" int dlnb_compute_mark(void) {
    NEED_INIT();
    if (G.capturing) return 0;
    CU_TRY(cudaEventRecord(G.ev_mark, G.s_compute));
    G.mark_set = true;
    return 0;
}
extern "C" int dlnb_slot_exposed(dlnb_comm_t c, int slot, float *ms) {
    NEED_INIT();
    if (!c || !ms) return fail("null argument");
    *ms = 0.f;
    if (G.capturing || !G.mark_set) return 0;
    if (slot == -1) slot = c->num_slots;
    if (slot < 0 || slot > c->num_slots || !c->slot_used[slot] || !c->slot_timed[slot]) return 0;
    CU_TRY(cudaEventSynchronize(G.ev_mark));
    CU_TRY(cudaEventSynchronize(c->ev_end[slot]));
    float d = 0.f;
    CU_TRY(cudaEventElapsedTime(&d, G.ev_mark, c->ev_end[slot]));
    *ms = d > 0.f ? d : 0.f;
    return 0;
}
extern "C" int dlnb_spin_calibrate(double *overhead_us, double *slope) {
    NEED_INIT();
    // Two isolated spins, device-timed with events: achieved = overhead + slope * requested.
    // slope is the ratio between the event clock and %globaltimer (measured ~1.0006 on B200);
    // overhead is the launch-to-first-instruction latency of an isolated spin (reported only).
    const double reqs[2] = {200.0, 20000.0};
    double got[2];
    CU_TRY(cudaStreamSynchronize(G.s_compute));
    for (int i = 0; i < 2; ++i) {
        double best = 1e30;
        for (int rep = 0; rep < 4; ++rep) {
            CU_TRY(cudaEventRecord(G.ev_tmp0, G.s_compute));
            if (launch_spin(reqs[i] * 1000.0)) return 1;
            CU_TRY(cudaEventRecord(G.ev_tmp1, G.s_compute));
            CU_TRY(cudaEventSynchronize(G.ev_tmp1));
            float ms = 0;
            CU_TRY(cudaEventElapsedTime(&ms, G.ev_tmp0, G.ev_tmp1));
            best = std::min(best, (double)ms * 1000.0);
            usleep(50); // break the chain: the next spin must start from its own clock
        }
        got[i] = best;
    }
    double b = (got[1] - got[0]) / (reqs[1] - reqs[0]);
    double a = got[0] - b * reqs[0];
    if (!(b > 0.9 && b < 1.1)) { a = 0; b = 1; }
    if (a < 0) a = 0;
    G.spin_overhead_us = a;
    G.spin_slope = b;
    if (overhead_us) *overhead_us = a;
    if (slope) *slope = b;
    return 0;
}

// ---------------------------------------------------------------------------
// collective launch
// ---------------------------------------------------------------------------
static cudaError_t launch_dtype(int dtype, int op, int nr, dim3 grid, int threads, cudaStream_t s,
                                const EmuViews &ev, size_t count, bool coop, bool pdl = false) {
    switch (dtype) {
    case DLNB_F32: return launch_coll_f32(op, nr, grid, threads, s, ev, count, coop, pdl);
    case DLNB_F16: return launch_coll_f16(op, nr, grid, threads, s, ev, count, coop, pdl);
    case DLNB_BF16: return launch_coll_bf16(op, nr, grid, threads, s, ev, count, coop, pdl);
    }
    return cudaErrorInvalidValue;
}
static int occupancy_dtype(int dtype, int op, int nr, int threads) {
    switch (dtype) {
    case DLNB_F32: return occupancy_coll_f32(op, nr, threads);
    case DLNB_F16: return occupancy_coll_f16(op, nr, threads);
    case DLNB_BF16: return occupancy_coll_bf16(op, nr, threads);
    }
    return 0;
}

// bytes the kernel's vector loop walks on this rank (sizes the grid)
static size_t work_bytes(int op, int n, size_t count, size_t esz) {
    switch (op) {
    case OP_AR1: return count * esz;
    case OP_AR2: return count * esz / (size_t)n;
    case OP_ARM: return count * esz / (size_t)n;
    case OP_RSM: return count * esz;
    case OP_AGM: return count * esz;
    case OP_A2A: return count * esz; // one block per peer, walked one after the other
    case OP_AG: return count * esz;
    default: return count * esz;
    }
}

static bool is_switch_op(int op) { return op == OP_ARM || op == OP_RSM || op == OP_AGM; }

static int pick_grid(long want, long want_mc, int dtype, int op, int n, int threads, size_t bytes, int *out) {
    int occ = occupancy_dtype(dtype, op, n, threads);
    if (occ <= 0) return fail("collective kernel does not fit on an SM with %d threads", threads);
    // n > 1: every CTA waits on its peers' CTA of the same index, so all must be resident;
    // n == 1: no barriers, and 4 CTAs per SM (two waves at occupancy 2) measured best for HBM
    int cap = n == 1 ? G.sms * 16 : std::min(G.sms * occ, MAX_FLAG_CTAS);
    long g;
    if (want > 0 && !(is_switch_op(op) && want_mc > 0)) {
        g = want;
    } else if (n == 1) {
        g = (long)G.sms * 4;
    } else if (is_switch_op(op)) {
        // the switch does the adding / replicating: few CTAs saturate it. Measured (profiles/r02/w4, w8): 16 CTAs
        // best on 4 and on 8 GPUs at every size >= 64 MiB (4 GPUs, 1 GiB: 695 / 677 / 616 GB/s with 16 / 32 / 48)
        (void)n;
        g = want_mc > 0 ? want_mc : 16;
    } else if (op == OP_AR2 && n == 2) {
        g = 64; // measured on 2 GPUs (profiles/r02): 632 / 666 / 676 GB/s at 64 MiB / 256 MiB / 1 GiB with 64 CTAs, 587 / 653 / 671 with 128
    } else {
        g = 128; // NVLink-bound peer loads/stores: measured best for >= 64 MB, harmless below
    }
    // no more CTAs than 16-byte vectors / threads
    const size_t vecs = bytes / 16 + 1;
    const long useful = (long)((vecs + threads - 1) / threads);
    g = std::min(g, std::max(1L, useful));
    g = std::min<long>(g, cap);
    *out = (int)std::max(1L, g);
    return 0;
}

static int join_streams() {
    // the collective may not start before compute / copies enqueued so far (both lanes wait)
    if (G.capturing && !G.comm_in_capture) { // fork the communication streams into the capture
        G.compute_dirty = true;
        G.comm_in_capture = true;
    }
    if (G.compute_dirty && !G.detach_compute) {
        CU_TRY(cudaEventRecord(G.ev_join, G.s_compute));
        CU_TRY(cudaStreamWaitEvent(G.s_comm, G.ev_join, 0));
        if (G.lanes > 1) {
            CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_join, 0));
            if (G.capturing) G.comm2_in_capture = true;
        }
        G.compute_dirty = false;
    }
    if (G.copy_dirty) {
        CU_TRY(cudaEventRecord(G.ev_copy, G.s_copy));
        CU_TRY(cudaStreamWaitEvent(G.s_comm, G.ev_copy, 0));
        if (G.lanes > 1) CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_copy, 0));
        G.copy_dirty = false;
    }
    return 0;
}

// Lane 0 is about to run something that must see (or must not overlap with) whatever lane 1 has
// in flight: make lane 0 wait for lane 1.
static int drain_lane1_into_lane0() {
    if (G.lane1_dirty) {
        CU_TRY(cudaEventRecord(G.ev_lane1, G.s_comm2));
        CU_TRY(cudaStreamWaitEvent(G.s_comm, G.ev_lane1, 0));
        G.lane1_dirty = false;
    }
    return 0;
}
// ... and whatever lane 1 runs next must come after that exclusive lane-0 work.
static int mark_exclusive_on_lane0() {
    if (G.lanes > 1) {
        CU_TRY(cudaEventRecord(G.ev_excl, G.s_comm));
        G.excl_pending = true;
    }
    return 0;
}

// what this rank sees of one collective on `lane` (peer mappings of the two buffers, flag rows, generations)
static int make_view(dlnb_comm *c, int op, int lane, const void *send, void *recv, const char *mc_s, char *mc_r, RankView &v) {
    memset(&v, 0, sizeof(v));
    v.rank = c->me;
    v.nranks = c->n;
    for (int i = 0; i < c->n; ++i) {
        char *s = translate(send, c->ranks[i]);
        char *r = translate(recv, c->ranks[i]);
        if (!s || !r) return fail("buffer %p/%p is not in the symmetric arena (use dlnb_symm_alloc)", send, recv);
        v.send[i] = s;
        v.recv[i] = r;
        v.flags[i] = (uint32_t *)translate(G.flag_area + (size_t)(c->id * 2 + lane) * FLAG_BYTES_PER_COMM, c->ranks[i]);
    }
    v.mc_send = is_switch_op(op) ? mc_s : nullptr;
    v.mc_recv = is_switch_op(op) ? mc_r : nullptr;
    v.strict = (int)G.strict_barriers;
    v.shift_overlap = (int)G.shift_overlap;
    v.ctl = G.ctl + (c->id * 2 + lane);
    v.host_error = G.host_error_dev;
    v.timeout_ns = (unsigned long long)c->timeout_ms * 1000000ull;
    return 0;
}

// A large allreduce split between the switch and the links: the first `hybrid_pct` percent (whole 8 KiB units) go
// through the in-switch kernel on lane 0, the rest through the fused two-shot kernel on lane 1, at the same time. The
// in-switch path stalls near 540 GB/s of wire rate per direction; the peer-to-peer path uses what the links have left.
// Both are exclusive: ordered after everything issued before, and everything issued later is ordered after both.
static int coll_hybrid(dlnb_comm *c, const void *send, void *recv, size_t count, int slot, const char *mc_s, char *mc_r) {
    const size_t esz = dtype_size(c->dtype);
    const size_t unit = 8192 / esz;
    const size_t c_mm = count * (size_t)c->hybrid_pct / 100 / unit * unit, c_pp = count - c_mm;
    const int threads = (int)c->threads;
    int g_mm = 1, g_pp = 1;
    if (pick_grid(0, c->mc_ctas, c->dtype, OP_ARM, c->n, threads, work_bytes(OP_ARM, c->n, c_mm, esz), &g_mm)) return 1;
    if (pick_grid(64, 0, c->dtype, OP_AR2, c->n, threads, work_bytes(OP_AR2, c->n, c_pp, esz), &g_pp)) return 1;
    if (join_streams()) return 1;
    if (drain_lane1_into_lane0()) return 1;
    CU_TRY(cudaEventRecord(G.ev_excl, G.s_comm)); // lane 1 starts after everything lane 0 has, too
    CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_excl, 0));
    G.excl_pending = false;
    const bool timed = G.slot_timing && !G.capturing;
    if (timed) CU_TRY(cudaEventRecord(c->ev_start[slot], G.s_comm));
    EmuViews ev;
    if (make_view(c, OP_ARM, 0, send, recv, mc_s, mc_r, ev.v[0])) return 1;
    cudaError_t e = launch_dtype(c->dtype, OP_ARM, c->n, dim3(g_mm, 1, 1), threads, G.s_comm, ev, c_mm, false);
    if (e != cudaSuccess) return fail("collective launch failed: %s", cudaGetErrorString(e));
    if (make_view(c, OP_AR2, 1, (const char *)send + c_mm * esz, (char *)recv + c_mm * esz, nullptr, nullptr, ev.v[0])) return 1;
    e = launch_dtype(c->dtype, OP_AR2, c->n, dim3(g_pp, 1, 1), threads, G.s_comm2, ev, c_pp, false);
    if (e != cudaSuccess) return fail("collective launch failed: %s", cudaGetErrorString(e));
    G.launches += 2;
    if (G.capturing) G.comm2_in_capture = true;
    CU_TRY(cudaEventRecord(G.ev_lane1, G.s_comm2)); // the slot is complete when both halves are
    CU_TRY(cudaStreamWaitEvent(G.s_comm, G.ev_lane1, 0));
    G.lane1_dirty = false;
    if (mark_exclusive_on_lane0()) return 1;
    if (timed) CU_TRY(cudaEventRecord(c->ev_end[slot], G.s_comm));
    CU_TRY(cudaEventRecord(c->ev_done[slot], G.s_comm));
    c->slot_timed[slot] = timed ? 1 : 0;
    c->slot_used[slot] = 1;
    return 0;
}

static int coll(dlnb_comm *c, int op, const void *send, void *recv, size_t count, int slot) {
    NEED_INIT();
    if (!c) return fail("null communicator");
    if (slot == -1) slot = c->num_slots; // internal slot of the blocking calls
    else if (slot < 0 || slot >= c->num_slots) return fail("slot %d out of range (communicator has %d)", slot, c->num_slots);
    if (count && (!send || !recv)) return fail("null buffer");
    const size_t esz = dtype_size(c->dtype);
    char *mc_s = nullptr, *mc_r = nullptr;
    const bool mc_possible = c->n > 1 && c->n == G.world && G.mc_enabled; // the multicast object spans every rank
    if (op == OP_AR1 || op == OP_AR2) {
        const size_t bytes = count * esz;
        // measured cross-over (B200): one-shot wins up to ~1 MiB on 2 GPUs, ~256 KiB on 8
        const size_t two_min = c->twoshot_min_bytes >= 0 ? (size_t)c->twoshot_min_bytes : (size_t)(2 << 20) / (size_t)c->n;
        const bool small = two_min > bytes;
        // in-switch reduction needs the multicast view of both buffers and 16-byte aligned buffers
        if (mc_possible && (c->algo == DLNB_ALGO_MULTIMEM || c->algo == DLNB_ALGO_AUTO) &&
            (((uintptr_t)send | (uintptr_t)recv) & 15) == 0) {
            mc_s = translate_mc(send);
            mc_r = translate_mc(recv);
        }
        const bool mc_ok = mc_s && mc_r;
        if (c->algo == DLNB_ALGO_MULTIMEM && !mc_ok)
            return fail("multimem allreduce requested but unavailable (multicast %s, group %d of world %d)",
                        G.mc_enabled ? "on" : "off", c->n, G.world);
        if (c->algo == DLNB_ALGO_ONESHOT) op = OP_AR1;
        else if (c->algo == DLNB_ALGO_TWOSHOT) op = OP_AR2;
        else if (c->algo == DLNB_ALGO_MULTIMEM) op = OP_ARM;
        else if (mc_ok && c->n >= 4 && bytes >= (size_t)c->multimem_min_bytes) op = OP_ARM; // n < 4: P2P moves fewer bytes
        else op = (c->n == 1 || small) ? OP_AR1 : OP_AR2;
        // In place (MPI_IN_PLACE / ncclAllReduce with sendbuff == recvbuff): the one-shot kernel has every
        // rank read all of its peers' send buffers while each rank writes its own recv, so a peer could
        // read an element that is already a sum. The two-shot and in-switch kernels read and write every
        // element through its one owner: they are in-place safe.
        if (op == OP_AR1 && c->n > 1) {
            const char *s0 = (const char *)send, *r0 = (const char *)recv;
            if (s0 < r0 + bytes && r0 < s0 + bytes) op = OP_AR2;
        }
    } else if ((op == OP_AG || op == OP_RS) && c->algo == DLNB_ALGO_MULTIMEM && c->n > 1) {
        // in-switch variants on request only (AUTO stays on the P2P kernels, see DESIGN: the byte count on
        // the busier direction does not shrink). Chunks that are not 16-byte aligned keep the P2P path.
        if (!mc_possible)
            return fail("multimem %s requested but unavailable (multicast %s, group %d of world %d)", op == OP_AG ? "allgather" : "reduce-scatter",
                        G.mc_enabled ? "on" : "off", c->n, G.world);
        if ((((uintptr_t)send | (uintptr_t)recv) & 15) == 0 && (count * esz) % 16 == 0) {
            if (op == OP_AG) mc_r = translate_mc(recv); else mc_s = translate_mc(send);
            if (op == OP_AG && mc_r) op = OP_AGM;
            else if (op == OP_RS && mc_s) op = OP_RSM;
        }
    }
    if (op == OP_ARM && c->hybrid_pct > 0 && c->hybrid_pct < 100 && G.lanes > 1 && count * esz >= (size_t)c->hybrid_min_bytes)
        return coll_hybrid(c, send, recv, count, slot, mc_s, mc_r);
    // Grid first: it decides the lane. Small allreduces (<= 64 CTAs, so two of them plus the spin
    // kernel are always co-resident) alternate between the two lanes; everything else is
    // exclusive on lane 0.
    int grid = 1;
    const int threads = (int)c->threads;
    if (count && pick_grid(c->ctas, c->mc_ctas, c->dtype, op, c->n, threads, work_bytes(op, c->n, count, esz), &grid)) return 1;
    const bool is_ar = op == OP_AR1 || op == OP_AR2 || op == OP_ARM;
    // up to 8 MiB the grid makes no measurable difference (26.2 / 26.5 / 26.1 us at 4 MiB with 32 / 64 /
    // 128 CTAs on 8 GPUs): keep such allreduces at <= 64 CTAs so that they can use the lanes
    if (G.lanes > 1 && is_ar && c->ctas <= 0 && count * esz <= ((size_t)8 << 20) && grid > 64) grid = 64;
    const bool laned = G.lanes > 1 && is_ar && c->n > 1 && count && grid <= 64;
    const int lane = laned ? c->next_lane : 0;
    if (laned) c->next_lane ^= 1;
    cudaStream_t st = lane ? G.s_comm2 : G.s_comm;
    if (join_streams()) return 1;
    if (!laned) {
        if (drain_lane1_into_lane0()) return 1;
    } else if (lane == 1 && G.excl_pending) {
        CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_excl, 0));
        G.excl_pending = false;
    }
    const bool timed = G.slot_timing && !G.capturing;
    if (timed) CU_TRY(cudaEventRecord(c->ev_start[slot], st));
    if (count) {
        EmuViews ev;
        if (make_view(c, op, lane, send, recv, mc_s, mc_r, ev.v[0])) return 1;
        cudaError_t e = launch_dtype(c->dtype, op, c->n, dim3(grid, 1, 1), threads, st, ev, count, false, c->n == 1 && G.pdl && !G.capturing);
        if (e != cudaSuccess) return fail("collective launch failed: %s", cudaGetErrorString(e));
        G.launches++;
    }
    if (lane == 1) G.lane1_dirty = true;
    else if (!laned && mark_exclusive_on_lane0()) return 1;
    if (G.capturing) {
        // no host-side completion inside a capture; the event lets Wait / WaitAll / the blocking calls
        // order the COMPUTE stream behind this collective, as the host's blocking did (dlnb_wait)
        CU_TRY(cudaEventRecord(c->ev_done[slot], st));
        c->slot_used[slot] = 1;
        c->slot_timed[slot] = 0;
        return 0;
    }
    if (timed) CU_TRY(cudaEventRecord(c->ev_end[slot], st));
    CU_TRY(cudaEventRecord(c->ev_done[slot], st));
    c->slot_timed[slot] = timed ? 1 : 0;
    c->slot_used[slot] = 1;
    return 0;
}

// Hooks for work of ANOTHER library on the communication stream (the `nccl` comparison build, cpp/proxy_classes.hpp:
// CCLCommunicator): the same ordering and slot bookkeeping a collective of this library gets — after the compute
// enqueued so far, timed per slot, completion recorded for Wait / WaitAll — around a call the caller enqueues itself.
extern "C" int dlnb_foreign_begin(dlnb_comm_t c, int slot, void **cuda_stream) {
    NEED_INIT();
    if (!c || !cuda_stream) return fail("null argument");
    if (G.capturing) return fail("foreign work cannot be captured into a graph");
    if (slot == -1) slot = c->num_slots;
    else if (slot < 0 || slot >= c->num_slots) return fail("slot %d out of range (communicator has %d)", slot, c->num_slots);
    if (join_streams()) return 1;
    if (drain_lane1_into_lane0()) return 1;
    if (G.slot_timing) CU_TRY(cudaEventRecord(c->ev_start[slot], G.s_comm));
    *cuda_stream = (void *)G.s_comm;
    return 0;
}
extern "C" int dlnb_foreign_end(dlnb_comm_t c, int slot) {
    NEED_INIT();
    if (!c) return fail("null communicator");
    if (slot == -1) slot = c->num_slots;
    else if (slot < 0 || slot >= c->num_slots) return fail("slot %d out of range", slot);
    if (mark_exclusive_on_lane0()) return 1;
    if (G.slot_timing) CU_TRY(cudaEventRecord(c->ev_end[slot], G.s_comm));
    CU_TRY(cudaEventRecord(c->ev_done[slot], G.s_comm));
    c->slot_timed[slot] = G.slot_timing ? 1 : 0;
    c->slot_used[slot] = 1;
    c->recv_src[slot] = -1;
    G.launches++;
    return 0;
}

static int check_watchdog() {
    if (G.host_error && *(volatile uint32_t *)G.host_error) {
        *(volatile uint32_t *)G.host_error = 0; // reported once; the caller decides whether to go on
        return fail("cross-GPU flag barrier timed out: a peer rank is dead or out of order");
    }
    return 0;
}

extern "C" int dlnb_iallreduce(dlnb_comm_t c, const void *send, void *recv, size_t count, int slot) {
    return coll(c, OP_AR2, send, recv, count, slot);
}
extern "C" int dlnb_iallgather(dlnb_comm_t c, const void *send, void *recv, size_t sendcount, int slot) {
    return coll(c, OP_AG, send, recv, sendcount, slot);
}
extern "C" int dlnb_ireduce_scatter(dlnb_comm_t c, const void *send, void *recv, size_t recvcount, int slot) {
    return coll(c, OP_RS, send, recv, recvcount, slot);
}
extern "C" int dlnb_ialltoall(dlnb_comm_t c, const void *send, void *recv, size_t sendcount, int slot) {
    return coll(c, OP_A2A, send, recv, sendcount, slot);
}
extern "C" int dlnb_alltoall(dlnb_comm_t c, const void *send, void *recv, size_t sendcount) {
    if (coll(c, OP_A2A, send, recv, sendcount, -1)) return 1;
    return dlnb_wait(c, -1);
}
static int finish_recv(dlnb_comm *c, int slot);
extern "C" int dlnb_wait(dlnb_comm_t c, int slot) {
    NEED_INIT();
    if (!c) return fail("null communicator");
    if (slot == -1) slot = c->num_slots;
    else if (slot < 0 || slot >= c->num_slots) return fail("slot %d out of range", slot);
    if (!c->slot_used[slot]) return 0;
    if (G.capturing) {
        // The host cannot block inside a capture. What its blocking meant for the schedule — nothing
        // enqueued afterwards (the next emulated compute) starts before this collective is done —
        // becomes a device-side edge: the compute stream waits for the collective's completion event.
        CU_TRY(cudaStreamWaitEvent(G.s_compute, c->ev_done[slot], 0));
        c->slot_used[slot] = 0;
        return 0;
    }
    if (finish_recv(c, slot)) return 1;
    CU_TRY(cudaEventSynchronize(c->ev_done[slot]));
    return check_watchdog();
}
extern "C" int dlnb_waitall(dlnb_comm_t c, int nslots) {
    NEED_INIT();
    if (!c) return fail("null communicator");
    if (nslots > c->num_slots) nslots = c->num_slots;
    if (G.capturing) { // see dlnb_wait: the compute stream waits for every outstanding slot
        for (int i = 0; i < nslots; ++i)
            if (c->slot_used[i]) {
                CU_TRY(cudaStreamWaitEvent(G.s_compute, c->ev_done[i], 0));
                c->slot_used[i] = 0;
            }
        return 0;
    }
    // the communication stream is ordered: waiting for the newest used slot suffices,
    // but slots may have been recorded in any order, so wait for each.
    for (int i = 0; i < nslots; ++i)
        if (c->slot_used[i] && finish_recv(c, i)) return 1;
    for (int i = 0; i < nslots; ++i)
        if (c->slot_used[i]) CU_TRY(cudaEventSynchronize(c->ev_done[i]));
    return check_watchdog();
}
extern "C" int dlnb_barrier(dlnb_comm_t c) {
    NEED_INIT();
    if (!c) return fail("null communicator");
    if (drain_lane1_into_lane0()) return 1; // both lanes: lane 0 finishes after whatever lane 1 has in flight
    if (G.capturing) { // device-side form of "the host waits for the communication stream"
        if (!G.comm_in_capture) return 0; // nothing captured on it yet
        CU_TRY(cudaEventRecord(G.ev_join, G.s_comm));
        CU_TRY(cudaStreamWaitEvent(G.s_compute, G.ev_join, 0));
        return 0;
    }
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return check_watchdog();
}
extern "C" int dlnb_allreduce(dlnb_comm_t c, const void *send, void *recv, size_t count) {
    if (coll(c, OP_AR2, send, recv, count, -1)) return 1;
    return dlnb_wait(c, -1);
}
extern "C" int dlnb_allgather(dlnb_comm_t c, const void *send, void *recv, size_t sendcount) {
    if (coll(c, OP_AG, send, recv, sendcount, -1)) return 1;
    return dlnb_wait(c, -1);
}
extern "C" int dlnb_reduce_scatter(dlnb_comm_t c, const void *send, void *recv, size_t recvcount) {
    if (coll(c, OP_RS, send, recv, recvcount, -1)) return 1;
    return dlnb_wait(c, -1);
}
extern "C" int dlnb_slot_elapsed(dlnb_comm_t c, int slot, float *ms) {
    NEED_INIT();
    if (!c || !ms) return fail("null argument");
    if (slot == -1) slot = c->num_slots; // the most recent blocking call
    if (slot < 0 || slot > c->num_slots || !c->slot_used[slot]) return fail("slot %d has no recorded collective", slot);
    if (c->recv_src[slot] >= 0) return fail("slot %d holds a receive that has not been waited for", slot);
    if (!c->slot_timed[slot]) return fail("slot %d was recorded with slot_timing off", slot);
    CU_TRY(cudaEventSynchronize(c->ev_end[slot]));
    CU_TRY(cudaEventElapsedTime(ms, c->ev_start[slot], c->ev_end[slot]));
    return 0;
}

// ---------------------------------------------------------------------------
// point-to-point (pipeline stages): rendezvous put over peer memory
// ---------------------------------------------------------------------------
static int p2p(dlnb_comm *c, bool is_send, const void *buf, size_t count, int peer, int slot) {
    NEED_INIT();
    if (!c) return fail("null communicator");
    if (slot == -1) slot = c->num_slots;
    else if (slot < 0 || slot >= c->num_slots) return fail("slot %d out of range (communicator has %d)", slot, c->num_slots);
    if (peer < 0 || peer >= c->n) return fail("peer %d outside the communicator (size %d)", peer, c->n);
    if (peer == c->me) return fail("send/recv to oneself is not supported");
    if (G.kind != ARENA_VMM) return fail("send/recv need the VMM arena (one contiguous window per rank)");
    if (count && !buf) return fail("null buffer");
    const int wpeer = c->ranks[peer];
    const size_t bytes = count * dtype_size(c->dtype);
    const char *b = (const char *)buf;
    if (count && !(b >= (const char *)G.window[G.rank] && b + bytes <= (const char *)G.window[G.rank] + G.va_used))
        return fail("buffer %p is not in the symmetric arena (use dlnb_symm_alloc)", buf);
    if (G.capturing) return fail("send/recv cannot be captured into a graph");
    if (c->recv_src[slot] >= 0)
        return fail("slot %d still holds a posted receive that has not been waited for (one pending operation per slot)", slot);
    if (join_streams()) return 1;
    if (drain_lane1_into_lane0()) return 1; // ordered on lane 0 behind everything, like allgather / reduce-scatter
    P2PBox *mine = G.p2p_boxes + wpeer;                                // my box for this peer
    P2PBox *theirs = (P2PBox *)translate(G.p2p_boxes, wpeer) + G.rank; // the peer's box for me
    const unsigned long long timeout_ns = (unsigned long long)c->timeout_ms * 1000000ull;
    const bool timed = G.slot_timing != 0;
    if (timed) CU_TRY(cudaEventRecord(c->ev_start[slot], G.s_comm));
    if (is_send) {
        const size_t vecs = bytes / 16 + 1;
        long g = c->ctas > 0 ? c->ctas : 64;
        g = std::min<long>(g, std::max<long>(1, (long)((vecs + 511) / 512)));
        g = std::min<long>(g, G.sms);
        send_kernel<<<(unsigned)g, 512, 0, G.s_comm>>>(mine, theirs, (char *)G.window[wpeer], buf, bytes, G.p2p_state, wpeer,
                                                        ++G.p2p_sent[wpeer], timeout_ns, G.host_error_dev);
        CU_TRY(cudaGetLastError());
        G.launches++;
        if (mark_exclusive_on_lane0()) return 1;
        if (timed) CU_TRY(cudaEventRecord(c->ev_end[slot], G.s_comm));
        CU_TRY(cudaEventRecord(c->ev_done[slot], G.s_comm));
        c->recv_src[slot] = -1;
    } else {
        const unsigned long long off = (unsigned long long)(b - (const char *)G.window[G.rank]);
        // the sender polls post[seq % P2P_RING]: a receive posted P2P_RING ahead of one that has not been
        // waited for would overwrite the word its sender may still be looking for
        if (G.p2p_posted[wpeer] - G.p2p_waited[wpeer] >= (uint32_t)P2P_RING)
            return fail("Irecv: %d receives from rank %d are already outstanding (wait for one first)", P2P_RING, wpeer);
        const uint32_t seq = ++G.p2p_posted[wpeer];
        recv_post_kernel<<<1, 32, 0, G.s_comm>>>(theirs, seq, off);
        CU_TRY(cudaGetLastError());
        G.launches++;
        if (mark_exclusive_on_lane0()) return 1;
        // the completion wait is enqueued by dlnb_wait / dlnb_waitall (see recv_wait_kernel)
        c->recv_src[slot] = wpeer;
        c->recv_seq[slot] = seq;
    }
    c->slot_timed[slot] = timed ? 1 : 0;
    c->slot_used[slot] = 1;
    return 0;
}
// enqueue the completion wait of a posted receive (if this slot holds one)
static int finish_recv(dlnb_comm *c, int slot) {
    if (c->recv_src[slot] < 0) return 0;
    if (drain_lane1_into_lane0()) return 1;
    recv_wait_kernel<<<1, 32, 0, G.s_comm>>>(G.p2p_boxes + c->recv_src[slot], c->recv_seq[slot],
                                             (unsigned long long)c->timeout_ms * 1000000ull, G.host_error_dev);
    CU_TRY(cudaGetLastError());
    G.launches++;
    if (mark_exclusive_on_lane0()) return 1;
    if (c->slot_timed[slot]) CU_TRY(cudaEventRecord(c->ev_end[slot], G.s_comm));
    CU_TRY(cudaEventRecord(c->ev_done[slot], G.s_comm));
    G.p2p_waited[c->recv_src[slot]]++;
    c->recv_src[slot] = -1;
    return 0;
}
extern "C" int dlnb_isend(dlnb_comm_t c, const void *buf, size_t count, int dest, int slot) { return p2p(c, true, buf, count, dest, slot); }
extern "C" int dlnb_irecv(dlnb_comm_t c, void *buf, size_t count, int source, int slot) { return p2p(c, false, buf, count, source, slot); }
extern "C" int dlnb_send(dlnb_comm_t c, const void *buf, size_t count, int dest) {
    if (p2p(c, true, buf, count, dest, -1)) return 1;
    return dlnb_wait(c, -1);
}
extern "C" int dlnb_recv(dlnb_comm_t c, void *buf, size_t count, int source) {
    if (p2p(c, false, buf, count, source, -1)) return 1;
    return dlnb_wait(c, -1);
}

// ---------------------------------------------------------------------------
// device timing of one proxy iteration (both streams)
// ---------------------------------------------------------------------------
// One flag barrier between all world ranks, enqueued on `st` (dlnb_set "align_iterations", default on).
static int launch_rank_sync(cudaStream_t st) {
    if (G.world <= 1 || !G.align_iterations) return 0;
    RankView v;
    memset(&v, 0, sizeof(v));
    v.rank = G.rank;
    v.nranks = G.world;
    for (int r = 0; r < G.world; ++r) {
        v.flags[r] = (uint32_t *)translate(G.flag_area + (size_t)SYS_FLAG_SET * FLAG_BYTES_PER_COMM, r);
        if (!v.flags[r]) return fail("rank sync: flag rows of rank %d are not mapped", r);
    }
    v.ctl = G.ctl + SYS_FLAG_SET;
    v.host_error = G.host_error_dev;
    v.timeout_ns = 30ull * 1000000000ull;
    rank_sync_kernel<<<1, 32, 0, st>>>(v);
    CU_TRY(cudaGetLastError());
    G.launches++;
    return 0;
}

extern "C" int dlnb_iter_begin(void) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_iter_begin inside a graph capture (time the replay instead)");
    // the iteration starts when all streams are idle up to here, and nothing of it starts earlier
    if (drain_lane1_into_lane0()) return 1;
    CU_TRY(cudaEventRecord(G.ev_join, G.s_comm));
    CU_TRY(cudaStreamWaitEvent(G.s_compute, G.ev_join, 0));
    // ... on every rank at the same moment (one flag round trip apart): the reference's ranks leave an
    // MPI_Barrier together before each timed run (dp.cpp:259, fsdp.cpp:391); here the ranks' compute
    // streams do, so the spin chains of all ranks run in step and a collective is not charged for the
    // time a late rank needs to arrive
    if (launch_rank_sync(G.s_compute)) return 1;
    CU_TRY(cudaEventRecord(G.ev_iter0, G.s_compute));
    CU_TRY(cudaStreamWaitEvent(G.s_comm, G.ev_iter0, 0));
    CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_iter0, 0));
    CU_TRY(cudaStreamWaitEvent(G.s_copy, G.ev_iter0, 0));
    CU_TRY(cudaStreamWaitEvent(G.s_out, G.ev_iter0, 0));
    return 0;
}
extern "C" int dlnb_batch_begin(void) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_batch_begin inside a graph capture");
    // starts the device timer ON THE COMMUNICATION STREAM, behind everything enqueued so far on
    // any stream: with a long spin enqueued first, the host can queue a whole batch of
    // collectives before the first one runs, and dlnb_iter_end() then reads pure device time.
    G.compute_dirty = true;
    G.copy_dirty = true;
    if (join_streams()) return 1;
    if (drain_lane1_into_lane0()) return 1;
    if (launch_rank_sync(G.s_comm)) return 1; // all ranks start the batch together
    CU_TRY(cudaEventRecord(G.ev_iter0, G.s_comm));
    CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_iter0, 0));
    return 0;
}
// Contention measurements: a batch of collectives timed on the communication stream WHILE emulated compute that was
// enqueued before runs beside it — between these two calls collectives do not wait for the compute stream.
extern "C" int dlnb_batch_begin_concurrent(double gate_us) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_batch_begin_concurrent inside a graph capture");
    if (drain_lane1_into_lane0()) return 1;
    G.detach_compute = true;
    if (gate_us > 0) { // one sleeping warp holds the communication stream while the host queues the batch behind it
        spin_kernel<<<1, 32, 0, G.s_comm>>>((unsigned long long)(gate_us * 1000.0), G.spin_stamps + 2, 0u, 0ull, (int)SPIN_SLEEP, nullptr, 0, G.spin_sink);
        CU_TRY(cudaGetLastError());
        G.launches++;
    }
    if (launch_rank_sync(G.s_comm)) return 1;
    CU_TRY(cudaEventRecord(G.ev_iter0, G.s_comm));
    CU_TRY(cudaStreamWaitEvent(G.s_comm2, G.ev_iter0, 0));
    return 0;
}
extern "C" int dlnb_batch_end(float *ms) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_batch_end inside a graph capture");
    if (drain_lane1_into_lane0()) return 1;
    CU_TRY(cudaEventRecord(G.ev_iter1, G.s_comm));
    CU_TRY(cudaEventSynchronize(G.ev_iter1));
    if (ms) CU_TRY(cudaEventElapsedTime(ms, G.ev_iter0, G.ev_iter1));
    G.detach_compute = false;
    return check_watchdog();
}
extern "C" int dlnb_iter_end(float *ms) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_iter_end inside a graph capture (time the replay instead)");
    G.compute_dirty = true;
    G.copy_dirty = true;
    if (join_streams()) return 1;
    if (drain_lane1_into_lane0()) return 1;
    if (G.out_dirty) { // results on their way to the host belong to the iteration
        CU_TRY(cudaEventRecord(G.ev_out, G.s_out));
        CU_TRY(cudaStreamWaitEvent(G.s_comm, G.ev_out, 0));
        G.out_dirty = false;
    }
    CU_TRY(cudaEventRecord(G.ev_iter1, G.s_comm));
    CU_TRY(cudaEventSynchronize(G.ev_iter1));
    if (ms) CU_TRY(cudaEventElapsedTime(ms, G.ev_iter0, G.ev_iter1));
    return check_watchdog();
}

// ---------------------------------------------------------------------------
// whole-iteration CUDA graphs (launch-bound configurations: thousands of small buckets)
// ---------------------------------------------------------------------------
struct dlnb_graph {
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
};

extern "C" int dlnb_graph_begin(void) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_graph_begin: a capture is already open");
    CU_TRY(cudaStreamSynchronize(G.s_compute));
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    CU_TRY(cudaStreamSynchronize(G.s_comm2));
    CU_TRY(cudaStreamSynchronize(G.s_copy));
    G.compute_dirty = G.copy_dirty = false;
    G.lane1_dirty = G.excl_pending = false;
    CU_TRY(cudaStreamBeginCapture(G.s_compute, cudaStreamCaptureModeRelaxed));
    G.capturing = true;
    G.comm_in_capture = G.comm2_in_capture = false;
    return 0;
}
extern "C" int dlnb_graph_end(dlnb_graph_t *out) {
    NEED_INIT();
    if (!G.capturing) return fail("dlnb_graph_end without dlnb_graph_begin");
    if (!out) return fail("dlnb_graph_end: null out pointer");
    cudaError_t e = cudaSuccess;
    if (G.comm2_in_capture) { // join lane 1 into lane 0 ...
        e = cudaEventRecord(G.ev_lane1, G.s_comm2);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(G.s_comm, G.ev_lane1, 0);
    }
    if (e == cudaSuccess && G.comm_in_capture) { // ... and the communication stream back into the origin stream
        e = cudaEventRecord(G.ev_join, G.s_comm);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(G.s_compute, G.ev_join, 0);
    }
    dlnb_graph *g = new dlnb_graph;
    cudaError_t e2 = cudaStreamEndCapture(G.s_compute, &g->graph);
    G.capturing = G.comm_in_capture = G.comm2_in_capture = false;
    G.compute_dirty = G.copy_dirty = false;
    G.lane1_dirty = G.excl_pending = false;
    for (dlnb_comm *c : G.comms) // slots recorded inside the capture have no host-side completion
        if (c) std::fill(c->slot_used.begin(), c->slot_used.end(), 0);
    if (e != cudaSuccess || e2 != cudaSuccess || !g->graph) {
        delete g;
        return fail("graph capture failed: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
    }
    e = cudaGraphInstantiate(&g->exec, g->graph, 0);
    if (e != cudaSuccess) {
        cudaGraphDestroy(g->graph);
        delete g;
        return fail("cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
    }
    *out = g;
    return 0;
}
extern "C" int dlnb_graph_launch(dlnb_graph_t g) {
    NEED_INIT();
    if (!g || !g->exec) return fail("dlnb_graph_launch: null graph");
    if (G.capturing) return fail("dlnb_graph_launch inside a capture");
    CU_TRY(cudaGraphLaunch(g->exec, G.s_compute));
    G.launches++; // one launch call; the kernels inside are counted by dlnb_graph_kernel_nodes()
    G.compute_dirty = true;
    return 0;
}
extern "C" int dlnb_graph_kernel_nodes(dlnb_graph_t g) {
    if (!g || !g->graph) return -1;
    size_t n = 0;
    if (cudaGraphGetNodes(g->graph, nullptr, &n) != cudaSuccess) return -1;
    std::vector<cudaGraphNode_t> nodes(n);
    if (n && cudaGraphGetNodes(g->graph, nodes.data(), &n) != cudaSuccess) return -1;
    int k = 0;
    for (auto nd : nodes) {
        cudaGraphNodeType t;
        if (cudaGraphNodeGetType(nd, &t) == cudaSuccess && t == cudaGraphNodeTypeKernel) ++k;
    }
    return k;
}
extern "C" int dlnb_graph_destroy(dlnb_graph_t g) {
    if (!g) return 0;
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->graph) cudaGraphDestroy(g->graph);
    delete g;
    return 0;
}

// ---------------------------------------------------------------------------
// buffer helpers
// ---------------------------------------------------------------------------
extern "C" int dlnb_memcpy_h2d(void *dst, const void *src, size_t bytes) {
    NEED_INIT();
    // On the communication stream and waited for: a plain cudaMemcpy from pageable memory may
    // return while the DMA is still in flight, and our streams are non-blocking (they do not
    // synchronise with the legacy default stream), so a kernel launched next could read stale data.
    if (drain_lane1_into_lane0()) return 1;
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, G.s_comm));
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return 0;
}
extern "C" int dlnb_memcpy_d2h(void *dst, const void *src, size_t bytes) {
    NEED_INIT();
    if (drain_lane1_into_lane0()) return 1;
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, G.s_comm));
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return 0;
}
extern "C" int dlnb_memcpy_h2d_async(void *dst, const void *src, size_t bytes) {
    NEED_INIT();
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, G.s_copy));
    G.copy_dirty = true;
    return 0;
}
extern "C" int dlnb_memcpy_d2h_async(void *dst, const void *src, size_t bytes) {
    NEED_INIT();
    // results are produced on the communication streams: order the copy after both lanes
    if (drain_lane1_into_lane0()) return 1;
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, G.s_comm));
    if (mark_exclusive_on_lane0()) return 1;
    return 0;
}
// End-to-end path of the bench: a bucket's gradients come from pinned host memory once "its" backward slice
// (the compute enqueued so far) has finished — not earlier, so the copies are not prefetched under the forward —
// and the reduced bucket goes back to the host as soon as its collective (slot) is done, on a stream of its own
// so that the copy engines and the communication stream do not queue behind one another.
extern "C" int dlnb_memcpy_h2d_after_compute(void *dst, const void *src, size_t bytes) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_memcpy_h2d_after_compute inside a graph capture");
    CU_TRY(cudaEventRecord(G.ev_gate, G.s_compute));
    CU_TRY(cudaStreamWaitEvent(G.s_copy, G.ev_gate, 0));
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, G.s_copy));
    G.copy_dirty = true;
    return 0;
}
extern "C" int dlnb_memcpy_d2h_after_slot(dlnb_comm_t c, int slot, void *dst, const void *src, size_t bytes) {
    NEED_INIT();
    if (G.capturing) return fail("dlnb_memcpy_d2h_after_slot inside a graph capture");
    if (!c) return fail("null communicator");
    if (slot < 0 || slot >= c->num_slots || !c->slot_used[slot]) return fail("slot %d has no recorded collective", slot);
    if (finish_recv(c, slot)) return 1;
    CU_TRY(cudaStreamWaitEvent(G.s_out, c->ev_done[slot], 0));
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, G.s_out));
    G.out_dirty = true;
    return 0;
}
extern "C" int dlnb_host_alloc(size_t bytes, void **ptr) {
    NEED_INIT();
    CU_TRY(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
    return 0;
}
extern "C" int dlnb_host_free(void *ptr) {
    if (ptr) cudaFreeHost(ptr);
    return 0;
}

static int aux_grid(size_t n) {
    size_t g = (n + 255) / 256;
    return (int)std::max<size_t>(1, std::min<size_t>(g, (size_t)G.sms * 8));
}

extern "C" int dlnb_fill_pattern(void *ptr, size_t count, int dtype, int kind, uint64_t seed, int as_rank) {
    NEED_INIT();
    if (drain_lane1_into_lane0()) return 1; // helper kernels run on lane 0 and must see lane 1's results
    if (!count) return 0;
    const int g = aux_grid(count);
    if (dtype == DLNB_F32) fill_kernel<float><<<g, 256, 0, G.s_comm>>>((float *)ptr, count, kind, seed, as_rank);
    else if (dtype == DLNB_F16) fill_kernel<__half><<<g, 256, 0, G.s_comm>>>((__half *)ptr, count, kind, seed, as_rank);
    else if (dtype == DLNB_BF16)
        fill_kernel<__nv_bfloat16><<<g, 256, 0, G.s_comm>>>((__nv_bfloat16 *)ptr, count, kind, seed, as_rank);
    else return fail("unknown dtype %d", dtype);
    CU_TRY(cudaGetLastError());
    G.launches++;
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return 0;
}

extern "C" int dlnb_verify_sum_pattern(const void *ptr, size_t count, size_t elem_offset, int dtype, int kind,
                                       uint64_t seed, const int *ranks, int nranks, uint64_t *mismatches) {
    NEED_INIT();
    if (drain_lane1_into_lane0()) return 1; // helper kernels run on lane 0 and must see lane 1's results
    if (!mismatches || !ranks || nranks < 1 || nranks > MAXR) return fail("bad argument");
    RankList rl;
    rl.n = nranks;
    for (int i = 0; i < nranks; ++i) rl.r[i] = ranks[i];
    CU_TRY(cudaMemsetAsync(G.d_scalar, 0, sizeof(unsigned long long), G.s_comm));
    if (count) {
        const int g = aux_grid(count);
        if (dtype == DLNB_F32)
            verify_sum_kernel<float><<<g, 256, 0, G.s_comm>>>((const float *)ptr, count, elem_offset, kind, seed, rl, G.d_scalar);
        else if (dtype == DLNB_F16)
            verify_sum_kernel<__half><<<g, 256, 0, G.s_comm>>>((const __half *)ptr, count, elem_offset, kind, seed, rl, G.d_scalar);
        else if (dtype == DLNB_BF16)
            verify_sum_kernel<__nv_bfloat16><<<g, 256, 0, G.s_comm>>>((const __nv_bfloat16 *)ptr, count, elem_offset, kind, seed, rl, G.d_scalar);
        else return fail("unknown dtype %d", dtype);
        CU_TRY(cudaGetLastError());
        G.launches++;
    }
    unsigned long long h = 0;
    CU_TRY(cudaMemcpyAsync(&h, G.d_scalar, sizeof(h), cudaMemcpyDeviceToHost, G.s_comm));
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    *mismatches = h;
    return 0;
}

extern "C" int dlnb_verify_sum_pattern_tol(const void *ptr, size_t count, size_t elem_offset, int dtype, int kind, uint64_t seed,
                                           const int *ranks, int nranks, double rel_tol, uint64_t *mismatches, double *worst_rel_err) {
    NEED_INIT();
    if (drain_lane1_into_lane0()) return 1;
    if (!mismatches || !ranks || nranks < 1 || nranks > MAXR) return fail("bad argument");
    RankList rl;
    rl.n = nranks;
    for (int i = 0; i < nranks; ++i) rl.r[i] = ranks[i];
    if (!G.d_worst) CU_TRY(cudaMalloc((void **)&G.d_worst, sizeof(unsigned)));
    unsigned *d_worst = G.d_worst;
    CU_TRY(cudaMemsetAsync(G.d_scalar, 0, sizeof(unsigned long long), G.s_comm));
    CU_TRY(cudaMemsetAsync(d_worst, 0, sizeof(unsigned), G.s_comm));
    if (count) {
        const int g = aux_grid(count);
        const float tol = (float)rel_tol;
        if (dtype == DLNB_F32)
            verify_sum_tol_kernel<float><<<g, 256, 0, G.s_comm>>>((const float *)ptr, count, elem_offset, kind, seed, rl, tol, G.d_scalar, d_worst);
        else if (dtype == DLNB_F16)
            verify_sum_tol_kernel<__half><<<g, 256, 0, G.s_comm>>>((const __half *)ptr, count, elem_offset, kind, seed, rl, tol, G.d_scalar, d_worst);
        else if (dtype == DLNB_BF16)
            verify_sum_tol_kernel<__nv_bfloat16><<<g, 256, 0, G.s_comm>>>((const __nv_bfloat16 *)ptr, count, elem_offset, kind, seed, rl, tol, G.d_scalar, d_worst);
        else return fail("unknown dtype %d", dtype);
        CU_TRY(cudaGetLastError());
        G.launches++;
    }
    unsigned long long h = 0;
    unsigned w = 0;
    CU_TRY(cudaMemcpyAsync(&h, G.d_scalar, sizeof(h), cudaMemcpyDeviceToHost, G.s_comm));
    CU_TRY(cudaMemcpyAsync(&w, d_worst, sizeof(w), cudaMemcpyDeviceToHost, G.s_comm));
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    *mismatches = h;
    if (worst_rel_err) {
        float f;
        memcpy(&f, &w, sizeof(f));
        *worst_rel_err = f;
    }
    return 0;
}

extern "C" int dlnb_checksum(const void *ptr, size_t bytes, uint64_t *out) {
    NEED_INIT();
    if (drain_lane1_into_lane0()) return 1; // helper kernels run on lane 0 and must see lane 1's results
    if (!out) return fail("null out pointer");
    CU_TRY(cudaMemsetAsync(G.d_scalar, 0, sizeof(unsigned long long), G.s_comm));
    const size_t nwords = bytes / 4;
    if (nwords) {
        checksum_kernel<<<aux_grid(nwords), 256, 0, G.s_comm>>>((const uint32_t *)ptr, nwords, G.d_scalar);
        CU_TRY(cudaGetLastError());
        G.launches++;
    }
    unsigned long long h = 0;
    CU_TRY(cudaMemcpyAsync(&h, G.d_scalar, sizeof(h), cudaMemcpyDeviceToHost, G.s_comm));
    uint32_t tail = 0;
    if (bytes % 4) CU_TRY(cudaMemcpyAsync(&tail, (const char *)ptr + nwords * 4, bytes % 4, cudaMemcpyDeviceToHost, G.s_comm));
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    *out = h + (uint64_t)mix32(tail + (uint32_t)nwords * 0x9e3779b9u) * (bytes % 4 ? 1 : 0);
    return 0;
}

extern "C" int dlnb_flush_l2(void) {
    NEED_INIT();
    if (!G.flush_buf) {
        G.flush_bytes = 256u << 20; // 2x the 126 MB L2
        CU_TRY(cudaMalloc(&G.flush_buf, G.flush_bytes));
    }
    static uint32_t tag = 0;
    flush_kernel<<<G.sms * 4, 512, 0, G.s_comm>>>((uint4 *)G.flush_buf, G.flush_bytes / 16, ++tag);
    CU_TRY(cudaGetLastError());
    G.launches++;
    return 0;
}

// ---------------------------------------------------------------------------
// single-GPU emulation of an n-rank collective (test vehicle)
// ---------------------------------------------------------------------------
static int emu(int op, int nranks, const void *const *send, void *const *recv, size_t count, int dtype, int ctas) {
    NEED_INIT();
    if (G.capturing) return fail("emu calls synchronise the stream: not usable between dlnb_graph_begin and dlnb_graph_end");
    if (nranks < 1 || nranks > MAXR) return fail("emu: nranks %d not in 1..%d", nranks, MAXR);
    if (!send || !recv) return fail("emu: null buffer list");
    if (!count) return 0;
    if (!G.emu_flags) {
        CU_TRY(cudaMalloc((void **)&G.emu_flags, MAXR * FLAG_BYTES_PER_COMM));
        CU_TRY(cudaMemset(G.emu_flags, 0, MAXR * FLAG_BYTES_PER_COMM));
        CU_TRY(cudaMalloc((void **)&G.emu_ctl, MAXR * sizeof(CommCtl)));
    }
    // A call with another nranks leaves the virtual ranks at different generations:
    // restart all of them from generation 0 when the group size changes (same-size
    // calls keep counting, which is what the multi-GPU path does).
    if (G.emu_last_n != nranks) {
        CU_TRY(cudaMemsetAsync(G.emu_flags, 0, MAXR * FLAG_BYTES_PER_COMM, G.s_comm));
        CU_TRY(cudaMemsetAsync(G.emu_ctl, 0, MAXR * sizeof(CommCtl), G.s_comm));
        G.emu_last_n = nranks;
    }
    EmuViews ev;
    memset(&ev, 0, sizeof(ev));
    for (int r = 0; r < nranks; ++r) {
        RankView &v = ev.v[r];
        v.rank = r;
        v.nranks = nranks;
        for (int p = 0; p < nranks; ++p) {
            v.send[p] = send[p];
            v.recv[p] = recv[p];
            v.flags[p] = (uint32_t *)((char *)G.emu_flags + (size_t)p * FLAG_BYTES_PER_COMM);
        }
        v.ctl = G.emu_ctl + r;
        v.host_error = G.host_error_dev;
        v.timeout_ns = 10ull * 1000000000ull;
        v.strict = (int)G.strict_barriers;
        v.shift_overlap = (int)G.shift_overlap;
    }
    const int threads = 512;
    int occ = occupancy_dtype(dtype, op, nranks, threads);
    if (occ <= 0) return fail("emu: kernel does not fit");
    // all virtual ranks' CTAs must be co-resident: they wait on one another
    int cap = std::min((G.sms * occ) / nranks, MAX_FLAG_CTAS);
    int grid = ctas > 0 ? std::min(ctas, cap) : std::min(16, cap);
    if (grid < 1) return fail("emu: device too small for %d virtual ranks", nranks);
    cudaError_t e = launch_dtype(dtype, op, nranks, dim3(grid, nranks, 1), threads, G.s_comm, ev, count, true);
    if (e != cudaSuccess) return fail("emu launch failed: %s", cudaGetErrorString(e));
    G.launches++;
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return check_watchdog();
}

extern "C" int dlnb_emu_allreduce(int nranks, const void *const *send, void *const *recv, size_t count, int dtype,
                                  int algo, int ctas) {
    int op = (algo == DLNB_ALGO_ONESHOT) ? OP_AR1 : (algo == DLNB_ALGO_TWOSHOT) ? OP_AR2
             : (count * dtype_size(dtype) < 256 * 1024 ? OP_AR1 : OP_AR2);
    return emu(op, nranks, send, recv, count, dtype, ctas);
}
extern "C" int dlnb_emu_allgather(int nranks, const void *const *send, void *const *recv, size_t sendcount, int dtype,
                                  int ctas) {
    return emu(OP_AG, nranks, send, recv, sendcount, dtype, ctas);
}
extern "C" int dlnb_emu_reduce_scatter(int nranks, const void *const *send, void *const *recv, size_t recvcount,
                                       int dtype, int ctas) {
    return emu(OP_RS, nranks, send, recv, recvcount, dtype, ctas);
}
extern "C" int dlnb_emu_alltoall(int nranks, const void *const *send, void *const *recv, size_t sendcount, int dtype,
                                 int ctas) {
    return emu(OP_A2A, nranks, send, recv, sendcount, dtype, ctas);
}

// Point-to-point on ONE GPU: the receive-post, send and receive-wait kernels of dlnb_irecv / dlnb_isend /
// dlnb_wait run in stream order against one mailbox that plays both ends (the receiver posts into the
// "sender's" box, the sender publishes "done" into the "receiver's": the fields are disjoint). The posted
// offset is the absolute address of `recv` against a null window base. Same device code as the
// multi-GPU path, so the one-GPU test box exercises the rendezvous, the shifted mover and the last-CTA
// completion count.
extern "C" int dlnb_emu_sendrecv(const void *send, void *recv, size_t count, int dtype, int ctas) {
    NEED_INIT();
    if (G.capturing) return fail("emu calls synchronise the stream: not usable between dlnb_graph_begin and dlnb_graph_end");
    if (dtype < DLNB_F32 || dtype > DLNB_BF16) return fail("emu_sendrecv: unknown dtype %d", dtype);
    if (count && (!send || !recv)) return fail("emu_sendrecv: null buffer");
    if (G.world != 1) return fail("emu_sendrecv is the one-GPU test vehicle (world size 1)");
    const size_t bytes = count * dtype_size(dtype);
    static uint32_t seq = 0;
    ++seq;
    P2PBox *box = G.p2p_boxes; // slot 0: unused by a world of one
    if (drain_lane1_into_lane0()) return 1;
    recv_post_kernel<<<1, 32, 0, G.s_comm>>>(box, seq, (unsigned long long)(uintptr_t)recv);
    CU_TRY(cudaGetLastError());
    const size_t vecs = bytes / 16 + 1;
    long g = ctas > 0 ? ctas : 64;
    g = std::min<long>(g, std::max<long>(1, (long)((vecs + 511) / 512)));
    g = std::min<long>(g, G.sms);
    send_kernel<<<(unsigned)g, 512, 0, G.s_comm>>>(box, box, (char *)nullptr, send, bytes, G.p2p_state, 0, seq, 10ull * 1000000000ull,
                                                    G.host_error_dev);
    CU_TRY(cudaGetLastError());
    recv_wait_kernel<<<1, 32, 0, G.s_comm>>>(box, seq, 10ull * 1000000000ull, G.host_error_dev);
    CU_TRY(cudaGetLastError());
    G.launches += 3;
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return check_watchdog();
}

// ---------------------------------------------------------------------------
// dlnb_emu_allreduce_noalloc — the emulated n-rank allreduce for callers that may not allocate
// ---------------------------------------------------------------------------
// Who this is for: a host program that has already handed every free byte of HBM to its own
// pool (a framework's caching allocator, or a proxy run sized to fill the card), or one that is
// in the middle of overlapped compute/communication where cudaMalloc's device-wide
// synchronisation would stall the other streams. dlnb_emu_allreduce() above takes its flag rows
// and per-row generations from cudaMalloc on first use; this variant takes them from a fixed
// array that is part of the library image instead, so the call itself never asks the driver for
// memory. The price is a hard upper bound on the launch shape and that all calls share the array.
//
//   scratch layout (18 KiB, resident on every GPU that loads the library, until exit):
//     [0, 16 KiB)        flag rows   : 8 virtual ranks x 64 CTA rows x 8 source ranks x uint32
//     [16 KiB, 18 KiB)   generations : 8 virtual ranks x 64 CTA rows x uint32
#define DLNB_NOALLOC_MAX_CTAS 64
#define DLNB_NOALLOC_FLAG_BYTES (MAXR * DLNB_NOALLOC_MAX_CTAS * FLAG_ROW * 4)
#define DLNB_NOALLOC_EPOCH_BYTES (MAXR * DLNB_NOALLOC_MAX_CTAS * 4)
static __device__ __align__(256) unsigned char dlnb_noalloc_scratch[DLNB_NOALLOC_FLAG_BYTES + DLNB_NOALLOC_EPOCH_BYTES];

// In host code the name above is only a handle for the runtime, not a device address: the
// address each device gave the array is looked up once per device and kept here.
static unsigned char *g_noalloc_dev_ptr[64] = {};

extern "C" int dlnb_emu_allreduce_noalloc(int nranks, const void *const *send, void *const *recv, size_t count,
                                          int dtype, int algo, int ctas) {
    NEED_INIT();
    if (G.capturing)
        return fail("emu_allreduce_noalloc synchronises the stream: not usable between dlnb_graph_begin and dlnb_graph_end");
    if (nranks < 1 || nranks > MAXR) return fail("emu_allreduce_noalloc: nranks %d not in 1..%d", nranks, MAXR);
    if (!send || !recv) return fail("emu_allreduce_noalloc: null buffer list");
    if (dtype < DLNB_F32 || dtype > DLNB_BF16) return fail("emu_allreduce_noalloc: unknown dtype %d", dtype);
    if (ctas > DLNB_NOALLOC_MAX_CTAS)
        return fail("emu_allreduce_noalloc: %d CTAs per virtual rank requested, the built-in scratch holds flag rows for "
                    "at most %d (use dlnb_emu_allreduce for larger grids)", ctas, DLNB_NOALLOC_MAX_CTAS);
    if (!count) return 0;

    // set-up, once per device, with that device current
    int dev = -1;
    CU_TRY(cudaGetDevice(&dev));
    if (dev != G.device) return fail("emu_allreduce_noalloc: device %d is current, the library was initialised on %d", dev, G.device);
    if (dev < 0 || dev >= 64) return fail("emu_allreduce_noalloc: device index %d not supported", dev);
    if (!g_noalloc_dev_ptr[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, dlnb_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail("cudaGetSymbolAddress(dlnb_noalloc_scratch) failed: %s", cudaGetErrorString(e));
        g_noalloc_dev_ptr[dev] = (unsigned char *)p;
    }
    unsigned char *scratch = g_noalloc_dev_ptr[dev];

    // the array still holds the flags and generations of whatever call came last (maybe with another
    // group size): every call starts all rows from generation 0
    CU_TRY(cudaMemsetAsync(scratch, 0, DLNB_NOALLOC_FLAG_BYTES + DLNB_NOALLOC_EPOCH_BYTES, G.s_comm));

    const int op = (algo == DLNB_ALGO_ONESHOT) ? OP_AR1 : (algo == DLNB_ALGO_TWOSHOT) ? OP_AR2
                   : (count * dtype_size(dtype) < 256 * 1024 ? OP_AR1 : OP_AR2);
    EmuViews ev;
    memset(&ev, 0, sizeof(ev));
    const size_t flags_per_rank = (size_t)DLNB_NOALLOC_MAX_CTAS * FLAG_ROW * 4;
    for (int r = 0; r < nranks; ++r) {
        RankView &v = ev.v[r];
        v.rank = r;
        v.nranks = nranks;
        for (int p = 0; p < nranks; ++p) {
            v.send[p] = send[p];
            v.recv[p] = recv[p];
            v.flags[p] = (uint32_t *)(scratch + (size_t)p * flags_per_rank);
        }
        // only rows < grid <= DLNB_NOALLOC_MAX_CTAS of the generation table are ever touched
        v.ctl = (CommCtl *)(scratch + DLNB_NOALLOC_FLAG_BYTES + (size_t)r * DLNB_NOALLOC_MAX_CTAS * 4);
        v.host_error = G.host_error_dev;
        v.timeout_ns = 10ull * 1000000000ull;
        v.shift_overlap = (int)G.shift_overlap;
    }
    const int threads = 512;
    const int occ = occupancy_dtype(dtype, op, nranks, threads);
    if (occ <= 0) return fail("emu_allreduce_noalloc: kernel does not fit");
    const int cap = std::min((G.sms * occ) / nranks, DLNB_NOALLOC_MAX_CTAS); // all virtual ranks co-resident
    const int grid = ctas > 0 ? std::min(ctas, cap) : std::min(16, cap);
    if (grid < 1) return fail("emu_allreduce_noalloc: device too small for %d virtual ranks", nranks);
    cudaError_t e = launch_dtype(dtype, op, nranks, dim3(grid, nranks, 1), threads, G.s_comm, ev, count, true);
    if (e != cudaSuccess) return fail("emu_allreduce_noalloc launch failed: %s", cudaGetErrorString(e));
    G.launches++;
    CU_TRY(cudaStreamSynchronize(G.s_comm));
    return check_watchdog();
}
