/*
 * dlnb.h — C ABI of the B200-native data-parallel hot path of DLNetBench.
 *
 * This is the drop-in boundary: plain C, plain pointers and sizes. The C++ host
 * side (cpp/proxy_classes.hpp: class B200Communicator : ProxyCommunicator) and
 * the Python test/bench harness (dlnetbench_b200/capi.py, ctypes) both bind
 * exactly these symbols from libdlnb.so.
 *
 * Each entry point cites the reference interface it replaces
 * (paths relative to the reference tree, HicrestLaboratory/DLNetBench).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on error; the message is
 *     available from dlnb_last_error() (the C++ wrapper prints "[rank r] msg"
 *     and exits, like the reference's CCUTILS_GPU_CHECK / MPI_ERRORS_ARE_FATAL).
 *   - one process per GPU, one host thread; not re-entrant.
 *   - `count` arguments are in ELEMENTS of the communicator's dtype and are
 *     size_t here (the reference's `int count` is widened by the C++ wrapper,
 *     which rejects negative values).
 *   - all data buffers passed to collectives must come from dlnb_symm_alloc()
 *     (the symmetric arena): kernels address the peers' copies of the same
 *     allocation directly over NVLink.
 *   - ranks of a communicator must issue collectives in the same order.
 */
#ifndef DLNB_H
#define DLNB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLNB_MAX_RANKS 8

typedef struct dlnb_comm *dlnb_comm_t;
typedef struct dlnb_graph *dlnb_graph_t;

/* element types; reference: cpp/data_types.hpp:36-51 (_FLOAT / NCCL_FLOAT_TYPE) */
enum dlnb_dtype { DLNB_F32 = 0, DLNB_F16 = 1, DLNB_BF16 = 2 };

/* allreduce algorithm selector (DLNB_ALGO_AUTO picks by message size and availability).
 * ONESHOT / TWOSHOT: peer loads and stores, fp32 accumulation in rank order (bit-identical to
 * the CPU oracle). MULTIMEM: in-switch reduction over the NVSwitch multicast mapping of the
 * arena (multimem.ld_reduce / multimem.st); summation order is the switch's, so results are
 * exact for integer-valued data and within 2(n-1)eps otherwise; needs the whole world in the
 * communicator; dlnb_get("multicast") tells whether it is available. */
enum dlnb_algo { DLNB_ALGO_AUTO = 0, DLNB_ALGO_ONESHOT = 1, DLNB_ALGO_TWOSHOT = 2, DLNB_ALGO_MULTIMEM = 3 };

/* buffer fill patterns used by tests/bench (SURVEY §8c pins) */
enum dlnb_pattern { DLNB_PAT_ZERO = 0, DLNB_PAT_INT = 1, DLNB_PAT_UNIFORM = 2 };

/* ---- process / device set-up ------------------------------------------------
 * Replaces MPI_Init + set_local_device + ncclGetUniqueId/MPI_Bcast/ncclCommInitRank
 * (cpp/data_parallel/dp.cpp:166-189, cpp/utils.hpp:74-117).
 * `session` names the rendezvous (abstract unix socket) shared by the `world`
 * processes of one box; rank 0 is the hub. world==1 needs no rendezvous. */
int dlnb_init(int rank, int world, int device, const char *session);
/* Replaces MPI_Finalize / CCLCommunicator::finalize (proxy_classes.hpp:229-234). */
int dlnb_finalize(void);
int dlnb_rank(void);
int dlnb_world(void);
int dlnb_device(void);
const char *dlnb_last_error(void);
const char *dlnb_version(void);
/* "vmm" (cuMem + POSIX fd export) or "ipc" (cudaIpc, diagnostic only) or "local" (world==1) */
const char *dlnb_arena_kind(void);

/* ---- host-side rendezvous helpers -------------------------------------------
 * Replace the MPI calls the mains make outside the hot path: MPI_Barrier
 * (dp.cpp:234), MPI_Allgather of strings (netcommunicators.hpp:19-77),
 * MPI_Allreduce/MPI_Bcast in estimate_runs (utils.hpp:127-132). */
int dlnb_host_barrier(void);
int dlnb_host_allgather(const void *in, size_t bytes, void *out);

/* ---- symmetric arena ---------------------------------------------------------
 * Replaces Tensor<T,Device::GPU>'s cudaMalloc + cudaMemset(0)
 * (cpp/proxy_classes.hpp:401-405). Collective: every rank calls it in the same
 * order with the same size. Memory is zero-filled. */
int dlnb_symm_alloc(size_t bytes, void **ptr);
/* Replaces CCUTILS_CUDA_FREE_SAFE (proxy_classes.hpp:434). Collective like the
 * allocation (same order on every rank): the block goes back to the arena and a later
 * dlnb_symm_alloc may reuse it (it is then cleared again, after all ranks went idle).
 * The physical memory returns to the driver at dlnb_finalize(). NULL is ignored; a
 * pointer that is not a live block is an error. */
int dlnb_symm_free(void *ptr);
size_t dlnb_symm_bytes_in_use(void);

/* ---- communicators -----------------------------------------------------------
 * Replaces `new CCLCommunicator(comm, num_streams)` / `new MPICommunicator(comm,
 * MPI_FLOAT, num_requests)` and MPI_Comm_split / ncclCommSplit
 * (proxy_classes.hpp:55-64,138-147; fsdp.cpp:258-283). `ranks` lists the world
 * ranks of the group in group order; must contain the caller. Collective over the
 * WORLD (every world rank creates and destroys its communicators in the same order).
 * At most 64 communicators are alive at a time: each owns a set of flag rows. Once 64
 * have been created, dlnb_comm_create reuses the rows of a destroyed communicator after
 * resetting them on every rank (all ranks idle -> clear -> all ranks cleared). */
int dlnb_comm_create(const int *ranks, int nranks, int num_slots, int dtype, dlnb_comm_t *out);
int dlnb_comm_destroy(dlnb_comm_t c);
int dlnb_comm_rank(dlnb_comm_t c);
int dlnb_comm_size(dlnb_comm_t c);
/* tuning knobs: "ctas" (CTAs per collective kernel), "mc_ctas" (the same for the in-switch
 * kernels; 0 = 16), "threads", "algo" (dlnb_algo; DLNB_ALGO_MULTIMEM also
 * sends allgather / reduce-scatter through the switch where the chunks are 16-byte aligned),
 * "twoshot_min_bytes" / "multimem_min_bytes" (AUTO thresholds), "timeout_ms" (flag-spin watchdog),
 * "hybrid_pct" (0..100, default 0 = off) / "hybrid_min_bytes" (default 32 MiB): an in-switch allreduce of at least
 * hybrid_min_bytes sends only the first hybrid_pct percent through the switch and the rest through the peer-to-peer
 * two-shot kernel on the second communication lane, both at the same time. */
int dlnb_comm_set(dlnb_comm_t c, const char *key, long value);
long dlnb_comm_get(dlnb_comm_t c, const char *key);

/* ---- emulated compute --------------------------------------------------------
 * Replaces usleep(fwd) / usleep(bwd) (dp.cpp:93,98; fsdp.cpp:103,120,145) with a
 * calibrated %globaltimer spin kernel on the compute stream. Asynchronous. */
int dlnb_compute_spin(double usec);
int dlnb_compute_sync(void);
/* dlnb_compute_mark: timestamp "the compute enqueued so far ends here" on the compute stream. dlnb_slot_exposed: how long
 * the collective recorded in `slot` kept running after that mark (0 if it finished earlier) — the wait that is LEFT after
 * the compute, which is what the reference's timers around Wait() hold (its host sleeps through the compute first:
 * fsdp.cpp:103-107,120-139). Both are no-ops inside a graph capture. */
int dlnb_compute_mark(void);
int dlnb_slot_exposed(dlnb_comm_t c, int slot, float *ms);
/* Consecutive spins chain: one that starts within "spin_slack_ns" (default 8000) of the
 * previous spin's nominal end counts from that end, so a chain of k spins lasts the sum of
 * the requests, not k launch gaps more.
 * dlnb_spin_calibrate measures two isolated spins with CUDA events: `slope` = event clock vs
 * %globaltimer rate (applied to every request), `overhead_us` = start-up latency of an
 * isolated spin (reported, not subtracted). */
int dlnb_spin_calibrate(double *overhead_us, double *slope);
/* Footprint of the emulated compute — what a collective has to share the GPU with:
 *   "spin_ctas" (0 = one CTA per SM, the default), "spin_threads" (default 256),
 *   "spin_mode" 0 = sleep (__nanosleep back-off: resident, issues nothing), 1 = FMA-busy
 *   (default: four dependent FMA chains per thread), 2 = HBM-busy (streams a 1 GiB buffer),
 *   "spin_smem" (dynamic shared memory per CTA, ballast only).
 *   dlnb_get("spin_coresident") = 1 when one spin CTA per SM leaves room for a collective CTA
 *   (512 threads x 96 registers) on the same SM; dlnb_get("spin_regs") = registers per thread.
 * "align_iterations" (default 1): dlnb_iter_begin / dlnb_batch_begin run one flag barrier between
 * all ranks on the device, so that every rank's iteration starts within a flag round trip.
 * "strict_barriers" (default 0): 1 = every flag barrier uses st.release.sys / ld.acquire.sys
 * (0: the barriers that publish no peer writes use relaxed flag traffic).
 * "spin_slack_ns": chaining window; "slot_timing": record a start event per slot;
 * "lanes" (1 or 2, default 2; all ranks alike): with 2, consecutive small allreduces (grid
 * <= 64 CTAs) alternate between two ordered communication streams with their own flag rows so
 * that their barrier latencies overlap; allgather, reduce-scatter and large allreduces are
 * ordered after and before everything else, as with one stream. */
int dlnb_set(const char *key, long value);
long dlnb_get(const char *key);

/* ---- collectives (asynchronous; ordered on the communication stream) ----------
 * Each first waits (device-side event) for the compute enqueued so far, so a
 * collective cannot start before "its" backward finished — this replaces the host
 * ordering usleep → Iallreduce of the reference. `slot` is the reference's
 * `index` (request / stream id). */
/* ProxyCommunicator::Iallreduce (proxy_classes.hpp:32,66-68,149-152). send == recv (in place, as
 * MPI_IN_PLACE / ncclAllReduce allow) is supported: overlapping buffers take the two-shot or the
 * in-switch kernel, where each element is read and written by one owner. Allgather,
 * reduce-scatter and all-to-all buffers must not overlap. */
int dlnb_iallreduce(dlnb_comm_t c, const void *send, void *recv, size_t count, int slot);
/* ProxyCommunicator::Iallgather (proxy_classes.hpp:34-35,91-93,202-204); recvcount==sendcount */
int dlnb_iallgather(dlnb_comm_t c, const void *send, void *recv, size_t sendcount, int slot);
/* non-blocking form of Reduce_Scatter_block, same semantics as below */
int dlnb_ireduce_scatter(dlnb_comm_t c, const void *send, void *recv, size_t recvcount, int slot);
/* ProxyCommunicator::Allreduce (proxy_classes.hpp:33,70-72,154-158): blocking */
int dlnb_allreduce(dlnb_comm_t c, const void *send, void *recv, size_t count);
/* ProxyCommunicator::Allgather (proxy_classes.hpp:36-37,95-97,197-200): blocking */
int dlnb_allgather(dlnb_comm_t c, const void *send, void *recv, size_t sendcount);
/* ProxyCommunicator::Reduce_Scatter_block (proxy_classes.hpp:38,99-101,206-209): blocking */
int dlnb_reduce_scatter(dlnb_comm_t c, const void *send, void *recv, size_t recvcount);
/* ProxyCommunicator::Alltoall (proxy_classes.hpp:39-40,74-77,160-182): block p of `send`
 * (sendcount elements at p*sendcount) lands in block `rank` of rank p's `recv`; recvcount ==
 * sendcount. The expert-parallel token exchange of the MoE proxy (hybrid_3d_moe.cpp:161-165). */
int dlnb_alltoall(dlnb_comm_t c, const void *send, void *recv, size_t sendcount);
int dlnb_ialltoall(dlnb_comm_t c, const void *send, void *recv, size_t sendcount, int slot);
/* ProxyCommunicator::send / recv / Isend / Irecv (proxy_classes.hpp:45-48,103-117,211-227): the
 * stage-to-stage traffic of the pipeline proxies (hybrid_2d.cpp:108-161). `dest` / `source` are
 * ranks of the communicator. The k-th send from a to b matches the k-th recv on b from a (no
 * tags: the proxies always use 0). A rendezvous put: the receiver posts where its buffer is, the
 * sender stores straight into it over NVLink. The blocking forms return when the message has
 * been delivered (send) / has arrived (recv). dlnb_irecv only posts the receive; the kernel that
 * waits for the data is enqueued by dlnb_wait / dlnb_waitall, so a posted receive never sits in
 * front of a send the peer is waiting for (Irecv + Isend rings are safe). At most 8 receives
 * from one source may be outstanding (posted and not yet waited for): the 9th dlnb_irecv fails. Needs the VMM arena (world > 1); not capturable. */
int dlnb_send(dlnb_comm_t c, const void *buf, size_t count, int dest);
int dlnb_recv(dlnb_comm_t c, void *buf, size_t count, int source);
int dlnb_isend(dlnb_comm_t c, const void *buf, size_t count, int dest, int slot);
int dlnb_irecv(dlnb_comm_t c, void *buf, size_t count, int source, int slot);
/* ProxyCommunicator::Wait / WaitAll / Barrier (proxy_classes.hpp:41-43,79-89,184-195) */
int dlnb_wait(dlnb_comm_t c, int slot);
int dlnb_waitall(dlnb_comm_t c, int nslots);
int dlnb_barrier(dlnb_comm_t c);
/* number of kernels this library has launched so far (bench "gpu_launches") */
uint64_t dlnb_launch_count(void);

/* ---- device timing -------------------------------------------------------------
 * Added (no reference counterpart; the reference only has host timers,
 * ccutils CCUTILS_MPI_TIMER_*): cudaEvent pair spanning both streams. */
/* With more than one rank dlnb_iter_begin / dlnb_batch_begin are COLLECTIVE over the world (every rank must call them the same
 * number of times): unless dlnb_set("align_iterations", 0), they enqueue one device-side flag barrier between all ranks so
 * that the iteration starts on all GPUs within a flag round trip (a rank that never arrives trips the 30 s watchdog). */
int dlnb_iter_begin(void);
int dlnb_iter_end(float *ms);
/* like dlnb_iter_begin, but the timer starts on the communication stream behind everything
 * enqueued so far (e.g. a long spin): times a pre-queued batch of collectives on the device
 * with the host out of the loop. Stop with dlnb_iter_end(). */
int dlnb_batch_begin(void);
/* Contention measurements: like dlnb_batch_begin / dlnb_iter_end, but the collectives issued between the two
 * calls do NOT wait for the compute stream — emulated compute enqueued before keeps running beside them — and
 * the end does not wait for it either. Always pair the two calls: until dlnb_batch_end() collectives keep ignoring the
 * compute stream. */
int dlnb_batch_begin_concurrent(double gate_us); /* gate_us > 0: a sleeping warp holds the stream that long while the host queues */
int dlnb_batch_end(float *ms);
/* device time of the most recent collective recorded in `slot` (ms); slot -1 = the most
 * recent BLOCKING call (those use an internal slot, never one of the caller's) */
int dlnb_slot_elapsed(dlnb_comm_t c, int slot, float *ms);

/* ---- another library's collectives on the communication stream (comparison builds only) --------
 * No reference counterpart. The `nccl` build (cpp/proxy_classes.hpp: CCLCommunicator, what the reference's
 * CCLCommunicator :135-253 does) enqueues ncclAllReduce & co. itself; these two calls give that work the
 * ordering and the slot bookkeeping a collective of this library gets: begin = wait for the compute enqueued
 * so far, start the slot's timer, hand out the stream (a cudaStream_t); end = stop the timer and record the
 * completion dlnb_wait / dlnb_waitall look at. The product path never calls them. Not capturable. */
int dlnb_foreign_begin(dlnb_comm_t c, int slot, void **cuda_stream);
int dlnb_foreign_end(dlnb_comm_t c, int slot);

/* ---- whole-iteration CUDA graphs ---------------------------------------------------
 * Added (no reference counterpart). For launch-bound configurations (vit_h with 4000 buckets:
 * 9 us of compute per bucket against ~30 us of host launch work) one proxy iteration —
 * every dlnb_compute_spin and collective issued between begin and end — is captured once and
 * replayed with a single launch. Inside a capture the host cannot block, so dlnb_wait /
 * dlnb_waitall / dlnb_barrier and the blocking collectives turn their blocking into a
 * device-side edge: the compute stream waits for the collective's completion event, i.e.
 * compute issued afterwards starts only when that collective is done — the schedule the
 * host-blocking calls produce outside a capture (fsdp: the backward of unit u-1 starts after
 * the ReduceScatter of unit u). dlnb_compute_sync returns at once. After dlnb_graph_launch the
 * host waits with dlnb_compute_sync() or dlnb_iter_end(). The kernels keep their barrier
 * generations and spin stamps in device memory, so a replay needs no patched arguments.
 * All ranks must capture the same sequence and launch the same number of replays. */
int dlnb_graph_begin(void);
int dlnb_graph_end(dlnb_graph_t *out);
int dlnb_graph_launch(dlnb_graph_t g);
int dlnb_graph_kernel_nodes(dlnb_graph_t g); /* kernels inside the graph (for launch accounting) */
int dlnb_graph_destroy(dlnb_graph_t g);

/* ---- buffer helpers (tests, bench, end-to-end path) ----------------------------- */
int dlnb_memcpy_h2d(void *dst, const void *src, size_t bytes);
int dlnb_memcpy_d2h(void *dst, const void *src, size_t bytes);
/* asynchronous copies on the communication stream from/to pinned host memory */
int dlnb_memcpy_h2d_async(void *dst, const void *src, size_t bytes);
int dlnb_memcpy_d2h_async(void *dst, const void *src, size_t bytes);
/* End-to-end path (no reference counterpart: the reference never moves its buffers). H2D on the copy
 * stream once the compute enqueued so far has finished ("this bucket's gradients exist when its backward
 * slice ends"); the next collective waits for it. D2H of a result on a stream of its own as soon as the
 * collective recorded in `slot` is done; dlnb_iter_end() waits for it. Pinned host memory (dlnb_host_alloc). */
int dlnb_memcpy_h2d_after_compute(void *dst, const void *src, size_t bytes);
int dlnb_memcpy_d2h_after_slot(dlnb_comm_t c, int slot, void *dst, const void *src, size_t bytes);
int dlnb_host_alloc(size_t bytes, void **ptr);
int dlnb_host_free(void *ptr);
/* x[i] = pattern(kind, seed, as_rank, i), i in [0,count) */
int dlnb_fill_pattern(void *ptr, size_t count, int dtype, int kind, uint64_t seed, int as_rank);
/* counts elements of `ptr` that differ (bitwise) from the rank-order fp32 sum over
 * `ranks` of pattern(kind, seed, r, elem_offset + i), rounded once to dtype. */
int dlnb_verify_sum_pattern(const void *ptr, size_t count, size_t elem_offset, int dtype, int kind,
                            uint64_t seed, const int *ranks, int nranks, uint64_t *mismatches);
/* Tolerance form, for reductions whose summation order is the NVSwitch's and not ours (DLNB_ALGO_MULTIMEM):
 * counts elements with |got - ref| > rel_tol * sum_r |x_r| + half an ulp of the dtype at |ref|, where ref is
 * the rank-order fp32 sum; SURVEY 8c pin (2) sets rel_tol = 2 (n-1) eps. `worst_rel_err` (may be NULL)
 * receives the largest |got - ref| / sum_r |x_r| seen. */
int dlnb_verify_sum_pattern_tol(const void *ptr, size_t count, size_t elem_offset, int dtype, int kind,
                                uint64_t seed, const int *ranks, int nranks, double rel_tol,
                                uint64_t *mismatches, double *worst_rel_err);
/* order-independent 64-bit checksum of a device buffer (sum of mixed 32-bit words) */
int dlnb_checksum(const void *ptr, size_t bytes, uint64_t *out);
/* writes a buffer larger than L2 (timing hygiene between timed iterations) */
int dlnb_flush_l2(void);

/* ---- single-GPU emulation of an n-rank collective --------------------------------
 * Test vehicle (no reference counterpart): runs the SAME device code as the
 * multi-GPU path for `nranks` virtual ranks inside one cooperative kernel launch on
 * one GPU (gridDim.y = virtual rank), including the flag barriers. send[r]/recv[r]
 * are ordinary device pointers (dlnb_symm_alloc or cudaMalloc). */
int dlnb_emu_allreduce(int nranks, const void *const *send, void *const *recv, size_t count,
                       int dtype, int algo, int ctas);
int dlnb_emu_allgather(int nranks, const void *const *send, void *const *recv, size_t sendcount,
                       int dtype, int ctas);
int dlnb_emu_reduce_scatter(int nranks, const void *const *send, void *const *recv,
                            size_t recvcount, int dtype, int ctas);
int dlnb_emu_alltoall(int nranks, const void *const *send, void *const *recv, size_t sendcount,
                      int dtype, int ctas);
/* send / recv (proxy_classes.hpp:103-117,211-227) on one GPU: the receive-post, send and
 * receive-wait kernels of dlnb_irecv / dlnb_isend / dlnb_wait in stream order against one mailbox
 * that plays both ends; `count` elements of `dtype` go from `send` to `recv` (any alignment).
 * World size 1 only. */
int dlnb_emu_sendrecv(const void *send, void *recv, size_t count, int dtype, int ctas);

/* ---- the emulated allreduce for callers that cannot allocate -----------------------
 * No counterpart in the reference (it has no such constraint: every backend there allocates
 * freely). Same arguments, same kernel and same results as dlnb_emu_allreduce(), but the
 * flag rows and barrier generations live in an 18 KiB array built into the library image
 * instead of coming from cudaMalloc on first use — the call makes no cudaMalloc / cudaFree
 * and no allocating library call. Meant for hosts that have handed all free HBM to their own
 * pool, or that must not trigger cudaMalloc's device-wide synchronisation while compute and
 * communication streams are busy.
 *   - limit: at most 64 CTAs per virtual rank (`ctas` 0 = default 16); a larger request is
 *     refused with an error, nothing is launched. nranks 1..8 as elsewhere.
 *   - the first call on a device looks up the array's device address (cudaGetSymbolAddress);
 *     make that first call before starting any stream capture. Between dlnb_graph_begin and
 *     dlnb_graph_end this entry point (like every dlnb_emu_* call) returns an error: it
 *     synchronises the communication stream.
 *   - all calls share the one array and it keeps the last call's contents; each call clears it
 *     on the communication stream before launching. Calls that overlap on the GPU would race
 *     on it: issue them one after another (the entry point returns after its kernel has
 *     finished, so a single host thread cannot overlap two).
 *   - every GPU that loads the library holds the 18 KiB until the process exits. */
int dlnb_emu_allreduce_noalloc(int nranks, const void *const *send, void *const *recv, size_t count,
                               int dtype, int algo, int ctas);

#ifdef __cplusplus
}
#endif
#endif /* DLNB_H */
