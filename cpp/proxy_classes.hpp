// cpp/proxy_classes.hpp — communicator interface, the B200 backend and Tensor.
//
// Drop-in for the reference's cpp/proxy_classes.hpp: `ProxyCommunicator` has the same 15
// virtuals with the same names, argument meaning and (void) error behaviour (:30-51), and
// `Tensor<T, Device>` the same public fields and RAII contract (:381-444). The backends of
// the reference (MPICommunicator :53-133, CCLCommunicator :135-253, OneCCLCommunicator
// :255-342) are replaced by ONE: B200Communicator, every method a single call into the C
// ABI of libdlnb.so (include/dlnb.h), whose kernels read and write peer HBM directly.
#ifndef PROXY_CLASSES_HPP
#define PROXY_CLASSES_HPP

#include <dlnb.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

#include <stdexcept>
#include <string>
#include <vector>

class ProxyCommunicator {
  public:
    virtual void Iallreduce(const void *sendbuf, void *recvbuf, int count, int index) = 0;
    virtual void Allreduce(const void *sendbuf, void *recvbuf, int count) = 0;
    virtual void Iallgather(const void *sendbuf, int sendcount, void *recvbuf, int recvcount, int index) = 0;
    virtual void Allgather(const void *sendbuf, int sendcount, void *recvbuf, int recvcount) = 0;
    virtual void Reduce_Scatter_block(const void *sendbuf, void *recvbuf, int recvcount) = 0;
    virtual void Alltoall(const void *sendbuf, int sendcount, void *recvbuf, int recvcount) = 0;
    virtual void Barrier() = 0;
    virtual void WaitAll(int num_waits) = 0;
    virtual void Wait(int index) = 0;
    virtual void finalize() = 0;
    virtual void send(const void *buf, int count, int dest) = 0;
    virtual void recv(void *buf, int count, int source) = 0;
    virtual void Isend(const void *buf, int count, int dest, int index) = 0;
    virtual void Irecv(void *buf, int count, int source, int index) = 0;
    virtual std::string get_name() = 0;
    virtual ~ProxyCommunicator() {}
};

// Errors do not propagate through the void interface: print "[rank r] what" and exit, as the
// reference does through CCUTILS_GPU_CHECK / MPI_ERRORS_ARE_FATAL.
#define B200_CHECK(call)                                                                           \
    do {                                                                                           \
        if ((call) != 0) {                                                                         \
            fprintf(stderr, "[rank %d] %s failed: %s\n", dlnb_rank(), #call, dlnb_last_error());   \
            fflush(stderr);                                                                        \
            _exit(1);                                                                              \
        }                                                                                          \
    } while (0)

class B200Communicator : public ProxyCommunicator {
  public:
    // `group` = world ranks of this communicator in group order (what MPI_Comm_split /
    // ncclCommSplit produce in fsdp.cpp:258-283); `num_slots` = the reference's
    // num_requests / num_streams; `dtype` = a dlnb_dtype.
    B200Communicator(const std::vector<int> &group, int num_slots, int dtype) {
        B200_CHECK(dlnb_comm_create(group.data(), (int)group.size(), num_slots > 0 ? num_slots : 1, dtype, &comm));
        rank = dlnb_comm_rank(comm);
        comm_size = dlnb_comm_size(comm);
    }

    void Iallreduce(const void *sendbuf, void *recvbuf, int count, int index) override {
        B200_CHECK(dlnb_iallreduce(comm, sendbuf, recvbuf, elems(count), index));
    }
    void Allreduce(const void *sendbuf, void *recvbuf, int count) override {
        B200_CHECK(dlnb_allreduce(comm, sendbuf, recvbuf, elems(count)));
    }
    void Iallgather(const void *sendbuf, int sendcount, void *recvbuf, int recvcount, int index) override {
        same(sendcount, recvcount);
        B200_CHECK(dlnb_iallgather(comm, sendbuf, recvbuf, elems(sendcount), index));
    }
    void Allgather(const void *sendbuf, int sendcount, void *recvbuf, int recvcount) override {
        same(sendcount, recvcount);
        B200_CHECK(dlnb_allgather(comm, sendbuf, recvbuf, elems(sendcount)));
    }
    void Reduce_Scatter_block(const void *sendbuf, void *recvbuf, int recvcount) override {
        B200_CHECK(dlnb_reduce_scatter(comm, sendbuf, recvbuf, elems(recvcount)));
    }
    void WaitAll(int num_waits) override { B200_CHECK(dlnb_waitall(comm, num_waits)); }
    void Wait(int index) override { B200_CHECK(dlnb_wait(comm, index)); }
    void Barrier() override { B200_CHECK(dlnb_barrier(comm)); }
    void finalize() override {
        if (comm) B200_CHECK(dlnb_comm_destroy(comm));
        comm = nullptr;
    }
    std::string get_name() override { return "B200P2P"; }

    // Stage-to-stage traffic of the pipeline proxies: a rendezvous put over peer memory.
    void send(const void *buf, int count, int dest) override { B200_CHECK(dlnb_send(comm, buf, elems(count), dest)); }
    void recv(void *buf, int count, int source) override { B200_CHECK(dlnb_recv(comm, buf, elems(count), source)); }
    void Isend(const void *buf, int count, int dest, int index) override {
        B200_CHECK(dlnb_isend(comm, buf, elems(count), dest, index));
    }
    void Irecv(void *buf, int count, int source, int index) override {
        B200_CHECK(dlnb_irecv(comm, buf, elems(count), source, index));
    }
    // Expert-parallel token exchange of the MoE proxy.
    void Alltoall(const void *sendbuf, int sendcount, void *recvbuf, int recvcount) override {
        same(sendcount, recvcount);
        B200_CHECK(dlnb_alltoall(comm, sendbuf, recvbuf, elems(sendcount)));
    }

    // additions
    // 64-bit element counts (the C ABI takes size_t). The interface above counts in int, and the
    // reference narrows its uint64_t sizes into it (dp.cpp:99, hybrid_2d.cpp:162, hybrid_3d_moe.cpp:207),
    // which wraps beyond 2^31-1 elements — e.g. the per-stage gradients of mixtral_8x7b on one box.
    virtual void Allreduce64(const void *sendbuf, void *recvbuf, uint64_t count) {
        B200_CHECK(dlnb_allreduce(comm, sendbuf, recvbuf, (size_t)count));
    }
    virtual void Iallreduce64(const void *sendbuf, void *recvbuf, uint64_t count, int index) {
        B200_CHECK(dlnb_iallreduce(comm, sendbuf, recvbuf, (size_t)count, index));
    }
    virtual void set(const char *key, long v) { B200_CHECK(dlnb_comm_set(comm, key, v)); }
    float slot_ms(int index) {
        float ms = 0;
        B200_CHECK(dlnb_slot_elapsed(comm, index, &ms));
        return ms;
    }
    dlnb_comm_t handle() const { return comm; }
    int size() const { return comm_size; }

  protected:
    static size_t elems(int count) {
        if (count < 0) {
            // the reference narrows uint64_t sizes to int (dp.cpp:99, fsdp.cpp:88): diagnose the wrap
            fprintf(stderr, "[rank %d] element count %d is negative: the message exceeds INT_MAX elements; "
                            "use more buckets / units\n", dlnb_rank(), count);
            _exit(1);
        }
        return (size_t)count;
    }
    static void same(int a, int b) {
        if (a != b) {
            fprintf(stderr, "[rank %d] sendcount %d != recvcount %d is not supported\n", dlnb_rank(), a, b);
            _exit(1);
        }
    }
    dlnb_comm_t comm = nullptr;
    int comm_size = 1;
    int rank = 0;
};

#ifdef PROXY_ENABLE_NCCL
// ------------------------------------------------------------------------------------------------
// COMPARISON backend, `nccl` build only (Makefile.flags.mk): what the reference's CCLCommunicator does
// (cpp/proxy_classes.hpp:135-253) — every collective one NCCL call — inside the same proxies, beside the
// same emulated compute, on the same buffers and with the same slot / timing bookkeeping as the product
// (dlnb_foreign_begin / dlnb_foreign_end put the NCCL call on libdlnb's communication stream, ordered
// after the compute enqueued so far). Never part of the `b200` build; there is no run-time switch.
// ------------------------------------------------------------------------------------------------
#include <nccl.h>

#define NCCL_CHECK(call)                                                                          \
    do {                                                                                          \
        ncclResult_t r_ = (call);                                                                 \
        if (r_ != ncclSuccess) {                                                                  \
            fprintf(stderr, "[rank %d] %s failed: %s\n", dlnb_rank(), #call, ncclGetErrorString(r_)); \
            fflush(stderr);                                                                       \
            _exit(1);                                                                             \
        }                                                                                         \
    } while (0)

class CCLCommunicator : public B200Communicator {
  public:
    CCLCommunicator(const std::vector<int> &group, int num_slots, int dtype) : B200Communicator(group, num_slots, dtype) {
        nccl_dtype = dtype == DLNB_F32 ? ncclFloat : dtype == DLNB_F16 ? ncclHalf : ncclBfloat16;
        // one ncclUniqueId per group: every rank offers one, the group uses its first member's
        // (the reference broadcasts rank 0's over MPI, dp.cpp:181-186)
        ncclUniqueId mine;
        NCCL_CHECK(ncclGetUniqueId(&mine));
        std::vector<ncclUniqueId> all((size_t)dlnb_world());
        B200_CHECK(dlnb_host_allgather(&mine, sizeof(mine), all.data()));
        NCCL_CHECK(ncclCommInitRank(&nccl, (int)group.size(), all[(size_t)group[0]], dlnb_comm_rank(handle())));
    }
    std::string get_name() override { return "NCCL"; }
    void set(const char *, long) override {} // kernel knobs of the product do not apply

    void Iallreduce(const void *s, void *r, int count, int index) override { Iallreduce64(s, r, (uint64_t)elems(count), index); }
    void Allreduce(const void *s, void *r, int count) override { Allreduce64(s, r, (uint64_t)elems(count)); }
    void Iallreduce64(const void *s, void *r, uint64_t count, int index) override {
        run(index, [&](cudaStream_t st) { NCCL_CHECK(ncclAllReduce(s, r, (size_t)count, nccl_dtype, ncclSum, nccl, st)); });
    }
    void Allreduce64(const void *s, void *r, uint64_t count) override {
        Iallreduce64(s, r, count, -1);
        Wait(-1);
    }
    void Iallgather(const void *s, int sendcount, void *r, int recvcount, int index) override {
        same(sendcount, recvcount);
        run(index, [&](cudaStream_t st) { NCCL_CHECK(ncclAllGather(s, r, elems(sendcount), nccl_dtype, nccl, st)); });
    }
    void Allgather(const void *s, int sendcount, void *r, int recvcount) override {
        Iallgather(s, sendcount, r, recvcount, -1);
        Wait(-1);
    }
    void Reduce_Scatter_block(const void *s, void *r, int recvcount) override {
        run(-1, [&](cudaStream_t st) { NCCL_CHECK(ncclReduceScatter(s, r, elems(recvcount), nccl_dtype, ncclSum, nccl, st)); });
        Wait(-1);
    }
    void Alltoall(const void *s, int sendcount, void *r, int recvcount) override { // grouped send/recv, as the reference :160-182
        same(sendcount, recvcount);
        const size_t bytes = elems(sendcount) * (nccl_dtype == ncclFloat ? 4 : 2);
        run(-1, [&](cudaStream_t st) {
            NCCL_CHECK(ncclGroupStart());
            for (int p = 0; p < size(); ++p) {
                NCCL_CHECK(ncclSend((const char *)s + (size_t)p * bytes, elems(sendcount), nccl_dtype, p, nccl, st));
                NCCL_CHECK(ncclRecv((char *)r + (size_t)p * bytes, elems(recvcount), nccl_dtype, p, nccl, st));
            }
            NCCL_CHECK(ncclGroupEnd());
        });
        Wait(-1);
    }
    void send(const void *buf, int count, int dest) override { Isend(buf, count, dest, -1); Wait(-1); }
    void recv(void *buf, int count, int source) override { Irecv(buf, count, source, -1); Wait(-1); }
    void Isend(const void *buf, int count, int dest, int index) override {
        run(index, [&](cudaStream_t st) { NCCL_CHECK(ncclSend(buf, elems(count), nccl_dtype, dest, nccl, st)); });
    }
    void Irecv(void *buf, int count, int source, int index) override {
        run(index, [&](cudaStream_t st) { NCCL_CHECK(ncclRecv(buf, elems(count), nccl_dtype, source, nccl, st)); });
    }
    void finalize() override {
        if (nccl) NCCL_CHECK(ncclCommDestroy(nccl)); // (the reference never gets here: `#ifdef PROXY_ENABLE_CLL`, dp.cpp:298)
        nccl = nullptr;
        B200Communicator::finalize();
    }

  private:
    template <typename Fn> void run(int slot, Fn fn) {
        void *st = nullptr;
        B200_CHECK(dlnb_foreign_begin(handle(), slot, &st));
        fn((cudaStream_t)st);
        B200_CHECK(dlnb_foreign_end(handle(), slot));
    }
    ncclComm_t nccl = nullptr;
    ncclDataType_t nccl_dtype = ncclFloat;
};
using ProxyBackend = CCLCommunicator;
#else
using ProxyBackend = B200Communicator;
#endif

// Allreduce of `count` elements through the interface when the count fits its int, through the
// 64-bit entry of the B200 backend when it does not (another backend would wrap: refuse).
inline void proxy_allreduce(ProxyCommunicator *c, const void *sendbuf, void *recvbuf, uint64_t count) {
    if (count <= 0x7fffffffull) return c->Allreduce(sendbuf, recvbuf, (int)count);
    B200Communicator *b = dynamic_cast<B200Communicator *>(c);
    if (!b) {
        fprintf(stderr, "[rank %d] allreduce of %llu elements exceeds the interface's int count\n", dlnb_rank(), (unsigned long long)count);
        _exit(1);
    }
    b->Allreduce64(sendbuf, recvbuf, count);
}
inline void proxy_iallreduce(ProxyCommunicator *c, const void *sendbuf, void *recvbuf, uint64_t count, int index) {
    if (count <= 0x7fffffffull) return c->Iallreduce(sendbuf, recvbuf, (int)count, index);
    B200Communicator *b = dynamic_cast<B200Communicator *>(c);
    if (!b) {
        fprintf(stderr, "[rank %d] allreduce of %llu elements exceeds the interface's int count\n", dlnb_rank(), (unsigned long long)count);
        _exit(1);
    }
    b->Iallreduce64(sendbuf, recvbuf, count, index);
}

enum class Device { CPU, GPU };

// Contiguous zero-filled buffer. GPU tensors live in libdlnb's symmetric arena, so that every
// rank's kernels can address every other rank's copy (construction is collective: same order
// and sizes on all ranks, which dp.cpp:229-232 / fsdp.cpp:348-354 already guarantee).
template <typename T, Device device = Device::CPU> class Tensor {
  public:
    T *data = nullptr;
    uint64_t size = 0;

    explicit Tensor(uint64_t size_) : size(size_) {
        if constexpr (device == Device::CPU) {
            data = static_cast<T *>(calloc(size ? size : 1, sizeof(T)));
            if (!data) throw std::runtime_error("Failed to allocate CPU memory");
        } else {
            void *p = nullptr;
            B200_CHECK(dlnb_symm_alloc(size * sizeof(T), &p));
            data = static_cast<T *>(p);
        }
    }
    Tensor(const Tensor &) = delete;
    Tensor &operator=(const Tensor &) = delete;
    ~Tensor() {
        if (!data) return;
        if constexpr (device == Device::CPU) free(data);
        else dlnb_symm_free(data);
    }
};

#endif // PROXY_CLASSES_HPP
