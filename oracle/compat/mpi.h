/*
 * oracle/compat/mpi.h — a minimal MPI over POSIX shared memory.
 *
 * TEST INFRASTRUCTURE ONLY (part of the oracle). The target image has no MPI, so the
 * reference's own data-parallel sources (cpp/data_parallel/dp.cpp, fsdp.cpp,
 * cpp/proxy_classes.hpp MPICommunicator, cpp/utils.hpp, cpp/netcommunicators.hpp)
 * are compiled UNMODIFIED against this header to obtain the reference's `mpi_cpu`
 * path (oracle/Makefile `ref`). It provides exactly the symbols those files use.
 *
 * Semantics follow the MPI standard for the calls provided. Reductions are
 * MPI_SUM over MPI_FLOAT only, accumulated in fp32 in fixed rank order 0..n-1
 * (deterministic; MPI leaves the order to the implementation). Non-blocking
 * collectives make progress only inside MPI_Wait/MPI_Waitall, like an MPI without
 * an asynchronous progress thread. One process per rank, started by
 * oracle/_ref/dlnb_mpirun, which creates the shared segment and sets
 * DLNB_MPI_RANK / DLNB_MPI_SIZE / DLNB_MPI_SHM.
 */
#ifndef DLNB_COMPAT_MPI_H
#define DLNB_COMPAT_MPI_H

#include <fcntl.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <vector>

#define MPI_SUCCESS 0
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_THREAD_MULTIPLE 3
#define MPI_COMM_TYPE_SHARED 1
#define MPI_INFO_NULL 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)
#define MPI_UNDEFINED (-32766)

typedef int MPI_Datatype;
#define MPI_CHAR 1
#define MPI_BYTE 2
#define MPI_INT 3
#define MPI_FLOAT 4
#define MPI_DOUBLE 5
#define MPI_UINT64_T 6
typedef int MPI_Op;
#define MPI_SUM 1
typedef int MPI_Info;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

namespace dlnb_mpi {

constexpr int MAX_RANKS = 64;
constexpr int MAX_COMMS = 256;
constexpr size_t SLOT_BYTES = 8u << 20; /* per-rank staging: IN half and OUT half */
constexpr size_t HALF = SLOT_BYTES / 2;

struct alignas(64) BarrierState {
    std::atomic<int> count;
    std::atomic<int> sense;
    char pad[56];
};
/* one mailbox per ordered pair of world ranks: blocking Send/Recv hand chunks over through the
 * sender's staging slot (state 0 = empty, else = bytes of the chunk that is ready) */
struct alignas(64) Mailbox {
    std::atomic<uint64_t> ready;
    char pad[56];
};
struct Shared {
    std::atomic<int> next_comm;
    std::atomic<int> abort_flag;
    char pad[56];
    BarrierState bar[MAX_COMMS];
    Mailbox box[MAX_RANKS][MAX_RANKS]; /* [source][destination] */
    /* followed by MAX_RANKS-many slots of SLOT_BYTES (only world of them are touched) */
};
inline size_t shared_bytes(int world) { return sizeof(Shared) + (size_t)world * SLOT_BYTES + 4096; }

struct Comm {
    int id = 0;          /* index into Shared::bar */
    int size = 1, rank = 0;
    int world_rank[MAX_RANKS] = {};
    int local_sense = 0;
};

enum ReqKind { REQ_NONE = 0, REQ_IALLREDUCE, REQ_IALLGATHER };
struct Request {
    int kind = REQ_NONE;
    const void *send = nullptr;
    void *recv = nullptr;
    int count = 0;
    MPI_Datatype dt = MPI_FLOAT;
    Comm *comm = nullptr;
};

struct State {
    bool inited = false;
    int rank = 0, world = 1;
    Shared *sh = nullptr;
    char *slots = nullptr;
    Comm world_comm;
    std::vector<Comm *> comms;
};
inline State &S() {
    static State s;
    return s;
}
inline char *slot_in(int world_rank) { return S().slots + (size_t)world_rank * SLOT_BYTES; }
inline char *slot_out(int world_rank) { return S().slots + (size_t)world_rank * SLOT_BYTES + HALF; }

[[noreturn]] inline void die(const char *msg) {
    fprintf(stderr, "[mini-MPI rank %d] %s\n", S().rank, msg);
    if (S().sh) S().sh->abort_flag.store(1);
    _exit(3);
}

inline size_t dt_size(MPI_Datatype d) {
    switch (d) {
    case MPI_CHAR: case MPI_BYTE: return 1;
    case MPI_INT: case MPI_FLOAT: return 4;
    case MPI_DOUBLE: case MPI_UINT64_T: return 8;
    }
    die("unsupported datatype");
}

inline void barrier(Comm *c) {
    if (c->size == 1) return;
    BarrierState &b = S().sh->bar[c->id];
    const int my = c->local_sense ^= 1;
    if (b.count.fetch_add(1, std::memory_order_acq_rel) == c->size - 1) {
        b.count.store(0, std::memory_order_relaxed);
        b.sense.store(my, std::memory_order_release);
    } else {
        unsigned spins = 0;
        while (b.sense.load(std::memory_order_acquire) != my) {
            if (++spins > 2000) {
                sched_yield();
                if ((spins & 0xffff) == 0 && S().sh->abort_flag.load()) _exit(3);
            }
        }
    }
}

/* recv = sum over ranks of send, fp32, rank order; chunked through the staging slots */
inline void allreduce_f32(const float *send, float *recv, size_t count, Comm *c) {
    if (c->size == 1) {
        if (recv != send) memcpy(recv, send, count * sizeof(float));
        return;
    }
    const size_t CH = HALF / sizeof(float);
    const int n = c->size, me = c->rank;
    for (size_t base = 0; base < count; base += CH) {
        const size_t len = std::min(CH, count - base);
        memcpy(slot_in(c->world_rank[me]), send + base, len * sizeof(float));
        barrier(c);
        /* I reduce my 1/n of the chunk, into my OUT half at the same offsets */
        const size_t per = (len + n - 1) / n, p0 = std::min(per * me, len), p1 = std::min(p0 + per, len);
        float *out = (float *)slot_out(c->world_rank[me]);
        const float *in0 = (const float *)slot_in(c->world_rank[0]);
        for (size_t i = p0; i < p1; ++i) out[i] = in0[i];
        for (int r = 1; r < n; ++r) {
            const float *in = (const float *)slot_in(c->world_rank[r]);
            for (size_t i = p0; i < p1; ++i) out[i] += in[i];
        }
        barrier(c);
        for (int r = 0; r < n; ++r) {
            const size_t q0 = std::min(per * r, len), q1 = std::min(q0 + per, len);
            memcpy(recv + base + q0, (const float *)slot_out(c->world_rank[r]) + q0, (q1 - q0) * sizeof(float));
        }
        barrier(c);
    }
}

inline void allgather_bytes(const char *send, size_t sbytes, char *recv, Comm *c) {
    if (c->size == 1) {
        if (recv != send) memcpy(recv, send, sbytes);
        return;
    }
    const int n = c->size, me = c->rank;
    for (size_t base = 0; base < sbytes || base == 0; base += SLOT_BYTES) {
        const size_t len = std::min(SLOT_BYTES, sbytes - base);
        memcpy(slot_in(c->world_rank[me]), send + base, len);
        barrier(c);
        for (int r = 0; r < n; ++r) memcpy(recv + (size_t)r * sbytes + base, slot_in(c->world_rank[r]), len);
        barrier(c);
        if (sbytes == 0) break;
    }
}

inline void allgatherv_bytes(const char *send, size_t sbytes, char *recv, const int *counts, const int *displs,
                             size_t esz, Comm *c) {
    const int n = c->size, me = c->rank;
    if (n == 1) {
        memcpy(recv + (size_t)displs[0] * esz, send, sbytes);
        return;
    }
    if (sbytes > SLOT_BYTES) die("MPI_Allgatherv: message larger than the staging slot");
    memcpy(slot_in(c->world_rank[me]), send, sbytes);
    barrier(c);
    for (int r = 0; r < n; ++r) memcpy(recv + (size_t)displs[r] * esz, slot_in(c->world_rank[r]), (size_t)counts[r] * esz);
    barrier(c);
}

/* rank me receives sum over ranks of block `me` (recvcount floats) */
inline void reduce_scatter_block_f32(const float *send, float *recv, size_t recvcount, Comm *c) {
    if (c->size == 1) {
        if (recv != send) memcpy(recv, send, recvcount * sizeof(float));
        return;
    }
    const int n = c->size, me = c->rank;
    const size_t CH = SLOT_BYTES / sizeof(float) / n; /* elements per destination per round */
    for (size_t base = 0; base < recvcount; base += CH) {
        const size_t len = std::min(CH, recvcount - base);
        float *mine = (float *)slot_in(c->world_rank[me]);
        for (int d = 0; d < n; ++d) memcpy(mine + (size_t)d * CH, send + (size_t)d * recvcount + base, len * sizeof(float));
        barrier(c);
        const float *in0 = (const float *)slot_in(c->world_rank[0]) + (size_t)me * CH;
        for (size_t i = 0; i < len; ++i) recv[base + i] = in0[i];
        for (int r = 1; r < n; ++r) {
            const float *in = (const float *)slot_in(c->world_rank[r]) + (size_t)me * CH;
            for (size_t i = 0; i < len; ++i) recv[base + i] += in[i];
        }
        barrier(c);
    }
}

/* block d of my send buffer goes to rank d; I receive block `me` of everybody's, in rank order */
inline void alltoall_bytes(const char *send, size_t bbytes, char *recv, Comm *c) {
    if (c->size == 1) {
        if (recv != send) memcpy(recv, send, bbytes);
        return;
    }
    const int n = c->size, me = c->rank;
    const size_t CH = SLOT_BYTES / (size_t)n; /* bytes per destination per round */
    for (size_t base = 0; base < bbytes || base == 0; base += CH) {
        const size_t len = std::min(CH, bbytes - base);
        char *mine = slot_in(c->world_rank[me]);
        for (int d = 0; d < n; ++d) memcpy(mine + (size_t)d * CH, send + (size_t)d * bbytes + base, len);
        barrier(c);
        for (int r = 0; r < n; ++r) memcpy(recv + (size_t)r * bbytes + base, slot_in(c->world_rank[r]) + (size_t)me * CH, len);
        barrier(c);
        if (bbytes == 0) break;
    }
}

inline void bcast_bytes(char *buf, size_t bytes, int root, Comm *c) {
    if (c->size == 1) return;
    for (size_t base = 0; base < bytes; base += SLOT_BYTES) {
        const size_t len = std::min(SLOT_BYTES, bytes - base);
        if (c->rank == root) memcpy(slot_in(c->world_rank[root]), buf + base, len);
        barrier(c);
        if (c->rank != root) memcpy(buf + base, slot_in(c->world_rank[root]), len);
        barrier(c);
    }
}

/* Blocking point-to-point, in order per (source, destination) pair; tags are not needed by
 * the proxies (always 0). The payload goes through the sender's staging slot in chunks. */
inline void wait_box(std::atomic<uint64_t> &a, bool want_full) {
    unsigned spins = 0;
    while ((a.load(std::memory_order_acquire) != 0) != want_full) {
        if (++spins > 2000) {
            sched_yield();
            if ((spins & 0xffff) == 0 && S().sh->abort_flag.load()) _exit(3);
        }
    }
}
inline void send_bytes(const char *buf, size_t bytes, int dest, Comm *c) {
    if (!S().sh) die("MPI_Send with a single rank");
    const int me = c->world_rank[c->rank], to = c->world_rank[dest];
    Mailbox &mb = S().sh->box[me][to];
    size_t base = 0;
    do {
        const size_t len = std::min(SLOT_BYTES, bytes - base);
        wait_box(mb.ready, false);
        memcpy(slot_in(me), buf + base, len);
        mb.ready.store(len ? len : (uint64_t)-1, std::memory_order_release); /* zero-length message still signals */
        base += len;
    } while (base < bytes);
    wait_box(mb.ready, false); /* the receiver has copied the last chunk: the slot is free again */
}
inline void recv_bytes(char *buf, size_t bytes, int source, Comm *c) {
    if (!S().sh) die("MPI_Recv with a single rank");
    const int me = c->world_rank[c->rank], from = c->world_rank[source];
    Mailbox &mb = S().sh->box[from][me];
    size_t base = 0;
    do {
        wait_box(mb.ready, true);
        uint64_t len = mb.ready.load(std::memory_order_acquire);
        if (len == (uint64_t)-1) len = 0;
        if (base + len > bytes) die("MPI_Recv: message longer than the receive buffer");
        memcpy(buf + base, slot_in(from), len);
        mb.ready.store(0, std::memory_order_release);
        base += len;
    } while (base < bytes);
}

inline void run_request(Request *r) {
    switch (r->kind) {
    case REQ_IALLREDUCE: allreduce_f32((const float *)r->send, (float *)r->recv, (size_t)r->count, r->comm); break;
    case REQ_IALLGATHER:
        allgather_bytes((const char *)r->send, (size_t)r->count * dt_size(r->dt), (char *)r->recv, r->comm);
        break;
    default: break;
    }
    r->kind = REQ_NONE;
}

} // namespace dlnb_mpi

typedef dlnb_mpi::Comm *MPI_Comm;
typedef dlnb_mpi::Request MPI_Request;
#define MPI_COMM_WORLD (&dlnb_mpi::S().world_comm)
#define MPI_COMM_NULL ((MPI_Comm)0)

inline int MPI_Init(int *, char ***) {
    using namespace dlnb_mpi;
    State &s = S();
    const char *r = getenv("DLNB_MPI_RANK"), *w = getenv("DLNB_MPI_SIZE"), *name = getenv("DLNB_MPI_SHM");
    s.rank = r ? atoi(r) : 0;
    s.world = w ? atoi(w) : 1;
    if (s.world > MAX_RANKS) die("too many ranks");
    if (s.world > 1) {
        if (!name) die("DLNB_MPI_SHM not set: start the program with dlnb_mpirun");
        int fd = shm_open(name, O_RDWR, 0600);
        if (fd < 0) die("shm_open failed");
        void *p = mmap(nullptr, shared_bytes(s.world), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        close(fd);
        if (p == MAP_FAILED) die("mmap failed");
        s.sh = (Shared *)p;
        s.slots = (char *)(((uintptr_t)p + sizeof(Shared) + 4095) & ~(uintptr_t)4095);
    }
    s.world_comm.id = 0;
    s.world_comm.size = s.world;
    s.world_comm.rank = s.rank;
    for (int i = 0; i < s.world; ++i) s.world_comm.world_rank[i] = i;
    s.inited = true;
    barrier(&s.world_comm);
    return MPI_SUCCESS;
}
inline int MPI_Init_thread(int *a, char ***b, int, int *provided) {
    if (provided) *provided = MPI_THREAD_MULTIPLE;
    return MPI_Init(a, b);
}
inline int MPI_Finalize() {
    using namespace dlnb_mpi;
    if (S().inited) barrier(&S().world_comm);
    S().inited = false;
    return MPI_SUCCESS;
}
inline int MPI_Abort(MPI_Comm, int code) { dlnb_mpi::die("MPI_Abort"); return code; }
inline int MPI_Comm_size(MPI_Comm c, int *n) { *n = c->size; return MPI_SUCCESS; }
inline int MPI_Comm_rank(MPI_Comm c, int *r) { *r = c->rank; return MPI_SUCCESS; }
inline double MPI_Wtime() {
    timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}
inline int MPI_Get_processor_name(char *name, int *len) {
    if (gethostname(name, MPI_MAX_PROCESSOR_NAME) != 0) strcpy(name, "localhost");
    name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
    *len = (int)strlen(name);
    return MPI_SUCCESS;
}
inline int MPI_Barrier(MPI_Comm c) { dlnb_mpi::barrier(c); return MPI_SUCCESS; }

inline int MPI_Allgather(const void *s, int sc, MPI_Datatype sd, void *r, int, MPI_Datatype, MPI_Comm c) {
    dlnb_mpi::allgather_bytes((const char *)s, (size_t)sc * dlnb_mpi::dt_size(sd), (char *)r, c);
    return MPI_SUCCESS;
}
inline int MPI_Allgatherv(const void *s, int sc, MPI_Datatype sd, void *r, const int *counts, const int *displs,
                          MPI_Datatype, MPI_Comm c) {
    dlnb_mpi::allgatherv_bytes((const char *)s, (size_t)sc * dlnb_mpi::dt_size(sd), (char *)r, counts, displs,
                               dlnb_mpi::dt_size(sd), c);
    return MPI_SUCCESS;
}
inline int MPI_Bcast(void *buf, int count, MPI_Datatype d, int root, MPI_Comm c) {
    dlnb_mpi::bcast_bytes((char *)buf, (size_t)count * dlnb_mpi::dt_size(d), root, c);
    return MPI_SUCCESS;
}
inline int MPI_Allreduce(const void *s, void *r, int count, MPI_Datatype d, MPI_Op op, MPI_Comm c) {
    if (d != MPI_FLOAT || op != MPI_SUM) dlnb_mpi::die("MPI_Allreduce: only MPI_FLOAT/MPI_SUM");
    dlnb_mpi::allreduce_f32((const float *)s, (float *)r, (size_t)count, c);
    return MPI_SUCCESS;
}
inline int MPI_Reduce_scatter_block(const void *s, void *r, int recvcount, MPI_Datatype d, MPI_Op op, MPI_Comm c) {
    if (d != MPI_FLOAT || op != MPI_SUM) dlnb_mpi::die("MPI_Reduce_scatter_block: only MPI_FLOAT/MPI_SUM");
    dlnb_mpi::reduce_scatter_block_f32((const float *)s, (float *)r, (size_t)recvcount, c);
    return MPI_SUCCESS;
}
inline int MPI_Iallreduce(const void *s, void *r, int count, MPI_Datatype d, MPI_Op op, MPI_Comm c, MPI_Request *q) {
    if (d != MPI_FLOAT || op != MPI_SUM) dlnb_mpi::die("MPI_Iallreduce: only MPI_FLOAT/MPI_SUM");
    *q = MPI_Request{dlnb_mpi::REQ_IALLREDUCE, s, r, count, d, c};
    return MPI_SUCCESS;
}
inline int MPI_Iallgather(const void *s, int sc, MPI_Datatype sd, void *r, int, MPI_Datatype, MPI_Comm c,
                          MPI_Request *q) {
    *q = MPI_Request{dlnb_mpi::REQ_IALLGATHER, s, r, sc, sd, c};
    return MPI_SUCCESS;
}
inline int MPI_Wait(MPI_Request *q, MPI_Status *) { dlnb_mpi::run_request(q); return MPI_SUCCESS; }
inline int MPI_Waitall(int n, MPI_Request *q, MPI_Status *) {
    for (int i = 0; i < n; ++i) dlnb_mpi::run_request(&q[i]);
    return MPI_SUCCESS;
}

/* MPI_Comm_split: members of the parent exchange (color, key); groups are ordered by
 * (key, parent rank); the lowest parent rank of each group draws the new id. */
inline int MPI_Comm_split(MPI_Comm parent, int color, int key, MPI_Comm *out) {
    using namespace dlnb_mpi;
    const int n = parent->size;
    std::vector<int> mine = {color, key}, all((size_t)2 * n);
    allgather_bytes((const char *)mine.data(), 2 * sizeof(int), (char *)all.data(), parent);
    std::vector<int> members;
    for (int r = 0; r < n; ++r)
        if (all[2 * r] == color) members.push_back(r);
    std::stable_sort(members.begin(), members.end(), [&](int a, int b) { return all[2 * a + 1] < all[2 * b + 1]; });
    int leader = *std::min_element(members.begin(), members.end());
    int id = -1;
    if (S().sh && parent->rank == leader) {
        id = S().sh->next_comm.fetch_add(1) + 1;
        if (id >= MAX_COMMS) die("too many communicators");
    }
    std::vector<int> ids((size_t)n);
    allgather_bytes((const char *)&id, sizeof(int), (char *)ids.data(), parent);
    if (color == MPI_UNDEFINED) {
        *out = MPI_COMM_NULL;
        return MPI_SUCCESS;
    }
    Comm *c = new Comm;
    c->id = S().sh ? ids[leader] : 0;
    c->size = (int)members.size();
    for (int i = 0; i < c->size; ++i) {
        c->world_rank[i] = parent->world_rank[members[i]];
        if (members[i] == parent->rank) c->rank = i;
    }
    /* the barrier slot may have been used before by a freed communicator: start from its sense */
    c->local_sense = S().sh ? S().sh->bar[c->id].sense.load() : 0;
    S().comms.push_back(c);
    *out = c;
    return MPI_SUCCESS;
}
inline int MPI_Comm_split_type(MPI_Comm parent, int, int key, MPI_Info, MPI_Comm *out) {
    return MPI_Comm_split(parent, 0, key, out); /* one host: everybody shares memory */
}
inline int MPI_Comm_free(MPI_Comm *c) { *c = MPI_COMM_NULL; return MPI_SUCCESS; }

/* Present in the reference's MPICommunicator but never called by dp / fsdp
 * (SURVEY §2b "never called"): provided so the class compiles; abort if reached. */
inline int MPI_Alltoall(const void *s, int sc, MPI_Datatype sd, void *r, int, MPI_Datatype, MPI_Comm c) {
    dlnb_mpi::alltoall_bytes((const char *)s, (size_t)sc * dlnb_mpi::dt_size(sd), (char *)r, c);
    return MPI_SUCCESS;
}
inline int MPI_Send(const void *b, int count, MPI_Datatype d, int dest, int, MPI_Comm c) {
    dlnb_mpi::send_bytes((const char *)b, (size_t)count * dlnb_mpi::dt_size(d), dest, c);
    return MPI_SUCCESS;
}
inline int MPI_Recv(void *b, int count, MPI_Datatype d, int source, int, MPI_Comm c, MPI_Status *) {
    dlnb_mpi::recv_bytes((char *)b, (size_t)count * dlnb_mpi::dt_size(d), source, c);
    return MPI_SUCCESS;
}
inline int MPI_Isend(const void *, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request *) { dlnb_mpi::die("MPI_Isend not provided"); }
inline int MPI_Irecv(void *, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request *) { dlnb_mpi::die("MPI_Irecv not provided"); }

#endif /* DLNB_COMPAT_MPI_H */
