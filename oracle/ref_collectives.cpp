// oracle/ref_collectives.cpp — TEST INFRASTRUCTURE (oracle).
// Buffer-level harness around the REFERENCE's own communicator: includes the reference's
// cpp/proxy_classes.hpp unmodified (MPICommunicator :53-133, Tensor :381-444) over
// oracle/compat/mpi.h, fills the send buffers with a seeded pattern (the reference itself
// only ever uses zeros), calls the same virtuals dp.cpp / fsdp.cpp call, and writes each
// rank's receive buffer to <prefix>.r<rank>.bin (raw fp32 — the MPI path is fp32-only,
// proxy_classes.hpp:67,71,100).
//
//   dlnb_mpirun -n K ref_collectives <allreduce|iallreduce|allgather|iallgather|reduce_scatter|alltoall>
//                                   <count> <kind> <seed> <prefix>
#include <mpi.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#include "proxy_classes.hpp" // the reference's, via -I$(REF)/cpp

static inline uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
static float pattern_value(int kind, uint64_t seed, int r, uint64_t i) {
    if (kind == 1) {
        const uint64_t h = i * 2654435761ull + (uint64_t)r * 40503ull;
        return (float)((int)(h % 17ull) - 8);
    } else if (kind == 2) {
        uint32_t h = mix32((uint32_t)i ^ mix32((uint32_t)(i >> 32) + 0x9e3779b9u * (uint32_t)(r + 1) + (uint32_t)seed));
        return (float)((int)(h >> 8) - (1 << 23)) * (1.0f / 8388608.0f);
    }
    return 0.0f;
}

int main(int argc, char **argv) {
    MPI_Init(&argc, &argv);
    int rank, size;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    if (argc < 6) {
        if (!rank) fprintf(stderr, "usage: ref_collectives <op> <count> <kind> <seed> <prefix>\n");
        MPI_Finalize();
        return 2;
    }
    const std::string op = argv[1];
    const uint64_t count = strtoull(argv[2], nullptr, 10);
    const int kind = atoi(argv[3]);
    const uint64_t seed = strtoull(argv[4], nullptr, 10);
    const std::string prefix = argv[5];

    ProxyCommunicator *comm = new MPICommunicator(MPI_COMM_WORLD, MPI_FLOAT, 4);
    uint64_t send_n = count, recv_n = count;
    if (op == "allgather" || op == "iallgather") recv_n = count * size;
    if (op == "reduce_scatter") send_n = count * size;
    if (op == "alltoall") { send_n = count * size; recv_n = count * size; }
    Tensor<float, Device::CPU> send(send_n ? send_n : 1), recv(recv_n ? recv_n : 1);
    for (uint64_t i = 0; i < send_n; ++i) send.data[i] = pattern_value(kind, seed, rank, i);

    if (op == "allreduce") comm->Allreduce(send.data, recv.data, (int)count);
    else if (op == "iallreduce") { comm->Iallreduce(send.data, recv.data, (int)count, 2); comm->WaitAll(4); }
    else if (op == "allgather") comm->Allgather(send.data, (int)count, recv.data, (int)count);
    else if (op == "iallgather") { comm->Iallgather(send.data, (int)count, recv.data, (int)count, 1); comm->Wait(1); }
    else if (op == "reduce_scatter") comm->Reduce_Scatter_block(send.data, recv.data, (int)count);
    else if (op == "alltoall") comm->Alltoall(send.data, (int)count, recv.data, (int)count);
    else { if (!rank) fprintf(stderr, "unknown op %s\n", op.c_str()); MPI_Finalize(); return 2; }

    const std::string path = prefix + ".r" + std::to_string(rank) + ".bin";
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { perror(path.c_str()); return 1; }
    fwrite(recv.data, sizeof(float), recv_n, f);
    fclose(f);
    MPI_Finalize();
    return 0;
}
