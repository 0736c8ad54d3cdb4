// bootstrap.hpp — host-side rendezvous of the `world` single-GPU processes of one box.
//
// Replaces what the reference gets from MPI outside the hot path: the launcher's
// process group, MPI_Barrier, MPI_Allgather of small blobs and the MPI_Bcast of the
// ncclUniqueId (cpp/data_parallel/dp.cpp:166-189). There is no MPI on the target
// image, so rank 0 is a hub on an abstract unix socket; POSIX file descriptors of
// cuMem allocations travel over the same sockets with SCM_RIGHTS.
#pragma once
#include <stddef.h>
#include <string>
#include <vector>

namespace dlnb {

class Bootstrap {
  public:
    Bootstrap() = default;
    ~Bootstrap();
    // returns empty string on success, else an error message
    std::string connect(int rank, int world, const std::string &session, int timeout_s);
    void close_all();
    bool active() const { return world_ > 1; }
    // out holds world*bytes, in rank order
    std::string allgather(const void *in, size_t bytes, void *out);
    std::string barrier();
    // every rank contributes one fd; every rank receives all `world` fds (its own
    // entry is a duplicate of what it sent). Caller closes them.
    std::string exchange_fds(int myfd, std::vector<int> &out);
    // rank 0's fd reaches every rank (returned in `out`; rank 0 gets a duplicate). Caller closes it.
    std::string bcast_fd(int root_fd, int &out);

  private:
    int rank_ = 0, world_ = 1;
    int listen_fd_ = -1;
    std::vector<int> peers_; // hub: fd per rank (self = -1); leaf: peers_[0] = hub fd
};

} // namespace dlnb
