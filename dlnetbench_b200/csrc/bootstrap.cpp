// bootstrap.cpp — see bootstrap.hpp
#include "bootstrap.hpp"

#include <errno.h>
#include <poll.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

namespace dlnb {

static std::string errstr(const char *what) {
    return std::string(what) + ": " + strerror(errno);
}

static bool send_all(int fd, const void *buf, size_t n) {
    const char *p = (const char *)buf;
    while (n) {
        ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
        if (k < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        p += k;
        n -= (size_t)k;
    }
    return true;
}

static bool recv_all(int fd, void *buf, size_t n) {
    char *p = (char *)buf;
    while (n) {
        ssize_t k = ::recv(fd, p, n, 0);
        if (k < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        if (k == 0) return false; // peer died
        p += k;
        n -= (size_t)k;
    }
    return true;
}

static socklen_t make_addr(sockaddr_un &a, const std::string &session) {
    memset(&a, 0, sizeof(a));
    a.sun_family = AF_UNIX;
    // abstract namespace: leading NUL, no file to clean up
    std::string name = "dlnb." + session;
    if (name.size() > sizeof(a.sun_path) - 2) name.resize(sizeof(a.sun_path) - 2);
    memcpy(a.sun_path + 1, name.data(), name.size());
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
}

Bootstrap::~Bootstrap() { close_all(); }

void Bootstrap::close_all() {
    for (int fd : peers_)
        if (fd >= 0) ::close(fd);
    peers_.clear();
    if (listen_fd_ >= 0) ::close(listen_fd_);
    listen_fd_ = -1;
    world_ = 1;
}

std::string Bootstrap::connect(int rank, int world, const std::string &session, int timeout_s) {
    rank_ = rank;
    world_ = world;
    if (world <= 1) return "";
    sockaddr_un addr;
    socklen_t alen = make_addr(addr, session);
    if (rank == 0) {
        listen_fd_ = ::socket(AF_UNIX, SOCK_STREAM, 0);
        if (listen_fd_ < 0) return errstr("socket");
        if (::bind(listen_fd_, (sockaddr *)&addr, alen) < 0) return errstr("bind (session name in use?)");
        if (::listen(listen_fd_, world) < 0) return errstr("listen");
        peers_.assign(world, -1);
        for (int k = 1; k < world; ++k) {
            pollfd pf{listen_fd_, POLLIN, 0};
            int pr;
            do { // a signal (profilers deliver them) interrupts poll: that is not a timeout
                pr = ::poll(&pf, 1, timeout_s * 1000);
            } while (pr < 0 && errno == EINTR);
            if (pr <= 0) return "rendezvous timeout: not all ranks connected to the hub";
            int fd;
            do {
                fd = ::accept(listen_fd_, nullptr, nullptr);
            } while (fd < 0 && errno == EINTR);
            if (fd < 0) return errstr("accept");
            int r = -1;
            if (!recv_all(fd, &r, sizeof(r)) || r <= 0 || r >= world || peers_[r] != -1) {
                ::close(fd);
                return "rendezvous: bad hello from a peer";
            }
            peers_[r] = fd;
        }
    } else {
        int fd = -1;
        timespec t0;
        clock_gettime(CLOCK_MONOTONIC, &t0);
        for (;;) {
            fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
            if (fd < 0) return errstr("socket");
            if (::connect(fd, (sockaddr *)&addr, alen) == 0) break;
            ::close(fd);
            timespec t1;
            clock_gettime(CLOCK_MONOTONIC, &t1);
            if (t1.tv_sec - t0.tv_sec > timeout_s) return "rendezvous timeout: hub (rank 0) not reachable";
            usleep(20000);
        }
        if (!send_all(fd, &rank, sizeof(rank))) {
            ::close(fd);
            return errstr("hello");
        }
        peers_.assign(1, fd);
    }
    return barrier();
}

std::string Bootstrap::allgather(const void *in, size_t bytes, void *out) {
    if (world_ <= 1) {
        if (out != in) memcpy(out, in, bytes);
        return "";
    }
    if (rank_ == 0) {
        char *o = (char *)out;
        memcpy(o, in, bytes);
        for (int r = 1; r < world_; ++r)
            if (!recv_all(peers_[r], o + (size_t)r * bytes, bytes)) return "allgather: lost a peer";
        for (int r = 1; r < world_; ++r)
            if (!send_all(peers_[r], o, bytes * (size_t)world_)) return "allgather: lost a peer";
    } else {
        if (!send_all(peers_[0], in, bytes)) return "allgather: lost the hub";
        if (!recv_all(peers_[0], out, bytes * (size_t)world_)) return "allgather: lost the hub";
    }
    return "";
}

std::string Bootstrap::barrier() {
    std::vector<char> tmp((size_t)world_);
    char c = 1;
    return allgather(&c, 1, tmp.data());
}

static bool send_fds(int sock, const int *fds, int n) {
    char dummy = 'F';
    iovec iov{&dummy, 1};
    std::vector<char> ctrl(CMSG_SPACE(sizeof(int) * n));
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl.data();
    msg.msg_controllen = ctrl.size();
    cmsghdr *cm = CMSG_FIRSTHDR(&msg);
    cm->cmsg_level = SOL_SOCKET;
    cm->cmsg_type = SCM_RIGHTS;
    cm->cmsg_len = CMSG_LEN(sizeof(int) * n);
    memcpy(CMSG_DATA(cm), fds, sizeof(int) * n);
    for (;;) {
        ssize_t k = ::sendmsg(sock, &msg, MSG_NOSIGNAL);
        if (k < 0 && errno == EINTR) continue;
        return k == 1;
    }
}

static bool recv_fds(int sock, int *fds, int n) {
    char dummy = 0;
    iovec iov{&dummy, 1};
    std::vector<char> ctrl(CMSG_SPACE(sizeof(int) * n));
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl.data();
    msg.msg_controllen = ctrl.size();
    for (;;) {
        ssize_t k = ::recvmsg(sock, &msg, 0);
        if (k < 0 && errno == EINTR) continue;
        if (k != 1) return false;
        break;
    }
    cmsghdr *cm = CMSG_FIRSTHDR(&msg);
    if (!cm || cm->cmsg_level != SOL_SOCKET || cm->cmsg_type != SCM_RIGHTS) return false;
    if (cm->cmsg_len != CMSG_LEN(sizeof(int) * n)) return false;
    memcpy(fds, CMSG_DATA(cm), sizeof(int) * n);
    return true;
}

std::string Bootstrap::exchange_fds(int myfd, std::vector<int> &out) {
    out.assign((size_t)world_, -1);
    if (world_ <= 1) {
        out[0] = ::dup(myfd);
        return "";
    }
    if (rank_ == 0) {
        out[0] = ::dup(myfd);
        for (int r = 1; r < world_; ++r)
            if (!recv_fds(peers_[r], &out[r], 1)) return "fd exchange: receive from a peer failed";
        for (int r = 1; r < world_; ++r)
            if (!send_fds(peers_[r], out.data(), world_)) return "fd exchange: send to a peer failed";
    } else {
        if (!send_fds(peers_[0], &myfd, 1)) return "fd exchange: send to the hub failed";
        if (!recv_fds(peers_[0], out.data(), world_)) return "fd exchange: receive from the hub failed";
    }
    return "";
}

std::string Bootstrap::bcast_fd(int root_fd, int &out) {
    out = -1;
    if (world_ <= 1 || rank_ == 0) {
        out = ::dup(root_fd);
        if (world_ > 1)
            for (int r = 1; r < world_; ++r)
                if (!send_fds(peers_[r], &root_fd, 1)) return "fd broadcast: send to a peer failed";
        return "";
    }
    if (!recv_fds(peers_[0], &out, 1)) return "fd broadcast: receive from the hub failed";
    return "";
}

} // namespace dlnb
