// Rendezvous and out-of-band exchange between the ranks of one node.
//
// Star topology over abstract-namespace unix-domain sockets: rank 0 listens on "\0dlbb-<job>",
// every other rank keeps one connection to it.  Unix sockets are used (rather than TCP) because
// CUDA VMM handles travel between processes as POSIX file descriptors, which only SCM_RIGHTS
// ancillary messages can carry.  Replaces what mpirun's PMI / torch's TCPStore do for the
// reference (run_mpi.py:29-43, collectives/1d/dsccl.py:47-57).
#pragma once
#include <stddef.h>

#include <string>
#include <vector>

namespace dlbb {

class Bootstrap {
  public:
    Bootstrap(int rank, int world, const std::string& job);
    ~Bootstrap();
    // Establish the star; returns false (with error()) on timeout or socket failure.
    bool connect(double timeout_s);
    // Every rank contributes `bytes` bytes; everyone receives world*bytes in rank order.
    bool allgather(const void* in, void* out, size_t bytes);
    // Every rank contributes one open file descriptor; everyone receives `world` descriptors
    // (duplicates owned by the receiver, in rank order; the caller closes them).
    bool allgather_fds(int fd, std::vector<int>& out);
    bool barrier();
    const std::string& error() const { return err_; }
    int rank() const { return rank_; }
    int world() const { return world_; }

  private:
    bool send_all(int fd, const void* p, size_t n);
    bool recv_all(int fd, void* p, size_t n);
    bool send_fds(int sock, const int* fds, int n);
    bool recv_fds(int sock, int* fds, int n);
    int rank_, world_;
    std::string job_, err_;
    int listen_fd_ = -1;
    int root_fd_ = -1;            // non-root: connection to rank 0
    std::vector<int> peer_fd_;    // root: connection to each rank (index = rank; [0] unused)
};

}  // namespace dlbb
