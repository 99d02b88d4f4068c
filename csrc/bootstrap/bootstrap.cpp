#include "bootstrap.h"

#include <errno.h>
#include <poll.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

namespace dlbb {

static double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}

// Abstract-namespace address (leading NUL): no file to clean up, vanishes with the process.
static socklen_t make_addr(const std::string& job, sockaddr_un* a) {
    memset(a, 0, sizeof(*a));
    a->sun_family = AF_UNIX;
    std::string name = "dlbb-" + job;
    if (name.size() > sizeof(a->sun_path) - 2) name.resize(sizeof(a->sun_path) - 2);
    memcpy(a->sun_path + 1, name.data(), name.size());
    return static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + name.size());
}

Bootstrap::Bootstrap(int rank, int world, const std::string& job) : rank_(rank), world_(world), job_(job) {}

Bootstrap::~Bootstrap() {
    if (listen_fd_ >= 0) close(listen_fd_);
    if (root_fd_ >= 0) close(root_fd_);
    for (int fd : peer_fd_)
        if (fd >= 0) close(fd);
}

bool Bootstrap::send_all(int fd, const void* p, size_t n) {
    const char* c = static_cast<const char*>(p);
    while (n) {
        ssize_t k = ::send(fd, c, n, MSG_NOSIGNAL);
        if (k < 0) {
            if (errno == EINTR) continue;
            err_ = std::string("bootstrap send: ") + strerror(errno);
            return false;
        }
        c += k;
        n -= static_cast<size_t>(k);
    }
    return true;
}

bool Bootstrap::recv_all(int fd, void* p, size_t n) {
    char* c = static_cast<char*>(p);
    while (n) {
        ssize_t k = ::recv(fd, c, n, 0);
        if (k < 0) {
            if (errno == EINTR) continue;
            err_ = std::string("bootstrap recv: ") + strerror(errno);
            return false;
        }
        if (k == 0) {
            err_ = "bootstrap recv: peer closed the connection (a rank died?)";
            return false;
        }
        c += k;
        n -= static_cast<size_t>(k);
    }
    return true;
}

bool Bootstrap::send_fds(int sock, const int* fds, int n) {
    char payload = 'F';
    iovec iov{&payload, 1};
    std::vector<char> ctrl(CMSG_SPACE(sizeof(int) * n), 0);
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl.data();
    msg.msg_controllen = ctrl.size();
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    cm->cmsg_level = SOL_SOCKET;
    cm->cmsg_type = SCM_RIGHTS;
    cm->cmsg_len = CMSG_LEN(sizeof(int) * n);
    memcpy(CMSG_DATA(cm), fds, sizeof(int) * n);
    for (;;) {
        ssize_t k = ::sendmsg(sock, &msg, MSG_NOSIGNAL);
        if (k == 1) return true;
        if (k < 0 && errno == EINTR) continue;
        err_ = std::string("bootstrap sendmsg(SCM_RIGHTS): ") + strerror(errno);
        return false;
    }
}

bool Bootstrap::recv_fds(int sock, int* fds, int n) {
    char payload = 0;
    iovec iov{&payload, 1};
    std::vector<char> ctrl(CMSG_SPACE(sizeof(int) * n), 0);
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl.data();
    msg.msg_controllen = ctrl.size();
    for (;;) {
        ssize_t k = ::recvmsg(sock, &msg, 0);
        if (k < 0 && errno == EINTR) continue;
        if (k != 1) {
            err_ = std::string("bootstrap recvmsg(SCM_RIGHTS): ") + (k == 0 ? "peer closed" : strerror(errno));
            return false;
        }
        break;
    }
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    if (!cm || cm->cmsg_level != SOL_SOCKET || cm->cmsg_type != SCM_RIGHTS ||
        cm->cmsg_len != CMSG_LEN(sizeof(int) * n)) {
        err_ = "bootstrap recvmsg: expected file descriptors were not delivered";
        return false;
    }
    memcpy(fds, CMSG_DATA(cm), sizeof(int) * n);
    return true;
}

bool Bootstrap::connect(double timeout_s) {
    if (world_ == 1) return true;
    sockaddr_un addr;
    const socklen_t alen = make_addr(job_, &addr);
    const double deadline = now_s() + timeout_s;
    if (rank_ == 0) {
        listen_fd_ = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (listen_fd_ < 0 || ::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), alen) < 0 ||
            ::listen(listen_fd_, world_ + 8) < 0) {
            err_ = std::string("bootstrap: cannot listen on dlbb-") + job_ + ": " + strerror(errno);
            return false;
        }
        peer_fd_.assign(world_, -1);
        for (int got = 1; got < world_;) {
            pollfd pfd{listen_fd_, POLLIN, 0};
            const double left = deadline - now_s();
            if (left <= 0 || ::poll(&pfd, 1, static_cast<int>(left * 1000)) <= 0) {
                err_ = "bootstrap: timed out waiting for " + std::to_string(world_ - got) + " rank(s) to connect";
                return false;
            }
            int fd = ::accept4(listen_fd_, nullptr, nullptr, SOCK_CLOEXEC);
            if (fd < 0) continue;
            int r = -1;
            if (!recv_all(fd, &r, sizeof(r)) || r <= 0 || r >= world_ || peer_fd_[r] >= 0) {
                close(fd);
                err_ = "bootstrap: bad hello from a connecting rank";
                return false;
            }
            peer_fd_[r] = fd;
            ++got;
        }
        return true;
    }
    // non-root: retry until rank 0 is listening
    for (;;) {
        int fd = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (fd < 0) {
            err_ = std::string("bootstrap socket: ") + strerror(errno);
            return false;
        }
        if (::connect(fd, reinterpret_cast<sockaddr*>(&addr), alen) == 0) {
            root_fd_ = fd;
            return send_all(fd, &rank_, sizeof(rank_));
        }
        close(fd);
        if (now_s() > deadline) {
            err_ = "bootstrap: timed out connecting to rank 0 (dlbb-" + job_ + ")";
            return false;
        }
        usleep(20000);
    }
}

bool Bootstrap::allgather(const void* in, void* out, size_t bytes) {
    char* o = static_cast<char*>(out);
    if (world_ == 1) {
        memcpy(o, in, bytes);
        return true;
    }
    if (rank_ == 0) {
        memcpy(o, in, bytes);
        for (int r = 1; r < world_; ++r)
            if (!recv_all(peer_fd_[r], o + r * bytes, bytes)) return false;
        for (int r = 1; r < world_; ++r)
            if (!send_all(peer_fd_[r], o, bytes * world_)) return false;
        return true;
    }
    return send_all(root_fd_, in, bytes) && recv_all(root_fd_, o, bytes * world_);
}

bool Bootstrap::allgather_fds(int fd, std::vector<int>& out) {
    out.assign(world_, -1);
    if (world_ == 1) {
        out[0] = dup(fd);
        return out[0] >= 0;
    }
    if (rank_ == 0) {
        out[0] = dup(fd);
        for (int r = 1; r < world_; ++r)
            if (!recv_fds(peer_fd_[r], &out[r], 1)) return false;
        for (int r = 1; r < world_; ++r)
            if (!send_fds(peer_fd_[r], out.data(), world_)) return false;
        return true;
    }
    return send_fds(root_fd_, &fd, 1) && recv_fds(root_fd_, out.data(), world_);
}

bool Bootstrap::barrier() {
    char c = 1;
    std::vector<char> all(world_);
    return allgather(&c, all.data(), 1);
}

}  // namespace dlbb
