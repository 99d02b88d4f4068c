// Host-side all-to-all exchange of one small record per rank through shared memory.
//
// Every rank owns one memfd-backed region with a few slots; all regions are mapped by everyone at
// creation (the descriptors travel over the bootstrap sockets).  Exchange number n: a rank writes its
// record into slot n % kSlots of its own region, publishes n with a release store and spins (acquire)
// until every peer's slot n % kSlots carries n.  A rank cannot finish exchange n+1 before every peer
// has published n+1, which a peer only does after it has read everything of exchange n, so a slot is
// never overwritten while somebody still reads it (kSlots = 2 would do; 4 is slack).
//
// Used by peer_registry.cpp to hand CUDA IPC handles around once per large collective (about a
// microsecond when the peers are already there, against tens of microseconds for a socket round trip).
// CUDA-free on purpose: tests/test_mailbox_cpu.py exercises it with forked processes on any host.
#pragma once
#include <sched.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <vector>

#include "bootstrap.h"

namespace dlbb {

class HostMailbox {
  public:
    static constexpr int kSlots = 4;
    static constexpr size_t kPayload = 112;
    static constexpr int kMaxRanks = 16;

    ~HostMailbox() {
        for (int p = 0; p < world_; ++p)
            if (regions_[p]) munmap(regions_[p], sizeof(Region));
    }

    // Collective over the bootstrap.  Returns false (on every rank alike) if any rank failed.
    bool create(Bootstrap& bs) {
        rank_ = bs.rank(); world_ = bs.world();
        if (world_ > kMaxRanks) return false;
        int fd = memfd_create("dlbb-mailbox", 0);
        bool ok = fd >= 0 && ftruncate(fd, sizeof(Region)) == 0;
        std::vector<int> fds;
        if (!bs.allgather_fds(ok ? fd : STDIN_FILENO, fds)) ok = false;   // every rank must take part
        if (ok) {
            for (int p = 0; p < world_; ++p) {
                void* m = mmap(nullptr, sizeof(Region), PROT_READ | PROT_WRITE, MAP_SHARED, fds[p], 0);
                if (m == MAP_FAILED) { ok = false; break; }
                regions_[p] = static_cast<Region*>(m);
            }
        }
        for (int f : fds)
            if (f >= 0) close(f);
        if (fd >= 0) close(fd);
        char mine = ok ? 1 : 0;
        std::vector<char> all(world_);
        if (!bs.allgather(&mine, all.data(), 1)) ok = false;
        for (char v : all) ok = ok && v;
        ready_ = ok;
        return ok;
    }

    bool ready() const { return ready_; }
    uint64_t sequence() const { return seq_; }

    // Publish `mine` (<= kPayload bytes) and wait for everybody's record of the same exchange.
    // out[p] points into peer p's region and stays valid until this rank's next-but-one exchange.
    // On timeout returns false and *missing names the first rank that did not arrive.
    bool exchange(const void* mine, size_t bytes, const void* out[], double timeout_s, int* missing) {
        const uint64_t seq = ++seq_;
        Slot& s = regions_[rank_]->slot[seq % kSlots];
        memcpy(s.payload, mine, bytes < kPayload ? bytes : kPayload);
        s.seq.store(seq, std::memory_order_release);
        const double t0 = now_s();
        for (int p = 0; p < world_; ++p) {
            Slot& q = regions_[p]->slot[seq % kSlots];
            unsigned spins = 0;
            while (q.seq.load(std::memory_order_acquire) != seq) {
                if ((++spins & 0xff) == 0) {
                    sched_yield();
                    if (now_s() - t0 > timeout_s) {
                        if (missing) *missing = p;
                        return false;
                    }
                }
            }
            out[p] = q.payload;
        }
        return true;
    }

  private:
    struct Slot {
        std::atomic<uint64_t> seq;
        unsigned char payload[kPayload];
    };
    struct Region {
        Slot slot[kSlots];
    };
    static double now_s() {
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec + 1e-9 * ts.tv_nsec;
    }
    int rank_ = 0, world_ = 1;
    bool ready_ = false;
    uint64_t seq_ = 0;
    Region* regions_[kMaxRanks] = {};
};

}  // namespace dlbb
