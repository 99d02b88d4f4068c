// CPU-only exercise of HostMailbox: W forked ranks rendezvous over the bootstrap, then run N exchanges with
// randomly delayed ranks; every record must carry the sender's rank and the current exchange number.
// Usage: mailbox_test <world> <exchanges> <job>     exit code 0 = all ranks agree.
#include <stdio.h>
#include <stdlib.h>
#include <sys/wait.h>

#include "../mailbox.h"

struct Record { int rank; uint64_t iter; uint64_t check; };

static int run_rank(int rank, int world, long iters, const char* job) {
    dlbb::Bootstrap bs(rank, world, job);
    if (!bs.connect(30.0)) { fprintf(stderr, "rank %d: %s\n", rank, bs.error().c_str()); return 2; }
    dlbb::HostMailbox mb;
    if (!mb.create(bs)) { fprintf(stderr, "rank %d: mailbox create failed\n", rank); return 3; }
    unsigned seed = 1234u + rank;
    const void* out[dlbb::HostMailbox::kMaxRanks];
    for (long i = 1; i <= iters; ++i) {
        if (rand_r(&seed) % 97 == 0) usleep(rand_r(&seed) % 300);   // make ranks run ahead of / behind each other
        Record r{rank, static_cast<uint64_t>(i), static_cast<uint64_t>(i) * 2654435761u + rank};
        int missing = -1;
        if (!mb.exchange(&r, sizeof(r), out, 20.0, &missing)) {
            fprintf(stderr, "rank %d: exchange %ld timed out waiting for rank %d\n", rank, i, missing);
            return 4;
        }
        for (int p = 0; p < world; ++p) {
            Record q;
            memcpy(&q, out[p], sizeof(q));
            if (q.rank != p || q.iter != static_cast<uint64_t>(i) || q.check != static_cast<uint64_t>(i) * 2654435761u + p) {
                fprintf(stderr, "rank %d: exchange %ld: bad record from rank %d (rank %d iter %llu)\n", rank, i, p, q.rank,
                        (unsigned long long)q.iter);
                return 5;
            }
        }
    }
    return mb.sequence() == static_cast<uint64_t>(iters) ? 0 : 6;
}

int main(int argc, char** argv) {
    const int world = argc > 1 ? atoi(argv[1]) : 4;
    const long iters = argc > 2 ? atol(argv[2]) : 20000;
    const char* job = argc > 3 ? argv[3] : "mailbox-test";
    std::vector<pid_t> kids;
    for (int r = 0; r < world; ++r) {
        pid_t pid = fork();
        if (pid == 0) _exit(run_rank(r, world, iters, job));
        kids.push_back(pid);
    }
    int bad = 0;
    for (pid_t k : kids) {
        int st = 0;
        waitpid(k, &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad = 1;
    }
    printf(bad ? "MAILBOX FAILED\n" : "MAILBOX OK %d ranks x %ld exchanges\n", world, iters);
    return bad;
}
