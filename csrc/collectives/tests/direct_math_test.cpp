// Host-side check of the two-hop kernels' partition arithmetic (direct_math.h), including world sizes that the 2 / 4 / 8-GPU
// boxes never exercise (3, 5, 6, 7) and messages smaller than the number of CTAs.  The test replays the schedule of
// bcast_relay_kernel and reduce_pull_kernel (direct_kernels.cuh) on index sets: who writes which vector where, which flags
// are signalled and waited for, and checks that
//   * every rank ends up with every vector of the message exactly once (broadcast) / the root reduces every vector once,
//   * a waited flag always has exactly one signaller with the same epoch, and no flag is signalled that nobody waits for
//     in a way that would desynchronise epochs (all ranks advance by the same nchunks),
//   * relay numbering is a bijection for every root.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>
#include <tuple>
#include <vector>

#include "../direct_math.h"

using namespace dlbb;

static int g_fail = 0;
#define CHECK(cond, ...)                                                  \
    do {                                                                  \
        if (!(cond)) {                                                    \
            if (g_fail < 20) { std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } \
            ++g_fail;                                                     \
        }                                                                 \
    } while (0)

// flag identity: (receiving rank, row, slot) -> epoch value written
using FlagKey = std::tuple<int, int, int>;

static void check_numbering(int P) {
    for (int root = 0; root < P; ++root) {
        std::set<int> seen;
        for (int r = 0; r < P; ++r) {
            const int j = relay_index(r, root, P);
            CHECK(j >= 0 && j < P, "relay_index range P=%d root=%d r=%d j=%d", P, root, r, j);
            CHECK((j == P - 1) == (r == root), "root must be index P-1 (P=%d root=%d r=%d j=%d)", P, root, r, j);
            if (r != root) {
                CHECK(relay_rank(j, root, P) == r, "relay_rank inverse P=%d root=%d r=%d", P, root, r);
                seen.insert(j);
            }
        }
        CHECK((int)seen.size() == P - 1, "relay indices not distinct P=%d root=%d", P, root);
        // groups 1..P-2 of relay `me` visit every other relay exactly once
        for (int me = 0; me < P - 1; ++me) {
            std::set<int> others;
            for (int g = 1; g <= P - 2; ++g) {
                const int j = other_relay(me, g, P);
                CHECK(j != me && j >= 0 && j < P - 1, "other_relay P=%d me=%d g=%d j=%d", P, me, g, j);
                others.insert(j);
            }
            CHECK((int)others.size() == P - 2, "other_relay not a bijection P=%d me=%d", P, me);
        }
    }
}

// bcast_relay_kernel (P2P form): grid = (P-1) groups x rows on every non-root rank.
//   group 0 of relay `me`: for each chunk of row-range of share me: copy root -> local, then signal flag (peer q, row, slot me)
//                          on every other relay q.
//   group g >= 1: share j = other_relay(me, g): wait flag (me, row, slot owner-of-j) == base + k + 1, copy owner -> local.
static void check_relay(int P, int64_t nvec, int rows, int chunk) {
    const int nchunks = share_nchunks(nvec, P - 1, rows, chunk);
    for (int root = 0; root < P; root += (P > 4 ? 3 : 1)) {
        std::vector<std::vector<int>> have(P, std::vector<int>(nvec, 0));
        std::map<FlagKey, std::vector<int>> signalled;   // epochs written per flag
        // pass 1: group 0 of every relay
        for (int rank = 0; rank < P; ++rank) {
            if (rank == root) continue;
            const int me = relay_index(rank, root, P);
            for (int row = 0; row < rows; ++row) {
                int64_t s_lo, r_lo, r_hi;
                share_row_range(nvec, P - 1, rows, me, row, s_lo, r_lo, r_hi);
                CHECK(s_lo <= r_lo && r_lo <= r_hi && r_hi <= nvec, "range order");
                for (int k = 0; k < nchunks; ++k) {
                    int64_t c_lo, c_hi;
                    chunk_range(r_lo, r_hi, k, chunk, c_lo, c_hi);
                    CHECK(r_lo <= c_lo && c_lo <= c_hi && c_hi <= r_hi, "chunk order");
                    for (int64_t v = c_lo; v < c_hi; ++v) have[rank][v] += 1;
                    for (int q = 0; q < P; ++q)
                        if (q != root && q != rank) signalled[{q, row, rank}].push_back(k + 1);
                }
                // the chunks must cover the row exactly
                int64_t c_lo, c_hi;
                chunk_range(r_lo, r_hi, nchunks - 1, chunk, c_lo, c_hi);
                CHECK(nchunks == 0 || c_hi == r_hi, "chunks do not reach the row end: P=%d nvec=%lld rows=%d chunk=%d", P, (long long)nvec, rows, chunk);
            }
        }
        // pass 2: groups >= 1
        for (int rank = 0; rank < P; ++rank) {
            if (rank == root) continue;
            const int me = relay_index(rank, root, P);
            for (int g = 1; g <= P - 2; ++g) {
                const int j = other_relay(me, g, P);
                const int owner = relay_rank(j, root, P);
                for (int row = 0; row < rows; ++row) {
                    int64_t s_lo, r_lo, r_hi;
                    share_row_range(nvec, P - 1, rows, j, row, s_lo, r_lo, r_hi);
                    for (int k = 0; k < nchunks; ++k) {
                        int64_t c_lo, c_hi;
                        chunk_range(r_lo, r_hi, k, chunk, c_lo, c_hi);
                        auto it = signalled.find({rank, row, owner});
                        CHECK(it != signalled.end() && (int)it->second.size() == nchunks && it->second[k] == k + 1,
                              "wait without matching signal: P=%d root=%d rank=%d owner=%d row=%d k=%d", P, root, rank, owner, row, k);
                        // the owner must hold the vectors this chunk reads
                        for (int64_t v = c_lo; v < c_hi; ++v) {
                            CHECK(have[owner][v] >= 1, "owner %d lacks vec %lld", owner, (long long)v);
                            have[rank][v] += 1;
                        }
                    }
                }
            }
        }
        for (int rank = 0; rank < P; ++rank) {
            if (rank == root) continue;
            for (int64_t v = 0; v < nvec; ++v)
                CHECK(have[rank][v] == 1, "bcast coverage: P=%d root=%d rank=%d vec=%lld count=%d (nvec=%lld rows=%d chunk=%d)", P, root, rank,
                      (long long)v, have[rank][v], (long long)nvec, rows, chunk);
        }
    }
}

// reduce_pull_kernel.  P2P form: nred = P-1 reducers (every rank but the root), reducer j = relay_index(rank); the root pulls
// share j from rank relay_rank(j) in group j.  MC form: nred = P reducers (reducer j = rank j); the root reduces its own share
// in group 0 and pulls share (root+g)%P in group g.
static void check_pull(int P, int64_t nvec, int rows, int chunk, bool mc) {
    const int nred = mc ? P : P - 1;
    const int nchunks = share_nchunks(nvec, nred, rows, chunk);
    for (int root = 0; root < P; root += (P > 4 ? 3 : 1)) {
        std::vector<int> reduced(nvec, 0), at_root(nvec, 0);
        std::map<FlagKey, int> sig;   // (root, row, from-rank) -> number of chunk signals
        for (int rank = 0; rank < P; ++rank) {
            const int my_j = mc ? rank : relay_index(rank, root, P);
            if (my_j >= nred) continue;                       // P2P: the root reduces nothing itself
            for (int row = 0; row < rows; ++row) {
                int64_t s_lo, r_lo, r_hi;
                share_row_range(nvec, nred, rows, my_j, row, s_lo, r_lo, r_hi);
                for (int k = 0; k < nchunks; ++k) {
                    int64_t c_lo, c_hi;
                    chunk_range(r_lo, r_hi, k, chunk, c_lo, c_hi);
                    for (int64_t v = c_lo; v < c_hi; ++v) {
                        reduced[v] += 1;
                        if (rank == root) at_root[v] += 1;   // MC: the root's own share lands in place
                    }
                    if (rank != root) sig[{root, row, rank}] += 1;
                }
            }
        }
        // the root's pull groups
        std::set<int> pulled_from;
        const int g_lo = mc ? 1 : 0, g_hi = mc ? P - 1 : P - 2;
        for (int g = g_lo; g <= g_hi; ++g) {
            const int j = mc ? (root + g) % P : g;
            const int q = mc ? j : relay_rank(j, root, P);
            CHECK(q != root, "root pulls from itself: P=%d root=%d g=%d", P, root, g);
            pulled_from.insert(q);
            for (int row = 0; row < rows; ++row) {
                int64_t s_lo, r_lo, r_hi;
                share_row_range(nvec, nred, rows, j, row, s_lo, r_lo, r_hi);
                CHECK((sig[{root, row, q}]) == nchunks, "signal count P=%d root=%d q=%d row=%d: %d vs %d", P, root, q, row, sig[{root, row, q}], nchunks);
                for (int k = 0; k < nchunks; ++k) {
                    int64_t c_lo, c_hi;
                    chunk_range(r_lo, r_hi, k, chunk, c_lo, c_hi);
                    for (int64_t v = c_lo; v < c_hi; ++v) at_root[v] += 1;
                }
            }
        }
        CHECK((int)pulled_from.size() == P - 1, "root must pull from every peer once: P=%d root=%d", P, root);
        for (int64_t v = 0; v < nvec; ++v) {
            CHECK(reduced[v] == 1, "reduce coverage P=%d root=%d mc=%d vec=%lld count=%d", P, root, (int)mc, (long long)v, reduced[v]);
            CHECK(at_root[v] == 1, "root coverage P=%d root=%d mc=%d vec=%lld count=%d", P, root, (int)mc, (long long)v, at_root[v]);
        }
    }
}

int main() {
    long cases = 0;
    for (int P = 3; P <= 8; ++P) {
        check_numbering(P);
        const int64_t sizes[] = {1, 2, 5, 7, 63, 64, 255, 1000, 4097, 16384, 65537};
        const int rowss[] = {1, 2, 3, 7, 16, 42, 49};
        const int chunks[] = {1, 8, 64, 8192};
        for (int64_t n : sizes)
            for (int rows : rowss)
                for (int ch : chunks) {
                    if (n > 5000 && ch == 1) continue;   // keep the run short
                    check_relay(P, n, rows, ch);
                    check_pull(P, n, rows, ch, false);
                    check_pull(P, n, rows, ch, true);
                    cases += 3;
                }
    }
    // P = 2: the relay kernels are not used (one relay = plain copy), but the arithmetic must still hold
    check_numbering(2);
    check_pull(2, 1000, 4, 64, false);
    check_pull(2, 1000, 4, 64, true);
    if (g_fail) { std::printf("DIRECT MATH FAILED (%d failures)\n", g_fail); return 1; }
    std::printf("DIRECT MATH OK (%ld schedules replayed, worlds 2-8)\n", cases);
    return 0;
}
