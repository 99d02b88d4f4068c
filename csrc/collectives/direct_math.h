// Partition arithmetic of the two-hop kernels (bcast_relay_kernel, reduce_pull_kernel in direct_kernels.cuh), shared with
// a host-side unit test (tests/direct_math_test.cpp): the message is cut into shares (one per relay / reducer), a share
// into rows (one per CTA of a group), a row into chunks (one flag each).  Every rank, every CTA and the host launcher must
// derive the same numbers from (nvec, shares, rows, chunk), and the pieces must tile [0, nvec) exactly once.
#pragma once
#include <stdint.h>

#if !defined(DLBB_HD)
#if defined(__CUDACC__)
#define DLBB_HD __host__ __device__ __forceinline__
#else
#define DLBB_HD inline
#endif
#endif

namespace dlbb {

DLBB_HD int64_t dm_min(int64_t a, int64_t b) { return a < b ? a : b; }

// length of one share (the last one may be shorter or empty)
DLBB_HD int64_t share_len(int64_t nvec, int nshares) { return (nvec + nshares - 1) / nshares; }

// rows [r_lo, r_hi) of share j handled by CTA `row` of a group; s_lo = first vector of the share
DLBB_HD void share_row_range(int64_t nvec, int nshares, int rows, int j, int row, int64_t& s_lo, int64_t& r_lo, int64_t& r_hi) {
    const int64_t share = share_len(nvec, nshares);
    s_lo = dm_min((int64_t)j * share, nvec);
    const int64_t s_hi = dm_min(s_lo + share, nvec);
    const int64_t per = (s_hi - s_lo + rows - 1) / rows;
    r_lo = dm_min(s_lo + (int64_t)row * per, s_hi);
    r_hi = dm_min(r_lo + per, s_hi);
}

// chunks per row - the same for every share and row (shorter rows simply have empty trailing chunks), so that every rank
// advances its flag epochs by the same amount
DLBB_HD int share_nchunks(int64_t nvec, int nshares, int rows, int chunk_vecs) {
    const int64_t share = share_len(nvec, nshares);
    return static_cast<int>((((share + rows - 1) / rows) + chunk_vecs - 1) / chunk_vecs);
}

DLBB_HD void chunk_range(int64_t r_lo, int64_t r_hi, int k, int chunk_vecs, int64_t& c_lo, int64_t& c_hi) {
    c_lo = dm_min(r_lo + (int64_t)k * chunk_vecs, r_hi);
    c_hi = dm_min(c_lo + chunk_vecs, r_hi);
}

// relays of a broadcast / peer reducers of a reduce: the P-1 ranks other than the root, numbered from root+1
DLBB_HD int relay_index(int rank, int root, int P) { return (rank - root - 1 + P) % P; }   // P-1 = the root itself
DLBB_HD int relay_rank(int j, int root, int P) { return (root + 1 + j) % P; }
// the relay whose share group g >= 1 of relay `me` pulls (g = 1 .. P-2 visits every other relay once)
DLBB_HD int other_relay(int me, int group, int P) { return (me + group) % (P - 1); }

}  // namespace dlbb
