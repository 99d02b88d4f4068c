// The fusion functor of the GEMM + all-reduce kernel (gemm_allreduce.cu): what the epilogue leader does when a partial
// tile has been stored (tile_stored), the reducer warps (extra_role) and the closing barrier (finish).  Kept apart from the
// GEMM body so that the protocol - per-tile flags, tile ownership, the addressing of the partial / output buffers, multimem
// or peer forms - also compiles for the host emulation (emu/, tests/emu_collectives_test.cpp) with a stand-in producer.
// Residency: a reducer waits for the partial tiles of every CTA of every rank, so the whole grid must be resident - the
// launcher takes min(tiles, SM count) CTAs, and the GEMM's shared memory keeps it at one CTA per SM.
#pragma once
#include "../kernels/gemm_tiles.h"
#if !defined(DLBB_HOST_EMU)
#include "../kernels/ptx.cuh"
#endif
#include "coll_device.cuh"

namespace dlbb {

#if defined(DLBB_HOST_EMU)
// the two PTX helpers of kernels/ptx.cuh that the functor uses, and the system-scope fence
inline void named_bar_sync(uint32_t id, uint32_t nthreads) { group_bar_sync(id, nthreads); }
template <int N> inline void tma_wait_group() {}   // the stand-in producer stores with plain stores: nothing in flight
inline void __threadfence_system() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
#endif

constexpr int kArBank = 2;
constexpr int kRedThreads = 256;   // 8 reducer warps: enough 16-byte requests in flight to cover NVLink latency
constexpr int kArThreads = kNumThreads + kRedThreads;
constexpr int kRU = 4;   // 16-byte vectors per reducer thread per batch

struct ArFusion {
    static constexpr int kExtraWarps = kRedThreads / 32;
    CommDev c;
    const char* part[DLBB_MAX_RANKS];
    char* out[DLBB_MAX_RANKS];
    const char* part_mc;
    char* out_mc;
    const __nv_bfloat16* residual;
    int64_t ld_res;
    int M, N, bn, num_m, num_n, num_tiles, group_m;
    uint32_t epoch;

    __device__ __forceinline__ uint32_t* flag(int owner, int tile, int src) const {
        return c.tile_flags[owner] + (size_t)tile * c.world + src;
    }

    // epilogue leader: tile t is complete in my partial buffer -> tell its owner
    __device__ __forceinline__ void tile_stored(int t, int, int) {
        tma_wait_group<0>();        // the bulk stores of this tile have been performed
        __threadfence_system();
        st_release_sys(flag(t % c.world, t, c.rank), epoch);
    }

    __device__ __forceinline__ void extra_role(int, int) {
        const int rt = threadIdx.x - kNumThreads;   // 0..kRedThreads-1
        for (int idx = blockIdx.x;; idx += gridDim.x) {
            const int t = idx * c.world + c.rank;     // my owned tiles, in production order
            if (t >= num_tiles) break;
            int m_blk, n_blk;
            tile_coords(t, num_m, num_n, group_m, m_blk, n_blk);
            const int m0 = m_blk * kBM, n0 = n_blk * bn;
            if (rt < c.world) wait_flag(flag(c.rank, t, rt), epoch, rt, c.watchdog_ns);
            named_bar_sync(2, kRedThreads);
            const int rows = min(kBM, M - m0);
            const int vpr = min(bn, N - n0) >> 3;     // 16-byte vectors per tile row
            const int total = rows * vpr;
            for (int v0 = rt; v0 < total; v0 += kRedThreads * kRU) {
                int64_t eoff[kRU];
                float acc[kRU][8];
#pragma unroll
                for (int u = 0; u < kRU; ++u) {
                    const int v = v0 + u * kRedThreads;
                    const int r = v / vpr, cv = v - r * vpr;
                    eoff[u] = (int64_t)(m0 + r) * N + n0 + cv * 8;
                }
                if (part_mc != nullptr) {
                    uint4 s[kRU];
#pragma unroll
                    for (int u = 0; u < kRU; ++u)
                        if (v0 + u * kRedThreads < total) s[u] = mm_ld_reduce_add_bf16(part_mc + eoff[u] * 2);
#pragma unroll
                    for (int u = 0; u < kRU; ++u) {
                        const uint32_t w4[4] = {s[u].x, s[u].y, s[u].z, s[u].w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpack_bf16x2(w4[k]);
                            acc[u][2 * k] = f.x; acc[u][2 * k + 1] = f.y;
                        }
                    }
                } else {
#pragma unroll
                    for (int u = 0; u < kRU; ++u)
#pragma unroll
                        for (int k = 0; k < 8; ++k) acc[u][k] = 0.f;
                    for (int p = 0; p < c.world; ++p) {   // rank order: identical sums on every owner
                        uint4 buf[kRU];
#pragma unroll
                        for (int u = 0; u < kRU; ++u)
                            if (v0 + u * kRedThreads < total) buf[u] = ld_peer_v4(part[p] + eoff[u] * 2);
#pragma unroll
                        for (int u = 0; u < kRU; ++u) {
                            if (v0 + u * kRedThreads < total) {
                                const uint32_t w4[4] = {buf[u].x, buf[u].y, buf[u].z, buf[u].w};
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const float2 f = unpack_bf16x2(w4[k]);
                                    acc[u][2 * k] += f.x; acc[u][2 * k + 1] += f.y;
                                }
                            }
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < kRU; ++u) {
                    const int v = v0 + u * kRedThreads;
                    if (v >= total) continue;
                    if (residual != nullptr) {
                        const int r = v / vpr, cv = v - r * vpr;
                        const uint4 x = ld_stream_v4(residual + (int64_t)(m0 + r) * ld_res + n0 + cv * 8);
                        const uint32_t w4[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpack_bf16x2(w4[k]);
                            acc[u][2 * k] += f.x; acc[u][2 * k + 1] += f.y;
                        }
                    }
                    const uint4 o = make_uint4(pack_bf16x2(acc[u][0], acc[u][1]), pack_bf16x2(acc[u][2], acc[u][3]),
                                               pack_bf16x2(acc[u][4], acc[u][5]), pack_bf16x2(acc[u][6], acc[u][7]));
                    if (out_mc != nullptr) {
                        mm_st_v4(out_mc + eoff[u] * 2, o);
                    } else {
#pragma unroll 1
                        for (int p = 0; p < c.world; ++p) {
                            const int q = (c.rank + p) % c.world;
                            st_v4(out[q] + eoff[u] * 2, o);
                        }
                    }
                }
            }
        }
    }

    // every thread: all of this CTA's result stores precede the barrier; the kernel ends only when every
    // CTA has met its peers, i.e. when every rank's output buffer is complete
    __device__ __forceinline__ void finish() { row_barrier(c, kArBank, blockIdx.x, epoch); }
};

// Per-tile flags that no call has used for a long time (a larger GEMM after many smaller ones) must stay within 2^31 of
// the epoch counter (comm_internal.h: refresh_flag_bank_if_due): between two row barriers every rank rewrites its own
// flag array with the current epoch.
__global__ void refresh_tile_flags_kernel(CommDev c, uint32_t epoch) {
    row_barrier(c, kArBank, blockIdx.x, epoch);
    const size_t n = kMaxFusedTiles * (size_t)c.world;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        st_release_sys(c.tile_flags[c.rank] + i, epoch + 1);
    row_barrier(c, kArBank, blockIdx.x, epoch + 1);
}

}  // namespace dlbb
