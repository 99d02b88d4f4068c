// Tile geometry shared by the GEMM kernels (gemm_core.cuh), the fused GEMM + all-reduce functor
// (csrc/collectives/gemm_allreduce_fusion.cuh) and its host emulation: plain arithmetic, host and device.
#pragma once

#if defined(__CUDACC__)
#define DLBB_TILES_HD __host__ __device__ __forceinline__
#else
#define DLBB_TILES_HD inline
#endif

namespace dlbb {

constexpr int kBM = 128;         // rows of an output tile (one CTA)
constexpr int kNumThreads = 256; // threads of the GEMM roles (TMA, MMA, epilogue warps)

// Tile t of the persistent schedule -> (m block, n block): groups of `group_m` m-blocks are walked column by column so
// that the CTAs working at the same time share operand panels in L2.
DLBB_TILES_HD void tile_coords(int t, int num_m, int num_n, int group_m, int& m_blk, int& n_blk) {
    const int per_group = group_m * num_n;
    const int g = t / per_group;
    const int first_m = g * group_m;
    const int gsize = group_m < num_m - first_m ? group_m : num_m - first_m;
    const int r = t - g * per_group;
    m_blk = first_m + r % gsize;
    n_blk = r / gsize;
}

}  // namespace dlbb
