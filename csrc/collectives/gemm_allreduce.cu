// K16: tensor-parallel row-parallel linear as ONE kernel - tcgen05 GEMM + all-reduce + residual.
//
//     out = sum over ranks( A_r[M, K/ws] . B_r[K/ws, N] ) + residual          (ref models.py:73-100,173,188)
//
// The reference does matmul -> bf16->fp32 -> host numpy -> MPI_Allreduce -> bf16 -> add.  Here every
// CTA runs the persistent GEMM of gemm_core.cuh and four extra "reducer" warps:
//   * the epilogue TMA-stores each finished partial tile into this rank's symmetric partial buffer and
//     raises a per-tile flag on the tile's owner rank (tile t is owned by rank t mod ws);
//   * the reducer warps of the owner wait until all ws partials of a tile have landed, sum them straight
//     out of the peers' memory - one multimem.ld_reduce per 16 bytes (the NVSwitch adds) or ws peer
//     loads - add the residual and write the bf16 result into every rank's output buffer
//     (multimem.st / peer stores).
// Tiles are reduced while later tiles are still in the tensor cores, so the NVLink traffic
// (2 (ws-1)/ws . M.N bf16 per rank) hides under the MMA time instead of following it.
#include "../../include/dlbb_comm.h"
#include "../kernels/gemm_core.cuh"
#include "coll_device.cuh"
#include "gemm_allreduce_fusion.cuh"

namespace dlbb {

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kArThreads, 1)
gemm_allreduce_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmC, const GemmParams p, ArFusion f) {
    gemm_cta_body<BN, A_MN, B_MN, ArFusion>(tmA, tmB, tmC, tmC, p, f);
}

template <bool A_MN, bool B_MN>
static int launch_ar(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmParams& p,
                     const ArFusion& f, int grid, cudaStream_t s) {
    using Cfg = GemmCfg<256>;
    auto kern = gemm_allreduce_kernel<256, A_MN, B_MN>;
    static bool configured = false;
    if (!configured) {
        DLBB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
        configured = true;
    }
    kern<<<grid, kArThreads, Cfg::kSmemBytes, s>>>(tmA, tmB, tmC, p, f);
    DLBB_CHECK_LAUNCH("gemm_allreduce_kernel");
    return 0;
}

}  // namespace dlbb

extern "C" int dlbb_gemm_allreduce_bf16(dlbb_comm_t* c, const void* A, const void* B, int M, int N, int K,
                                        int64_t lda, int64_t ldb, int a_mn_major, int b_mn_major, int part_seg,
                                        int64_t part_elem_off, int out_seg, int64_t out_elem_off, const void* residual,
                                        int64_t ld_res, int use_multicast, void* stream) {
    using namespace dlbb;
    constexpr int BN = 256;
    DLBB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_allreduce: empty problem");
    DLBB_REQUIRE(N % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && part_elem_off % 8 == 0 && out_elem_off % 8 == 0 &&
                     (residual == nullptr || ld_res % 8 == 0),
                 "gemm_allreduce: N, leading dimensions and offsets must be multiples of 8 elements");
    DLBB_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(residual)) & 15) == 0,
                 "gemm_allreduce: operands must be 16-byte aligned");
    const size_t need = (size_t)M * N * 2;
    DLBB_REQUIRE(dlbb_symm_ptr(c, part_seg, 0) && (size_t)part_elem_off * 2 + need <= dlbb_symm_size(c, part_seg),
                 "gemm_allreduce: partial buffer too small");
    DLBB_REQUIRE(dlbb_symm_ptr(c, out_seg, 0) && (size_t)out_elem_off * 2 + need <= dlbb_symm_size(c, out_seg),
                 "gemm_allreduce: output buffer too small");
    const int num_m = (M + kBM - 1) / kBM, num_n = (N + BN - 1) / BN;
    DLBB_REQUIRE((size_t)num_m * num_n <= kMaxFusedTiles, "gemm_allreduce: too many output tiles (%d)", num_m * num_n);
    int sms = 0, dev = 0;
    DLBB_CHECK_CUDA(cudaGetDevice(&dev));
    DLBB_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    char* part_local = static_cast<char*>(dlbb_symm_ptr(c, part_seg, c->rank)) + part_elem_off * 2;
    CUtensorMap tmA, tmB, tmC;
    int rc;
    if (!a_mn_major) rc = make_map_2d(&tmA, A, K, M, lda, kBK, kBM, CU_TENSOR_MAP_SWIZZLE_128B);
    else             rc = make_map_2d(&tmA, A, M, K, lda, 64, kBK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (!b_mn_major) rc = make_map_2d(&tmB, B, K, N, ldb, kBK, BN, CU_TENSOR_MAP_SWIZZLE_128B);
    else             rc = make_map_2d(&tmB, B, N, K, ldb, 64, kBK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    rc = make_map_2d(&tmC, part_local, N, M, N, kEpiCols, kBM, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
    GemmParams p{};
    p.M = M; p.N = N; p.K = K; p.epi = DLBB_EPI_NONE; p.aux_out = 0;
    p.a_policy = p.b_policy = kL2EvictNormal;
    const double a_bytes = 2.0 * M * K, b_bytes = 2.0 * N * K;
    if (b_bytes <= a_bytes) p.group_m = std::max(1, std::min(num_m, (sms + num_n - 1) / num_n));
    else                    p.group_m = (a_bytes <= 64e6) ? num_m : 8;
    ArFusion f{};
    f.c = c->dev;
    for (int q = 0; q < c->world; ++q) {
        f.part[q] = static_cast<const char*>(dlbb_symm_ptr(c, part_seg, q)) + part_elem_off * 2;
        f.out[q] = static_cast<char*>(dlbb_symm_ptr(c, out_seg, q)) + out_elem_off * 2;
    }
    const char* pmc = static_cast<const char*>(dlbb_symm_mc_ptr(c, part_seg));
    char* omc = static_cast<char*>(dlbb_symm_mc_ptr(c, out_seg));
    const bool nvls = use_multicast && pmc && omc && c->world > 1;
    f.part_mc = nvls ? pmc + part_elem_off * 2 : nullptr;
    f.out_mc = nvls ? omc + out_elem_off * 2 : nullptr;
    f.residual = static_cast<const __nv_bfloat16*>(residual);
    f.ld_res = ld_res;
    f.M = M; f.N = N; f.bn = BN; f.num_m = num_m; f.num_n = num_n; f.num_tiles = num_m * num_n; f.group_m = p.group_m;
    {
        static const uint32_t every = getenv("DLBB_FLAG_REFRESH_EVERY") ? static_cast<uint32_t>(strtoul(getenv("DLBB_FLAG_REFRESH_EVERY"), nullptr, 0))
                                                                        : (1u << 28);
        if (c->world > 1 && c->epoch[kArBank] - c->last_refresh[kArBank] >= every) {
            // rows of the bank first (a barrier over all of them), then the tile flags
            c->last_refresh[kArBank] = c->epoch[kArBank] - every;   // make the bank refresh due now, whatever its own history
            if (const int rg = refresh_flag_bank_if_due(c, kArBank, stream)) return rg;
            const uint32_t ep = c->epoch[kArBank] + 1;
            c->epoch[kArBank] += 2;
            refresh_tile_flags_kernel<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(c->dev, ep);
            DLBB_CHECK_LAUNCH("refresh_tile_flags_kernel");
            c->last_refresh[kArBank] = c->epoch[kArBank];
        }
    }
    f.epoch = ++c->epoch[kArBank];
    // every rank must launch the same grid: CTA b meets CTA b of the peers at the end
    const int grid = std::min(std::min(num_m * num_n, sms), kMaxCtas);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!a_mn_major && !b_mn_major) return launch_ar<false, false>(tmA, tmB, tmC, p, f, grid, s);
    if (!a_mn_major && b_mn_major)  return launch_ar<false, true>(tmA, tmB, tmC, p, f, grid, s);
    if (a_mn_major && !b_mn_major)  return launch_ar<true, false>(tmA, tmB, tmC, p, f, grid, s);
    return launch_ar<true, true>(tmA, tmB, tmC, p, f, grid, s);
}
