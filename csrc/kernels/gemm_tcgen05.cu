// Persistent warp-specialised bf16 GEMM for sm_100a.
//
//   TMA (cp.async.bulk.tensor, 128B swizzle)  ->  shared-memory ring (4-6 stages)
//   tcgen05.mma (M=128, N=128|256, K=16, one elected thread)  ->  fp32 accumulators in TMEM
//   tcgen05.ld  ->  registers  ->  fused epilogue  ->  swizzled smem  ->  TMA store
//
// One CTA per SM, 8 warps: warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM
// allocator, warp 3 idle, warps 4-7 = epilogue (one TMEM lane quarter each).  The accumulator
// is double buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps the MMAs of
// tile i+1.  Both operands may be K-major or MN-major (the transposes needed by dgrad/wgrad),
// selected through the instruction descriptor and the shape of the TMA boxes, so no transpose
// kernel ever runs.  Partial tiles rely on TMA zero-fill (loads) and clipping (stores).
//
// Covers the matmuls of the reference model: models.py:47 (column parallel), models.py:81
// (row parallel) plus the GELU (models.py:182) and residual (models.py:173,188) epilogues.
#include "gemm_core.cuh"

namespace dlbb {

// ------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------
static int g_force_bn = 0;
static int g_num_sms = 0;

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                       const CUtensorMap& tmAux, const GemmParams& p, int grid, cudaStream_t stream) {
    using Cfg = GemmCfg<BN>;
    auto kern = gemm_bf16_tcgen05_kernel<BN, A_MN, B_MN>;
    static bool configured = false;
    if (!configured) {
        DLBB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes));
        configured = true;
    }
    kern<<<grid, kNumThreads, Cfg::kSmemBytes, stream>>>(tmA, tmB, tmC, tmAux, p);
    DLBB_CHECK_LAUNCH("gemm_bf16_tcgen05_kernel");
    return 0;
}

template <int BN>
static int dispatch_gemm(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda,
                         int64_t ldb, int64_t ldc, int a_mn, int b_mn, int epi, void* aux,
                         int64_t ld_aux, cudaStream_t stream) {
    CUtensorMap tmA, tmB, tmC, tmAux;
    int rc;
    if (!a_mn) rc = make_map_2d(&tmA, A, K, M, lda, kBK, kBM, CU_TENSOR_MAP_SWIZZLE_128B);
    else       rc = make_map_2d(&tmA, A, M, K, lda, 64, kBK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (!b_mn) rc = make_map_2d(&tmB, B, K, N, ldb, kBK, BN, CU_TENSOR_MAP_SWIZZLE_128B);
    else       rc = make_map_2d(&tmB, B, N, K, ldb, 64, kBK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    rc = make_map_2d(&tmC, C, N, M, ldc, kEpiCols, kBM, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
    if (aux != nullptr) {
        rc = make_map_2d(&tmAux, aux, N, M, ld_aux, kEpiCols, kBM, CU_TENSOR_MAP_SWIZZLE_64B);
        if (rc) return rc;
    } else {
        tmAux = tmC;
    }
    if (g_num_sms == 0) {
        int dev = 0;
        DLBB_CHECK_CUDA(cudaGetDevice(&dev));
        DLBB_CHECK_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.epi = epi;
    p.aux_out = (epi == DLBB_EPI_GELU && aux != nullptr) ? 1 : 0;
    const int num_m = (M + kBM - 1) / kBM, num_n = (N + BN - 1) / BN;
    // Rasterisation: keep the smaller operand L2-resident and stream the other exactly once.
    const double a_bytes = 2.0 * M * K, b_bytes = 2.0 * N * K;
    if (b_bytes <= a_bytes) {
        p.group_m = std::max(1, std::min(num_m, (g_num_sms + num_n - 1) / num_n));
        p.a_policy = kL2EvictFirst;   // A is streamed once, B is re-read by every wave
        p.b_policy = kL2EvictLast;
    } else {
        p.group_m = (a_bytes <= 64e6) ? num_m : 8;
        p.a_policy = kL2EvictLast;
        p.b_policy = kL2EvictFirst;
    }
    // Measured: evict-first on the streamed operand costs ~5% (its lines are shared by the CTAs of one wave
    // and get dropped between their reads), so the hints are opt-in.
    if (!getenv("DLBB_GEMM_L2_HINTS")) p.a_policy = p.b_policy = kL2EvictNormal;
    const int grid = std::min(num_m * num_n, g_num_sms);
    if (!a_mn && !b_mn) return launch_gemm<BN, false, false>(tmA, tmB, tmC, tmAux, p, grid, stream);
    if (!a_mn && b_mn)  return launch_gemm<BN, false, true>(tmA, tmB, tmC, tmAux, p, grid, stream);
    if (a_mn && !b_mn)  return launch_gemm<BN, true, false>(tmA, tmB, tmC, tmAux, p, grid, stream);
    return launch_gemm<BN, true, true>(tmA, tmB, tmC, tmAux, p, grid, stream);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace dlbb

extern "C" {

void dlbb_gemm_set_tile_n(int bn) { dlbb::g_force_bn = (bn == 128 || bn == 256) ? bn : 0; }

int dlbb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda,
                   int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major, int epilogue,
                   void* aux, int64_t ld_aux, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
    DLBB_REQUIRE(epilogue >= DLBB_EPI_NONE && epilogue <= DLBB_EPI_MUL_DGELU, "gemm: bad epilogue %d",
                 epilogue);
    DLBB_REQUIRE(!((epilogue == DLBB_EPI_ADD || epilogue == DLBB_EPI_MUL_DGELU) && aux == nullptr),
                 "gemm: epilogue %d needs aux", epilogue);
    const bool tma_ok = aligned16(A) && aligned16(B) && aligned16(C) && aligned16(aux) &&
                        lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0 &&
                        (aux == nullptr || ld_aux % 8 == 0);
    if (!tma_ok)
        return dlbb_gemm_bf16_simt(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue,
                                   aux, ld_aux, stream);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int bn = g_force_bn;
    if (bn == 0) {
        // Measured on B200 (profiles/gemm_shapes.md): the 128-wide tile is shared-memory-bandwidth
        // bound (A and B are re-read at 128 B/clk/SM) and loses ~30% to the 256-wide tile on every
        // model shape, even where 256 wastes part of the last wave.  Use 128 only for narrow outputs.
        bn = (N <= 128) ? 128 : 256;
    }
    if (bn == 256)
        return dispatch_gemm<256>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue,
                                  aux, ld_aux, s);
    return dispatch_gemm<128>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue, aux,
                              ld_aux, s);
}
}
