// Persistent warp-specialised bf16 GEMM for sm_100a.
//
//   TMA (cp.async.bulk.tensor, 128B swizzle)  ->  shared-memory ring (4-6 stages)
//   tcgen05.mma (M=128, N=128|256, K=16, one elected thread)  ->  fp32 accumulators in TMEM
//   tcgen05.ld  ->  registers  ->  fused epilogue  ->  swizzled smem  ->  TMA store
//
// Default for outputs larger than one 128-row block: CTA pairs.  A cluster of two CTAs (the two SMs of a
// TPC) owns one 256 x 256 tile; tcgen05.mma.cta_group::2 (M = 256) is issued by the pair's leader, each
// CTA stages its own 128 rows of A and only half of B, holds its 128 accumulator lanes in its own TMEM
// and runs its own epilogue.  Shared-memory traffic per MMA drops by a third, which is what limited the
// one-CTA 128 x 256 kernel; measured 4-14% faster on the 7B model's shapes (profiles/gemm_shapes.md).
// The one-CTA kernel (below) remains for narrow outputs, DLBB_GEMM_PAIR=0 and the fused GEMM+all-reduce.
// The default is the pair kernel built with __maxnreg__(168): a cluster launch of the 198-register build allocates
// 230 registers per thread (ncu launch__registers_per_thread), which keeps the data-parallel bucket kernel
// (128 threads x 112 registers) off the GEMM's SMs; capped, both fit (200 x 256 + 112 x 128 = 65536) and the
// 2-GPU training step drops from 156.7 to 145.1 ms (one-CTA kernel: 160.8 ms; profiles/raw/dp_pairmode_n2.json).
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
static int g_pair_mode = -1;   // -1: from DLBB_GEMM_PAIR (default 2), 0: one CTA per tile, 1: CTA pairs,
                               // 2: CTA pairs capped at 168 registers (leaves room for a co-resident kernel)
static int g_pair_clusters = 0;

template <bool A_MN, bool B_MN, bool LOWREG>
static int launch_gemm_pair(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                            const CUtensorMap& tmAux, const GemmParams& p, int clusters, cudaStream_t stream) {
    using Cfg = GemmCfg<256, 2>;
    auto kern = LOWREG ? gemm_bf16_tcgen05_pair_lowreg_kernel<A_MN, B_MN> : gemm_bf16_tcgen05_pair_kernel<A_MN, B_MN>;
    static bool configured = false;
    if (!configured) {
        DLBB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes));
        configured = true;
    }
    kern<<<2 * clusters, kNumThreads, Cfg::kSmemBytes, stream>>>(tmA, tmB, tmC, tmAux, p);
    DLBB_CHECK_LAUNCH("gemm_bf16_tcgen05_pair_kernel");
    return 0;
}

// 256 x 256 tiles on CTA pairs (tcgen05.mma.cta_group::2): each SM stages half of B, so shared-memory
// traffic per MMA drops by a third against the one-CTA 128 x 256 tile.
template <bool LOWREG>
static int dispatch_gemm_pair(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda,
                              int64_t ldb, int64_t ldc, int a_mn, int b_mn, int epi, void* aux,
                              int64_t ld_aux, cudaStream_t stream) {
    constexpr int BN = 256, kHalfN = 128;
    CUtensorMap tmA, tmB, tmC, tmAux;
    int rc;
    if (!a_mn) rc = make_map_2d(&tmA, A, K, M, lda, kBK, kBM, CU_TENSOR_MAP_SWIZZLE_128B);
    else       rc = make_map_2d(&tmA, A, M, K, lda, 64, kBK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (!b_mn) rc = make_map_2d(&tmB, B, K, N, ldb, kBK, kHalfN, CU_TENSOR_MAP_SWIZZLE_128B);
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
    if (g_pair_clusters == 0) {
        int dev = 0;
        DLBB_CHECK_CUDA(cudaGetDevice(&dev));
        DLBB_CHECK_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
        g_pair_clusters = std::max(1, g_num_sms / 2);
    }
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.epi = epi;
    p.aux_out = (epi == DLBB_EPI_GELU && aux != nullptr) ? 1 : 0;
    const int num_m = (M + 2 * kBM - 1) / (2 * kBM), num_n = (N + BN - 1) / BN;
    const double a_bytes = 2.0 * M * K, b_bytes = 2.0 * N * K;
    if (b_bytes <= a_bytes) p.group_m = std::max(1, std::min(num_m, (g_pair_clusters + num_n - 1) / num_n));
    else p.group_m = (a_bytes <= 64e6) ? num_m : 4;
    p.a_policy = p.b_policy = kL2EvictNormal;
    const int clusters = std::min(num_m * num_n, g_pair_clusters);
    if (!a_mn && !b_mn) return launch_gemm_pair<false, false, LOWREG>(tmA, tmB, tmC, tmAux, p, clusters, stream);
    if (!a_mn && b_mn)  return launch_gemm_pair<false, true, LOWREG>(tmA, tmB, tmC, tmAux, p, clusters, stream);
    if (a_mn && !b_mn)  return launch_gemm_pair<true, false, LOWREG>(tmA, tmB, tmC, tmAux, p, clusters, stream);
    return launch_gemm_pair<true, true, LOWREG>(tmA, tmB, tmC, tmAux, p, clusters, stream);
}

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
void dlbb_gemm_set_pair_mode(int mode) { dlbb::g_pair_mode = (mode >= 0 && mode <= 2) ? mode : -1; }

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
    if (g_pair_mode < 0) {
        const char* e = getenv("DLBB_GEMM_PAIR");
        g_pair_mode = e ? std::max(0, std::min(2, atoi(e))) : 2;
    }
    if (g_pair_mode >= 1 && bn == 0 && M > kBM && N > 128) {
        if (g_pair_mode == 2)
            return dispatch_gemm_pair<true>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue, aux,
                                            ld_aux, s);
        return dispatch_gemm_pair<false>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue, aux,
                                         ld_aux, s);
    }
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
