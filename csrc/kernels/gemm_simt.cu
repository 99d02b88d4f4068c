// SIMT fallback GEMM: same contract as dlbb_gemm_bf16 for shapes TMA cannot address (row
// pitches that are not multiples of 16 bytes, e.g. the 10->20->5 toy layers of the reference's
// DeepSpeed smoke tests, test/ccl.py:65-69).  32x32 output tile per CTA, fp32 accumulation,
// operands staged through padded shared memory.  Also used by the tests as an independent
// implementation to cross-check the tcgen05 kernel.
#include "../../include/dlbb.h"
#include "common.cuh"

namespace dlbb {

constexpr int kT = 32;

__global__ void __launch_bounds__(kT* kT)
gemm_bf16_simt_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B,
                      __nv_bfloat16* __restrict__ C, int M, int N, int K, int64_t lda, int64_t ldb,
                      int64_t ldc, int a_mn, int b_mn, int epi, __nv_bfloat16* aux, int64_t ld_aux) {
    __shared__ float sA[kT][kT + 1];  // [m][k]
    __shared__ float sB[kT][kT + 1];  // [k][n]
    const int tx = threadIdx.x % kT, ty = threadIdx.x / kT;
    const int m0 = blockIdx.y * kT, n0 = blockIdx.x * kT;
    float acc = 0.f;
    for (int k0 = 0; k0 < K; k0 += kT) {
        // Each operand is loaded with its contiguous dimension along threadIdx.x.
        {
            int m, k;
            if (a_mn) { m = m0 + tx; k = k0 + ty; } else { m = m0 + ty; k = k0 + tx; }
            float v = 0.f;
            if (m < M && k < K) v = __bfloat162float(a_mn ? A[(int64_t)k * lda + m] : A[(int64_t)m * lda + k]);
            sA[m - m0][k - k0] = v;
        }
        {
            int n, k;
            if (b_mn) { n = n0 + tx; k = k0 + ty; } else { n = n0 + ty; k = k0 + tx; }
            float v = 0.f;
            if (n < N && k < K) v = __bfloat162float(b_mn ? B[(int64_t)k * ldb + n] : B[(int64_t)n * ldb + k]);
            sB[k - k0][n - n0] = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kT; ++k) acc = fmaf(sA[ty][k], sB[k][tx], acc);
        __syncthreads();
    }
    const int m = m0 + ty, n = n0 + tx;
    if (m >= M || n >= N) return;
    float out = acc;
    if (epi == DLBB_EPI_GELU) {
        if (aux) aux[(int64_t)m * ld_aux + n] = __float2bfloat16(acc);
        out = gelu_exact(acc);
    } else if (epi == DLBB_EPI_ADD) {
        out = acc + __bfloat162float(aux[(int64_t)m * ld_aux + n]);
    } else if (epi == DLBB_EPI_MUL_DGELU) {
        out = acc * dgelu_exact(__bfloat162float(aux[(int64_t)m * ld_aux + n]));
    }
    C[(int64_t)m * ldc + n] = __float2bfloat16(out);
}

}  // namespace dlbb

extern "C" int dlbb_gemm_bf16_simt(const void* A, const void* B, void* C, int M, int N, int K,
                                   int64_t lda, int64_t ldb, int64_t ldc, int a_mn_major,
                                   int b_mn_major, int epilogue, void* aux, int64_t ld_aux,
                                   void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_simt: empty problem %dx%dx%d", M, N, K);
    DLBB_REQUIRE(!((epilogue == DLBB_EPI_ADD || epilogue == DLBB_EPI_MUL_DGELU) && aux == nullptr),
                 "gemm_simt: epilogue %d needs aux", epilogue);
    dim3 grid((N + kT - 1) / kT, (M + kT - 1) / kT);
    gemm_bf16_simt_kernel<<<grid, kT * kT, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(A), static_cast<const __nv_bfloat16*>(B),
        static_cast<__nv_bfloat16*>(C), M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue,
        static_cast<__nv_bfloat16*>(aux), ld_aux);
    DLBB_CHECK_LAUNCH("gemm_bf16_simt_kernel");
    return 0;
}
