/* dlbb — B200-native kernels and collectives, C ABI.
 *
 * Every entry point takes raw device pointers and a CUDA stream (as void*), launches
 * asynchronously on that stream and returns 0 on success or a negative error code;
 * dlbb_last_error() returns a human-readable description of the last failure in the
 * calling thread.  All kernels are compiled for sm_100a only.
 *
 * Reference citations ("ref:") name the file:line in
 * hardik-jinda/distributed-llm-backend-benchmark whose behaviour the entry point covers.
 */
#ifndef DLBB_H_
#define DLBB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dlbb_last_error(void);
/* Number of kernel launches issued through this library by the calling process so far. */
uint64_t dlbb_launch_count(void);
/* Build-time description: arch, compiler, feature flags. */
const char* dlbb_build_info(void);

/* ------------------------------------------------------------------------------------------
 * GEMM (bf16 inputs, fp32 accumulation in TMEM, bf16 output) — tcgen05.mma fed by TMA.
 * ref: models.py:47 (ColumnParallelLinear matmul), models.py:81 (RowParallelLinear matmul).
 *
 *   C[m,n] = epi( sum_k A(m,k) * B(k,n) )
 *   A(m,k) = a_mn_major ? A[k*lda + m] : A[m*lda + k]
 *   B(k,n) = b_mn_major ? B[k*ldb + n] : B[n*ldb + k]
 *   C row-major with leading dimension ldc.
 *
 * Epilogues:
 *   DLBB_EPI_NONE       C = acc
 *   DLBB_EPI_GELU       C = gelu(acc) (exact erf, ref models.py:182); if aux != NULL also
 *                       writes the pre-activation acc to aux (saved for backward)
 *   DLBB_EPI_ADD        C = acc + aux            (residual add, ref models.py:173,188)
 *   DLBB_EPI_MUL_DGELU  C = acc * gelu'(aux)     (backward of the GELU epilogue)
 * aux has the shape of C with leading dimension ld_aux.
 * Alignment: all base pointers 16 B aligned; lda, ldb, ldc, ld_aux multiples of 8 elements.
 * Shapes that violate this are routed to a SIMT fallback kernel (correct, slow).
 */
enum { DLBB_EPI_NONE = 0, DLBB_EPI_GELU = 1, DLBB_EPI_ADD = 2, DLBB_EPI_MUL_DGELU = 3 };

int dlbb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda,
                   int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major, int epilogue,
                   void* aux, int64_t ld_aux, void* stream);
/* Same contract, always the SIMT fallback (used by tests as an independent implementation). */
int dlbb_gemm_bf16_simt(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda,
                        int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major, int epilogue,
                        void* aux, int64_t ld_aux, void* stream);
/* Tuning knob: 0 = auto, 128 or 256 = force the N tile. */
void dlbb_gemm_set_tile_n(int bn);
/* Tuning knob: 2 = 256 x 256 tiles on CTA pairs (tcgen05.mma.cta_group::2, cluster of two) built with at most 168
 * registers per thread so that a second kernel can share the SMs (the default), 1 = the same without the register cap,
 * 0 = one CTA per 128 x 256 tile, -1 = take DLBB_GEMM_PAIR from the environment (default 2). */
void dlbb_gemm_set_pair_mode(int mode);

/* ------------------------------------------------------------------------------------------
 * LayerNorm over the last dimension (bf16 in/out, fp32 statistics).
 * ref: models.py:122,135,222 (nn.LayerNorm(hidden, bf16)), used at models.py:156,177,236.
 *   fwd: y = (x - mean) * rstd * gamma + beta ; mean/rstd (fp32 [rows]) saved when non-NULL.
 *   bwd: dx = LN'(dy) (+ dres if non-NULL) ; dgamma/dbeta (fp32 [cols]) accumulated from
 *        per-CTA partial sums held in `scratch` (fp32, dlbb_layernorm_bwd_scratch_bytes()).
 */
int dlbb_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                       float* rstd, int rows, int cols, float eps, void* stream);
size_t dlbb_layernorm_bwd_scratch_bytes(int rows, int cols);
int dlbb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd,
                       const void* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                       void* scratch, int rows, int cols, void* stream);
/* Same, with optional bf16 copies of the parameter gradients (dgamma_bf16 / dbeta_bf16, [cols]) written by the fold
 * kernel itself - the training engine points them into its flat gradient bucket, so no cast-and-copy follows. */
int dlbb_layernorm_bwd_ex(const void* dy, const void* x, const float* mean, const float* rstd,
                          const void* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                          void* dgamma_bf16, void* dbeta_bf16, void* scratch, int rows, int cols, void* stream);

/* LayerNorm backward that allocates nothing and needs no caller scratch.  It has no counterpart
 * in the reference; it exists for callers that may not call cudaMalloc where they need the op (a
 * process whose device memory is owned by its own pool, a plugin inside another runtime, a loop
 * that cannot take cudaMalloc's device-wide synchronisation).  Same arguments and results as
 * dlbb_layernorm_bwd minus `scratch`: the dgamma/dbeta partial sums go to an 8 MiB array built
 * into the library (held by every GPU that loads it, until process exit).
 *   - cols must be <= 8192, otherwise the call fails with an error and launches nothing;
 *   - the first call on each device looks up the array's device address; make that call before
 *     starting any stream capture;
 *   - all calls on a device share the array: calls that may overlap on the GPU race on it and must
 *     be serialised by the caller (same stream, or an event between them).  Nothing in the array
 *     needs clearing between calls - each call overwrites every element it later reads. */
int dlbb_layernorm_bwd_noalloc(const void* dy, const void* x, const float* mean, const float* rstd,
                               const void* gamma, const void* dres, void* dx, float* dgamma,
                               float* dbeta, int rows, int cols, void* stream);

/* ------------------------------------------------------------------------------------------
 * Elementwise (bf16).  ref: models.py:182 (gelu), models.py:173,188 (residual add).
 */
int dlbb_gelu_fwd(const void* x, void* y, int64_t n, void* stream);
int dlbb_gelu_bwd(const void* dy, const void* x, void* dx, int64_t n, void* stream);
int dlbb_add(const void* a, const void* b, void* out, int64_t n, void* stream);
/* Strided 2-D copy (slice materialisation, ref models.py:166-167). */
int dlbb_copy2d(const void* src, int64_t ld_src, void* dst, int64_t ld_dst, int rows, int cols,
                void* stream);
/* dtype conversion between fp16/bf16/fp32/fp64 (ref models.py:84,98; collectives/3d/openmpi.py:43).
 * dtype codes: DLBB_F16, DLBB_BF16, DLBB_F32, DLBB_F64. */
enum { DLBB_F16 = 0, DLBB_BF16 = 1, DLBB_F32 = 2, DLBB_F64 = 3, DLBB_U8 = 4, DLBB_I32 = 5, DLBB_I64 = 6 };
int dlbb_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------
 * Mean-squared-error loss and its gradient in one pass.
 * ref: test/ccl.py:110, test/ds_mpi_test.py:44 (F.mse_loss / mean on toy models); the LLM
 * training loss is new semantics (the reference model is forward-only).
 *   loss[0] = mean((y - t)^2) ; dy = 2 (y - t) / n * grad_scale   (dy may be NULL)
 * `scratch` holds per-block partial sums and a ticket counter: dlbb_mse_scratch_bytes(n)
 * bytes, zero-initialised once by the caller (the kernel leaves the counter at zero).
 */
size_t dlbb_mse_scratch_bytes(int64_t n);
int dlbb_mse_loss(const void* y, const void* target, void* dy, float* loss, void* scratch,
                  int64_t n, float grad_scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * Adam (ref: DeepSpeed engine.step with optimizer{type: Adam}, test/ccl.py:78-84,115).
 * bf16 parameter + bf16 gradient, fp32 moments and fp32 master copy (master may be NULL:
 * then the bf16 parameter is the only copy).  grad_scale multiplies the gradient first
 * (1/world_size, loss-scale removal).  step is 1-based.
 */
int dlbb_adam_bf16(void* param, const void* grad, float* m, float* v, float* master, int64_t n,
                   float lr, float beta1, float beta2, float eps, float weight_decay,
                   float grad_scale, int step, void* stream);
/* Cap the grid of the Adam kernels (0 = default).  With at most one 128-thread CTA per SM the update
 * co-resides with a running persistent GEMM instead of displacing it (used by the DP engine while the
 * backward pass is still in flight). */
void dlbb_adam_set_max_ctas(int n);
/* fp32 parameter variant for the toy DeepSpeed-style models (ref test/ds_mpi_test.py:8). */
int dlbb_adam_f32(float* param, const float* grad, float* m, float* v, int64_t n, float lr,
                  float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                  int step, void* stream);

/* L2 flush helper for benchmarks: streams `bytes` of writes through `buf`. */
int dlbb_fill_u32(void* buf, uint32_t value, int64_t n_words, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLBB_H_ */
