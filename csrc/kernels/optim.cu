// Adam update kernels (local, single GPU).  The multi-GPU variant that fuses the gradient
// reduce-scatter, 1/N scale, Adam and parameter all-gather lives in csrc/collectives/fused_adam.cu
// and shares adam_update() below.
// ref: DeepSpeed engine.step() with optimizer {type: Adam, params: {lr}} — test/ccl.py:78-84,115,
// test/ds_mpi_test.py:19-22,47.
#include "../../include/dlbb.h"
#include "adam_math.cuh"
#include "common.cuh"

namespace dlbb {

// Measured on B200 (tools/gpu_probe_corun.py, profiles/optimizer_overlap.md): running this kernel beside the
// persistent GEMM does not buy time - the GEMM slows down by about what the optimizer gains (both are limited by
// HBM / the 1 kW power cap) - so it is sized to run alone at full HBM speed: 256-thread CTAs, 16 per SM.
constexpr int kAdamThreads = 256;

__global__ void __launch_bounds__(kAdamThreads)
adam_bf16_kernel(__nv_bfloat16* __restrict__ param, const __nv_bfloat16* __restrict__ grad,
                 float* __restrict__ m, float* __restrict__ v, float* __restrict__ master, int64_t n,
                 AdamHyper h, int vec_ok) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t nvec = n >> 3;
        for (int64_t i = tid; i < nvec; i += nthreads) {
            const uint4 ug = reinterpret_cast<const uint4*>(grad)[i];
            const uint32_t gw[4] = {ug.x, ug.y, ug.z, ug.w};
            float g[8], w[8], mm[8], vv[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(gw[j]);
                g[2 * j] = f.x; g[2 * j + 1] = f.y;
            }
            float4* m4 = reinterpret_cast<float4*>(m) + 2 * i;
            float4* v4 = reinterpret_cast<float4*>(v) + 2 * i;
            const float4 ma = m4[0], mb = m4[1];
            const float4 va = v4[0], vb = v4[1];
            mm[0] = ma.x; mm[1] = ma.y; mm[2] = ma.z; mm[3] = ma.w;
            mm[4] = mb.x; mm[5] = mb.y; mm[6] = mb.z; mm[7] = mb.w;
            vv[0] = va.x; vv[1] = va.y; vv[2] = va.z; vv[3] = va.w;
            vv[4] = vb.x; vv[5] = vb.y; vv[6] = vb.z; vv[7] = vb.w;
            if (master) {
                const float4 wa = reinterpret_cast<const float4*>(master)[2 * i];
                const float4 wb = reinterpret_cast<const float4*>(master)[2 * i + 1];
                w[0] = wa.x; w[1] = wa.y; w[2] = wa.z; w[3] = wa.w;
                w[4] = wb.x; w[5] = wb.y; w[6] = wb.z; w[7] = wb.w;
            } else {
                const uint4 up = reinterpret_cast<const uint4*>(param)[i];
                const uint32_t pw[4] = {up.x, up.y, up.z, up.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(pw[j]);
                    w[2 * j] = f.x; w[2 * j + 1] = f.y;
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) adam_update(w[j], mm[j], vv[j], g[j], h);
            m4[0] = make_float4(mm[0], mm[1], mm[2], mm[3]);
            m4[1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
            v4[0] = make_float4(vv[0], vv[1], vv[2], vv[3]);
            v4[1] = make_float4(vv[4], vv[5], vv[6], vv[7]);
            if (master) {
                reinterpret_cast<float4*>(master)[2 * i] = make_float4(w[0], w[1], w[2], w[3]);
                reinterpret_cast<float4*>(master)[2 * i + 1] = make_float4(w[4], w[5], w[6], w[7]);
            }
            reinterpret_cast<uint4*>(param)[i] =
                make_uint4(pack_bf16x2(w[0], w[1]), pack_bf16x2(w[2], w[3]), pack_bf16x2(w[4], w[5]),
                           pack_bf16x2(w[6], w[7]));
        }
        done = nvec << 3;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) {
        float w = master ? master[i] : __bfloat162float(param[i]);
        float mm = m[i], vv = v[i];
        adam_update(w, mm, vv, __bfloat162float(grad[i]), h);
        m[i] = mm; v[i] = vv;
        if (master) master[i] = w;
        param[i] = __float2bfloat16(w);
    }
}

__global__ void __launch_bounds__(kAdamThreads)
adam_f32_kernel(float* __restrict__ param, const float* __restrict__ grad, float* __restrict__ m,
                float* __restrict__ v, int64_t n, AdamHyper h) {
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nthreads) {
        float w = param[i], mm = m[i], vv = v[i];
        adam_update(w, mm, vv, grad[i], h);
        param[i] = w; m[i] = mm; v[i] = vv;
    }
}

static int g_adam_max_ctas = 0;

static int adam_grid(int64_t items) {
    int64_t b = (items + kAdamThreads - 1) / kAdamThreads;
    const int64_t cap = g_adam_max_ctas > 0 ? g_adam_max_ctas : 148 * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

}  // namespace dlbb

extern "C" {

void dlbb_adam_set_max_ctas(int n) { dlbb::g_adam_max_ctas = n; }

int dlbb_adam_bf16(void* param, const void* grad, float* m, float* v, float* master, int64_t n,
                   float lr, float beta1, float beta2, float eps, float weight_decay,
                   float grad_scale, int step, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(step >= 1, "adam: step must be >= 1 (got %d)", step);
    if (n <= 0) return 0;
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    const uintptr_t al = reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) |
                         reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
                         reinterpret_cast<uintptr_t>(master);
    const int vec_ok = (al & 15) == 0;
    // Same shared-memory carve-out as the GEMM (max shared): an SM can only switch its L1/smem split
    // when it is idle, so a CTA of this kernel with the default split would keep the next GEMM CTA
    // (226 KB of smem) off its SM until it exits - and serialise the two kernels.
    static bool carveout_set = false;
    if (!carveout_set) {
        cudaFuncSetAttribute(adam_bf16_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        carveout_set = true;
    }
    adam_bf16_kernel<<<adam_grid((n + 7) / 8), kAdamThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<__nv_bfloat16*>(param), static_cast<const __nv_bfloat16*>(grad), m, v, master, n, h,
        vec_ok);
    DLBB_CHECK_LAUNCH("adam_bf16");
    return 0;
}

int dlbb_adam_f32(float* param, const float* grad, float* m, float* v, int64_t n, float lr,
                  float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                  int step, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(step >= 1, "adam: step must be >= 1 (got %d)", step);
    if (n <= 0) return 0;
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    adam_f32_kernel<<<adam_grid(n), kAdamThreads, 0, static_cast<cudaStream_t>(stream)>>>(param, grad, m,
                                                                                        v, n, h);
    DLBB_CHECK_LAUNCH("adam_f32");
    return 0;
}
}
