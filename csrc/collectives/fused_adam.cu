// K20: gradient reduce-scatter + 1/N scale + Adam + parameter all-gather in ONE kernel, over
// symmetric memory.  This is the data-parallel "all-reduce the bucket, then step the optimizer"
// of DeepSpeed ZeRO (ref test/ccl.py:86-88,114-115; test/ds_mpi_test.py:13-47) collapsed so that
// the reduced gradient never exists in memory: each rank pulls its 1/world shard of the bucket
// from every peer (NVLink loads, or one multimem.ld_reduce that the NVSwitch reduces), updates
// the shard with its local fp32 moments / master weights and pushes the new bf16 parameters to
// every rank (peer stores, or one multimem.st).  Optimizer state is therefore sharded 1/world per
// rank and Adam runs once per element instead of once per rank.
//
// CTA b of every rank only synchronises with CTA b of the peers (row barriers on pad bank 1):
//   entry barrier - every rank has reached this bucket, i.e. its gradients are final
//   exit barrier  - every store into this rank's parameters has landed
#include "../../include/dlbb.h"
#include "../../include/dlbb_comm.h"
#include "../kernels/adam_math.cuh"
#include "coll_device.cuh"

namespace dlbb {

constexpr int kFusedThreads = 128;   // small CTAs: they must fit beside a resident GEMM CTA (regs, no smem)
constexpr int kFusedBank = 1;
constexpr int kFU = 4;               // vectors per thread per iteration (x world peer loads in flight)

struct PeerPtrs {
    const char* grads[DLBB_MAX_RANKS];
    char* params[DLBB_MAX_RANKS];
    const char* grads_mc;
    char* params_mc;
};

template <bool NVLS>
__global__ void __maxnreg__(112)
fused_reduce_adam_kernel(CommDev c, PeerPtrs pp, int64_t lo, int64_t hi, float* __restrict__ m,
                         float* __restrict__ v, float* __restrict__ master, AdamHyper h, uint32_t epoch) {
    row_barrier(c, kFusedBank, blockIdx.x, epoch);
    // [lo, hi) is this rank's shard in elements relative to the buffers; lo is a multiple of 8.
    const int64_t n = hi - lo;
    const int64_t nvec = n >> 3;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = tid; i0 < nvec; i0 += nth * kFU) {
        float g[kFU][8];
        if (NVLS) {
            uint4 s[kFU];
#pragma unroll
            for (int u = 0; u < kFU; ++u)
                if (i0 + u * nth < nvec) s[u] = mm_ld_reduce_add_bf16(pp.grads_mc + (lo + ((i0 + u * nth) << 3)) * 2);
#pragma unroll
            for (int u = 0; u < kFU; ++u) {
                const uint32_t w4[4] = {s[u].x, s[u].y, s[u].z, s[u].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(w4[j]);
                    g[u][2 * j] = f.x; g[u][2 * j + 1] = f.y;
                }
            }
        } else {
#pragma unroll
            for (int u = 0; u < kFU; ++u)
#pragma unroll
                for (int j = 0; j < 8; ++j) g[u][j] = 0.f;
            for (int p = 0; p < c.world; ++p) {   // rank order: every rank would compute identical sums
                uint4 buf[kFU];
#pragma unroll
                for (int u = 0; u < kFU; ++u)
                    if (i0 + u * nth < nvec) buf[u] = ld_peer_v4(pp.grads[p] + (lo + ((i0 + u * nth) << 3)) * 2);
#pragma unroll
                for (int u = 0; u < kFU; ++u) {
                    if (i0 + u * nth < nvec) {
                        const uint32_t w4[4] = {buf[u].x, buf[u].y, buf[u].z, buf[u].w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpack_bf16x2(w4[k]);
                            g[u][2 * k] += f.x; g[u][2 * k + 1] += f.y;
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kFU; ++u) {
            const int64_t i = i0 + u * nth;
            if (i >= nvec) continue;
            const int64_t boff = (lo + (i << 3)) * 2;  // byte offset of this 16 B vector
            float w[8], mm[8], vv[8];
            float4* m4 = reinterpret_cast<float4*>(m) + 2 * i;
            float4* v4 = reinterpret_cast<float4*>(v) + 2 * i;
            const float4 ma = m4[0], mb = m4[1], va = v4[0], vb = v4[1];
            mm[0] = ma.x; mm[1] = ma.y; mm[2] = ma.z; mm[3] = ma.w; mm[4] = mb.x; mm[5] = mb.y; mm[6] = mb.z; mm[7] = mb.w;
            vv[0] = va.x; vv[1] = va.y; vv[2] = va.z; vv[3] = va.w; vv[4] = vb.x; vv[5] = vb.y; vv[6] = vb.z; vv[7] = vb.w;
            if (master) {
                const float4 wa = reinterpret_cast<const float4*>(master)[2 * i], wb = reinterpret_cast<const float4*>(master)[2 * i + 1];
                w[0] = wa.x; w[1] = wa.y; w[2] = wa.z; w[3] = wa.w; w[4] = wb.x; w[5] = wb.y; w[6] = wb.z; w[7] = wb.w;
            } else {
                const uint4 up = *reinterpret_cast<const uint4*>(pp.params[c.rank] + boff);
                const uint32_t pw[4] = {up.x, up.y, up.z, up.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(pw[j]);
                    w[2 * j] = f.x; w[2 * j + 1] = f.y;
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) adam_update(w[j], mm[j], vv[j], g[u][j], h);
            m4[0] = make_float4(mm[0], mm[1], mm[2], mm[3]); m4[1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
            v4[0] = make_float4(vv[0], vv[1], vv[2], vv[3]); v4[1] = make_float4(vv[4], vv[5], vv[6], vv[7]);
            if (master) {
                reinterpret_cast<float4*>(master)[2 * i] = make_float4(w[0], w[1], w[2], w[3]);
                reinterpret_cast<float4*>(master)[2 * i + 1] = make_float4(w[4], w[5], w[6], w[7]);
            }
            const uint4 o = make_uint4(pack_bf16x2(w[0], w[1]), pack_bf16x2(w[2], w[3]), pack_bf16x2(w[4], w[5]),
                                       pack_bf16x2(w[6], w[7]));
            if (NVLS) {
                mm_st_v4(pp.params_mc + boff, o);
            } else {
#pragma unroll 1
                for (int p = 0; p < c.world; ++p) {
                    const int q = (c.rank + p) % c.world;
                    st_v4(pp.params[q] + boff, o);
                }
            }
        }
    }
    // ragged tail (< 8 elements) of the shard: scalar, always through peer pointers
    for (int64_t e = (nvec << 3) + tid; e < n; e += nth) {
        const int64_t idx = lo + e;
        float g = 0.f;
        for (int p = 0; p < c.world; ++p)
            g += __bfloat162float(ld_peer_scalar(reinterpret_cast<const __nv_bfloat16*>(pp.grads[p]) + idx));
        float w = master ? master[e] : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(pp.params[c.rank])[idx]);
        float mm = m[e], vv = v[e];
        adam_update(w, mm, vv, g, h);
        m[e] = mm; v[e] = vv;
        if (master) master[e] = w;
        for (int p = 0; p < c.world; ++p) reinterpret_cast<__nv_bfloat16*>(pp.params[p])[idx] = __float2bfloat16(w);
    }
    row_barrier(c, kFusedBank, blockIdx.x, epoch + 1);
}

}  // namespace dlbb

extern "C" int dlbb_fused_reduce_adam_bf16(dlbb_comm_t* c, int grads_seg, int params_seg, int64_t elem_off, int64_t n,
                                           float* m, float* v, float* master, float lr, float beta1, float beta2,
                                           float eps, float weight_decay, float grad_scale, int step,
                                           int use_multicast, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(step >= 1, "fused_reduce_adam: step must be >= 1");
    DLBB_REQUIRE(n > 0 && elem_off >= 0 && elem_off % 8 == 0, "fused_reduce_adam: offset must be a multiple of 8 elements");
    DLBB_REQUIRE(dlbb_symm_ptr(c, grads_seg, 0) && dlbb_symm_ptr(c, params_seg, 0), "fused_reduce_adam: bad segment");
    DLBB_REQUIRE((size_t)(elem_off + n) * 2 <= dlbb_symm_size(c, grads_seg) &&
                     (size_t)(elem_off + n) * 2 <= dlbb_symm_size(c, params_seg),
                 "fused_reduce_adam: bucket exceeds the symmetric buffers");
    const int64_t q = (int64_t)c->world * 8;
    const int64_t per = (n + q - 1) / q * 8;  // shard length, multiple of 8
    const int64_t lo = std::min<int64_t>((int64_t)c->rank * per, n), hi = std::min<int64_t>(lo + per, n);
    PeerPtrs pp{};
    for (int p = 0; p < c->world; ++p) {
        pp.grads[p] = static_cast<const char*>(dlbb_symm_ptr(c, grads_seg, p)) + elem_off * 2;
        pp.params[p] = static_cast<char*>(dlbb_symm_ptr(c, params_seg, p)) + elem_off * 2;
    }
    const char* gmc = static_cast<const char*>(dlbb_symm_mc_ptr(c, grads_seg));
    char* pmc = static_cast<char*>(dlbb_symm_mc_ptr(c, params_seg));
    const bool nvls = use_multicast && gmc && pmc && c->world > 1;
    pp.grads_mc = nvls ? gmc + elem_off * 2 : nullptr;
    pp.params_mc = nvls ? pmc + elem_off * 2 : nullptr;
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    int ctas = c->ctas > 0 ? c->ctas : 120;
    const int64_t want = ((hi - lo) / 8 + kFusedThreads - 1) / kFusedThreads;
    if (want < ctas) ctas = static_cast<int>(std::max<int64_t>(1, want));
    ctas = std::min(ctas, kMaxCtas);
    // every rank must launch the same grid (row barriers pair CTA b with CTA b): derive from `per`
    {
        const int64_t want_all = (per / 8 + kFusedThreads - 1) / kFusedThreads;
        ctas = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(c->ctas > 0 ? c->ctas : 120, want_all)));
        ctas = std::min(ctas, kMaxCtas);
    }
    if (const int rg = refresh_flag_bank_if_due(c, kFusedBank, stream)) return rg;   // (comm_internal.h: flags stay within 2^31 of the epoch)
    const uint32_t ep = c->epoch[kFusedBank] + 1;
    c->epoch[kFusedBank] += 2;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
#if !defined(DLBB_HOST_EMU)
    static bool carveout_set = false;
    if (!carveout_set) {  // co-reside with the GEMM CTAs: see csrc/kernels/optim.cu
        cudaFuncSetAttribute(fused_reduce_adam_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(fused_reduce_adam_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        carveout_set = true;
    }
#endif
    if (nvls)
        fused_reduce_adam_kernel<true><<<ctas, kFusedThreads, 0, s>>>(c->dev, pp, lo, hi, m, v, master, h, ep);
    else
        fused_reduce_adam_kernel<false><<<ctas, kFusedThreads, 0, s>>>(c->dev, pp, lo, hi, m, v, master, h, ep);
    DLBB_CHECK_LAUNCH("fused_reduce_adam_kernel");
    return 0;
}
