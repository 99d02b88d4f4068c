// LayerNorm forward / backward (bf16 activations, fp32 statistics), one CTA per row.
// ref: models.py:122,135,222 (nn.LayerNorm(hidden, dtype=bf16)) used at models.py:156,177,236.
//
// The row is read once with 16-byte loads and kept in registers for both passes (mean, then
// centred variance) so HBM traffic is exactly read-x + write-y.  Backward fuses the residual
// branch gradient (dx = LN'(dy) + dres) and produces dgamma/dbeta through per-CTA register
// partials that a second tiny kernel folds together (deterministic, no atomics).
#include <mutex>

#include "../../include/dlbb.h"
#include "common.cuh"

namespace dlbb {

constexpr int kLnThreads = 256;

__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();  // protect sh from the previous use
    if (l == 0) sh[w] = v;
    __syncthreads();
    float t = (l < kLnThreads / 32) ? sh[l] : 0.f;
    return warp_sum(t);
}

// Two sums at once: half the barriers of two block_sum calls (the backward needs both per row).
__device__ __forceinline__ float2 block_sum2(float a, float b, float2* sh) {
    a = warp_sum(a);
    b = warp_sum(b);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = make_float2(a, b);
    __syncthreads();
    float2 t = (l < kLnThreads / 32) ? sh[l] : make_float2(0.f, 0.f);
    return make_float2(warp_sum(t.x), warp_sum(t.y));
}

// ---------------------------------------------------------------- vector path (cols % 8 == 0)
template <int VPT>
__global__ void __launch_bounds__(kLnThreads)
ln_fwd_vec_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gamma,
                  const uint4* __restrict__ beta, uint4* __restrict__ y, float* __restrict__ mean_out,
                  float* __restrict__ rstd_out, int rows, int cols, float eps) {
    __shared__ float sh[32];
    const int nvec = cols >> 3;
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const uint4* xr = x + (size_t)row * nvec;
        float v[VPT][8];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * kLnThreads;
            if (c < nvec) {
                const uint4 u = xr[c];
                const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z),
                             d = unpack_bf16x2(u.w);
                v[i][0] = a.x; v[i][1] = a.y; v[i][2] = b.x; v[i][3] = b.y;
                v[i][4] = cc.x; v[i][5] = cc.y; v[i][6] = d.x; v[i][7] = d.y;
#pragma unroll
                for (int j = 0; j < 8; ++j) s += v[i][j];
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
            }
        }
        const float mean = block_sum(s, sh) / cols;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * kLnThreads;
            if (c < nvec) {
#pragma unroll
                for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; q += d * d; }
            }
        }
        const float rstd = rsqrtf(block_sum(q, sh) / cols + eps);
        if (threadIdx.x == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
        uint4* yr = y + (size_t)row * nvec;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * kLnThreads;
            if (c < nvec) {
                const uint4 g = gamma[c], b = beta[c];
                const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, bw[4] = {b.x, b.y, b.z, b.w};
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 gg = unpack_bf16x2(gw[j]), bb = unpack_bf16x2(bw[j]);
                    o[j] = pack_bf16x2((v[i][2 * j] - mean) * rstd * gg.x + bb.x,
                                       (v[i][2 * j + 1] - mean) * rstd * gg.y + bb.y);
                }
                yr[c] = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
    }
}

// Two CTAs per SM (<= 128 registers): the normalised input and the scaled gradient are recomputed in the second
// pass instead of being held across the two block-wide sums - at 146 registers one CTA per SM left the kernel at 41 %
// of the HBM rate (ncu: 50 us for [4096, 4096], 12 % achieved occupancy).
template <int VPT>
__global__ void __launch_bounds__(kLnThreads, 2)
ln_bwd_vec_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ x,
                  const float* __restrict__ mean, const float* __restrict__ rstd,
                  const uint4* __restrict__ gamma, const uint4* __restrict__ dres,
                  uint4* __restrict__ dx, float* __restrict__ partials, int rows, int cols) {
    __shared__ float2 sh2[32];
    const int nvec = cols >> 3;
    float pg[VPT][8], pb[VPT][8];  // per-thread column partials of dgamma / dbeta
    float gam[VPT][8];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * kLnThreads;
        uint4 g = make_uint4(0, 0, 0, 0);
        if (c < nvec) g = gamma[c];
        const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16x2(gw[j]);
            gam[i][2 * j] = f.x; gam[i][2 * j + 1] = f.y;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { pg[i][j] = 0.f; pb[i][j] = 0.f; }
    }
    // Software pipeline over this CTA's rows: the loads of row r+1 are issued before the two block-wide
    // reductions of row r, so HBM latency hides behind the barriers instead of adding to them.
    uint4 nd[VPT], nx[VPT], nr[VPT];
    auto fetch = [&](int row) {
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * kLnThreads;
            if (c < nvec) {
                nd[i] = dy[(size_t)row * nvec + c];
                nx[i] = x[(size_t)row * nvec + c];
                if (dres) nr[i] = dres[(size_t)row * nvec + c];
            }
        }
    };
    if ((int)blockIdx.x < rows) fetch(blockIdx.x);
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const float mu = mean[row], rs = rstd[row];
        uint4 cd[VPT], cx[VPT], cr[VPT];
#pragma unroll
        for (int i = 0; i < VPT; ++i) { cd[i] = nd[i]; cx[i] = nx[i]; cr[i] = nr[i]; }
        if (row + (int)gridDim.x < rows) fetch(row + gridDim.x);
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * kLnThreads;
            if (c < nvec) {
                const uint32_t dw[4] = {cd[i].x, cd[i].y, cd[i].z, cd[i].w}, xw[4] = {cx[i].x, cx[i].y, cx[i].z, cx[i].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 d = unpack_bf16x2(dw[j]), xx = unpack_bf16x2(xw[j]);
                    const float h0 = (xx.x - mu) * rs, h1 = (xx.y - mu) * rs;
                    pg[i][2 * j] += d.x * h0; pg[i][2 * j + 1] += d.y * h1;
                    pb[i][2 * j] += d.x; pb[i][2 * j + 1] += d.y;
                    const float g0 = d.x * gam[i][2 * j], g1 = d.y * gam[i][2 * j + 1];
                    s1 += g0 + g1;
                    s2 += g0 * h0 + g1 * h1;
                }
            }
        }
        const float2 cs = block_sum2(s1, s2, reinterpret_cast<float2*>(sh2));
        const float c1 = cs.x / cols, c2 = cs.y / cols;
        uint4* dxr = dx + (size_t)row * nvec;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * kLnThreads;
            if (c < nvec) {
                const uint32_t dw[4] = {cd[i].x, cd[i].y, cd[i].z, cd[i].w}, xw[4] = {cx[i].x, cx[i].y, cx[i].z, cx[i].w};
                float r[8];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 d = unpack_bf16x2(dw[j]), xx = unpack_bf16x2(xw[j]);
                    const float h0 = (xx.x - mu) * rs, h1 = (xx.y - mu) * rs;
                    r[2 * j] = rs * (d.x * gam[i][2 * j] - c1 - h0 * c2);
                    r[2 * j + 1] = rs * (d.y * gam[i][2 * j + 1] - c1 - h1 * c2);
                }
                if (dres) {
                    const uint32_t rw[4] = {cr[i].x, cr[i].y, cr[i].z, cr[i].w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 f = unpack_bf16x2(rw[j]);
                        r[2 * j] += f.x; r[2 * j + 1] += f.y;
                    }
                }
                dxr[c] = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]),
                                    pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
            }
        }
    }
    // partials[blockIdx.x][0][cols] = dgamma part, [1][cols] = dbeta part
    float* pgo = partials + (size_t)blockIdx.x * 2 * cols;
    float* pbo = pgo + cols;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * kLnThreads;
        if (c < nvec) {
            float4* g4 = reinterpret_cast<float4*>(pgo + c * 8);
            float4* b4 = reinterpret_cast<float4*>(pbo + c * 8);
            g4[0] = make_float4(pg[i][0], pg[i][1], pg[i][2], pg[i][3]);
            g4[1] = make_float4(pg[i][4], pg[i][5], pg[i][6], pg[i][7]);
            b4[0] = make_float4(pb[i][0], pb[i][1], pb[i][2], pb[i][3]);
            b4[1] = make_float4(pb[i][4], pb[i][5], pb[i][6], pb[i][7]);
        }
    }
}

// ---------------------------------------------------------------- generic path (any cols)
__global__ void __launch_bounds__(kLnThreads)
ln_fwd_generic_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                      const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                      float* mean_out, float* rstd_out, int rows, int cols, float eps) {
    __shared__ float sh[32];
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const __nv_bfloat16* xr = x + (size_t)row * cols;
        float s = 0.f;
        for (int c = threadIdx.x; c < cols; c += kLnThreads) s += __bfloat162float(xr[c]);
        const float mean = block_sum(s, sh) / cols;
        float q = 0.f;
        for (int c = threadIdx.x; c < cols; c += kLnThreads) {
            const float d = __bfloat162float(xr[c]) - mean;
            q += d * d;
        }
        const float rstd = rsqrtf(block_sum(q, sh) / cols + eps);
        if (threadIdx.x == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
        for (int c = threadIdx.x; c < cols; c += kLnThreads)
            y[(size_t)row * cols + c] =
                __float2bfloat16((__bfloat162float(xr[c]) - mean) * rstd * __bfloat162float(gamma[c]) +
                                 __bfloat162float(beta[c]));
    }
}

__global__ void __launch_bounds__(kLnThreads)
ln_bwd_generic_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                      const float* __restrict__ mean, const float* __restrict__ rstd,
                      const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ dres,
                      __nv_bfloat16* __restrict__ dx, float* __restrict__ partials, int rows, int cols) {
    __shared__ float sh[32];
    float* pgo = partials + (size_t)blockIdx.x * 2 * cols;
    float* pbo = pgo + cols;
    for (int c = threadIdx.x; c < cols; c += kLnThreads) { pgo[c] = 0.f; pbo[c] = 0.f; }
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const float mu = mean[row], rs = rstd[row];
        const size_t off = (size_t)row * cols;
        float s1 = 0.f, s2 = 0.f;
        for (int c = threadIdx.x; c < cols; c += kLnThreads) {
            const float d = __bfloat162float(dy[off + c]);
            const float h = (__bfloat162float(x[off + c]) - mu) * rs;
            const float g = d * __bfloat162float(gamma[c]);
            pgo[c] += d * h;  // each thread owns its columns of this CTA's slice
            pbo[c] += d;
            s1 += g;
            s2 += g * h;
        }
        const float c1 = block_sum(s1, sh) / cols;
        const float c2 = block_sum(s2, sh) / cols;
        for (int c = threadIdx.x; c < cols; c += kLnThreads) {
            const float d = __bfloat162float(dy[off + c]);
            const float h = (__bfloat162float(x[off + c]) - mu) * rs;
            float r = rs * (d * __bfloat162float(gamma[c]) - c1 - h * c2);
            if (dres) r += __bfloat162float(dres[off + c]);
            dx[off + c] = __float2bfloat16(r);
        }
    }
}

// Fold the per-CTA partials.  A 256-thread block owns 32 columns: lane = column, warp = slice of the
// partial rows (8 slices), so every load is a coalesced 128 B line; each thread keeps 8 independent loads per array in
// flight (the dependent-add version took 30 us for 296 x 4096 partials, ~600 ns per step of its chain); the slices are
// combined through shared memory in a fixed order (deterministic).  The sums go to fp32 outputs and / or, as bf16,
// straight into a gradient bucket (dgamma16 / dbeta16) so that no cast-and-copy kernel has to follow.
__global__ void __launch_bounds__(256)
ln_bwd_reduce_kernel(const float* __restrict__ partials, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     int nparts, int cols, __nv_bfloat16* __restrict__ dgamma16 = nullptr,
                     __nv_bfloat16* __restrict__ dbeta16 = nullptr) {
    __shared__ float sg[8][33], sb[8][33];
    const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float g = 0.f, b = 0.f;
    if (c < cols) {
        constexpr int kIlp = 8;
        int p = slice;
        for (; p + (kIlp - 1) * 8 < nparts; p += kIlp * 8) {
            float tg[kIlp], tb[kIlp];
#pragma unroll
            for (int k = 0; k < kIlp; ++k) {
                tg[k] = partials[(size_t)(p + k * 8) * 2 * cols + c];
                tb[k] = partials[(size_t)(p + k * 8) * 2 * cols + cols + c];
            }
#pragma unroll
            for (int k = 0; k < kIlp; ++k) { g += tg[k]; b += tb[k]; }
        }
        for (; p < nparts; p += 8) {
            g += partials[(size_t)p * 2 * cols + c];
            b += partials[(size_t)p * 2 * cols + cols + c];
        }
    }
    sg[slice][lane] = g;
    sb[slice][lane] = b;
    __syncthreads();
    if (slice == 0 && c < cols) {
#pragma unroll
        for (int k = 1; k < 8; ++k) { g += sg[k][lane]; b += sb[k][lane]; }
        if (dgamma) dgamma[c] = g;
        if (dbeta) dbeta[c] = b;
        if (dgamma16) dgamma16[c] = __float2bfloat16(g);
        if (dbeta16) dbeta16[c] = __float2bfloat16(b);
    }
}

static int ln_bwd_parts(int rows) {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            sms = 148;
    }
    return rows < 2 * sms ? rows : 2 * sms;
}

}  // namespace dlbb

extern "C" {

int dlbb_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                       float* rstd, int rows, int cols, float eps, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(rows > 0 && cols > 0, "layernorm_fwd: empty input");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool vec = (cols % 8 == 0) && cols <= 8 * kLnThreads * 4 &&
                     ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) |
                       reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15) == 0;
    const int grid = rows;
    if (vec) {
        const int vpt = (cols / 8 + kLnThreads - 1) / kLnThreads;
#define LN_FWD(V)                                                                                  \
    ln_fwd_vec_kernel<V><<<grid, kLnThreads, 0, s>>>(                                              \
        static_cast<const uint4*>(x), static_cast<const uint4*>(gamma),                            \
        static_cast<const uint4*>(beta), static_cast<uint4*>(y), mean, rstd, rows, cols, eps)
        if (vpt == 1) LN_FWD(1); else if (vpt == 2) LN_FWD(2); else if (vpt == 3) LN_FWD(3); else LN_FWD(4);
#undef LN_FWD
    } else {
        ln_fwd_generic_kernel<<<grid, kLnThreads, 0, s>>>(
            static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(gamma),
            static_cast<const __nv_bfloat16*>(beta), static_cast<__nv_bfloat16*>(y), mean, rstd, rows,
            cols, eps);
    }
    DLBB_CHECK_LAUNCH("layernorm_fwd");
    return 0;
}

size_t dlbb_layernorm_bwd_scratch_bytes(int rows, int cols) {
    return (size_t)dlbb::ln_bwd_parts(rows) * 2 * cols * sizeof(float);
}

int dlbb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd,
                       const void* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                       void* scratch, int rows, int cols, void* stream) {
    return dlbb_layernorm_bwd_ex(dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, nullptr, nullptr, scratch, rows, cols,
                                 stream);
}

int dlbb_layernorm_bwd_ex(const void* dy, const void* x, const float* mean, const float* rstd,
                          const void* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                          void* dgamma_bf16, void* dbeta_bf16, void* scratch, int rows, int cols, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(rows > 0 && cols > 0, "layernorm_bwd: empty input");
    DLBB_REQUIRE(scratch != nullptr, "layernorm_bwd: scratch is required");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int parts = ln_bwd_parts(rows);
    uintptr_t al = reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x) |
                   reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(dx) |
                   reinterpret_cast<uintptr_t>(dres) | reinterpret_cast<uintptr_t>(scratch);
    const bool vec = (cols % 8 == 0) && cols <= 8 * kLnThreads * 4 && (al & 15) == 0;
    float* partials = static_cast<float*>(scratch);
    if (vec) {
        const int vpt = (cols / 8 + kLnThreads - 1) / kLnThreads;
#define LN_BWD(V)                                                                                  \
    ln_bwd_vec_kernel<V><<<parts, kLnThreads, 0, s>>>(                                             \
        static_cast<const uint4*>(dy), static_cast<const uint4*>(x), mean, rstd,                   \
        static_cast<const uint4*>(gamma), static_cast<const uint4*>(dres), static_cast<uint4*>(dx), \
        partials, rows, cols)
        if (vpt == 1) LN_BWD(1); else if (vpt == 2) LN_BWD(2); else if (vpt == 3) LN_BWD(3); else LN_BWD(4);
#undef LN_BWD
    } else {
        ln_bwd_generic_kernel<<<parts, kLnThreads, 0, s>>>(
            static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x), mean, rstd,
            static_cast<const __nv_bfloat16*>(gamma), static_cast<const __nv_bfloat16*>(dres),
            static_cast<__nv_bfloat16*>(dx), partials, rows, cols);
    }
    DLBB_CHECK_LAUNCH("layernorm_bwd");
    if (dgamma || dbeta || dgamma_bf16 || dbeta_bf16) {
        ln_bwd_reduce_kernel<<<(cols + 31) / 32, 256, 0, s>>>(partials, dgamma, dbeta, parts, cols,
                                                              static_cast<__nv_bfloat16*>(dgamma_bf16),
                                                              static_cast<__nv_bfloat16*>(dbeta_bf16));
        DLBB_CHECK_LAUNCH("layernorm_bwd_reduce");
    }
    return 0;
}
}

// ---------------------------------------------------------------------------------------------
// LayerNorm backward for callers that cannot allocate device memory.
//
// Some hosts of this library are not allowed to call cudaMalloc at the point where they need a
// LayerNorm backward: a server that hands the whole device to its own memory pool at start-up and
// forbids driver allocations afterwards, a plugin running inside someone else's runtime with no
// access to that runtime's allocator, or a latency-critical loop that cannot afford the implicit
// device-wide synchronisation of cudaMalloc / cudaFree.  dlbb_layernorm_bwd() needs such a caller
// to bring a scratch buffer; this variant brings its own instead: the per-CTA dgamma / dbeta
// partial sums live in the fixed array below, which is part of the library image.
//
// Cost and limits of that choice:
//   * every GPU on which this library is loaded holds the array (8 MiB) until the process exits,
//     whether or not the entry point is ever called;
//   * the array fits kNoAllocMaxParts CTAs x 2 x kNoAllocMaxCols floats, so cols is limited to
//     kNoAllocMaxCols (8192) - larger inputs are rejected, not truncated; rows is unlimited
//     (at most kNoAllocMaxParts CTAs walk the rows);
//   * all calls on a device share the one array and it keeps whatever the previous call left in
//     it.  Nothing in it has to start at zero: each CTA overwrites its whole slice (columns
//     < cols) before the fold kernel, launched behind it on the same stream, reads it.  Two calls
//     that can overlap on the GPU (different streams, no event between them) race on the array
//     and corrupt each other's dgamma / dbeta; serialise them.
//
// Host code cannot use the array's name as a pointer - on the host it is only a handle for the
// runtime.  The device address is looked up once per device with cudaGetSymbolAddress (with that
// device current) and cached; the first call on a device therefore does a little set-up work and
// should be made before any stream capture is started.
// ---------------------------------------------------------------------------------------------
namespace dlbb {

constexpr int kNoAllocMaxParts = 128;
constexpr int kNoAllocMaxCols = 8192;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_ln_noalloc_scratch[(size_t)kNoAllocMaxParts * 2 * kNoAllocMaxCols];

static float* g_ln_noalloc_ptr[kNoAllocMaxDevices];  // device address of the array, per device
static std::mutex g_ln_noalloc_mu;

// Returns the device address of the scratch array on the current device (nullptr + error set on failure).
static float* ln_noalloc_scratch_for_current_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error("layernorm_bwd_noalloc: cudaGetDevice failed: %s", cudaGetErrorString(e));
        return nullptr;
    }
    if (dev < 0 || dev >= kNoAllocMaxDevices) {
        set_error("layernorm_bwd_noalloc: device ordinal %d is outside the supported range [0, %d)", dev,
                  kNoAllocMaxDevices);
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_ln_noalloc_mu);
    if (g_ln_noalloc_ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_ln_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("layernorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                      cudaGetErrorString(e));
            return nullptr;
        }
        g_ln_noalloc_ptr[dev] = static_cast<float*>(p);
    }
    return g_ln_noalloc_ptr[dev];
}

}  // namespace dlbb

extern "C" int dlbb_layernorm_bwd_noalloc(const void* dy, const void* x, const float* mean, const float* rstd,
                                          const void* gamma, const void* dres, void* dx, float* dgamma,
                                          float* dbeta, int rows, int cols, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(rows > 0 && cols > 0, "layernorm_bwd_noalloc: empty input (%d x %d)", rows, cols);
    DLBB_REQUIRE(cols <= kNoAllocMaxCols,
                 "layernorm_bwd_noalloc: cols=%d exceeds the built-in scratch (at most %d columns); "
                 "use dlbb_layernorm_bwd with a caller-provided scratch buffer",
                 cols, kNoAllocMaxCols);
    float* partials = ln_noalloc_scratch_for_current_device();
    if (partials == nullptr) return -1;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int parts = rows < kNoAllocMaxParts ? rows : kNoAllocMaxParts;
    const uintptr_t al = reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x) |
                         reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(dx) |
                         reinterpret_cast<uintptr_t>(dres);
    const bool vec = (cols % 8 == 0) && (al & 15) == 0;  // cols <= 8192 always fits VPT <= 4
    if (vec) {
        const int vpt = (cols / 8 + kLnThreads - 1) / kLnThreads;
#define LN_BWD_NA(V)                                                                               \
    ln_bwd_vec_kernel<V><<<parts, kLnThreads, 0, s>>>(                                             \
        static_cast<const uint4*>(dy), static_cast<const uint4*>(x), mean, rstd,                   \
        static_cast<const uint4*>(gamma), static_cast<const uint4*>(dres), static_cast<uint4*>(dx), \
        partials, rows, cols)
        if (vpt == 1) LN_BWD_NA(1); else if (vpt == 2) LN_BWD_NA(2); else if (vpt == 3) LN_BWD_NA(3); else LN_BWD_NA(4);
#undef LN_BWD_NA
    } else {
        ln_bwd_generic_kernel<<<parts, kLnThreads, 0, s>>>(
            static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x), mean, rstd,
            static_cast<const __nv_bfloat16*>(gamma), static_cast<const __nv_bfloat16*>(dres),
            static_cast<__nv_bfloat16*>(dx), partials, rows, cols);
    }
    DLBB_CHECK_LAUNCH("layernorm_bwd_noalloc");
    if (dgamma || dbeta) {
        ln_bwd_reduce_kernel<<<(cols + 31) / 32, 256, 0, s>>>(partials, dgamma, dbeta, parts, cols);
        DLBB_CHECK_LAUNCH("layernorm_bwd_noalloc_reduce");
    }
    return 0;
}
