// Elementwise / reduction kernels: GELU fwd/bwd, add, strided copy, dtype casts, MSE loss,
// buffer fill.  All are HBM-bound: 16-byte vector accesses, grid-stride loops sized to a few
// waves of 148 SMs.
// ref: models.py:182 (F.gelu, exact erf), models.py:173,188 (residual add), models.py:166-167
// (slice), models.py:84,98 (bf16<->fp32 casts), test/ccl.py:110 (mse_loss).
#include "../../include/dlbb.h"
#include "common.cuh"

namespace dlbb {

constexpr int kEwThreads = 256;

static int ew_grid(int64_t work_items) {
    int64_t blocks = (work_items + kEwThreads - 1) / kEwThreads;
    const int64_t cap = 148 * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

enum { OP_GELU_FWD = 0, OP_GELU_BWD = 1, OP_ADD = 2 };

template <int OP>
__device__ __forceinline__ float ew_apply(float a, float b) {
    if (OP == OP_GELU_FWD) return gelu_exact(a);
    if (OP == OP_GELU_BWD) return a * dgelu_exact(b);  // a = dy, b = x
    return a + b;
}

// out = f(a, b) over bf16 arrays; b unused for unary ops.
template <int OP>
__global__ void __launch_bounds__(kEwThreads)
ew_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
               __nv_bfloat16* __restrict__ out, int64_t n, int vec_ok) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t nvec = n >> 3;
        const uint4* a4 = reinterpret_cast<const uint4*>(a);
        const uint4* b4 = reinterpret_cast<const uint4*>(b);
        uint4* o4 = reinterpret_cast<uint4*>(out);
        for (int64_t i = tid; i < nvec; i += nthreads) {
            const uint4 ua = a4[i];
            uint4 ub = make_uint4(0, 0, 0, 0);
            if (OP != OP_GELU_FWD) ub = b4[i];
            const uint32_t aw[4] = {ua.x, ua.y, ua.z, ua.w}, bw[4] = {ub.x, ub.y, ub.z, ub.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 fa = unpack_bf16x2(aw[j]), fb = unpack_bf16x2(bw[j]);
                o[j] = pack_bf16x2(ew_apply<OP>(fa.x, fb.x), ew_apply<OP>(fa.y, fb.y));
            }
            o4[i] = make_uint4(o[0], o[1], o[2], o[3]);
        }
        done = nvec << 3;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) {
        const float fa = __bfloat162float(a[i]);
        const float fb = (OP != OP_GELU_FWD) ? __bfloat162float(b[i]) : 0.f;
        out[i] = __float2bfloat16(ew_apply<OP>(fa, fb));
    }
}

__global__ void __launch_bounds__(kEwThreads)
copy2d_bf16_kernel(const __nv_bfloat16* __restrict__ src, int64_t ld_src,
                   __nv_bfloat16* __restrict__ dst, int64_t ld_dst, int rows, int cols, int vec_ok) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    if (vec_ok) {
        const int cv = cols >> 3;
        const int64_t total = (int64_t)rows * cv;
        for (int64_t i = tid; i < total; i += nthreads) {
            const int64_t r = i / cv;
            const int c = static_cast<int>(i - r * cv);
            reinterpret_cast<uint4*>(dst + r * ld_dst)[c] =
                reinterpret_cast<const uint4*>(src + r * ld_src)[c];
        }
    } else {
        const int64_t total = (int64_t)rows * cols;
        for (int64_t i = tid; i < total; i += nthreads) {
            const int64_t r = i / cols;
            const int c = static_cast<int>(i - r * cols);
            dst[r * ld_dst + c] = src[r * ld_src + c];
        }
    }
}

template <typename T> __device__ __forceinline__ double load_as_double(const T* p, int64_t i);
template <> __device__ __forceinline__ double load_as_double<__half>(const __half* p, int64_t i) { return __half2float(p[i]); }
template <> __device__ __forceinline__ double load_as_double<__nv_bfloat16>(const __nv_bfloat16* p, int64_t i) { return __bfloat162float(p[i]); }
template <> __device__ __forceinline__ double load_as_double<float>(const float* p, int64_t i) { return p[i]; }
template <> __device__ __forceinline__ double load_as_double<double>(const double* p, int64_t i) { return p[i]; }
template <typename T> __device__ __forceinline__ void store_from_double(T* p, int64_t i, double v);
template <> __device__ __forceinline__ void store_from_double<__half>(__half* p, int64_t i, double v) { p[i] = __float2half(static_cast<float>(v)); }
template <> __device__ __forceinline__ void store_from_double<__nv_bfloat16>(__nv_bfloat16* p, int64_t i, double v) { p[i] = __float2bfloat16(static_cast<float>(v)); }
template <> __device__ __forceinline__ void store_from_double<float>(float* p, int64_t i, double v) { p[i] = static_cast<float>(v); }
template <> __device__ __forceinline__ void store_from_double<double>(double* p, int64_t i, double v) { p[i] = v; }

template <typename S, typename D>
__global__ void __launch_bounds__(kEwThreads) cast_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t n) {
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nthreads)
        store_from_double<D>(dst, i, load_as_double<S>(src, i));
}

template <typename S>
static int cast_dispatch_dst(const void* src, void* dst, int dd, int64_t n, cudaStream_t s) {
    const int grid = ew_grid(n);
    const S* sp = static_cast<const S*>(src);
    switch (dd) {
        case DLBB_F16: cast_kernel<S, __half><<<grid, kEwThreads, 0, s>>>(sp, static_cast<__half*>(dst), n); break;
        case DLBB_BF16: cast_kernel<S, __nv_bfloat16><<<grid, kEwThreads, 0, s>>>(sp, static_cast<__nv_bfloat16*>(dst), n); break;
        case DLBB_F32: cast_kernel<S, float><<<grid, kEwThreads, 0, s>>>(sp, static_cast<float*>(dst), n); break;
        case DLBB_F64: cast_kernel<S, double><<<grid, kEwThreads, 0, s>>>(sp, static_cast<double*>(dst), n); break;
        default: set_error("cast: bad destination dtype %d", dd); return -2;
    }
    return 0;
}

// ---------------------------------------------------------------- MSE loss (+ gradient)
// Pass 1: every block writes its partial sum; the last block to finish (ticket counter) folds
// the partials in index order, so the result is deterministic without a second launch.
__global__ void __launch_bounds__(kEwThreads)
mse_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ t,
           __nv_bfloat16* __restrict__ dy, float* __restrict__ loss, float* __restrict__ partials,
           unsigned int* __restrict__ ticket, int64_t n, float inv_n, float gscale, int vec_ok) {
    __shared__ float sh[32];
    __shared__ bool is_last;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    float acc = 0.f;
    int64_t done = 0;
    const float g = 2.f * inv_n * gscale;
    if (vec_ok) {
        const int64_t nvec = n >> 3;
        const uint4* y4 = reinterpret_cast<const uint4*>(y);
        const uint4* t4 = reinterpret_cast<const uint4*>(t);
        uint4* d4 = reinterpret_cast<uint4*>(dy);
        for (int64_t i = tid; i < nvec; i += nthreads) {
            const uint4 uy = y4[i], ut = t4[i];
            const uint32_t yw[4] = {uy.x, uy.y, uy.z, uy.w}, tw[4] = {ut.x, ut.y, ut.z, ut.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 fy = unpack_bf16x2(yw[j]), ft = unpack_bf16x2(tw[j]);
                const float d0 = fy.x - ft.x, d1 = fy.y - ft.y;
                acc += d0 * d0 + d1 * d1;
                o[j] = pack_bf16x2(d0 * g, d1 * g);
            }
            if (dy) d4[i] = make_uint4(o[0], o[1], o[2], o[3]);
        }
        done = nvec << 3;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) {
        const float d = __bfloat162float(y[i]) - __bfloat162float(t[i]);
        acc += d * d;
        if (dy) dy[i] = __float2bfloat16(d * g);
    }
    acc = warp_sum(acc);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = acc;
    __syncthreads();
    if (w == 0) {
        float v = (l < kEwThreads / 32) ? sh[l] : 0.f;
        v = warp_sum(v);
        if (l == 0) {
            partials[blockIdx.x] = v;
            __threadfence();
            const unsigned int prev = atomicAdd(ticket, 1u);
            is_last = (prev == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (is_last && w == 0) {
        __threadfence();
        float v = 0.f;
        for (int i = l; i < (int)gridDim.x; i += 32) v += __ldcg(partials + i);
        // fixed order inside each lane + fixed shuffle tree => run-to-run deterministic
        v = warp_sum(v);
        if (l == 0) {
            loss[0] = v * inv_n;
            *ticket = 0;  // ready for the next call
        }
    }
}

__global__ void __launch_bounds__(kEwThreads) fill_u32_kernel(uint32_t* buf, uint32_t v, int64_t n) {
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nthreads) buf[i] = v;
}

constexpr int kMseMaxBlocks = 1024;

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int OP>
static int launch_ew(const void* a, const void* b, void* out, int64_t n, void* stream, const char* name) {
    DLBB_REQUIRE(n >= 0, "%s: negative length", name);
    if (n == 0) return 0;
    const int vec_ok = al16(a) && al16(b) && al16(out);
    ew_bf16_kernel<OP><<<ew_grid((n + 7) / 8), kEwThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b),
        static_cast<__nv_bfloat16*>(out), n, vec_ok);
    DLBB_CHECK_LAUNCH(name);
    return 0;
}

}  // namespace dlbb

extern "C" {

int dlbb_gelu_fwd(const void* x, void* y, int64_t n, void* stream) {
    return dlbb::launch_ew<dlbb::OP_GELU_FWD>(x, nullptr, y, n, stream, "gelu_fwd");
}
int dlbb_gelu_bwd(const void* dy, const void* x, void* dx, int64_t n, void* stream) {
    return dlbb::launch_ew<dlbb::OP_GELU_BWD>(dy, x, dx, n, stream, "gelu_bwd");
}
int dlbb_add(const void* a, const void* b, void* out, int64_t n, void* stream) {
    return dlbb::launch_ew<dlbb::OP_ADD>(a, b, out, n, stream, "add");
}

int dlbb_copy2d(const void* src, int64_t ld_src, void* dst, int64_t ld_dst, int rows, int cols,
                void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(rows >= 0 && cols >= 0, "copy2d: negative extent");
    if (rows == 0 || cols == 0) return 0;
    const int vec_ok = al16(src) && al16(dst) && cols % 8 == 0 && ld_src % 8 == 0 && ld_dst % 8 == 0;
    const int64_t items = vec_ok ? (int64_t)rows * (cols / 8) : (int64_t)rows * cols;
    copy2d_bf16_kernel<<<ew_grid(items), kEwThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(src), ld_src, static_cast<__nv_bfloat16*>(dst), ld_dst, rows,
        cols, vec_ok);
    DLBB_CHECK_LAUNCH("copy2d");
    return 0;
}

int dlbb_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, void* stream) {
    using namespace dlbb;
    if (n == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc;
    switch (src_dtype) {
        case DLBB_F16: rc = cast_dispatch_dst<__half>(src, dst, dst_dtype, n, s); break;
        case DLBB_BF16: rc = cast_dispatch_dst<__nv_bfloat16>(src, dst, dst_dtype, n, s); break;
        case DLBB_F32: rc = cast_dispatch_dst<float>(src, dst, dst_dtype, n, s); break;
        case DLBB_F64: rc = cast_dispatch_dst<double>(src, dst, dst_dtype, n, s); break;
        default: set_error("cast: bad source dtype %d", src_dtype); return -2;
    }
    if (rc) return rc;
    DLBB_CHECK_LAUNCH("cast");
    return 0;
}

size_t dlbb_mse_scratch_bytes(int64_t n) {
    (void)n;
    return (dlbb::kMseMaxBlocks + 4) * sizeof(float);
}

int dlbb_mse_loss(const void* y, const void* target, void* dy, float* loss, void* scratch,
                  int64_t n, float grad_scale, void* stream) {
    using namespace dlbb;
    DLBB_REQUIRE(n > 0, "mse_loss: empty input");
    DLBB_REQUIRE(scratch != nullptr && loss != nullptr, "mse_loss: scratch and loss are required");
    int grid = ew_grid((n + 7) / 8);
    if (grid > kMseMaxBlocks) grid = kMseMaxBlocks;
    float* partials = static_cast<float*>(scratch);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(partials + kMseMaxBlocks);
    const int vec_ok = al16(y) && al16(target) && al16(dy);
    mse_kernel<<<grid, kEwThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(target),
        static_cast<__nv_bfloat16*>(dy), loss, partials, ticket, n, 1.0f / static_cast<float>(n),
        grad_scale, vec_ok);
    DLBB_CHECK_LAUNCH("mse_loss");
    return 0;
}

int dlbb_fill_u32(void* buf, uint32_t value, int64_t n_words, void* stream) {
    using namespace dlbb;
    if (n_words <= 0) return 0;
    fill_u32_kernel<<<ew_grid(n_words), kEwThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<uint32_t*>(buf), value, n_words);
    DLBB_CHECK_LAUNCH("fill_u32");
    return 0;
}
}
