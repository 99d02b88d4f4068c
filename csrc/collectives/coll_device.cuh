// Device-side building blocks of the collectives: system-scope flag protocol on the signal pads,
// coherent peer loads, NVLS multimem instructions, typed reduction operators.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "../bootstrap/comm_internal.h"
#include "../kernels/common.cuh"

namespace dlbb {

#ifndef DLBB_WATCHDOG
#define DLBB_WATCHDOG 1
#endif

#if defined(DLBB_HOST_EMU)
}  // namespace dlbb
#include "emu/cuda_emu.h"
namespace dlbb {
#include "emu/device_emu.h"   // host versions of the PTX helpers (flags, peer accesses, multimem)
#else
__device__ __forceinline__ void group_bar_sync(uint32_t id, uint32_t nthreads) {   // named barrier of a warp group
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ flags
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Wait until *flag has reached `epoch` (monotonic, wrap-safe).  A peer that never arrives traps
// after `limit_ns` (CommDev::watchdog_ns: 60 s unless DLBB_WATCHDOG_SECONDS says otherwise) instead of
// hanging the GPU (watchdog on device-side flag waits).  The message names the missing rank.
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t epoch, int peer, long long limit_ns) {
#if DLBB_WATCHDOG
    long long t0 = 0;
    uint32_t it = 0;
#endif
    while (static_cast<int32_t>(ld_acquire_sys(flag) - epoch) < 0) {
#if DLBB_WATCHDOG
        if ((++it & 0xfff) == 0) {
            long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > limit_ns) {
                printf("dlbb: device barrier timeout: block %d waited %lld ms for rank %d (epoch %u, flag %u)\n",
                       blockIdx.x, limit_ns / 1000000, peer, epoch, *flag);
                __trap();
            }
        }
#endif
    }
}

#endif  // DLBB_HOST_EMU

__device__ __forceinline__ uint32_t* pad_slot(const CommDev& c, int owner, int bank, int row, int col) {
    return c.pads[owner] + bank * kPadWordsPerBank + row * kPadStride + col;
}

// Row barrier: CTA `row` of every rank meets CTA `row` of every other rank.  All earlier global
// writes of this CTA (to local or peer memory) are visible to the peers' CTA after it returns.
__device__ __forceinline__ void row_barrier(const CommDev& c, int bank, int row, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x < c.world) {
        const int p = threadIdx.x;
        st_release_sys(pad_slot(c, p, bank, row, c.rank), epoch);
        wait_flag(pad_slot(c, c.rank, bank, row, p), epoch, p, c.watchdog_ns);
    }
    __syncthreads();
}

// Grid barrier: every CTA of every rank.  CTA b announces itself on row b; everyone waits for all
// `nrows` rows of all ranks (needed when a CTA reads peer data that other CTAs produced).
__device__ __forceinline__ void grid_barrier(const CommDev& c, int bank, int nrows, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x < c.world) st_release_sys(pad_slot(c, threadIdx.x, bank, blockIdx.x, c.rank), epoch);
    for (int i = threadIdx.x; i < nrows * c.world; i += blockDim.x) {
        const int row = i / c.world, p = i - row * c.world;
        wait_flag(pad_slot(c, c.rank, bank, row, p), epoch, p, c.watchdog_ns);
    }
    __syncthreads();
}

#if !defined(DLBB_HOST_EMU)
// ------------------------------------------------------------------ coherent 16-byte accesses
// Peer data produced during this kernel must not come from a stale L1 line: strong (relaxed,
// system-scope) loads always go to the owner's L2.
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
template <typename T>
__device__ __forceinline__ T ld_peer_scalar(const T* p) {
    return *reinterpret_cast<const volatile T*>(p);
}
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {  // local, read-once
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void st_v4(void* p, const uint4& v) {
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// ------------------------------------------------------------------ NVLS (multimem)
// In-switch reduction of one 16-byte vector across every rank bound to the multicast object.
__device__ __forceinline__ uint4 mm_ld_reduce_add_bf16(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ uint4 mm_ld_reduce_add_f16(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ uint4 mm_ld_reduce_add_f32(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
// Broadcast store of one 16-byte vector into every rank's copy.
__device__ __forceinline__ void mm_st_v4(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

#endif  // !DLBB_HOST_EMU

// ------------------------------------------------------------------ typed reduction
template <typename T> struct Acc { using type = float; };
template <> struct Acc<double> { using type = double; };
template <> struct Acc<int32_t> { using type = int32_t; };
template <> struct Acc<int64_t> { using type = int64_t; };

template <typename T> __device__ __forceinline__ typename Acc<T>::type to_acc(T v);
template <> __device__ __forceinline__ float to_acc<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_acc<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_acc<float>(float v) { return v; }
template <> __device__ __forceinline__ double to_acc<double>(double v) { return v; }
template <> __device__ __forceinline__ int32_t to_acc<int32_t>(int32_t v) { return v; }
template <> __device__ __forceinline__ int64_t to_acc<int64_t>(int64_t v) { return v; }
template <typename T> __device__ __forceinline__ T from_acc(typename Acc<T>::type v);
template <> __device__ __forceinline__ __half from_acc<__half>(float v) { return __float2half(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16>(float v) { return __float2bfloat16(v); }
template <> __device__ __forceinline__ float from_acc<float>(float v) { return v; }
template <> __device__ __forceinline__ double from_acc<double>(double v) { return v; }
template <> __device__ __forceinline__ int32_t from_acc<int32_t>(int32_t v) { return v; }
template <> __device__ __forceinline__ int64_t from_acc<int64_t>(int64_t v) { return v; }

template <int OP, typename A>
__device__ __forceinline__ A red_op(A a, A b) {
    if (OP == DLBB_SUM) return a + b;
    if (OP == DLBB_MAX) return a > b ? a : b;
    if (OP == DLBB_MIN) return a < b ? a : b;
    return a * b;
}

}  // namespace dlbb
