// Host versions of the PTX helpers of coll_device.cuh (included by it, inside namespace dlbb, when DLBB_HOST_EMU is
// defined).  See cuda_emu.h for the model.
#pragma once

inline void group_bar_sync(uint32_t id, uint32_t nthreads) { emu::tl_cta->bar[id & 15].sync(nthreads); }

// ------------------------------------------------------------------ flags
// (EMU_WEAK_FLAGS: negative control - flags without release / acquire semantics; the sanitizer must then report the data
//  accesses that only the flags were ordering)
#if defined(EMU_WEAK_FLAGS)
inline void st_release_sys(uint32_t* p, uint32_t v) { __atomic_store_n(p, v, __ATOMIC_RELAXED); }
inline uint32_t ld_acquire_sys(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
#else
inline void st_release_sys(uint32_t* p, uint32_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
inline uint32_t ld_acquire_sys(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
#endif

inline void wait_flag(const uint32_t* flag, uint32_t epoch, int peer, long long limit_ns) {
    long long t0 = 0;
    uint32_t it = 0;
    while (static_cast<int32_t>(ld_acquire_sys(flag) - epoch) < 0) {
        sched_yield();
        if ((++it & 0x3ff) == 0) {
            usleep(50);
            const long long now = emu::now_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > limit_ns) {
                fprintf(stderr, "dlbb: device barrier timeout: block %d waited %lld ms for rank %d (epoch %u, flag %u)\n",
                        blockIdx.x, limit_ns / 1000000, peer, epoch, __atomic_load_n(flag, __ATOMIC_RELAXED));
                abort();
            }
        }
    }
}

// ------------------------------------------------------------------ 16-byte accesses (plain: ordering must come from the flags)
inline uint4 ld_peer_v4(const void* p) {
    uint4 v;
    memcpy(&v, p, 16);
    return v;
}
template <typename T>
inline T ld_peer_scalar(const T* p) {
    T v;
    memcpy(&v, p, sizeof(T));
    return v;
}
inline uint4 ld_stream_v4(const void* p) { return ld_peer_v4(p); }
inline void st_v4(void* p, const uint4& v) {
    if (reinterpret_cast<uintptr_t>(p) & 15) {
        fprintf(stderr, "emu: misaligned 16-byte store to %p\n", p);
        abort();
    }
    memcpy(p, &v, 16);
}

// ------------------------------------------------------------------ multimem
namespace emu_detail {
inline void check_aligned(const void* p) {
    if (reinterpret_cast<uintptr_t>(p) & 15) {
        fprintf(stderr, "emu: misaligned multimem access at %p\n", p);
        abort();
    }
}
}  // namespace emu_detail

inline uint4 mm_ld_reduce_add_bf16(const void* mc) {
    emu_detail::check_aligned(mc);
    const emu::McWindow& w = emu::window_of(mc);
    const size_t off = static_cast<const char*>(mc) - w.base;
    float acc[8] = {};
    for (int r = 0; r < w.world; ++r) {
        __nv_bfloat16 x[8];
        memcpy(x, w.rank_base[r] + off, 16);
        for (int e = 0; e < 8; ++e) acc[e] += __bfloat162float(x[e]);
    }
    __nv_bfloat16 o[8];
    for (int e = 0; e < 8; ++e) o[e] = __float2bfloat16(acc[e]);
    uint4 v;
    memcpy(&v, o, 16);
    return v;
}
inline uint4 mm_ld_reduce_add_f16(const void* mc) {
    emu_detail::check_aligned(mc);
    const emu::McWindow& w = emu::window_of(mc);
    const size_t off = static_cast<const char*>(mc) - w.base;
    float acc[8] = {};
    for (int r = 0; r < w.world; ++r) {
        __half x[8];
        memcpy(x, w.rank_base[r] + off, 16);
        for (int e = 0; e < 8; ++e) acc[e] += __half2float(x[e]);
    }
    __half o[8];
    for (int e = 0; e < 8; ++e) o[e] = __float2half(acc[e]);
    uint4 v;
    memcpy(&v, o, 16);
    return v;
}
inline uint4 mm_ld_reduce_add_f32(const void* mc) {
    emu_detail::check_aligned(mc);
    const emu::McWindow& w = emu::window_of(mc);
    const size_t off = static_cast<const char*>(mc) - w.base;
    float acc[4] = {};
    for (int r = 0; r < w.world; ++r) {
        float x[4];
        memcpy(x, w.rank_base[r] + off, 16);
        for (int e = 0; e < 4; ++e) acc[e] += x[e];
    }
    uint4 v;
    memcpy(&v, acc, 16);
    return v;
}
inline void mm_st_v4(void* mc, const uint4& v) {
    emu_detail::check_aligned(mc);
    const emu::McWindow& w = emu::window_of(mc);
    const size_t off = static_cast<char*>(mc) - w.base;
    for (int r = 0; r < w.world; ++r) memcpy(w.rank_base[r] + off, &v, 16);
}
