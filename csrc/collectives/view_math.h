// Index arithmetic of strided views (K12), shared by the collective kernels (collectives.cu) and a host-side unit test
// (tests/view_math_test.cpp, run by tests/test_view_math_cpu.py): logical element / byte index -> offset inside a view of
// up to four dimensions, and the conditions under which 16-byte vectors may be moved with one index decomposition each.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define DLBB_HD __host__ __device__ __forceinline__
#else
#define DLBB_HD inline
#endif

namespace dlbb {

struct ViewDev {
    int ndim;            // 0 = contiguous
    int64_t size[4];
    int64_t stride[4];
    int64_t base;        // logical element offset added to every index (a launch that covers a later part of the view)
};


DLBB_HD int64_t view_offset(const ViewDev& v, int64_t idx) {
    if (v.ndim == 0) return idx;
    idx += v.base;
    int64_t off = 0;
#pragma unroll
    for (int d = 3; d >= 0; --d) {
        if (d < v.ndim) {
            const int64_t q = idx / v.size[d];
            off += (idx - q * v.size[d]) * v.stride[d];
            idx = q;
        }
    }
    return off;
}

// A strided destination (K12, unpack side): can 16-byte vectors of T be stored with one index decomposition each?
// (innermost dimension dense and a whole number of vectors, every other stride and the base pointer vector aligned)
template <typename T>
DLBB_HD bool view_vec_ok(const ViewDev& v, const void* base) {
    constexpr int VE = 16 / sizeof(T);
    if (v.ndim == 0) return (reinterpret_cast<uintptr_t>(base) & 15) == 0;
    const int last = v.ndim - 1;
    bool ok = v.stride[last] == 1 && v.size[last] % VE == 0 && v.base % VE == 0 && (reinterpret_cast<uintptr_t>(base) & 15) == 0;
    for (int d = 0; d < last; ++d) ok = ok && (v.stride[d] % VE == 0);
    return ok;
}
DLBB_HD int64_t view_byte_off(const ViewDev& v, int es, int64_t b) {   // logical byte -> byte in the view
    if (v.ndim == 0) return b;
    const int64_t e = b / es;
    return view_offset(v, e) * es + (b - e * es);
}
DLBB_HD bool view_vec_ok_bytes(const ViewDev& v, int es, const void* base) {
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return false;
    if (v.ndim == 0) return true;
    if (es > 16 || 16 % es != 0) return false;
    const int VE = 16 / es, last = v.ndim - 1;
    bool ok = v.stride[last] == 1 && v.size[last] % VE == 0 && v.base % VE == 0;
    for (int d = 0; d < last; ++d) ok = ok && (v.stride[d] % VE == 0);
    return ok;
}

}  // namespace dlbb
