// Every collective entry point of collectives.cu, run on the host emulation (../emu): P ranks as host threads, every
// CUDA thread a host thread, flags as release / acquire atomics - built plain, with ThreadSanitizer (races between data
// accesses that the flag protocol does not order) or AddressSanitizer (out-of-bounds staging / view arithmetic).
// World sizes that no box offers (3, 5, 6, 7) run here like 2, 4 and 8.  Each case repeats the call several times back
// to back WITHOUT a host barrier in between and refills its input right after each call, so a protocol that lets a rank
// overwrite staging / tensors that a slower peer is still reading fails here (wrong numbers, or a TSan report).
//
//   emu_collectives_test <world> [mc|nomc] [case-filter]
// Environment: the DLBB_* knobs of the library (thresholds are lowered by tests/test_emu_collectives_cpu.py so that the
// large-message paths run on kilobytes), EMU_MAX_BLOCK, EMU_CTAS, EMU_STAGE_KB, EMU_ZC_MIN.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../../include/dlbb.h"
#include "../emu/emu_world.h"

using emu::World;

static std::atomic<int> g_fail{0};
static std::atomic<long> g_checks{0};
static const char* g_case = "";

#define FAIL(...)                                                   \
    do {                                                            \
        if (g_fail.fetch_add(1) < 25) {                             \
            fprintf(stderr, "FAIL [%s] ", g_case);                  \
            fprintf(stderr, __VA_ARGS__);                           \
            fprintf(stderr, "\n");                                  \
        }                                                           \
    } while (0)
#define CALL(expr)                                                                   \
    do {                                                                             \
        const int _rc = (expr);                                                      \
        if (_rc != 0) FAIL("%s -> %d (%s)", #expr, _rc, emu::last_error().c_str());  \
    } while (0)

// ------------------------------------------------------------------ typed helpers
template <typename T> struct DT;
template <> struct DT<float> { static constexpr int code = DLBB_F32; static constexpr const char* name = "f32"; };
template <> struct DT<double> { static constexpr int code = DLBB_F64; static constexpr const char* name = "f64"; };
template <> struct DT<int32_t> { static constexpr int code = DLBB_I32; static constexpr const char* name = "i32"; };
template <> struct DT<int64_t> { static constexpr int code = DLBB_I64; static constexpr const char* name = "i64"; };
template <> struct DT<__nv_bfloat16> { static constexpr int code = DLBB_BF16; static constexpr const char* name = "bf16"; };
template <> struct DT<__half> { static constexpr int code = DLBB_F16; static constexpr const char* name = "f16"; };

template <typename T> inline T from_d(double v) { return static_cast<T>(v); }
template <> inline __nv_bfloat16 from_d<__nv_bfloat16>(double v) { return __float2bfloat16(static_cast<float>(v)); }
template <> inline __half from_d<__half>(double v) { return __float2half(static_cast<float>(v)); }
template <typename T> inline double to_d(T v) { return static_cast<double>(v); }
template <> inline double to_d<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> inline double to_d<__half>(__half v) { return __half2float(v); }

// small integers: sums / products over <= 8 ranks stay exact in every dtype (bf16 has 8 bits of mantissa)
static inline double val(int rank, int64_t i, int rep) { return static_cast<double>(((i * 7 + rank * 3 + rep * 5) % 5) - 2); }

static double reduce_expected(int P, int64_t i, int rep, int op) {
    double a = val(0, i, rep);
    for (int r = 1; r < P; ++r) {
        const double x = val(r, i, rep);
        a = op == DLBB_SUM ? a + x : op == DLBB_MAX ? std::max(a, x) : op == DLBB_MIN ? std::min(a, x) : a * x;
    }
    return a;
}
static const char* op_name(int op) { return op == DLBB_SUM ? "sum" : op == DLBB_MAX ? "max" : op == DLBB_MIN ? "min" : "prod"; }

// per-rank buffers, plain ("cudaMalloc") or in one symmetric segment
struct Bufs {
    char* p[16] = {};
    bool symmetric = false;
    Bufs(World& w, size_t bytes, bool symm) : symmetric(symm) {
        if (symm) {
            emu::SymmSeg& s = w.alloc_symm((bytes + 255) / 256 * 256 + 256);
            for (int r = 0; r < w.P; ++r) p[r] = s.base[r];
        } else {
            for (int r = 0; r < w.P; ++r) p[r] = emu::alloc_zero(bytes + 256);
        }
    }
    template <typename T> T* at(int r, int64_t elem_off = 0) const { return reinterpret_cast<T*>(p[r]) + elem_off; }
};

template <typename T>
static void fill(T* dst, int rank, int64_t n, int rep) {
    for (int64_t i = 0; i < n; ++i) dst[i] = from_d<T>(val(rank, i, rep));
}

static constexpr int kReps = 4;

// ------------------------------------------------------------------ all-reduce
template <typename T>
static void case_allreduce(World& w, int64_t n, int op, int algo, bool symm, bool inplace, int64_t misalign = 0) {
    char name[160];
    snprintf(name, sizeof(name), "allreduce %s %s n=%lld algo=%d %s %s off=%lld", DT<T>::name, op_name(op), (long long)n, algo,
             symm ? "symm" : "plain", inplace ? "inplace" : "outofplace", (long long)misalign);
    g_case = name;
    Bufs in(w, (n + misalign) * sizeof(T), symm), out(w, (n + misalign) * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            T* src = in.at<T>(r, misalign);
            T* dst = inplace ? src : out.at<T>(r, misalign);
            fill(src, r, n, rep);
            CALL(dlbb_allreduce(c, src, dst, n, DT<T>::code, op, algo, nullptr, nullptr));
            for (int64_t i = 0; i < n; ++i) {
                const double e = reduce_expected(w.P, i, rep, op), g = to_d(dst[i]);
                if (e != g) { FAIL("rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)i, g, e); break; }
            }
            g_checks.fetch_add(1);
        }
    });
}

// ------------------------------------------------------------------ driver
static bool want(const char* filter, const char* group) { return filter == nullptr || strstr(group, filter) != nullptr; }

int main(int argc, char** argv) {
    const int P = argc > 1 ? atoi(argv[1]) : 2;
    const bool mc = argc > 2 ? strcmp(argv[2], "mc") == 0 : true;
    const char* filter = argc > 3 ? argv[3] : nullptr;
    const int ctas = getenv("EMU_CTAS") ? atoi(getenv("EMU_CTAS")) : 3;
    const size_t stage = (getenv("EMU_STAGE_KB") ? atol(getenv("EMU_STAGE_KB")) : 64) << 10;
    const size_t zc_min = getenv("EMU_ZC_MIN") ? atol(getenv("EMU_ZC_MIN")) : 8192;
    World w(P, stage, mc && P > 1, ctas, P > 1 ? zc_min : 0);
    emu::world() = &w;

    const int64_t sizes[] = {1, 7, 300, 4099, 20000, 70001};
    if (want(filter, "allreduce")) {
        for (int64_t n : sizes) {
            for (int algo : {DLBB_ALGO_AUTO, DLBB_ALGO_ONESHOT, DLBB_ALGO_TWOSHOT, DLBB_ALGO_NVLS}) {
                case_allreduce<float>(w, n, DLBB_SUM, algo, false, true);
                case_allreduce<__nv_bfloat16>(w, n, DLBB_SUM, algo, false, false);
            }
            case_allreduce<float>(w, n, DLBB_SUM, DLBB_ALGO_AUTO, true, true);
            case_allreduce<__nv_bfloat16>(w, n, DLBB_SUM, DLBB_ALGO_AUTO, true, true);
            case_allreduce<__half>(w, n, DLBB_SUM, DLBB_ALGO_AUTO, true, false);
            case_allreduce<int32_t>(w, n, DLBB_MAX, DLBB_ALGO_AUTO, false, true);
            case_allreduce<int64_t>(w, n, DLBB_MIN, DLBB_ALGO_AUTO, true, true);
            case_allreduce<double>(w, n, DLBB_PROD, DLBB_ALGO_AUTO, false, true);
            case_allreduce<float>(w, n, DLBB_MAX, DLBB_ALGO_AUTO, true, true);
            case_allreduce<float>(w, n, DLBB_SUM, DLBB_ALGO_DIRECT, false, true);
            case_allreduce<float>(w, n, DLBB_SUM, DLBB_ALGO_AUTO, false, true, 1);   // 4-byte aligned only
        }
    }
    if (g_fail.load()) {
        printf("EMU COLLECTIVES FAILED: %d failures (world %d, %s)\n", g_fail.load(), P, mc ? "multicast" : "no multicast");
        return 1;
    }
    printf("EMU COLLECTIVES OK: world %d, %s, %ld checks, %ld kernel launches emulated\n", P, mc ? "multicast" : "no multicast",
           g_checks.load(), emu::launches().load());
    return 0;
}
