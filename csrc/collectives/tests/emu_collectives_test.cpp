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
#include <memory>
#include <string>
#include <vector>

#include "../../../include/dlbb.h"
#include "../emu/emu_world.h"
#include "../gemm_allreduce_fusion.cuh"

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
            for (int r = 0; r < w.P; ++r) p[r] = emu::alloc_zero(bytes);   // exact: AddressSanitizer sees a one-byte overrun
        }
    }
    template <typename T> T* at(int r, int64_t elem_off = 0) const { return reinterpret_cast<T*>(p[r]) + elem_off; }
};

template <typename T>
static void fill(T* dst, int rank, int64_t n, int rep) {
    for (int64_t i = 0; i < n; ++i) dst[i] = from_d<T>(val(rank, i, rep));
}

static const int kReps = getenv("EMU_REPS") ? atoi(getenv("EMU_REPS")) : 4;   // back-to-back repetitions of every call

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

// data-movement payloads: depend on the source rank, the index and the repetition; exact in bf16 (|v| <= 100)
static inline double mval(int rank, int64_t i, int rep) { return static_cast<double>(((i * 7 + rank * 31 + rep * 3) % 201) - 100); }
template <typename T>
static void mfill(T* dst, int rank, int64_t n, int rep, int64_t i0 = 0) {
    for (int64_t i = 0; i < n; ++i) dst[i] = from_d<T>(mval(rank, i0 + i, rep));
}
template <typename T>
static bool expect_eq(const T* got, int64_t n, int src_rank, int rep, int64_t i0, int rank, const char* what) {
    for (int64_t i = 0; i < n; ++i) {
        const double e = mval(src_rank, i0 + i, rep), g = to_d(got[i]);
        if (e != g) {
            FAIL("%s: rank %d rep %d elem %lld (from rank %d): got %g expected %g", what, rank, rep, (long long)i, src_rank, g, e);
            return false;
        }
    }
    return true;
}

// ------------------------------------------------------------------ reduce-scatter / reduce
template <typename T>
static void case_reduce_scatter(World& w, int64_t cpr, int op, bool symm) {
    char name[160];
    snprintf(name, sizeof(name), "reduce_scatter %s %s cpr=%lld %s", DT<T>::name, op_name(op), (long long)cpr, symm ? "symm" : "plain");
    g_case = name;
    Bufs in(w, cpr * w.P * sizeof(T), symm), out(w, cpr * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            fill(in.at<T>(r), r, cpr * w.P, rep);
            CALL(dlbb_reduce_scatter(c, in.at<T>(r), out.at<T>(r), cpr, DT<T>::code, op, nullptr));
            for (int64_t i = 0; i < cpr; ++i) {
                const double e = reduce_expected(w.P, r * cpr + i, rep, op), g = to_d(out.at<T>(r)[i]);
                if (e != g) { FAIL("rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)i, g, e); break; }
            }
            g_checks.fetch_add(1);
        }
    });
}

template <typename T>
static void case_reduce(World& w, int64_t n, int op, int root, bool symm, bool inplace) {
    char name[160];
    snprintf(name, sizeof(name), "reduce %s %s n=%lld root=%d %s %s", DT<T>::name, op_name(op), (long long)n, root,
             symm ? "symm" : "plain", inplace ? "inplace" : "outofplace");
    g_case = name;
    Bufs in(w, n * sizeof(T), symm), out(w, n * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            T* src = in.at<T>(r);
            T* dst = inplace ? src : out.at<T>(r);
            fill(src, r, n, rep);
            CALL(dlbb_reduce(c, src, dst, n, DT<T>::code, op, root, nullptr));
            if (r == root)
                for (int64_t i = 0; i < n; ++i) {
                    const double e = reduce_expected(w.P, i, rep, op), g = to_d(dst[i]);
                    if (e != g) { FAIL("root rep %d elem %lld: got %g expected %g", rep, (long long)i, g, e); break; }
                }
            g_checks.fetch_add(1);
        }
    });
}

// ------------------------------------------------------------------ all-gather / gather / scatter / broadcast / all-to-all
template <typename T>
static void case_allgather(World& w, int64_t cpr, bool symm) {
    char name[160];
    snprintf(name, sizeof(name), "allgather %s cpr=%lld %s", DT<T>::name, (long long)cpr, symm ? "symm" : "plain");
    g_case = name;
    Bufs in(w, cpr * sizeof(T), symm), out(w, cpr * w.P * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            mfill(in.at<T>(r), r, cpr, rep);
            CALL(dlbb_allgather(c, in.at<T>(r), out.at<T>(r), cpr, DT<T>::code, nullptr, nullptr));
            for (int p = 0; p < w.P; ++p)
                if (!expect_eq(out.at<T>(r) + p * cpr, cpr, p, rep, 0, r, "allgather")) break;
            g_checks.fetch_add(1);
        }
    });
}

template <typename T>
static void case_gather_scatter(World& w, int64_t n, int root, bool symm) {
    char name[160];
    snprintf(name, sizeof(name), "gather+scatter %s n=%lld root=%d %s", DT<T>::name, (long long)n, root, symm ? "symm" : "plain");
    g_case = name;
    Bufs small(w, n * sizeof(T), symm), big(w, n * w.P * sizeof(T), symm), back(w, n * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            mfill(small.at<T>(r), r, n, rep);
            CALL(dlbb_gather(c, small.at<T>(r), big.at<T>(r), n, DT<T>::code, root, nullptr));
            if (r == root)
                for (int p = 0; p < w.P; ++p)
                    if (!expect_eq(big.at<T>(r) + p * n, n, p, rep, 0, r, "gather")) break;
            // ... and back: the root scatters what it gathered
            CALL(dlbb_scatter(c, big.at<T>(r), back.at<T>(r), n, DT<T>::code, root, nullptr));
            expect_eq(back.at<T>(r), n, r, rep, 0, r, "scatter");
            g_checks.fetch_add(2);
        }
    });
}

template <typename T>
static void case_broadcast(World& w, int64_t n, int root, bool symm) {
    char name[160];
    snprintf(name, sizeof(name), "broadcast %s n=%lld root=%d %s", DT<T>::name, (long long)n, root, symm ? "symm" : "plain");
    g_case = name;
    Bufs buf(w, n * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            if (r == root) mfill(buf.at<T>(r), root, n, rep);
            else memset(buf.at<T>(r), 0x5a, n * sizeof(T));
            CALL(dlbb_broadcast(c, buf.at<T>(r), n, DT<T>::code, root, nullptr));
            expect_eq(buf.at<T>(r), n, root, rep, 0, r, "broadcast");
            g_checks.fetch_add(1);
        }
    });
}

template <typename T>
static void case_alltoall(World& w, int64_t cpr, bool symm) {
    char name[160];
    snprintf(name, sizeof(name), "alltoall %s cpr=%lld %s", DT<T>::name, (long long)cpr, symm ? "symm" : "plain");
    g_case = name;
    Bufs in(w, cpr * w.P * sizeof(T), symm), out(w, cpr * w.P * sizeof(T), symm);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            // chunk p of my input: payload "from r" at logical offset p * cpr
            mfill(in.at<T>(r), r, cpr * w.P, rep);
            CALL(dlbb_alltoall(c, in.at<T>(r), out.at<T>(r), cpr, DT<T>::code, nullptr));
            for (int p = 0; p < w.P; ++p)
                if (!expect_eq(out.at<T>(r) + p * cpr, cpr, p, rep, r * cpr, r, "alltoall")) break;
            g_checks.fetch_add(1);
        }
    });
}

// ------------------------------------------------------------------ point-to-point, fused residual, barrier
template <typename T>
static void case_ring_sendrecv(World& w, int64_t n) {
    char name[160];
    snprintf(name, sizeof(name), "sendrecv ring %s n=%lld", DT<T>::name, (long long)n);
    g_case = name;
    if (w.P < 2) return;
    Bufs snd(w, n * sizeof(T), false), rcv(w, n * sizeof(T), false);
    w.run([&](int r, dlbb_comm* c) {
        const int next = (r + 1) % w.P, prev = (r + w.P - 1) % w.P;
        for (int rep = 0; rep < kReps; ++rep) {
            mfill(snd.at<T>(r), r, n, rep);
            if (w.P == 2 && r == 1) {   // two ranks: one send and one receive each, in matching order
                CALL(dlbb_sendrecv(c, nullptr, 0, -1, rcv.at<T>(r), n, prev, DT<T>::code, nullptr));
                CALL(dlbb_sendrecv(c, snd.at<T>(r), n, next, nullptr, 0, -1, DT<T>::code, nullptr));
            } else if (w.P == 2) {
                CALL(dlbb_sendrecv(c, snd.at<T>(r), n, next, nullptr, 0, -1, DT<T>::code, nullptr));
                CALL(dlbb_sendrecv(c, nullptr, 0, -1, rcv.at<T>(r), n, prev, DT<T>::code, nullptr));
            } else {
                CALL(dlbb_sendrecv(c, snd.at<T>(r), n, next, rcv.at<T>(r), n, prev, DT<T>::code, nullptr));
            }
            expect_eq(rcv.at<T>(r), n, prev, rep, 0, r, "sendrecv");
            g_checks.fetch_add(1);
        }
    });
}

static void case_residual(World& w, int64_t n) {
    char name[160];
    snprintf(name, sizeof(name), "allreduce_residual bf16 n=%lld", (long long)n);
    g_case = name;
    using T = __nv_bfloat16;
    Bufs in(w, n * sizeof(T), false), res(w, n * sizeof(T), false), out(w, n * sizeof(T), false);
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            fill(in.at<T>(r), r, n, rep);
            for (int64_t i = 0; i < n; ++i) res.at<T>(r)[i] = from_d<T>(static_cast<double>((i + rep) % 3));
            CALL(dlbb_allreduce_residual_bf16(c, in.at<T>(r), res.at<T>(r), out.at<T>(r), n, nullptr));
            for (int64_t i = 0; i < n; ++i) {
                const double e = reduce_expected(w.P, i, rep, DLBB_SUM) + static_cast<double>((i + rep) % 3), g = to_d(out.at<T>(r)[i]);
                if (e != g) { FAIL("rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)i, g, e); break; }
            }
            g_checks.fetch_add(1);
        }
    });
}

static void case_barrier(World& w) {
    g_case = "barrier";
    // a barrier orders plain host-visible writes: every rank publishes a counter, barrier, reads all the others
    static int slots[16];
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 1; rep <= kReps; ++rep) {
            __atomic_store_n(&slots[r], rep, __ATOMIC_RELAXED);
            CALL(dlbb_barrier(c, nullptr));
            for (int p = 0; p < w.P; ++p)
                if (__atomic_load_n(&slots[p], __ATOMIC_RELAXED) < rep) FAIL("rank %d passed barrier %d before rank %d arrived", r, rep, p);
            CALL(dlbb_barrier(c, nullptr));
            g_checks.fetch_add(1);
        }
    });
}

// ------------------------------------------------------------------ lists of tensors (K13 and the list collectives)
template <typename T>
static void case_multi(World& w, int tensors, int64_t n_each, int op) {
    char name[160];
    snprintf(name, sizeof(name), "allreduce_multi %s %s %d x %lld", DT<T>::name, op_name(op), tensors, (long long)n_each);
    g_case = name;
    std::vector<Bufs> bufs;
    std::vector<int64_t> counts;
    for (int t = 0; t < tensors; ++t) {
        counts.push_back(n_each + (t % 3) * 5 + (t == 1 ? 0 : t));   // ragged sizes, some not a multiple of 16 bytes
        bufs.emplace_back(w, (counts.back() + 1) * sizeof(T), false);   // (+1: every third tensor starts one element in)
    }
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            std::vector<void*> ptrs;
            for (int t = 0; t < tensors; ++t) {
                T* b = bufs[t].at<T>(r, t % 3 == 1 ? 1 : 0);   // every third tensor is not 16-byte aligned
                for (int64_t i = 0; i < counts[t]; ++i) b[i] = from_d<T>(val(r, i + 11 * t, rep));
                ptrs.push_back(b);
            }
            CALL(dlbb_allreduce_multi(c, ptrs.data(), counts.data(), tensors, DT<T>::code, op, nullptr));
            for (int t = 0; t < tensors; ++t)
                for (int64_t i = 0; i < counts[t]; ++i) {
                    const double e = reduce_expected(w.P, i + 11 * t, rep, op), g = to_d(bufs[t].at<T>(r, t % 3 == 1 ? 1 : 0)[i]);
                    if (e != g) { FAIL("rank %d rep %d tensor %d elem %lld: got %g expected %g", r, rep, t, (long long)i, g, e); t = tensors; break; }
                }
            g_checks.fetch_add(1);
        }
    });
}

template <typename T>
static void case_ptrs(World& w, int64_t n, int root) {
    char name[160];
    snprintf(name, sizeof(name), "allgather_ptrs/alltoall_ptrs %s n=%lld root=%d", DT<T>::name, (long long)n, root);
    g_case = name;
    Bufs in(w, n * sizeof(T), false), dense_in(w, n * w.P * sizeof(T), false);
    std::vector<Bufs> outs, a2a_in, a2a_out;
    for (int p = 0; p < w.P; ++p) {
        outs.emplace_back(w, n * sizeof(T), false);
        a2a_in.emplace_back(w, n * sizeof(T), false);
        a2a_out.emplace_back(w, n * sizeof(T), false);
    }
    w.run([&](int r, dlbb_comm* c) {
        for (int rep = 0; rep < kReps; ++rep) {
            mfill(in.at<T>(r), r, n, rep);
            std::vector<void*> op(w.P), ao(w.P);
            std::vector<const void*> ai(w.P);
            for (int p = 0; p < w.P; ++p) {
                op[p] = outs[p].at<T>(r);
                memset(op[p], 0x11, n * sizeof(T));
                mfill(a2a_in[p].at<T>(r), r, n, rep, 1000 * p);   // what rank r sends to rank p
                ai[p] = a2a_in[p].at<T>(r);
                ao[p] = a2a_out[p].at<T>(r);
            }
            CALL(dlbb_allgather_ptrs(c, in.at<T>(r), op.data(), n, DT<T>::code, root, nullptr, nullptr));
            if (root < 0 || r == root)
                for (int p = 0; p < w.P; ++p)
                    if (!expect_eq(static_cast<T*>(op[p]), n, p, rep, 0, r, "allgather_ptrs")) break;
            CALL(dlbb_alltoall_ptrs(c, ai.data(), ao.data(), n * (int64_t)sizeof(T), nullptr));
            for (int p = 0; p < w.P; ++p)
                if (!expect_eq(static_cast<T*>(ao[p]), n, p, rep, 1000 * r, r, "alltoall_ptrs")) break;
            // the same with the input list being the chunks of ONE tensor (tensor.chunk(world)): zero-copy from 8 KiB
            T* flat = dense_in.at<T>(r);
            for (int p = 0; p < w.P; ++p) {
                mfill(flat + p * n, r, n, rep, 2000 * p);
                ai[p] = flat + p * n;
                memset(ao[p], 0x22, n * sizeof(T));
            }
            CALL(dlbb_alltoall_ptrs(c, ai.data(), ao.data(), n * (int64_t)sizeof(T), nullptr));
            for (int p = 0; p < w.P; ++p)
                if (!expect_eq(static_cast<T*>(ao[p]), n, p, rep, 2000 * r, r, "alltoall_ptrs (dense list)")) break;
            g_checks.fetch_add(3);
        }
    });
}

// ------------------------------------------------------------------ strided views (K12)
struct TestView {
    const char* name;
    dlbb_view_t v;
    int64_t count;      // logical elements
    int64_t extent;     // elements of the underlying allocation
};
// [b, s, h] views of the kinds the 3-D sweeps pass (collectives/3d/*.py with DLBB_SWEEP_STRIDED=1) plus an element-wise one
static std::vector<TestView> test_views(int64_t b, int64_t s, int64_t h) {
    std::vector<TestView> out;
    {   // last-dim slice of a wider tensor: [b, s, :h] of [b, s, h + 8]  (vector path)
        TestView t{"slice", {3, {b, s, h, 0}, {s * (h + 8), h + 8, 1, 0}}, b * s * h, b * s * (h + 8)};
        out.push_back(t);
    }
    {   // permuted batch / sequence: [s, b, h] view of a contiguous [b, s, h]  (vector path)
        TestView t{"permute", {3, {s, b, h, 0}, {h, s * h, 1, 0}}, b * s * h, b * s * h};
        out.push_back(t);
    }
    {   // transposed inner dims: [b, h, s] view of [b, s, h]  (element path)
        TestView t{"transpose", {3, {b, h, s, 0}, {s * h, 1, h, 0}}, b * s * h, b * s * h};
        out.push_back(t);
    }
    {   // every second row of a 2-D tensor, odd width (element path, 2 dims)
        TestView t{"rows2", {2, {b * s, h - 1, 0, 0}, {2 * h, 1, 0, 0}}, b * s * (h - 1), 2 * b * s * h};
        out.push_back(t);
    }
    return out;
}
static int64_t view_off(const dlbb_view_t& v, int64_t i) {   // independent of view_math.h: plain mixed-radix walk
    int64_t idx[4] = {0, 0, 0, 0}, off = 0;
    for (int d = v.ndim - 1; d >= 0; --d) { idx[d] = i % v.size[d]; i /= v.size[d]; }
    for (int d = 0; d < v.ndim; ++d) off += idx[d] * v.stride[d];
    return off;
}
static constexpr double kSentinel = 77.0;

template <typename T>
static void poison(T* base, int64_t extent) { for (int64_t i = 0; i < extent; ++i) base[i] = from_d<T>(kSentinel); }
template <typename T>
static void check_untouched(const T* base, const TestView& tv, int rank, const char* what) {
    std::vector<char> in_view(tv.extent, 0);
    for (int64_t i = 0; i < tv.count; ++i) in_view[view_off(tv.v, i)] = 1;
    for (int64_t e = 0; e < tv.extent; ++e)
        if (!in_view[e] && to_d(base[e]) != kSentinel) { FAIL("%s: rank %d wrote outside the view at element %lld", what, rank, (long long)e); return; }
}

template <typename T>
static void case_views(World& w, int64_t b, int64_t s, int64_t h, bool symm) {
    for (const TestView& tv : test_views(b, s, h)) {
        char name[200];
        snprintf(name, sizeof(name), "views %s %s [%lld,%lld,%lld] %s", DT<T>::name, tv.name, (long long)b, (long long)s, (long long)h,
                 symm ? "symm" : "plain");
        g_case = name;
        Bufs t(w, tv.extent * sizeof(T), symm), dense(w, tv.count * w.P * sizeof(T), symm);
        const int root = w.P - 1;
        w.run([&](int r, dlbb_comm* c) {
            T* base = t.at<T>(r);
            T* flat = dense.at<T>(r);
            for (int rep = 0; rep < 2; ++rep) {
                // all-reduce in place on the view
                poison(base, tv.extent);
                for (int64_t i = 0; i < tv.count; ++i) base[view_off(tv.v, i)] = from_d<T>(val(r, i, rep));
                CALL(dlbb_allreduce_v(c, base, base, tv.count, DT<T>::code, DLBB_SUM, DLBB_ALGO_AUTO, &tv.v, &tv.v, nullptr));
                for (int64_t i = 0; i < tv.count; ++i) {
                    const double e = reduce_expected(w.P, i, rep, DLBB_SUM), g = to_d(base[view_off(tv.v, i)]);
                    if (e != g) { FAIL("allreduce_v rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)i, g, e); break; }
                }
                check_untouched(base, tv, r, "allreduce_v");
                // strided -> contiguous all-reduce (pack only)
                for (int64_t i = 0; i < tv.count; ++i) base[view_off(tv.v, i)] = from_d<T>(val(r, i, rep + 1));
                CALL(dlbb_allreduce(c, base, flat, tv.count, DT<T>::code, DLBB_SUM, DLBB_ALGO_AUTO, &tv.v, nullptr));
                for (int64_t i = 0; i < tv.count; ++i) {
                    const double e = reduce_expected(w.P, i, rep + 1, DLBB_SUM), g = to_d(flat[i]);
                    if (e != g) { FAIL("allreduce(view in) rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)i, g, e); break; }
                }
                // broadcast of the view
                poison(base, tv.extent);
                if (r == root) for (int64_t i = 0; i < tv.count; ++i) base[view_off(tv.v, i)] = from_d<T>(mval(root, i, rep));
                CALL(dlbb_broadcast_v(c, base, tv.count, DT<T>::code, root, &tv.v, nullptr));
                for (int64_t i = 0; i < tv.count; ++i) {
                    const double e = mval(root, i, rep), g = to_d(base[view_off(tv.v, i)]);
                    if (e != g) { FAIL("broadcast_v rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)i, g, e); break; }
                }
                check_untouched(base, tv, r, "broadcast_v");
                // reduce to the root, in place on the view
                poison(base, tv.extent);
                for (int64_t i = 0; i < tv.count; ++i) base[view_off(tv.v, i)] = from_d<T>(val(r, i, rep));
                CALL(dlbb_reduce_v(c, base, base, tv.count, DT<T>::code, DLBB_SUM, root, &tv.v, &tv.v, nullptr));
                if (r == root) {
                    for (int64_t i = 0; i < tv.count; ++i) {
                        const double e = reduce_expected(w.P, i, rep, DLBB_SUM), g = to_d(base[view_off(tv.v, i)]);
                        if (e != g) { FAIL("reduce_v rep %d elem %lld: got %g expected %g", rep, (long long)i, g, e); break; }
                    }
                    check_untouched(base, tv, r, "reduce_v");
                }
                // all-gather and gather with a strided input
                for (int64_t i = 0; i < tv.count; ++i) base[view_off(tv.v, i)] = from_d<T>(mval(r, i, rep));
                CALL(dlbb_allgather(c, base, flat, tv.count, DT<T>::code, &tv.v, nullptr));
                for (int p = 0; p < w.P; ++p)
                    if (!expect_eq(flat + p * tv.count, tv.count, p, rep, 0, r, "allgather(view in)")) break;
                memset(flat, 0, tv.count * w.P * sizeof(T));
                CALL(dlbb_gather_v(c, base, flat, tv.count, DT<T>::code, root, &tv.v, nullptr));
                if (r == root)
                    for (int p = 0; p < w.P; ++p)
                        if (!expect_eq(flat + p * tv.count, tv.count, p, rep, 0, r, "gather_v")) break;
                g_checks.fetch_add(6);
            }
        });
    }
}

// ------------------------------------------------------------------ K20: fused gradient reduce + Adam + parameter broadcast
static void ref_adam(float& w, float& m, float& v, float g_raw, float lr, float b1, float b2, float eps, float wd, float gs, int step) {
    const float bc1_inv = static_cast<float>(1.0 / (1.0 - pow(static_cast<double>(b1), step)));
    const float bc2_inv_sqrt = static_cast<float>(1.0 / sqrt(1.0 - pow(static_cast<double>(b2), step)));
    float g = g_raw * gs;
    if (wd != 0.f) g = fmaf(wd, w, g);
    m = fmaf(b1, m, (1.f - b1) * g);
    v = fmaf(b2, v, (1.f - b2) * g * g);
    w -= lr * (m * bc1_inv) / (sqrtf(v) * bc2_inv_sqrt + eps);
}

static void case_fused_adam(World& w, int64_t n, int64_t elem_off, bool use_mc, bool with_master) {
    char name[160];
    snprintf(name, sizeof(name), "fused_adam n=%lld off=%lld %s %s", (long long)n, (long long)elem_off, use_mc ? "multimem" : "peer",
             with_master ? "master" : "nomaster");
    g_case = name;
    using T = __nv_bfloat16;
    const int P = w.P;
    const size_t bytes = (elem_off + n) * sizeof(T);
    w.alloc_symm((bytes + 255) / 256 * 256);
    const int gseg = static_cast<int>(w.segs.size()) - 1;
    w.alloc_symm((bytes + 255) / 256 * 256);
    const int pseg = static_cast<int>(w.segs.size()) - 1;
    const int64_t per = (n + P * 8 - 1) / (P * 8) * 8;
    const float lr = 0.01f, b1 = 0.9f, b2 = 0.999f, eps = 1e-8f, wd = 0.01f, gs = 1.0f / P;
    // expected trajectory of every element (fp32 state lives on the owning rank only; the arithmetic is the same everywhere)
    std::vector<float> ew(n), em(n, 0.f), ev(n, 0.f);
    for (int64_t i = 0; i < n; ++i) ew[i] = to_d(from_d<T>(0.25 * ((i % 9) - 4)));
    std::vector<std::vector<float>> m(P), v(P), master(P);
    for (int r = 0; r < P; ++r) {
        const int64_t lo = std::min<int64_t>(r * per, n), hi = std::min<int64_t>(lo + per, n);
        m[r].assign(per + 8, 0.f);
        v[r].assign(per + 8, 0.f);
        master[r].assign(per + 8, 0.f);
        for (int64_t i = lo; i < hi; ++i) master[r][i - lo] = ew[i];
        T* pr = reinterpret_cast<T*>(w.segs[pseg].base[r]) + elem_off;
        for (int64_t i = 0; i < n; ++i) pr[i] = from_d<T>(ew[i]);
    }
    const int steps = 3;
    std::vector<std::vector<float>> expect(steps, std::vector<float>(n));
    for (int st = 1; st <= steps; ++st)
        for (int64_t i = 0; i < n; ++i) {
            float g = 0.f;
            for (int r = 0; r < P; ++r) g += static_cast<float>(val(r, i, st));
            // without fp32 master weights the update starts from the bf16 parameter
            float wv = with_master ? ew[i] : static_cast<float>(to_d(from_d<T>(ew[i])));
            ref_adam(wv, em[i], ev[i], g, lr, b1, b2, eps, wd, gs, st);
            ew[i] = wv;
            expect[st - 1][i] = static_cast<float>(to_d(from_d<T>(wv)));
        }
    w.run([&](int r, dlbb_comm* c) {
        T* gr = reinterpret_cast<T*>(w.segs[gseg].base[r]) + elem_off;
        const T* pr = reinterpret_cast<const T*>(w.segs[pseg].base[r]) + elem_off;
        for (int st = 1; st <= steps; ++st) {
            for (int64_t i = 0; i < n; ++i) gr[i] = from_d<T>(val(r, i, st));
            CALL(dlbb_fused_reduce_adam_bf16(c, gseg, pseg, elem_off, n, m[r].data(), v[r].data(), with_master ? master[r].data() : nullptr,
                                             lr, b1, b2, eps, wd, gs, st, use_mc ? 1 : 0, nullptr));
            for (int64_t i = 0; i < n; ++i) {
                const float g = static_cast<float>(to_d(pr[i])), e = expect[st - 1][i];
                if (fabsf(g - e) > 0.008f * fabsf(e) + 1e-6f) {   // at most one bf16 ulp (rounding of an fp32 value on a tie)
                    FAIL("rank %d step %d elem %lld: param %g expected %g", r, st, (long long)i, g, e);
                    break;
                }
            }
            g_checks.fetch_add(1);
        }
    });
}

// ------------------------------------------------------------------ K16: the all-reduce half of the fused GEMM + all-reduce
// The tensor-core half cannot run here; a stand-in producer plays the GEMM roles of gemm_cta_body: CTA b "computes" tiles
// b, b + grid, ... of the persistent schedule by writing rank- and element-dependent values into this rank's partial
// buffer, and its leader calls ArFusion::tile_stored for each; threads >= kNumThreads are the real reducer warps
// (ArFusion::extra_role), and everybody ends in ArFusion::finish - the functor is the one gemm_allreduce.cu compiles.
static inline double pval(int rank, int64_t e, int rep) { return static_cast<double>(((e * 5 + rank * 7 + rep * 3) % 7) - 3); }

static void emu_gemm_allreduce_kernel(dlbb::ArFusion f, __nv_bfloat16* my_part, int rep) {
    using namespace dlbb;
    if (threadIdx.x < (unsigned)kNumThreads) {
        for (int t = blockIdx.x; t < f.num_tiles; t += gridDim.x) {
            int m_blk, n_blk;
            tile_coords(t, f.num_m, f.num_n, f.group_m, m_blk, n_blk);
            const int m0 = m_blk * kBM, n0 = n_blk * f.bn;
            const int rows = std::min(kBM, f.M - m0), cols = std::min(f.bn, f.N - n0);
            for (int i = threadIdx.x; i < rows * cols; i += kNumThreads) {
                const int64_t e = (int64_t)(m0 + i / cols) * f.N + n0 + i % cols;
                my_part[e] = __float2bfloat16(static_cast<float>(pval(f.c.rank, e, rep)));
            }
            group_bar_sync(1, kNumThreads);          // the whole tile is written ...
            if (threadIdx.x == 0) f.tile_stored(t, m0, n0);   // ... before the leader raises its flag
        }
    } else {
        f.extra_role(0, 0);
    }
    f.finish();
}

static void case_gemm_allreduce(World& w, int M, int N, int group_m, int grid, bool use_mc, bool with_residual) {
    char name[160];
    snprintf(name, sizeof(name), "gemm_allreduce M=%d N=%d group_m=%d grid=%d %s %s", M, N, group_m, grid, use_mc ? "multimem" : "peer",
             with_residual ? "residual" : "noresidual");
    g_case = name;
    using T = __nv_bfloat16;
    using namespace dlbb;
    const int P = w.P, BN = 256;
    const size_t bytes = ((size_t)M * N * sizeof(T) + 255) / 256 * 256;
    w.alloc_symm(bytes);
    const int part_seg = static_cast<int>(w.segs.size()) - 1;
    w.alloc_symm(bytes);
    const int out_seg = static_cast<int>(w.segs.size()) - 1;
    Bufs res(w, (size_t)M * N * sizeof(T), false);
    const int num_m = (M + kBM - 1) / kBM, num_n = (N + BN - 1) / BN;
    grid = std::min(grid, std::max(2, 24 / P));   // 512 host threads per CTA and rank: stay below the process's thread limit
    // like the launcher (grid = min(tiles, SM count), one CTA per SM): the reducers wait for tiles of every CTA, so the grid
    // must be resident as a whole
    if (emu::resident_ctas() > 0) grid = std::min(grid, emu::resident_ctas());
    w.run([&](int r, dlbb_comm* c) {
        T* rs = res.at<T>(r);
        const T* out = reinterpret_cast<const T*>(w.segs[out_seg].base[r]);
        for (int rep = 0; rep < kReps; ++rep) {
            if (with_residual)
                for (int64_t e = 0; e < (int64_t)M * N; ++e) rs[e] = from_d<T>(static_cast<double>((e + rep) % 3));
            if (rep % 2 == 1) {   // the launcher's periodic flag refresh (every 2^28 calls there, every other call here)
                c->last_refresh[kArBank] = c->epoch[kArBank] - (1u << 28);
                CALL(refresh_flag_bank_if_due(c, kArBank, nullptr));
                const uint32_t ep = c->epoch[kArBank] + 1;
                c->epoch[kArBank] += 2;
                EMU_LAUNCH(4, 256, refresh_tile_flags_kernel)(c->dev, ep);
            }
            ArFusion f{};   // as dlbb_gemm_allreduce_bf16 fills it
            f.c = c->dev;
            for (int q = 0; q < P; ++q) {
                f.part[q] = w.segs[part_seg].base[q];
                f.out[q] = w.segs[out_seg].base[q];
            }
            const bool nvls = use_mc && w.segs[part_seg].mc && P > 1;
            f.part_mc = nvls ? w.segs[part_seg].mc : nullptr;
            f.out_mc = nvls ? w.segs[out_seg].mc : nullptr;
            f.residual = with_residual ? rs : nullptr;
            f.ld_res = N;
            f.M = M; f.N = N; f.bn = BN; f.num_m = num_m; f.num_n = num_n; f.num_tiles = num_m * num_n; f.group_m = group_m;
            f.epoch = ++c->epoch[kArBank];
            EMU_LAUNCH(std::min(grid, num_m * num_n), kArThreads, emu_gemm_allreduce_kernel)(f, reinterpret_cast<T*>(w.segs[part_seg].base[r]), rep);
            for (int64_t e = 0; e < (int64_t)M * N; ++e) {
                double x = 0;
                for (int q = 0; q < P; ++q) x += pval(q, e, rep);
                if (with_residual) x += static_cast<double>((e + rep) % 3);
                const double g = to_d(out[e]);
                if (g != x) { FAIL("rank %d rep %d elem %lld: got %g expected %g", r, rep, (long long)e, g, x); break; }
            }
            g_checks.fetch_add(1);
        }
    });
}

// ------------------------------------------------------------------ mixed sequences
// Different collectives back to back on one communicator, no host synchronisation in between: what one kernel leaves
// behind in the staging halves / flag banks is what the next one starts from.  (Cases above join their rank threads
// between calls of different kinds; here the only ordering between ranks is the kernels' own flag protocol.)
static void case_mixed(World& w, unsigned seed, int length) {
    char name[96];
    snprintf(name, sizeof(name), "mixed seed=%u len=%d", seed, length);
    g_case = name;
    const int P = w.P;
    const int64_t sizes[] = {5, 640, 4100, 20000};
    struct Slot { Bufs a, b; Slot(World& w, size_t n, bool symm) : a(w, n, symm), b(w, n, symm) {} };
    std::vector<std::unique_ptr<Slot>> plain, symm;
    for (int64_t n : sizes) {
        plain.emplace_back(new Slot(w, n * P * sizeof(float), false));
        symm.emplace_back(new Slot(w, n * P * sizeof(float), true));
    }
    struct Step { int kind, size, symm, arg; };
    std::vector<Step> seq;
    unsigned x = seed * 2654435761u + 12345u;
    auto rnd = [&](int m) { x = x * 1664525u + 1013904223u; return static_cast<int>((x >> 10) % m); };
    for (int i = 0; i < length; ++i) seq.push_back(Step{rnd(9), rnd(4), rnd(2), rnd(64)});
    w.run([&](int r, dlbb_comm* c) {
        int rep = 0;
        for (const Step& st : seq) {
            ++rep;
            const int64_t n = sizes[st.size];
            Slot& sl = st.symm ? *symm[st.size] : *plain[st.size];
            float* a = sl.a.at<float>(r);
            float* b = sl.b.at<float>(r);
            const int root = st.arg % P;
            switch (st.kind) {
                case 0: {   // all-reduce, in place, algorithm by argument
                    const int algos[] = {DLBB_ALGO_AUTO, DLBB_ALGO_ONESHOT, DLBB_ALGO_TWOSHOT, DLBB_ALGO_NVLS};
                    fill(a, r, n, rep);
                    CALL(dlbb_allreduce(c, a, a, n, DLBB_F32, DLBB_SUM, algos[st.arg % 4], nullptr, nullptr));
                    for (int64_t i = 0; i < n; ++i)
                        if (to_d(a[i]) != reduce_expected(P, i, rep, DLBB_SUM)) { FAIL("mixed allreduce rank %d step %d elem %lld", r, rep, (long long)i); break; }
                    break;
                }
                case 1: {   // all-reduce out of place
                    fill(a, r, n, rep);
                    CALL(dlbb_allreduce(c, a, b, n, DLBB_F32, DLBB_MAX, DLBB_ALGO_AUTO, nullptr, nullptr));
                    for (int64_t i = 0; i < n; ++i)
                        if (to_d(b[i]) != reduce_expected(P, i, rep, DLBB_MAX)) { FAIL("mixed allreduce(max) rank %d step %d elem %lld", r, rep, (long long)i); break; }
                    break;
                }
                case 2: {
                    if (r == root) mfill(a, root, n, rep);
                    CALL(dlbb_broadcast(c, a, n, DLBB_F32, root, nullptr));
                    expect_eq(a, n, root, rep, 0, r, "mixed broadcast");
                    break;
                }
                case 3: {
                    fill(a, r, n, rep);
                    CALL(dlbb_reduce(c, a, a, n, DLBB_F32, DLBB_SUM, root, nullptr));
                    if (r == root)
                        for (int64_t i = 0; i < n; ++i)
                            if (to_d(a[i]) != reduce_expected(P, i, rep, DLBB_SUM)) { FAIL("mixed reduce step %d elem %lld", rep, (long long)i); break; }
                    break;
                }
                case 4: {
                    mfill(a, r, n, rep);
                    CALL(dlbb_allgather(c, a, b, n, DLBB_F32, nullptr, nullptr));
                    for (int p = 0; p < P; ++p)
                        if (!expect_eq(b + p * n, n, p, rep, 0, r, "mixed allgather")) break;
                    break;
                }
                case 5: {
                    const int64_t cpr = (n + 3) / 4 * 4;   // (buffers hold n * P floats; keep shards 16-byte multiples half of the time)
                    const int64_t k = (st.arg & 1) ? cpr : n;
                    if (k * P > n * P + 0 && k != n) { /* cpr * P may exceed the slot: fall back to n */ }
                    const int64_t use = (k * P <= n * P) ? k : n;
                    fill(a, r, use * P, rep);
                    CALL(dlbb_reduce_scatter(c, a, b, use, DLBB_F32, DLBB_SUM, nullptr));
                    for (int64_t i = 0; i < use; ++i)
                        if (to_d(b[i]) != reduce_expected(P, r * use + i, rep, DLBB_SUM)) { FAIL("mixed reduce_scatter rank %d step %d elem %lld", r, rep, (long long)i); break; }
                    break;
                }
                case 6: {
                    mfill(a, r, n * P, rep);
                    CALL(dlbb_alltoall(c, a, b, n, DLBB_F32, nullptr));
                    for (int p = 0; p < P; ++p)
                        if (!expect_eq(b + p * n, n, p, rep, r * n, r, "mixed alltoall")) break;
                    break;
                }
                case 7: {
                    mfill(a, r, n, rep);
                    CALL(dlbb_gather(c, a, b, n, DLBB_F32, root, nullptr));
                    if (r == root)
                        for (int p = 0; p < P; ++p)
                            if (!expect_eq(b + p * n, n, p, rep, 0, r, "mixed gather")) break;
                    break;
                }
                default: {
                    if (P > 2) {
                        mfill(a, r, n, rep);
                        CALL(dlbb_sendrecv(c, a, n, (r + 1) % P, b, n, (r + P - 1) % P, DLBB_F32, nullptr));
                        expect_eq(b, n, (r + P - 1) % P, rep, 0, r, "mixed sendrecv");
                    } else {
                        CALL(dlbb_barrier(c, nullptr));
                    }
                    break;
                }
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

    if (filter && strcmp(filter, "watchdog") == 0) {
        // the last rank never issues the collective: the others' flag waits must give up after DLBB_WATCHDOG_SECONDS and
        // name it (the process aborts, as the GPU traps) - run on its own, expected to fail
        g_case = "watchdog";
        Bufs b(w, 4096, false);
        w.run([&](int r, dlbb_comm* c) {
            if (r != P - 1) dlbb_allreduce(c, b.at<float>(r), b.at<float>(r), 1024, DLBB_F32, DLBB_SUM, DLBB_ALGO_AUTO, nullptr, nullptr);
        });
        printf("watchdog did not fire\n");
        return 3;
    }
    const int64_t sizes[] = {0, 1, 7, 300, 4099, 20000, 70001};
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
    const int roots[] = {0, P - 1, P / 2};
    if (want(filter, "reduce_scatter"))
        for (int64_t n : {4, 12, 300, 4100, 20000}) {
            case_reduce_scatter<float>(w, n, DLBB_SUM, false);
            case_reduce_scatter<__nv_bfloat16>(w, n * 2, DLBB_SUM, true);
            case_reduce_scatter<int32_t>(w, n + 1, DLBB_MAX, false);
        }
    if (want(filter, "reduce "))
        for (int64_t n : sizes)
            for (int root : roots) {
                case_reduce<float>(w, n, DLBB_SUM, root, false, true);
                case_reduce<__nv_bfloat16>(w, n, DLBB_SUM, root, true, true);
                case_reduce<float>(w, n, DLBB_SUM, root, false, false);
                case_reduce<int64_t>(w, n, DLBB_MIN, root, false, true);
            }
    if (want(filter, "allgather"))
        for (int64_t n : sizes) {
            case_allgather<float>(w, n, false);
            case_allgather<__nv_bfloat16>(w, n, true);
            case_allgather<double>(w, n, false);
        }
    if (want(filter, "gather+scatter"))
        for (int64_t n : {1, 300, 4099, 20000})
            for (int root : roots) {
                case_gather_scatter<float>(w, n, root, false);
                case_gather_scatter<__nv_bfloat16>(w, n, root, true);
            }
    if (want(filter, "broadcast"))
        for (int64_t n : sizes)
            for (int root : roots) {
                case_broadcast<float>(w, n, root, false);
                case_broadcast<__nv_bfloat16>(w, n, root, true);
            }
    if (want(filter, "alltoall"))
        for (int64_t n : {1, 16, 301, 4096, 9000}) {
            case_alltoall<float>(w, n, false);
            case_alltoall<__nv_bfloat16>(w, n, true);
        }
    if (want(filter, "sendrecv"))
        for (int64_t n : sizes) case_ring_sendrecv<float>(w, n);
    if (want(filter, "residual"))
        for (int64_t n : {8, 4096, 70000}) case_residual(w, n);
    if (want(filter, "barrier")) case_barrier(w);
    if (want(filter, "multi")) {
        case_multi<float>(w, 5, 100, DLBB_SUM);
        case_multi<__nv_bfloat16>(w, 16, 2048, DLBB_SUM);
        case_multi<float>(w, 100, 40, DLBB_SUM);       // more tensors than one table holds
        case_multi<int32_t>(w, 7, 9000, DLBB_MAX);     // more bytes than one bucket holds
    }
    if (want(filter, "ptrs"))
        for (int64_t n : {5, 1024, 20000})
            for (int root : {-1, 0, P - 1}) case_ptrs<float>(w, n, root);
    if (want(filter, "gemm_allreduce")) {
        case_gemm_allreduce(w, 256, 512, 2, 3, false, true);            // 2 x 2 tiles, more tiles than CTAs
        case_gemm_allreduce(w, 200, 264, 1, 2, mc && P > 1, true);      // ragged last row / column of tiles
        case_gemm_allreduce(w, 640, 256, 5, 4, mc && P > 1, false);     // one column of five tiles in one group
        case_gemm_allreduce(w, 384, 768, 2, 9, false, false);           // as many CTAs as tiles, group of 2 + a group of 1
    }
    if (want(filter, "fused_adam"))
        for (int64_t n : {8, 1000, 4096 + 24, 30001}) {
            case_fused_adam(w, n, 0, false, true);
            case_fused_adam(w, n, 64, mc && P > 1, true);
            case_fused_adam(w, n, 8, false, false);
        }
    if (want(filter, "mixed"))
        for (unsigned seed = 1; seed <= (getenv("EMU_MIXED_SEEDS") ? (unsigned)atoi(getenv("EMU_MIXED_SEEDS")) : 6u); ++seed) case_mixed(w, seed, 60);
    if (want(filter, "views")) {
        case_views<float>(w, 2, 5, 16, false);
        case_views<__nv_bfloat16>(w, 3, 7, 24, false);
        case_views<__nv_bfloat16>(w, 4, 33, 64, true);   // 16.5 KiB per rank: beyond the zero-copy threshold
        case_views<float>(w, 4, 64, 96, false);          // 96 KiB per rank: sliced through the staging area
    }
    if (g_fail.load()) {
        printf("EMU COLLECTIVES FAILED: %d failures (world %d, %s)\n", g_fail.load(), P, mc ? "multicast" : "no multicast");
        return 1;
    }
    printf("kernels:");
    for (const auto& kv : emu::kernel_counts().n) printf(" %s=%ld", kv.first.c_str(), kv.second);
    printf("\n");
    printf("EMU COLLECTIVES OK: world %d, %s, %ld checks, %ld kernel launches emulated\n", P, mc ? "multicast" : "no multicast",
           g_checks.load(), emu::launches().load());
    return 0;
}
