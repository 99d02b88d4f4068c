// Host-side check of the strided-view index arithmetic used by the collective kernels (csrc/collectives/view_math.h).
// For random views of 1-4 dimensions (slices, permutations and padded rows of a base tensor) and element sizes 1-8:
//   * view_offset agrees with a nested-loop walk of the view;
//   * whenever view_vec_ok_bytes / view_vec_ok<T> allow the vector path, every 16-byte vector of the logical order is 16
//     contiguous, 16-byte aligned bytes of the view;
//   * a gather (view -> dense) followed by a scatter (dense -> view) written exactly like the kernels' loops - vector path
//     when allowed, byte path otherwise, starting at arbitrary logical offsets - reproduces a plain copy.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../view_math.h"

using dlbb::ViewDev;

static uint64_t rng_state = 88172645463325252ull;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }
static int64_t rint(int64_t lo, int64_t hi) { return lo + (int64_t)(rnd() % (uint64_t)(hi - lo + 1)); }

struct Case { ViewDev v; int64_t span; };   // span: elements of the base storage

static Case make_case(int es) {
    Case c{};
    const int nd = (int)rint(1, 4);
    int64_t base_size[4], order[4];
    for (int d = 0; d < nd; ++d) { base_size[d] = rint(1, 6) * (rnd() % 2 ? 16 / es > 0 ? (16 / es) : 1 : 1) + (rnd() % 3 == 0 ? 1 : 0); order[d] = d; }
    if (rnd() % 3 == 0) for (int d = nd - 1; d > 0; --d) { int j = (int)rint(0, d); std::swap(order[d], order[j]); }   // permute dims
    int64_t bstride[4], st = 1;
    for (int d = nd - 1; d >= 0; --d) { bstride[d] = st; st *= base_size[d] + (rnd() % 4 == 0 ? rint(0, 8) : 0); }   // padded rows
    c.span = st + 64;
    c.v.ndim = nd;
    for (int d = 0; d < nd; ++d) {
        const int o = (int)order[d];
        c.v.size[d] = rint(1, base_size[o]);        // a slice [0, size) of that base dimension
        c.v.stride[d] = bstride[o];
    }
    return c;
}

template <typename T>
static int check_typed(const Case& c) {
    constexpr int VE = 16 / sizeof(T);
    ViewDev v = c.v;
    int64_t n = 1;
    for (int d = 0; d < v.ndim; ++d) n *= v.size[d];
    v.base = (rnd() % 2) ? 0 : rint(0, n - 1) / VE * VE;
    static char aligned[16] __attribute__((aligned(16)));
    if (!dlbb::view_vec_ok<T>(v, aligned)) return 0;
    for (int64_t i = 0; (i + 1) * VE + v.base <= n; ++i) {
        const int64_t o0 = dlbb::view_offset(v, i * VE);
        if (o0 % VE) return 1;
        for (int e = 1; e < VE; ++e)
            if (dlbb::view_offset(v, i * VE + e) != o0 + e) return 2;
    }
    return 0;
}

int main(int argc, char** argv) {
    const int cases = argc > 1 ? atoi(argv[1]) : 4000;
    int vec_cases = 0;
    for (int it = 0; it < cases; ++it) {
        const int es = 1 << (int)rint(0, 3);
        Case c = make_case(es);
        const ViewDev& v = c.v;
        int64_t n = 1;
        for (int d = 0; d < v.ndim; ++d) n *= v.size[d];
        // 1. view_offset vs nested loops (row-major order over the view's sizes)
        {
            int64_t idx[4] = {0, 0, 0, 0};
            for (int64_t i = 0; i < n; ++i) {
                int64_t want = 0;
                for (int d = 0; d < v.ndim; ++d) want += idx[d] * v.stride[d];
                if (dlbb::view_offset(v, i) != want) { printf("FAIL view_offset case %d i %lld\n", it, (long long)i); return 1; }
                for (int d = v.ndim - 1; d >= 0; --d) { if (++idx[d] < v.size[d]) break; idx[d] = 0; }
            }
        }
        // 2. the vector conditions imply contiguous, aligned vectors
        std::vector<char> store((size_t)c.span * es + 64, 0), dense((size_t)n * es + 64, 0), back((size_t)c.span * es + 64, 0);
        char* src = store.data() + ((16 - ((uintptr_t)store.data() & 15)) & 15);
        char* dst = back.data() + ((16 - ((uintptr_t)back.data() & 15)) & 15);
        char* dn = dense.data() + ((16 - ((uintptr_t)dense.data() & 15)) & 15);
        for (int64_t i = 0; i < c.span * es; ++i) src[i] = (char)(rnd() & 0xff);
        const int64_t bytes_total = n * es;
        const bool vok = dlbb::view_vec_ok_bytes(v, es, src);
        if (vok) {
            ++vec_cases;
            for (int64_t b = 0; b + 16 <= bytes_total; b += 16) {
                const int64_t o0 = dlbb::view_byte_off(v, es, b);
                if (o0 & 15) { printf("FAIL alignment case %d\n", it); return 1; }
                for (int k = 1; k < 16; ++k)
                    if (dlbb::view_byte_off(v, es, b + k) != o0 + k) { printf("FAIL contiguity case %d b %lld\n", it, (long long)b); return 1; }
            }
        }
        if (check_typed<uint16_t>(c) || check_typed<uint32_t>(c) || check_typed<uint64_t>(c)) { printf("FAIL typed vectors case %d\n", it); return 1; }
        // 3. gather + scatter exactly like grid_gather_view / grid_scatter_view, in two launches split at a random offset
        const int64_t split = (rnd() % 2) ? rint(0, bytes_total) / 16 * 16 : rint(0, bytes_total);
        for (int part = 0; part < 2; ++part) {
            const int64_t b0 = part ? split : 0, len = part ? bytes_total - split : split;
            const bool vec = dlbb::view_vec_ok_bytes(v, es, src) && (b0 & 15) == 0;
            int64_t done = 0;
            if (vec) {
                const int64_t nvec = len >> 4;
                for (int64_t i = 0; i < nvec; ++i) memcpy(dn + b0 + (i << 4), src + dlbb::view_byte_off(v, es, b0 + (i << 4)), 16);
                done = nvec << 4;
            }
            for (int64_t b = done; b < len; ++b) dn[b0 + b] = src[dlbb::view_byte_off(v, es, b0 + b)];
            const bool svec = dlbb::view_vec_ok_bytes(v, es, dst) && (b0 & 15) == 0;
            done = 0;
            if (svec) {
                const int64_t nvec = len >> 4;
                for (int64_t i = 0; i < nvec; ++i) memcpy(dst + dlbb::view_byte_off(v, es, b0 + (i << 4)), dn + b0 + (i << 4), 16);
                done = nvec << 4;
            }
            for (int64_t b = done; b < len; ++b) dst[dlbb::view_byte_off(v, es, b0 + b)] = dn[b0 + b];
        }
        {   // reference: nested-loop copy of the view, everything outside the view stays zero
            std::vector<char> want((size_t)c.span * es, 0);
            int64_t idx[4] = {0, 0, 0, 0};
            for (int64_t i = 0; i < n; ++i) {
                int64_t off = 0;
                for (int d = 0; d < v.ndim; ++d) off += idx[d] * v.stride[d];
                memcpy(want.data() + off * es, src + off * es, es);
                if (memcmp(dn + i * es, src + off * es, es) != 0) { printf("FAIL gather case %d elem %lld\n", it, (long long)i); return 1; }
                for (int d = v.ndim - 1; d >= 0; --d) { if (++idx[d] < v.size[d]) break; idx[d] = 0; }
            }
            if (memcmp(want.data(), dst, (size_t)c.span * es) != 0) { printf("FAIL scatter case %d\n", it); return 1; }
        }
    }
    printf("VIEW MATH OK (%d cases, %d on the vector path)\n", cases, vec_cases);
    return 0;
}
