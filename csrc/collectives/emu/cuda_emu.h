// Host emulation of the small part of the CUDA execution model that the collective kernels use, so that the SAME kernel
// and launcher source (collectives.cu, direct_kernels.cuh, coll_device.cuh) can be compiled with g++ and run on a machine
// without a GPU - under ThreadSanitizer / AddressSanitizer when wanted (race detection).
//
//   * every CUDA thread is a host thread; threadIdx / blockIdx / blockDim / gridDim are thread_local
//   * __syncthreads() and named barriers (bar.sync id, n) are mutex + condition-variable barriers of the CTA
//   * system-scope release stores / acquire loads on the signal pads are __atomic release / acquire operations, so the
//     sanitizer knows exactly the ordering the PTX memory model guarantees - and reports every data access that is NOT
//     ordered by the flag protocol as a race
//   * peer loads / stores of data are plain 16-byte accesses (every "GPU" lives in this address space)
//   * a multicast address is an address inside a reserved, inaccessible window; multimem.st writes the vector at the same
//     offset of every rank's buffer, multimem.ld_reduce sums the ranks' vectors in fp32 (bf16 / f16) or in order (f32)
//   * a launch runs its grid to completion on the calling thread's behalf (the rank's "stream" is the rank's host thread);
//     ranks run independently, so kernel k+1 of one rank overlaps kernel k of another exactly as on the box
//
// What the emulation cannot see (these stay with the GPU tests): stale L1 lines (a cacheable load of data a peer wrote
// during the kernel - ld_stream_v4 where ld_peer_v4 is needed - reads fresh data here), warp-level primitives (none are
// used by these kernels), and of course time.  Limited residency IS modelled (EMU_RESIDENT_CTAS, see run_grid).
//
// tools/emu_build.py rewrites `kernel<<<grid, block, smem, stream>>>(args)` into EMU_LAUNCH(...) and compiles the
// result; nothing else in the sources differs between the two builds.
#ifndef DLBB_CUDA_EMU_H   // (reached both through -include and through relative includes: a guard, not #pragma once)
#define DLBB_CUDA_EMU_H
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

// ------------------------------------------------------------------ execution model
#undef __launch_bounds__
#define __launch_bounds__(...)
// (__noinline__ is rewritten by tools/emu_build.py: as a macro it would break libstdc++'s __attribute__((__noinline__)))
#undef __grid_constant__
#define __grid_constant__
inline thread_local uint3 threadIdx{0, 0, 0}, blockIdx{0, 0, 0};
inline thread_local dim3 blockDim(1), gridDim(1);

namespace emu {

struct NamedBarrier {
    std::mutex m;
    std::condition_variable cv;
    unsigned count = 0, gen = 0;
    void sync(unsigned n) {
        std::unique_lock<std::mutex> lk(m);
        const unsigned g = gen;
        if (++count == n) {
            count = 0;
            ++gen;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return gen != g; });
        }
    }
};
struct Cta {
    NamedBarrier bar[16];
};
inline thread_local Cta* tl_cta = nullptr;

// blocks of the generic kernels (they only need blockDim.x >= world) are shrunk to this many threads; kernels whose
// roles depend on the exact block size keep it (see launch())
inline int max_block() {
    static const int v = getenv("EMU_MAX_BLOCK") ? atoi(getenv("EMU_MAX_BLOCK")) : 8;
    return v;
}
inline int keep_block_at() {   // block sizes at or above this are kept as they are (role-partitioned kernels)
    static const int v = getenv("EMU_KEEP_BLOCK") ? atoi(getenv("EMU_KEEP_BLOCK")) : 512;
    return v;
}
inline std::atomic<long>& launches() {
    static std::atomic<long> n{0};
    return n;
}
// launches per kernel (the text of the launch expression), so that a test can tell which code paths a configuration took
struct KernelCounts {
    std::mutex m;
    std::map<std::string, long> n;
    void add(const char* name) {
        std::lock_guard<std::mutex> lk(m);
        ++n[name];
    }
};
inline KernelCounts& kernel_counts() {
    static KernelCounts k;
    return k;
}

// EMU_RESIDENT_CTAS = N: at most N CTAs of a grid run at a time, started in blockIdx order as slots free up - what the
// hardware does when a grid does not fit.  A kernel in which a CTA waits for a higher-numbered CTA of its own grid, or a
// grid barrier over more CTAs than slots, then stops making progress here too (and the watchdog reports it).
inline int resident_ctas() {
    static const int v = getenv("EMU_RESIDENT_CTAS") ? atoi(getenv("EMU_RESIDENT_CTAS")) : 0;
    return v;
}

template <typename F>
void run_grid(int grid, int block, F&& body) {
    if (block < keep_block_at()) block = std::min(block, max_block());
    launches().fetch_add(1, std::memory_order_relaxed);
    std::vector<Cta> ctas(grid);
    std::vector<std::thread> th;
    th.reserve(static_cast<size_t>(grid) * block);
    const int slots = resident_ctas() > 0 ? resident_ctas() : grid;
    std::mutex m;
    std::condition_variable cv;
    int running = 0;
    std::vector<int> left(grid, block);   // threads of CTA b still running (under m)
    for (int b = 0; b < grid; ++b) {
        {
            std::unique_lock<std::mutex> lk(m);
            cv.wait(lk, [&] { return running < slots; });
            ++running;
        }
        for (int t = 0; t < block; ++t)
            th.emplace_back([&, b, t] {
                threadIdx = uint3{static_cast<unsigned>(t), 0, 0};
                blockIdx = uint3{static_cast<unsigned>(b), 0, 0};
                blockDim = dim3(block);
                gridDim = dim3(grid);
                tl_cta = &ctas[b];
                body();
                std::lock_guard<std::mutex> lk(m);
                if (--left[b] == 0) {
                    --running;
                    cv.notify_all();
                }
            });
    }
    for (auto& x : th) x.join();
}

template <typename K>
struct Launcher {
    const char* name;
    int grid, block;
    K kernel;
    template <typename... A>
    void operator()(A&&... a) const {
        kernel_counts().add(name);
        run_grid(grid, block, [&] { kernel(a...); });
    }
};
template <typename K>
Launcher<K> launcher(const char* name, int grid, int block, K k) { return Launcher<K>{name, grid, block, k}; }

// ------------------------------------------------------------------ multicast windows
struct McWindow {
    char* base = nullptr;      // reserved, inaccessible
    size_t size = 0;
    int world = 0;
    char* rank_base[16] = {};
};
inline std::vector<McWindow>& windows() {
    static std::vector<McWindow> w;
    return w;
}
inline const McWindow& window_of(const void* mc) {
    const char* p = static_cast<const char*>(mc);
    for (const auto& w : windows())
        if (p >= w.base && p < w.base + w.size) return w;
    fprintf(stderr, "emu: %p is not a multicast address\n", mc);
    abort();
}

inline long long now_ns() {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace emu

#define EMU_LAUNCH(grid, block, ...) ::emu::launcher(#__VA_ARGS__, static_cast<int>(grid), static_cast<int>(block), __VA_ARGS__)

// the integer / floating-point min and max overloads that device code finds in the global namespace
#define EMU_MINMAX(T)                                  \
    inline T min(T a, T b) { return b < a ? b : a; }   \
    inline T max(T a, T b) { return a < b ? b : a; }
EMU_MINMAX(int)
EMU_MINMAX(unsigned)
EMU_MINMAX(long)
EMU_MINMAX(unsigned long)
EMU_MINMAX(long long)
EMU_MINMAX(unsigned long long)
EMU_MINMAX(float)
EMU_MINMAX(double)
#undef EMU_MINMAX

template <typename T>
inline T __shfl_xor_sync(unsigned, T, int) {   // not used by the collective kernels (warp reductions of the math kernels)
    fprintf(stderr, "emu: warp shuffles are not emulated\n");
    abort();
}
inline void __syncthreads() { emu::tl_cta->bar[0].sync(blockDim.x); }
[[noreturn]] inline void __trap() { abort(); }
#endif  // DLBB_CUDA_EMU_H
