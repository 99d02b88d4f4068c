// The "box" of the host emulation: P communicators in one process (one host thread per rank), their staging areas, signal
// pads and symmetric segments with emulated multicast windows, plus the few host-side functions that collectives.cu
// expects from the rest of libdlbb (error text, launch counter, zero-copy registry, multicast query).
// Include from exactly one translation unit of the test program.
#pragma once
#include <stdarg.h>
#include <sys/mman.h>

#include <deque>
#include <functional>
#include <memory>
#include <string>

#include "../../../include/dlbb_comm.h"
#include "../../bootstrap/comm_internal.h"
#include "cuda_emu.h"

namespace emu {

struct SymmSeg {
    size_t size = 0;
    char* base[16] = {};
    char* mc = nullptr;
};

inline char* alloc_zero(size_t bytes) {
    void* p = nullptr;
    if (posix_memalign(&p, 256, bytes ? bytes : 256) != 0) abort();
    memset(p, 0, bytes);
    return static_cast<char*>(p);
}

struct World {
    int P;
    bool multicast;
    std::vector<std::unique_ptr<dlbb_comm>> comm;
    std::deque<SymmSeg> segs;   // (references handed out by alloc_symm stay valid)
    // host rendezvous (what the shared-memory mailbox of peer_registry.cpp does between processes)
    std::mutex m;
    std::condition_variable cv;
    int arrived = 0;
    unsigned gen = 0;
    struct Rec { const void* ptr; size_t bytes; } recs[16] = {};

    World(int world, size_t stage_bytes, bool mc, int ctas, size_t zero_copy_min) : P(world), multicast(mc) {
        SymmSeg& pads = alloc_symm(dlbb::kPadBytes);
        SymmSeg& stage = alloc_symm(3 * stage_bytes);
        SymmSeg& flags = alloc_symm(dlbb::kTileFlagBytes);
        for (int r = 0; r < P; ++r) {
            auto c = std::make_unique<dlbb_comm>();
            c->rank = r;
            c->world = P;
            c->mc_supported = mc;
            c->stage_bytes = stage_bytes;
            c->ctas = ctas;
            // the emulated device has EMU_RESIDENT_CTAS slots (cuda_emu.h); EMU_NO_CLAMP keeps the library unaware of it
            // (negative control: a grid barrier over more CTAs than slots must then stall until the watchdog fires)
            if (const char* e = getenv("EMU_RESIDENT_CTAS"))
                if (!getenv("EMU_NO_CLAMP")) c->max_resident_ctas = atoi(e) > 0 ? atoi(e) : 128;
            c->zero_copy_min = zero_copy_min;
            c->debug_poison = getenv("DLBB_DEBUG_POISON") != nullptr && atoi(getenv("DLBB_DEBUG_POISON")) != 0;
            c->dev.rank = r;
            c->dev.world = P;
            for (int p = 0; p < P; ++p) {
                c->dev.pads[p] = reinterpret_cast<uint32_t*>(pads.base[p]);
                c->dev.stage[p] = stage.base[p];
                c->dev.tile_flags[p] = reinterpret_cast<uint32_t*>(flags.base[p]);
            }
            c->dev.stage_mc = mc ? stage.mc : nullptr;
            c->dev.pads_mc = mc ? reinterpret_cast<uint32_t*>(pads.mc) : nullptr;
            // EMU_EPOCH0: start every epoch counter (and every flag) just below 2^32, so that the run crosses the wrap-around
            if (const char* e0 = getenv("EMU_EPOCH0")) {
                const uint32_t v = static_cast<uint32_t>(strtoul(e0, nullptr, 0));
                for (uint32_t& e : c->epoch) e = v;
                for (int p = 0; p < P; ++p) { c->send_epoch[p] = v; c->recv_epoch[p] = v; }
                if (r == 0)
                    for (int p = 0; p < P; ++p) {
                        uint32_t* w = reinterpret_cast<uint32_t*>(pads.base[p]);
                        for (size_t i = 0; i < dlbb::kPadBytes / 4; ++i) w[i] = v;
                        uint32_t* tf = reinterpret_cast<uint32_t*>(flags.base[p]);   // per-tile flags of the fused GEMM + all-reduce
                        for (size_t i = 0; i < dlbb::kTileFlagBytes / 4; ++i) tf[i] = v;
                    }
            }
            const char* wd = getenv("DLBB_WATCHDOG_SECONDS");
            c->dev.watchdog_ns = (wd ? atoll(wd) : 300) * 1000000000ll;
            comm.push_back(std::move(c));
        }
    }

    // one buffer per rank at the same offset of one segment (+ a multicast window when the box has NVLS)
    SymmSeg& alloc_symm(size_t bytes) {
        SymmSeg s;
        s.size = bytes;
        for (int r = 0; r < P; ++r) s.base[r] = alloc_zero(bytes);
        if (multicast) {
            void* w = mmap(nullptr, bytes, PROT_NONE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
            if (w == MAP_FAILED) abort();
            s.mc = static_cast<char*>(w);
            McWindow mw;
            mw.base = s.mc;
            mw.size = bytes;
            mw.world = P;
            for (int r = 0; r < P; ++r) mw.rank_base[r] = s.base[r];
            windows().push_back(mw);
        }
        segs.push_back(s);
        return segs.back();
    }

    int segment_of(int rank, const void* ptr, size_t bytes) const {
        const char* p = static_cast<const char*>(ptr);
        for (size_t i = 0; i < segs.size(); ++i)
            if (p >= segs[i].base[rank] && p + bytes <= segs[i].base[rank] + segs[i].size) return static_cast<int>(i);
        return -1;
    }

    void host_barrier() {
        std::unique_lock<std::mutex> lk(m);
        const unsigned g = gen;
        if (++arrived == P) {
            arrived = 0;
            ++gen;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return gen != g; });
        }
    }

    // run f(rank, comm) on one host thread per rank
    void run(const std::function<void(int, dlbb_comm*)>& f) {
        std::vector<std::thread> th;
        for (int r = 0; r < P; ++r) th.emplace_back([&, r] { f(r, comm[r].get()); });
        for (auto& t : th) t.join();
    }
};

inline World*& world() {
    static World* w = nullptr;
    return w;
}
inline std::string& last_error() {
    static thread_local std::string e;
    return e;
}

}  // namespace emu

// ------------------------------------------------------------------ what collectives.cu links against
namespace dlbb {
std::atomic<uint64_t> g_launch_count{0};
void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    emu::last_error() = buf;
}
int symm_segment_of(const dlbb_comm* c, const void* ptr, size_t bytes) { return emu::world()->segment_of(c->rank, ptr, bytes); }

// Collective: every rank publishes its buffer, then builds the same PeerView.  A null buffer ("not usable on this rank")
// makes all ranks fall back together, as in peer_registry.cpp.
int peer_registry_resolve(dlbb_comm* c, PeerRegistry*, const void* ptr, size_t bytes, PeerView* pv, bool* all_ok) {
    emu::World& w = *emu::world();
    w.recs[c->rank] = {ptr, bytes};
    w.host_barrier();
    bool ok = true;
    int seg0 = -2;
    size_t off0 = 0;
    bool symmetric = true;
    for (int r = 0; r < w.P; ++r) {
        pv->ptr[r] = static_cast<char*>(const_cast<void*>(w.recs[r].ptr));
        ok = ok && w.recs[r].ptr != nullptr && w.recs[r].bytes == bytes;
        if (!w.recs[r].ptr) { symmetric = false; continue; }
        const int s = w.segment_of(r, w.recs[r].ptr, bytes);
        const size_t off = s >= 0 ? static_cast<size_t>(static_cast<const char*>(w.recs[r].ptr) - w.segs[s].base[r]) : 0;
        if (seg0 == -2) { seg0 = s; off0 = off; }
        symmetric = symmetric && s >= 0 && s == seg0 && off == off0;
    }
    pv->mc = (ok && symmetric && w.multicast) ? w.segs[seg0].mc + off0 : nullptr;
    *all_ok = ok;
    w.host_barrier();   // nobody overwrites a record before everyone has read it
    return 0;
}
}  // namespace dlbb

extern "C" {
int dlbb_comm_has_multicast(const dlbb_comm_t* c) { return c->mc_supported ? 1 : 0; }
// symmetric segments by id (the fused kernels take segment ids): index into the emulated box's segment list
void* dlbb_symm_ptr(dlbb_comm_t*, int seg, int peer) {
    emu::World& w = *emu::world();
    return seg >= 0 && seg < (int)w.segs.size() && peer >= 0 && peer < w.P ? w.segs[seg].base[peer] : nullptr;
}
void* dlbb_symm_mc_ptr(dlbb_comm_t*, int seg) {
    emu::World& w = *emu::world();
    return seg >= 0 && seg < (int)w.segs.size() ? w.segs[seg].mc : nullptr;
}
size_t dlbb_symm_size(dlbb_comm_t*, int seg) {
    emu::World& w = *emu::world();
    return seg >= 0 && seg < (int)w.segs.size() ? w.segs[seg].size : 0;
}
const char* cudaGetErrorString(cudaError_t) { return "emulated"; }
cudaError_t cudaGetLastError(void) { return cudaSuccess; }
cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) {   // the rank's stream is its host thread: runs in order
    memset(p, v, n);
    return cudaSuccess;
}
}
