// Internal layout of a communicator, shared by the bootstrap/VMM code (host C++) and the
// collective launchers (CUDA).
#pragma once
#include <cuda.h>
#include <stdint.h>

#include <vector>

#include "../../include/dlbb_comm.h"
#include "bootstrap.h"

namespace dlbb {

constexpr int kMaxCtas = 320;         // channels: one signal-pad row per CTA of a collective kernel (two CTAs per SM fit)
constexpr int kPadStride = DLBB_MAX_RANKS;  // uint32 slots per row
// Signal pad layout (uint32 words): [kMaxCtas rows][kPadStride] for kernel barriers, then
// a second bank of the same shape for the fused DP kernel (it runs on a side stream concurrently
// with collectives issued on the main stream, so it must not share rows or epochs with them).
constexpr int kPadWordsPerBank = kMaxCtas * kPadStride;
constexpr int kPadBanks = 6;          // 0: collectives, 1: fused DP kernel, 2: fused GEMM+all-reduce, 3: point-to-point,
                                      // 4/5: pack-done / reduce-done flags of the pipelined all-reduce
constexpr size_t kPadBytes = sizeof(uint32_t) * kPadWordsPerBank * kPadBanks;
// Per-tile arrival flags of the fused GEMM + all-reduce: [tile][source rank], epoch valued.
constexpr size_t kMaxFusedTiles = 16384;
constexpr size_t kTileFlagBytes = sizeof(uint32_t) * kMaxFusedTiles * DLBB_MAX_RANKS;

struct Segment {
    bool live = false;
    size_t size = 0;
    CUmemGenericAllocationHandle local = 0;
    std::vector<CUmemGenericAllocationHandle> handles;  // per rank (own = local)
    std::vector<CUdeviceptr> va;                          // per rank mapping in this process
    CUmemGenericAllocationHandle mc = 0;
    CUdeviceptr mc_va = 0;
};

// What kernels receive by value.
struct CommDev {
    int rank, world;
    uint32_t* pads[DLBB_MAX_RANKS];
    char* stage[DLBB_MAX_RANKS];
    char* stage_mc;       // multicast mapping of the staging segment (nullptr if unavailable)
    uint32_t* pads_mc;
    uint32_t* tile_flags[DLBB_MAX_RANKS];
    long long watchdog_ns;   // device-side flag waits trap after this long (DLBB_WATCHDOG_SECONDS, default 60)
};

// The same logical user buffer on every rank, as mapped in this process (peer_registry.cpp).
// `mc` is the NVLS multicast address of that buffer: only set when every rank's buffer sits at the same offset of
// the same symmetric segment and that segment is bound to a multicast object (multimem.* may then be used on it).
struct PeerView {
    char* ptr[DLBB_MAX_RANKS];
    char* mc;
};
struct PeerRegistry;

}  // namespace dlbb

struct dlbb_comm;
namespace dlbb {
PeerRegistry* peer_registry_create(dlbb_comm* c);
void peer_registry_destroy(PeerRegistry* r);
bool peer_registry_usable(const PeerRegistry* r);
int peer_registry_resolve(dlbb_comm* c, PeerRegistry* r, const void* ptr, size_t bytes, PeerView* pv, bool* all_ok);
// id of the live symmetric segment that holds [ptr, ptr + bytes) of this rank's memory, or -1
int symm_segment_of(const dlbb_comm* c, const void* ptr, size_t bytes);
// Flags are compared wrap-safely (signed 32-bit difference to the awaited epoch), which is only right while every flag is
// within 2^31 of the bank's epoch counter - but a pad row that no launch has used for a long time keeps its old value.
// Every 2^28 epochs (DLBB_FLAG_REFRESH_EVERY) the launchers therefore run a barrier over ALL rows of the bank, which
// rewrites every flag with the current epoch.  Collective, enqueued on `stream`; returns 0 or a negative error code.
int refresh_flag_bank_if_due(dlbb_comm* c, int bank, void* stream);
}  // namespace dlbb

struct dlbb_comm {
    int rank = 0, world = 1, device = 0;
    dlbb::Bootstrap* bs = nullptr;
    std::vector<dlbb::Segment> segs;
    bool mc_supported = false;
    size_t gran = 2u << 20;
    int pad_seg = -1, stage_seg = -1, flag_seg = -1;
    size_t stage_bytes = 0;        // bytes of each staging region: [A][B] alternate for collectives, [C] point-to-point
    uint32_t epoch[dlbb::kPadBanks] = {0, 0, 0, 0, 0, 0};  // last epoch value used on each pad bank
    uint32_t last_refresh[dlbb::kPadBanks] = {0, 0, 0, 0, 0, 0};   // epoch at which all flags of the bank were last rewritten
    uint32_t send_epoch[DLBB_MAX_RANKS] = {};      // point-to-point: messages sent to / received from each peer
    uint32_t recv_epoch[DLBB_MAX_RANKS] = {};
    unsigned parity = 0;           // alternates the staging half used by single-barrier kernels
    int ctas = 0;                  // 0 = auto
    int max_resident_ctas = 128;   // most CTAs a kernel with a GRID barrier may launch: they must all be resident together
                                   // (min(128, SM count) unless DLBB_MAX_RESIDENT_CTAS says otherwise - SM-limited partitions)
    size_t oneshot_max = 0;        // 0 = auto
    int default_algo = 0;          // DLBB_ALGO: what DLBB_ALGO_AUTO resolves to (0 = size-based choice)
    dlbb::PeerRegistry* reg = nullptr;   // zero-copy access to the peers' ordinary allocations
    size_t zero_copy_min = 0;      // messages of at least this many bytes try the zero-copy kernels (0 = never)
    const dlbb_view_t* out_view = nullptr;   // set by the *_v entry points for the duration of one call: strided destination
    int64_t out_base = 0;                    // ... and the logical element offset of the launch being issued inside it
    bool debug_poison = false;     // DLBB_DEBUG_POISON=1: 0xFF-fill a staging region before the kernel that rewrites it
    dlbb::CommDev dev{};
};
