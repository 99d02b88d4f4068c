// Zero-copy collectives on ordinary device allocations ("user buffer registration").
//
// The staged collectives copy the user's tensor into symmetric staging memory and back; for large
// messages those two HBM passes cost as much as the NVLink phase.  This file lets the kernels read
// and write the peers' *own* tensors instead: each rank exports the cudaMalloc allocation behind its
// tensor as a CUDA IPC handle, the handles are exchanged, every rank opens the peers' allocations
// once (cached) and the collective receives one pointer per rank to the same logical buffer.
//
// The exchange happens on the host at enqueue time, once per large collective, through a tiny
// shared-memory mailbox (mailbox.h: one memfd per rank, mapped by everyone at communicator creation;
// rank r publishes {sequence, handle, offset} in its own slot and spins until every peer's slot carries
// the same sequence number - about a microsecond when the peers are already there).  All ranks see the
// same set of "ok" flags, so they take the same decision (zero-copy or staged) without another round.
//
// No counterpart in the reference (its tensors live in host memory and go through MPI / Gloo / oneCCL);
// the closest analogue is NCCL's user-buffer registration.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "comm_internal.h"
#include "mailbox.h"

namespace dlbb {
void set_error(const char* fmt, ...);  // common.cu

namespace {

// What a rank tells the others about its buffer (one HostMailbox record).
//   ok == 1: a cudaMalloc block, exported as a CUDA IPC handle
//   ok == 2: memory inside one of the communicator's symmetric segments (dlbb_symm_alloc): every peer already has
//            that segment mapped, so {segment id, offset} is all it needs - and when all ranks name the same
//            segment and offset the buffer can also be addressed through the segment's multicast mapping
struct BufferRecord {
    uint32_t ok;
    uint32_t segment;       // ok == 2: segment id
    uint64_t offset;        // of the tensor inside its allocation / segment
    uint64_t alloc_base;    // address of the allocation in the owner's address space (cache key)
    uint64_t alloc_size;
    unsigned char handle[sizeof(cudaIpcMemHandle_t)];
};
enum : uint32_t { kBufNone = 0, kBufIpc = 1, kBufSymm = 2 };
static_assert(sizeof(BufferRecord) <= HostMailbox::kPayload, "record must fit a mailbox slot");

struct PeerMapping {
    unsigned char handle[sizeof(cudaIpcMemHandle_t)];
    uint64_t remote_base, size;
    char* local;            // where the peer's allocation is mapped in this process
};

typedef CUresult (*GetAddressRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
GetAddressRangeFn get_address_range_fn() {
    static GetAddressRangeFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &sym, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<GetAddressRangeFn>(sym);
    }
    return fn;
}

}  // namespace

struct PeerRegistry {
    int rank = 0, world = 1;
    HostMailbox mailbox;
    std::vector<PeerMapping> cache[DLBB_MAX_RANKS];
    bool usable = false;
};

// Which live symmetric segment (if any) holds [ptr, ptr + bytes) of this rank's memory.
static int find_symm_segment(const dlbb_comm* c, const void* ptr, size_t bytes, uint64_t* offset) {
    const CUdeviceptr a = reinterpret_cast<CUdeviceptr>(ptr);
    for (int i = 0; i < static_cast<int>(c->segs.size()); ++i) {
        const Segment& s = c->segs[i];
        if (!s.live || static_cast<int>(s.va.size()) <= c->rank) continue;
        const CUdeviceptr base = s.va[c->rank];
        if (a >= base && a + bytes <= base + s.size) {
            *offset = a - base;
            return i;
        }
    }
    return -1;
}

int symm_segment_of(const dlbb_comm* c, const void* ptr, size_t bytes) {
    uint64_t off;
    return ptr ? find_symm_segment(c, ptr, bytes, &off) : -1;
}

// Creates the mailbox (collective over the bootstrap) and checks once, with a scratch allocation, that
// IPC export / import works between all ranks of this job; on any failure zero-copy stays off everywhere.
PeerRegistry* peer_registry_create(dlbb_comm* c) {
    PeerRegistry* r = new PeerRegistry();
    r->rank = c->rank; r->world = c->world;
    if (c->world == 1) return r;
    if (!r->mailbox.create(*c->bs)) return r;
    void* probe = nullptr;
    const bool good = cudaMalloc(&probe, 1 << 20) == cudaSuccess;
    PeerView pv;
    bool all_ok = false;
    r->usable = true;   // resolve() below needs it
    if (peer_registry_resolve(c, r, good ? probe : nullptr, 1 << 20, &pv, &all_ok) != 0) all_ok = false;
    char mine = all_ok ? 1 : 0;   // an import can fail on one rank only: agree once more
    std::vector<char> all(c->world);
    c->bs->allgather(&mine, all.data(), 1);
    for (char v : all) all_ok = all_ok && v;
    r->usable = all_ok;
    for (int p = 0; p < c->world; ++p) {   // drop the probe mappings again
        for (auto& m : r->cache[p]) cudaIpcCloseMemHandle(m.local);
        r->cache[p].clear();
    }
    c->bs->barrier();
    if (probe) cudaFree(probe);
    cudaGetLastError();
    return r;
}

void peer_registry_destroy(PeerRegistry* r) {
    if (!r) return;
    for (int p = 0; p < r->world; ++p)
        for (auto& m : r->cache[p]) cudaIpcCloseMemHandle(m.local);
    cudaGetLastError();
    delete r;
}

bool peer_registry_usable(const PeerRegistry* r) { return r && r->usable; }

// Collective and host-blocking: every rank calls it with its own buffer, in the same order.
// *all_ok is the same on every rank; when false, pv is not filled and the caller falls back to staging.
// Two mailbox rounds: the buffer records, then each rank's import status - an import (cudaIpcOpenMemHandle) can fail
// on one rank only, and that rank must not be the only one that falls back to staging (its peers would wait in the
// zero-copy kernel until the device watchdog fires).  An error is returned only when an exchange itself fails; the
// registry is then unusable (the sequence numbers of the ranks no longer agree).
int peer_registry_resolve(dlbb_comm* c, PeerRegistry* r, const void* ptr, size_t bytes, PeerView* pv, bool* all_ok) {
    *all_ok = false;
    if (!r || !r->usable || r->world == 1) return 0;
    pv->mc = nullptr;
    // ---- describe my buffer: a symmetric segment, or an exportable cudaMalloc block
    BufferRecord mine;
    memset(&mine, 0, sizeof(mine));
    cudaIpcMemHandle_t h;
    memset(&h, 0, sizeof(h));
    uint64_t symm_off = 0;
    const int seg = ptr ? find_symm_segment(c, ptr, bytes, &symm_off) : -1;
    if (seg >= 0) {
        mine.ok = kBufSymm;
        mine.segment = static_cast<uint32_t>(seg);
        mine.offset = symm_off;
    } else {
        CUdeviceptr base = 0;
        size_t asize = 0;
        bool ok = ptr != nullptr && get_address_range_fn() != nullptr &&
                  get_address_range_fn()(&base, &asize, reinterpret_cast<CUdeviceptr>(ptr)) == CUDA_SUCCESS &&
                  reinterpret_cast<CUdeviceptr>(ptr) + bytes <= base + asize;
        if (ok && cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(base)) != cudaSuccess) {
            cudaGetLastError();   // VMM / pool memory cannot be exported this way: not an error, just not zero-copy
            ok = false;
        }
        mine.ok = ok ? kBufIpc : kBufNone;
        mine.offset = ok ? reinterpret_cast<CUdeviceptr>(ptr) - base : 0;
        mine.alloc_base = base;
        mine.alloc_size = asize;
        memcpy(mine.handle, &h, sizeof(h));
    }
    // ---- round 1: everybody's record
    const void* recs[HostMailbox::kMaxRanks];
    int missing = -1;
    if (!r->mailbox.exchange(&mine, sizeof(mine), recs, 120.0, &missing)) {
        r->usable = false;
        set_error("zero-copy exchange %llu: rank %d did not arrive within 120 s (ranks must issue large collectives in "
                  "the same order); zero-copy is now off on this rank", (unsigned long long)r->mailbox.sequence(), missing);
        return -1;
    }
    BufferRecord all[HostMailbox::kMaxRanks];
    bool everyone = true, same_symm = true;
    for (int p = 0; p < r->world; ++p) {
        memcpy(&all[p], recs[p], sizeof(BufferRecord));
        everyone = everyone && all[p].ok != kBufNone;
        same_symm = same_symm && all[p].ok == kBufSymm && all[p].segment == all[0].segment && all[p].offset == all[0].offset;
    }
    if (!everyone) return 0;   // the same verdict on every rank: no second round needed
    // ---- map the peers' buffers (IPC mappings are cached by handle)
    bool import_ok = true;
    char import_err[256] = "";
    for (int p = 0; p < r->world && import_ok; ++p) {
        if (p == r->rank) { pv->ptr[p] = static_cast<char*>(const_cast<void*>(ptr)); continue; }
        const BufferRecord& s = all[p];
        if (s.ok == kBufSymm) {
            const bool known = s.segment < c->segs.size() && c->segs[s.segment].live &&
                               s.offset + bytes <= c->segs[s.segment].size;
            if (!known) { import_ok = false; snprintf(import_err, sizeof(import_err), "rank %d names an unknown symmetric segment", p); break; }
            pv->ptr[p] = reinterpret_cast<char*>(c->segs[s.segment].va[p]) + s.offset;
            continue;
        }
        char* local = nullptr;
        auto& cache = r->cache[p];
        for (size_t i = 0; i < cache.size(); ++i) {
            if (memcmp(cache[i].handle, s.handle, sizeof(s.handle)) == 0) {
                local = cache[i].local;
                if (i + 1 != cache.size()) {   // most recently used mapping goes to the back
                    PeerMapping pm = cache[i];
                    cache.erase(cache.begin() + i);
                    cache.push_back(pm);
                }
                break;
            }
            if (cache[i].remote_base == s.alloc_base) {
                // same address, different handle: the peer freed and re-allocated - drop the stale mapping
                cudaIpcCloseMemHandle(cache[i].local);
                cache.erase(cache.begin() + i);
                break;
            }
        }
        if (!local) {
            cudaIpcMemHandle_t ph;
            memcpy(&ph, s.handle, sizeof(ph));
            void* m = nullptr;
            const cudaError_t e = cudaIpcOpenMemHandle(&m, ph, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                import_ok = false;
                snprintf(import_err, sizeof(import_err), "cudaIpcOpenMemHandle of rank %d's buffer failed: %s", p, cudaGetErrorString(e));
                break;
            }
            // Bound the foreign mappings per peer, least recently used first (a mapping keeps the peer's block
            // alive after the peer has freed it).  Rare: drain the GPU before unmapping.
            size_t mapped = 0;
            for (auto& pm : cache) mapped += pm.size;
            while (!cache.empty() && (cache.size() >= 64 || mapped + s.alloc_size > (static_cast<size_t>(64) << 30))) {
                cudaDeviceSynchronize();
                mapped -= cache.front().size;
                cudaIpcCloseMemHandle(cache.front().local);
                cache.erase(cache.begin());
            }
            PeerMapping pm;
            memcpy(pm.handle, s.handle, sizeof(pm.handle));
            pm.remote_base = s.alloc_base; pm.size = s.alloc_size; pm.local = static_cast<char*>(m);
            cache.push_back(pm);
            local = pm.local;
        }
        pv->ptr[p] = local + s.offset;
    }
    // ---- round 2: did everybody manage to map everybody?
    BufferRecord status;
    memset(&status, 0, sizeof(status));
    status.ok = import_ok ? 1 : 0;
    if (!r->mailbox.exchange(&status, sizeof(status), recs, 120.0, &missing)) {
        r->usable = false;
        set_error("zero-copy status exchange %llu: rank %d did not arrive within 120 s; zero-copy is now off on this rank",
                  (unsigned long long)r->mailbox.sequence(), missing);
        return -1;
    }
    bool all_imported = true;
    for (int p = 0; p < r->world; ++p) all_imported = all_imported && static_cast<const BufferRecord*>(recs[p])->ok != 0;
    if (!all_imported) {
        if (!import_ok && getenv("DLBB_LOG_LEVEL")) fprintf(stderr, "dlbb: rank %d: zero-copy import failed (%s); all ranks stage this call\n", r->rank, import_err);
        return 0;   // everybody falls back to staging together
    }
    if (same_symm && c->segs[all[0].segment].mc_va)
        pv->mc = reinterpret_cast<char*>(c->segs[all[0].segment].mc_va) + all[0].offset;
    *all_ok = true;
    return 0;
}

}  // namespace dlbb
