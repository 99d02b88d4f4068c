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
#include <string.h>

#include "comm_internal.h"
#include "mailbox.h"

namespace dlbb {
void set_error(const char* fmt, ...);  // common.cu

namespace {

// What a rank tells the others about its buffer (one HostMailbox record).
struct BufferRecord {
    uint32_t ok;
    uint32_t reserved;
    uint64_t offset;        // of the tensor inside its allocation
    uint64_t alloc_base;    // address of the allocation in the owner's address space (cache key)
    uint64_t alloc_size;
    unsigned char handle[sizeof(cudaIpcMemHandle_t)];
};
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
int peer_registry_resolve(dlbb_comm* c, PeerRegistry* r, const void* ptr, size_t bytes, PeerView* pv, bool* all_ok) {
    (void)c;
    *all_ok = false;
    if (!r || !r->usable || r->world == 1) return 0;
    // ---- export my allocation
    BufferRecord mine;
    memset(&mine, 0, sizeof(mine));
    cudaIpcMemHandle_t h;
    memset(&h, 0, sizeof(h));
    CUdeviceptr base = 0;
    size_t asize = 0;
    bool ok = ptr != nullptr && get_address_range_fn() != nullptr &&
              get_address_range_fn()(&base, &asize, reinterpret_cast<CUdeviceptr>(ptr)) == CUDA_SUCCESS &&
              reinterpret_cast<CUdeviceptr>(ptr) + bytes <= base + asize;
    if (ok && cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(base)) != cudaSuccess) {
        cudaGetLastError();   // VMM / pool memory cannot be exported this way: not an error, just not zero-copy
        ok = false;
    }
    mine.ok = ok ? 1 : 0;
    mine.offset = ok ? reinterpret_cast<CUdeviceptr>(ptr) - base : 0;
    mine.alloc_base = base;
    mine.alloc_size = asize;
    memcpy(mine.handle, &h, sizeof(h));
    // ---- everybody's record of this exchange
    const void* recs[HostMailbox::kMaxRanks];
    int missing = -1;
    if (!r->mailbox.exchange(&mine, sizeof(mine), recs, 120.0, &missing)) {
        set_error("zero-copy exchange %llu: rank %d did not arrive within 120 s (ranks must issue large collectives in "
                  "the same order)", (unsigned long long)r->mailbox.sequence(), missing);
        return -1;
    }
    bool everyone = true;
    for (int p = 0; p < r->world; ++p) everyone = everyone && static_cast<const BufferRecord*>(recs[p])->ok;
    if (!everyone) return 0;
    // ---- map the peers' allocations (cached by handle)
    for (int p = 0; p < r->world; ++p) {
        if (p == r->rank) { pv->ptr[p] = static_cast<char*>(const_cast<void*>(ptr)); continue; }
        BufferRecord s;
        memcpy(&s, recs[p], sizeof(s));
        char* local = nullptr;
        auto& cache = r->cache[p];
        for (size_t i = 0; i < cache.size(); ++i) {
            if (memcmp(cache[i].handle, s.handle, sizeof(s.handle)) == 0) { local = cache[i].local; break; }
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
                set_error("zero-copy: cudaIpcOpenMemHandle of rank %d's buffer failed: %s", p, cudaGetErrorString(e));
                return -1;
            }
            if (cache.size() >= 64) {   // bound the number of foreign mappings (rare: drain the GPU first)
                cudaDeviceSynchronize();
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
    *all_ok = true;
    return 0;
}

}  // namespace dlbb
