// Communicator life cycle and symmetric memory over the CUDA virtual-memory-management API.
//
// Each symmetric segment is one cuMemCreate allocation per rank, exported as a POSIX file
// descriptor, passed to every peer over the bootstrap sockets (SCM_RIGHTS), imported and mapped
// into every process, so a kernel on rank r can dereference rank p's copy directly (NVLink P2P
// through NVSwitch).  When the devices support it the same physical memory is also bound to an
// NVLS multicast object whose mapping lets `multimem.ld_reduce` / `multimem.st` reduce and
// broadcast inside the switch.  (SURVEY.md 5.8 "Bootstrap"; no counterpart in the reference,
// whose substrate is libmpi / Gloo / oneCCL shared memory.)
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <fcntl.h>
#include <unistd.h>

#include <string>

#include "comm_internal.h"

namespace dlbb {
void set_error(const char* fmt, ...);  // common.cu
}
using dlbb::set_error;

namespace {

// ---- driver entry points, resolved through the runtime so libcuda is not a link dependency
struct Driver {
    decltype(&cuMemCreate) MemCreate = nullptr;
    decltype(&cuMemRelease) MemRelease = nullptr;
    decltype(&cuMemExportToShareableHandle) MemExport = nullptr;
    decltype(&cuMemImportFromShareableHandle) MemImport = nullptr;
    decltype(&cuMemAddressReserve) AddressReserve = nullptr;
    decltype(&cuMemAddressFree) AddressFree = nullptr;
    decltype(&cuMemMap) MemMap = nullptr;
    decltype(&cuMemUnmap) MemUnmap = nullptr;
    decltype(&cuMemSetAccess) MemSetAccess = nullptr;
    decltype(&cuMemGetAllocationGranularity) GetGranularity = nullptr;
    decltype(&cuMulticastCreate) McCreate = nullptr;
    decltype(&cuMulticastAddDevice) McAddDevice = nullptr;
    decltype(&cuMulticastBindMem) McBindMem = nullptr;
    decltype(&cuMulticastGetGranularity) McGetGranularity = nullptr;
    decltype(&cuMulticastUnbind) McUnbind = nullptr;
    decltype(&cuDeviceGet) DeviceGet = nullptr;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
    decltype(&cuGetErrorString) GetErrorString = nullptr;
    bool ok = false;
};

template <typename T>
bool resolve(const char* name, T& fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
        sym == nullptr)
        return false;
    fn = reinterpret_cast<T>(sym);
    return true;
}

Driver& drv() {
    static Driver d;
    static bool tried = false;
    if (!tried) {
        tried = true;
        bool ok = true;
        ok &= resolve("cuMemCreate", d.MemCreate);
        ok &= resolve("cuMemRelease", d.MemRelease);
        ok &= resolve("cuMemExportToShareableHandle", d.MemExport);
        ok &= resolve("cuMemImportFromShareableHandle", d.MemImport);
        ok &= resolve("cuMemAddressReserve", d.AddressReserve);
        ok &= resolve("cuMemAddressFree", d.AddressFree);
        ok &= resolve("cuMemMap", d.MemMap);
        ok &= resolve("cuMemUnmap", d.MemUnmap);
        ok &= resolve("cuMemSetAccess", d.MemSetAccess);
        ok &= resolve("cuMemGetAllocationGranularity", d.GetGranularity);
        ok &= resolve("cuDeviceGet", d.DeviceGet);
        ok &= resolve("cuDeviceGetAttribute", d.DeviceGetAttribute);
        ok &= resolve("cuGetErrorString", d.GetErrorString);
        // multicast is optional
        resolve("cuMulticastCreate", d.McCreate);
        resolve("cuMulticastAddDevice", d.McAddDevice);
        resolve("cuMulticastBindMem", d.McBindMem);
        resolve("cuMulticastGetGranularity", d.McGetGranularity);
        resolve("cuMulticastUnbind", d.McUnbind);
        d.ok = ok;
    }
    return d;
}

const char* cu_err(CUresult r) {
    const char* s = nullptr;
    if (drv().GetErrorString && drv().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "unknown CUDA driver error";
}

#define CU_TRY(call, what)                                                              \
    do {                                                                                \
        CUresult _r = (call);                                                           \
        if (_r != CUDA_SUCCESS) {                                                       \
            set_error("%s failed: %s (rank %d)", what, cu_err(_r), c->rank);            \
            return -1;                                                                  \
        }                                                                               \
    } while (0)

size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

// Agree on a boolean across ranks: true only if every rank says true.
bool all_agree(dlbb_comm* c, bool mine) {
    std::vector<char> all(c->world);
    char m = mine ? 1 : 0;
    if (!c->bs->allgather(&m, all.data(), 1)) return false;
    for (char v : all)
        if (!v) return false;
    return true;
}

int try_multicast(dlbb_comm* c, dlbb::Segment& s) {
    Driver& d = drv();
    CUmulticastObjectProp mp{};
    mp.numDevices = static_cast<unsigned>(c->world);
    mp.size = s.size;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mp.flags = 0;
    CUdevice cudev;
    bool ok = d.DeviceGet(&cudev, c->device) == CUDA_SUCCESS;
    // stage 1: root creates and exports, everyone imports
    int fd = -1;
    CUmemGenericAllocationHandle mc = 0;
    if (ok && c->rank == 0) {
        ok = d.McCreate(&mc, &mp) == CUDA_SUCCESS &&
             d.MemExport(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
    }
    if (fd < 0) fd = open("/dev/null", O_RDONLY | O_CLOEXEC);  // placeholder from ranks with nothing to share
    std::vector<int> fds;
    const bool xok = c->bs->allgather_fds(fd, fds);
    if (fd >= 0) close(fd);
    if (!xok) return -1;
    if (ok && c->rank != 0)
        ok = d.MemImport(&mc, reinterpret_cast<void*>(static_cast<intptr_t>(fds[0])),
                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
    for (int f : fds)
        if (f >= 0) close(f);
    if (!all_agree(c, ok)) { if (mc) d.MemRelease(mc); return 1; }
    // stage 2: every device joins, then binds its physical memory
    ok = d.McAddDevice(mc, cudev) == CUDA_SUCCESS;
    if (!all_agree(c, ok)) { d.MemRelease(mc); return 1; }
    ok = d.McBindMem(mc, 0, s.local, 0, s.size, 0) == CUDA_SUCCESS;
    if (!all_agree(c, ok)) { d.MemRelease(mc); return 1; }
    // stage 3: map the multicast object
    CUdeviceptr va = 0;
    ok = d.AddressReserve(&va, s.size, c->gran, 0, 0) == CUDA_SUCCESS;
    if (ok) ok = d.MemMap(va, s.size, 0, mc, 0) == CUDA_SUCCESS;
    if (ok) {
        CUmemAccessDesc ad{};
        ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ad.location.id = c->device;
        ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        ok = d.MemSetAccess(va, s.size, &ad, 1) == CUDA_SUCCESS;
    }
    if (!all_agree(c, ok)) return 1;  // leak the partial mapping; multicast simply stays off
    s.mc = mc;
    s.mc_va = va;
    return 0;
}

void fill_dev(dlbb_comm* c) {
    const char* wd = getenv("DLBB_WATCHDOG_SECONDS");
    const double wd_s = wd && atof(wd) > 0 ? atof(wd) : 60.0;
    c->dev.watchdog_ns = static_cast<long long>(wd_s * 1e9);
    c->dev.rank = c->rank;
    c->dev.world = c->world;
    for (int p = 0; p < DLBB_MAX_RANKS; ++p) {
        c->dev.pads[p] = nullptr;
        c->dev.stage[p] = nullptr;
    }
    if (c->pad_seg >= 0)
        for (int p = 0; p < c->world; ++p)
            c->dev.pads[p] = reinterpret_cast<uint32_t*>(c->segs[c->pad_seg].va[p]);
    if (c->stage_seg >= 0)
        for (int p = 0; p < c->world; ++p) c->dev.stage[p] = reinterpret_cast<char*>(c->segs[c->stage_seg].va[p]);
    for (int p = 0; p < DLBB_MAX_RANKS; ++p)
        c->dev.tile_flags[p] = (c->flag_seg >= 0 && p < c->world) ? reinterpret_cast<uint32_t*>(c->segs[c->flag_seg].va[p]) : nullptr;
    c->dev.pads_mc = c->pad_seg >= 0 ? reinterpret_cast<uint32_t*>(c->segs[c->pad_seg].mc_va) : nullptr;
    c->dev.stage_mc = c->stage_seg >= 0 ? reinterpret_cast<char*>(c->segs[c->stage_seg].mc_va) : nullptr;
}

}  // namespace

extern "C" {

int dlbb_symm_alloc(dlbb_comm_t* c, size_t bytes, int* seg_id) {
    Driver& d = drv();
    if (!d.ok) { set_error("CUDA VMM driver entry points are not available"); return -1; }
    if (bytes == 0) { set_error("symm_alloc: zero bytes"); return -2; }
    dlbb::Segment s;
    s.size = round_up(bytes, c->gran);
    CUmemAllocationProp prop{};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = c->device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CU_TRY(d.MemCreate(&s.local, s.size, &prop, 0), "cuMemCreate");
    int fd = -1;
    CU_TRY(d.MemExport(&fd, s.local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
    std::vector<int> fds;
    const bool xok = c->bs->allgather_fds(fd, fds);
    close(fd);
    if (!xok) { set_error("symm_alloc: fd exchange failed: %s", c->bs->error().c_str()); return -1; }
    s.handles.assign(c->world, 0);
    s.va.assign(c->world, 0);
    CUmemAccessDesc ad{};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = c->device;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    for (int p = 0; p < c->world; ++p) {
        if (p == c->rank) {
            s.handles[p] = s.local;
        } else {
            CU_TRY(d.MemImport(&s.handles[p], reinterpret_cast<void*>(static_cast<intptr_t>(fds[p])),
                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
                   "cuMemImportFromShareableHandle");
        }
        CU_TRY(d.AddressReserve(&s.va[p], s.size, c->gran, 0, 0), "cuMemAddressReserve");
        CU_TRY(d.MemMap(s.va[p], s.size, 0, s.handles[p], 0), "cuMemMap");
        CU_TRY(d.MemSetAccess(s.va[p], s.size, &ad, 1), "cuMemSetAccess (is NVLink peer access available?)");
    }
    for (int f : fds)
        if (f >= 0) close(f);
    if (cudaMemset(reinterpret_cast<void*>(s.va[c->rank]), 0, s.size) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
        set_error("symm_alloc: zero-fill failed: %s", cudaGetErrorString(cudaGetLastError()));
        return -1;
    }
    if (c->mc_supported && c->world > 1) {
        const int rc = try_multicast(c, s);
        if (rc < 0) { set_error("symm_alloc: multicast exchange failed: %s", c->bs->error().c_str()); return -1; }
        if (rc > 0) c->mc_supported = false;  // some rank could not set it up: P2P only from now on
    }
    if (!c->bs->barrier()) { set_error("symm_alloc: barrier failed: %s", c->bs->error().c_str()); return -1; }
    s.live = true;
    c->segs.push_back(s);
    *seg_id = static_cast<int>(c->segs.size()) - 1;
    return 0;
}

void* dlbb_symm_ptr(dlbb_comm_t* c, int seg, int peer) {
    if (seg < 0 || seg >= (int)c->segs.size() || !c->segs[seg].live || peer < 0 || peer >= c->world) return nullptr;
    return reinterpret_cast<void*>(c->segs[seg].va[peer]);
}

void* dlbb_symm_mc_ptr(dlbb_comm_t* c, int seg) {
    if (seg < 0 || seg >= (int)c->segs.size() || !c->segs[seg].live) return nullptr;
    return reinterpret_cast<void*>(c->segs[seg].mc_va);
}

size_t dlbb_symm_size(dlbb_comm_t* c, int seg) {
    if (seg < 0 || seg >= (int)c->segs.size() || !c->segs[seg].live) return 0;
    return c->segs[seg].size;
}

int dlbb_symm_free(dlbb_comm_t* c, int seg) {
    Driver& d = drv();
    if (seg < 0 || seg >= (int)c->segs.size() || !c->segs[seg].live) { set_error("symm_free: bad segment"); return -2; }
    dlbb::Segment& s = c->segs[seg];
    cudaDeviceSynchronize();
    c->bs->barrier();  // nobody may still be touching a peer's copy
    if (s.mc_va) {
        d.MemUnmap(s.mc_va, s.size);
        d.AddressFree(s.mc_va, s.size);
        CUdevice cudev;
        if (d.McUnbind && d.DeviceGet(&cudev, c->device) == CUDA_SUCCESS) d.McUnbind(s.mc, cudev, 0, s.size);
        d.MemRelease(s.mc);
    }
    for (int p = 0; p < c->world; ++p) {
        if (s.va[p]) {
            d.MemUnmap(s.va[p], s.size);
            d.AddressFree(s.va[p], s.size);
        }
        if (s.handles[p]) d.MemRelease(s.handles[p]);
    }
    s.live = false;
    c->bs->barrier();
    return 0;
}

int dlbb_comm_create(int rank, int world, int device, const char* job, size_t staging_bytes, dlbb_comm_t** out) {
    if (world < 1 || world > DLBB_MAX_RANKS || rank < 0 || rank >= world) {
        set_error("comm_create: bad rank/world %d/%d (max %d ranks)", rank, world, DLBB_MAX_RANKS);
        return -2;
    }
    if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) {
        set_error("comm_create: cannot initialise CUDA device %d", device);
        return -1;
    }
    Driver& d = drv();
    if (!d.ok) { set_error("comm_create: CUDA VMM driver entry points are not available"); return -1; }
    dlbb_comm* c = new dlbb_comm();
    c->rank = rank; c->world = world; c->device = device;
    c->bs = new dlbb::Bootstrap(rank, world, job ? job : "default");
    const char* to = getenv("DLBB_BOOTSTRAP_TIMEOUT");
    if (!c->bs->connect(to ? atof(to) : 120.0)) {
        set_error("comm_create: %s", c->bs->error().c_str());
        delete c->bs; delete c;
        return -1;
    }
    // allocation granularity (and multicast granularity if usable)
    CUmemAllocationProp prop{};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t g = 0;
    if (d.GetGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g) c->gran = g;
    bool mc = world > 1 && d.McCreate && d.McAddDevice && d.McBindMem && d.McGetGranularity &&
              getenv("DLBB_DISABLE_MULTICAST") == nullptr;
    if (mc) {
        CUdevice cudev;
        int sup = 0;
        mc = d.DeviceGet(&cudev, device) == CUDA_SUCCESS &&
             d.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) == CUDA_SUCCESS && sup;
        if (mc) {
            CUmulticastObjectProp mp{};
            mp.numDevices = static_cast<unsigned>(world);
            mp.size = c->gran;
            mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
            size_t mg = 0;
            if (d.McGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg)
                c->gran = c->gran > mg ? c->gran : mg;
            else
                mc = false;
        }
    }
    c->mc_supported = all_agree(c, mc);
    // agree on the granularity (max over ranks)
    {
        std::vector<uint64_t> all(world);
        uint64_t mine = c->gran;
        c->bs->allgather(&mine, all.data(), sizeof(mine));
        for (uint64_t v : all) if (v > c->gran) c->gran = v;
    }
    if (staging_bytes == 0) {
        const char* e = getenv("DLBB_STAGING_MB");
        staging_bytes = (e ? static_cast<size_t>(atol(e)) : 256) << 20;   // fewer, longer launches for large messages
    }
    c->stage_bytes = round_up(staging_bytes, 1 << 20);
    int rc = dlbb_symm_alloc(c, dlbb::kPadBytes, &c->pad_seg);
    if (rc == 0) rc = dlbb_symm_alloc(c, 3 * c->stage_bytes, &c->stage_seg);
    if (rc == 0) rc = dlbb_symm_alloc(c, dlbb::kTileFlagBytes, &c->flag_seg);
    if (rc != 0) { delete c->bs; delete c; return rc; }
    fill_dev(c);
    const char* e = getenv("DLBB_CTAS");
    if (e) c->ctas = atoi(e);
    {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0)
            c->max_resident_ctas = sms < 128 ? sms : 128;
        e = getenv("DLBB_MAX_RESIDENT_CTAS");
        if (e && atoi(e) > 0) c->max_resident_ctas = atoi(e);
    }
    e = getenv("DLBB_ONESHOT_MAX_BYTES");
    if (e) c->oneshot_max = static_cast<size_t>(atol(e));
    e = getenv("DLBB_ALGO");   // analogue of CCL_ALLREDUCE (collectives/3d/launch_dsccl.sh:46)
    if (e) {
        const std::string a(e);
        c->default_algo = a == "oneshot" ? DLBB_ALGO_ONESHOT : a == "twoshot" ? DLBB_ALGO_TWOSHOT
                        : a == "nvls" ? DLBB_ALGO_NVLS : a == "direct" ? DLBB_ALGO_DIRECT : DLBB_ALGO_AUTO;
    }
    c->debug_poison = getenv("DLBB_DEBUG_POISON") != nullptr && atoi(getenv("DLBB_DEBUG_POISON")) != 0;
    // Zero-copy path for large messages on ordinary allocations (peer_registry.cpp); DLBB_ZEROCOPY_MIN_BYTES=0 disables.
    e = getenv("DLBB_ZEROCOPY_MIN_BYTES");
    c->zero_copy_min = e ? static_cast<size_t>(atoll(e)) : (static_cast<size_t>(16) << 20);
    if (world > 1 && c->zero_copy_min > 0) {
        c->reg = dlbb::peer_registry_create(c);
        if (!dlbb::peer_registry_usable(c->reg)) c->zero_copy_min = 0;
    } else {
        c->zero_copy_min = 0;
    }
    if (getenv("DLBB_LOG_LEVEL") && rank == 0)
        fprintf(stderr, "dlbb: comm up, %d ranks, granularity %zu MiB, staging 2x%zu MiB, multicast %s, zero-copy %s\n",
                world, c->gran >> 20, c->stage_bytes >> 20, c->mc_supported ? "on" : "off",
                c->zero_copy_min ? "on" : "off");
    *out = c;
    return 0;
}

int dlbb_comm_destroy(dlbb_comm_t* c) {
    if (!c) return 0;
    for (int i = static_cast<int>(c->segs.size()) - 1; i >= 0; --i)
        if (c->segs[i].live) dlbb_symm_free(c, i);
    dlbb::peer_registry_destroy(c->reg);
    delete c->bs;
    delete c;
    return 0;
}

int dlbb_comm_rank(const dlbb_comm_t* c) { return c->rank; }
int dlbb_comm_world(const dlbb_comm_t* c) { return c->world; }
int dlbb_comm_has_multicast(const dlbb_comm_t* c) {
    return c->mc_supported && c->stage_seg >= 0 && c->segs[c->stage_seg].mc_va != 0;
}

int dlbb_comm_host_barrier(dlbb_comm_t* c) {
    if (!c->bs->barrier()) { set_error("host_barrier: %s", c->bs->error().c_str()); return -1; }
    return 0;
}

int dlbb_comm_host_allgather(dlbb_comm_t* c, const void* in, void* out, size_t bytes) {
    if (!c->bs->allgather(in, out, bytes)) { set_error("host_allgather: %s", c->bs->error().c_str()); return -1; }
    return 0;
}

void dlbb_comm_set_ctas(dlbb_comm_t* c, int ctas) { c->ctas = ctas < 0 ? 0 : (ctas > dlbb::kMaxCtas ? dlbb::kMaxCtas : ctas); }
void dlbb_comm_set_oneshot_max_bytes(dlbb_comm_t* c, size_t bytes) { c->oneshot_max = bytes; }
void dlbb_comm_set_zero_copy_min_bytes(dlbb_comm_t* c, size_t bytes) {
    c->zero_copy_min = dlbb::peer_registry_usable(c->reg) ? bytes : 0;
}
size_t dlbb_comm_zero_copy_min_bytes(const dlbb_comm_t* c) { return c->zero_copy_min; }

}  // extern "C"
