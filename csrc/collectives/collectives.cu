// Collective kernels over symmetric memory (K1-K12 of SURVEY.md section 8).
//
// Every collective is ONE kernel launch on the caller's stream.  Inside the kernel:
//   1. "pack":   each rank copies its contribution from the user tensor (optionally a strided
//                view of up to 4 dims - no .contiguous() pre-copy) into its staging region;
//   2. barrier:  system-scope release/acquire flags on the peers' signal pads;
//   3. "pull":   the rank reads what it needs straight out of the peers' staging regions over
//                NVLink (plain peer loads) and reduces / scatters it into the user output.
// Large all-reduces use the bandwidth-optimal two-shot form (reduce-scatter, then all-gather by
// peer stores or by NVLS multimem.ld_reduce / multimem.st) with a second barrier.
//
// ref (what these replace): collectives/1d/dsccl.py:59-163, collectives/1d/openmpi.py:55-198,
// collectives/3d/dsccl.py:46-118, models.py:95.
#include <algorithm>

#include "../../include/dlbb.h"
#include "../../include/dlbb_comm.h"
#include "coll_device.cuh"
#include "view_math.h"

namespace dlbb {

constexpr int kCollThreads = 256;
constexpr int kU = 8;   // independent 16 B requests a thread keeps in flight (NVLink latency ~2 us)

struct Piece {           // a contiguous run of elements / bytes
    int peer;            // pull: whose staging to read (pack: unused)
    int64_t a_off;       // pack: logical source element offset    pull: offset in the peer's staging
    int64_t b_off;       // pack: offset in my staging              pull: offset in my output
    int64_t n;           // elements (typed kernels) or bytes (move kernel)
};

struct PieceList {
    int count;
    Piece p[DLBB_MAX_RANKS];
};

// Store vector i (elements [i*VE, (i+1)*VE) of the logical, dense order) into a dense or strided destination.
// (the element-wise fall-back is kept out of line: it is the slow path, and inlined into every unrolled store site its
//  four 64-bit divisions per element dominated the size and the compile time of the reduction kernels)
template <typename T>
__device__ __noinline__ void store_view_elems(T* dst, const ViewDev& v, int64_t i, uint4 o) {
    constexpr int VE = 16 / sizeof(T);
    const T* ot = reinterpret_cast<const T*>(&o);
#pragma unroll 1
    for (int e = 0; e < VE; ++e) dst[view_offset(v, i * VE + e)] = ot[e];
}
template <typename T>
__device__ __forceinline__ void store_view_vec(T* __restrict__ dst, const ViewDev& v, int64_t i, const uint4& o, bool vec_ok) {
    constexpr int VE = 16 / sizeof(T);
    if (vec_ok) st_v4(dst + view_offset(v, i * VE), o);
    else store_view_elems<T>(dst, v, i, o);
}

// Copy `bytes` bytes with the whole grid.  16-byte vectors when everything is aligned.
template <bool PEER_SRC>
__device__ __forceinline__ void grid_copy_bytes(char* dst, const char* src, int64_t bytes) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * blockDim.x;
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
        const int64_t nvec = bytes >> 4;
        int64_t i = tid;
        for (; i + (kU - 1) * nth < nvec; i += kU * nth) {  // kU independent 16 B requests in flight per thread
            uint4 v[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u)
                v[u] = PEER_SRC ? ld_peer_v4(src + ((i + u * nth) << 4)) : ld_stream_v4(src + ((i + u * nth) << 4));
#pragma unroll
            for (int u = 0; u < kU; ++u) st_v4(dst + ((i + u * nth) << 4), v[u]);
        }
        for (; i < nvec; i += nth) st_v4(dst + (i << 4), PEER_SRC ? ld_peer_v4(src + (i << 4)) : ld_stream_v4(src + (i << 4)));
        for (int64_t b = (nvec << 4) + tid; b < bytes; b += nth)
            dst[b] = PEER_SRC ? ld_peer_scalar(src + b) : src[b];
    } else {
        for (int64_t b = tid; b < bytes; b += nth) dst[b] = PEER_SRC ? ld_peer_scalar(src + b) : src[b];
    }
}

// ---- byte-typed view access for the data-movement kernels (element size known only at run time)
// 16 logical bytes starting at logical byte offset b (a multiple of 16) of a dense or strided local buffer
__device__ __noinline__ uint4 ld_view_bytes(const char* base, const ViewDev& v, int es, int64_t b) {   // slow path, out of line
    uint4 r;
    char* rb = reinterpret_cast<char*>(&r);
#pragma unroll 1
    for (int k = 0; k < 16; ++k) rb[k] = base[view_byte_off(v, es, b + k)];
    return r;
}
__device__ __noinline__ void st_view_bytes(char* base, const ViewDev& v, int es, int64_t b, uint4 val) {
    const char* vb = reinterpret_cast<const char*>(&val);
#pragma unroll 1
    for (int k = 0; k < 16; ++k) base[view_byte_off(v, es, b + k)] = vb[k];
}
__device__ __forceinline__ uint4 ld_view_v4(const char* base, const ViewDev& v, int es, int64_t b, bool vec_ok) {
    if (vec_ok) return ld_stream_v4(base + view_byte_off(v, es, b));
    return ld_view_bytes(base, v, es, b);
}
__device__ __forceinline__ void st_view_v4(char* base, const ViewDev& v, int es, int64_t b, const uint4& val, bool vec_ok) {
    if (vec_ok) st_v4(base + view_byte_off(v, es, b), val);
    else st_view_bytes(base, v, es, b, val);
}
// Whole grid: copy `bytes` logical bytes from a peer's (dense) staging into a strided local destination, starting at
// logical byte offset b0 of the view.
__device__ __forceinline__ void grid_scatter_view(char* dst, const ViewDev& ov, int es, int64_t b0, const char* src, int64_t bytes) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * blockDim.x;
    const bool vec = view_vec_ok_bytes(ov, es, dst) && (b0 & 15) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    int64_t done = 0;
    if (vec) {
        const int64_t nvec = bytes >> 4;
        for (int64_t i0 = tid; i0 < nvec; i0 += nth * 4) {
            uint4 t[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (i0 + u * nth < nvec) t[u] = ld_peer_v4(src + ((i0 + u * nth) << 4));
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (i0 + u * nth < nvec) st_v4(dst + view_byte_off(ov, es, b0 + ((i0 + u * nth) << 4)), t[u]);
        }
        done = nvec << 4;
    }
    for (int64_t b = done + tid; b < bytes; b += nth) dst[view_byte_off(ov, es, b0 + b)] = ld_peer_scalar(src + b);
}

// Whole grid: gather `bytes` logical bytes of a strided local source, starting at logical byte offset b0 of the view,
// into dense (staging) memory.
__device__ __forceinline__ void grid_gather_view(char* dst, const char* src, const ViewDev& v, int es, int64_t b0, int64_t bytes) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * blockDim.x;
    const bool vec = view_vec_ok_bytes(v, es, src) && (b0 & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    int64_t done = 0;
    if (vec) {
        const int64_t nvec = bytes >> 4;
        for (int64_t i0 = tid; i0 < nvec; i0 += nth * 4) {
            uint4 t[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (i0 + u * nth < nvec) t[u] = ld_stream_v4(src + view_byte_off(v, es, b0 + ((i0 + u * nth) << 4)));
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (i0 + u * nth < nvec) st_v4(dst + ((i0 + u * nth) << 4), t[u]);
        }
        done = nvec << 4;
    }
    for (int64_t b = done + tid; b < bytes; b += nth) dst[b] = src[view_byte_off(v, es, b0 + b)];
}

// ------------------------------------------------------------------------------------------
// K1: barrier
__global__ void barrier_kernel(CommDev c, int bank, uint32_t epoch) { row_barrier(c, bank, blockIdx.x, epoch); }

// ------------------------------------------------------------------------------------------
// K6/K7/K9/K10/K11: data movement (all-gather, broadcast, gather, scatter, all-to-all, sendrecv)
//
// Zero-copy form (direct != 0, peer_registry.cpp): the pull pieces address the peers' own source tensors
// (pv.ptr[peer] + a_off), nothing is packed, and a second barrier keeps every rank inside the kernel - and its
// stream from overwriting the source - until all peers have read it.  CTA b only meets CTA b of the peers.
__global__ void __launch_bounds__(kCollThreads)
move_kernel(CommDev c, const char* __restrict__ src, char* __restrict__ dst, PieceList pack, PieceList pull,
            ViewDev view, int elem_size, int64_t stage_off, int bank, uint32_t epoch, const __grid_constant__ PeerView pv, int direct,
            ViewDev oview) {
    if (direct) {
        row_barrier(c, bank, blockIdx.x, epoch);   // the peers' kernels run, so their sources are final
        for (int i = 0; i < pull.count; ++i) {
            const Piece& pc = pull.p[i];
            const char* from = pv.ptr[pc.peer] + pc.a_off;
            if (from != dst + pc.b_off) grid_copy_bytes<true>(dst + pc.b_off, from, pc.n);
        }
        row_barrier(c, bank, blockIdx.x, epoch + 1);
        return;
    }
    char* my_stage = c.stage[c.rank] + stage_off;
    for (int i = 0; i < pack.count; ++i) {
        const Piece& pc = pack.p[i];
        if (view.ndim == 0) {
            grid_copy_bytes<false>(my_stage + pc.b_off, src + pc.a_off, pc.n);
        } else {
            // strided source: gathered here, 16 bytes at a time when the view's inner dimension allows it
            grid_gather_view(my_stage + pc.b_off, src, view, elem_size, pc.a_off, pc.n);
        }
    }
    grid_barrier(c, bank, gridDim.x, epoch);
    for (int i = 0; i < pull.count; ++i) {
        const Piece& pc = pull.p[i];
        if (oview.ndim == 0) grid_copy_bytes<true>(dst + pc.b_off, c.stage[pc.peer] + stage_off + pc.a_off, pc.n);
        else grid_scatter_view(dst, oview, elem_size, pc.b_off, c.stage[pc.peer] + stage_off + pc.a_off, pc.n);   // unpack in-kernel
    }
}

// ------------------------------------------------------------------------------------------
// K7 on more than two ranks: broadcast through the switch.  Pulling from the root costs the root's link
// (world-1) x the message; instead the root writes its tensor ONCE to the multicast mapping of the staging
// area (multimem.st - NVSwitch replicates it into every rank's staging), then every other rank copies its
// local staging copy into the user tensor.  CTA b owns byte range b and only meets CTA b of the peers.
constexpr int kBcChunkVecs = 16384;   // 256 KiB sub-chunks: receivers copy chunk k while the root multicasts chunk k+1
constexpr int kBcFlagBank = 4;        // one-directional "chunk k is in your staging" flags (shared with the pipelined all-reduce)

__global__ void __launch_bounds__(kCollThreads)
broadcast_nvls_kernel(CommDev c, const char* __restrict__ src, char* __restrict__ dst, int64_t bytes, int root,
                      int64_t stage_off, uint32_t flag_base, int nchunks, int bank, uint32_t end_epoch, ViewDev view, int es,
                      int64_t b0) {   // b0: logical byte offset of this launch inside the view (strided tensors only)
    const int64_t nvec = bytes >> 4;
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t v_lo = std::min<int64_t>((int64_t)blockIdx.x * per, nvec);
    const int64_t v_hi = std::min<int64_t>(v_lo + per, nvec);
    const int64_t tail_lo = nvec << 4;   // < 16 trailing bytes travel with chunk 0 of CTA 0 as plain peer stores
    const int row = blockIdx.x;
    const char* mine = c.stage[c.rank] + stage_off;
    // (a strided tensor - the same view on every rank - is gathered by the root and scattered by the receivers here)
    const bool src_vec = view_vec_ok_bytes(view, es, src), dst_vec = view_vec_ok_bytes(view, es, dst);
    // Everybody has entered this kernel - and so has left the previous one - before the root writes into their staging:
    // the region this launch uses was the OUTPUT half of a preceding two-shot all-reduce, which a slower rank may still be
    // unpacking (found by the host emulation under ThreadSanitizer: csrc/collectives/tests/emu_collectives_test.cpp).
    row_barrier(c, bank, row, end_epoch - 1);
    for (int k = 0; k < nchunks; ++k) {
        const int64_t c_lo = std::min<int64_t>(v_lo + (int64_t)k * kBcChunkVecs, v_hi);
        const int64_t c_hi = std::min<int64_t>(c_lo + kBcChunkVecs, v_hi);
        if (c.rank == root) {
            char* mc = c.stage_mc + stage_off;
            if (view.ndim != 0) {
                for (int64_t i = c_lo + threadIdx.x; i < c_hi; i += blockDim.x)
                    mm_st_v4(mc + (i << 4), ld_view_v4(src, view, es, b0 + (i << 4), src_vec && (b0 & 15) == 0));
            } else if (src_vec) {
                for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
                    uint4 v[kU];
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < c_hi) v[u] = ld_stream_v4(src + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < c_hi) mm_st_v4(mc + ((i0 + u * blockDim.x) << 4), v[u]);
                }
            } else {
                for (int64_t i = c_lo + threadIdx.x; i < c_hi; i += blockDim.x) {
                    uint4 v;
                    char* vb = reinterpret_cast<char*>(&v);
#pragma unroll
                    for (int b = 0; b < 16; ++b) vb[b] = src[(i << 4) + b];
                    mm_st_v4(mc + (i << 4), v);
                }
            }
            if (row == 0 && k == 0)
                for (int64_t b = tail_lo + threadIdx.x; b < bytes; b += blockDim.x)
                    for (int p = 0; p < c.world; ++p) c.stage[p][stage_off + b] = src[view_byte_off(view, es, b0 + b)];
            __syncthreads();
            if (threadIdx.x < c.world && threadIdx.x != root)
                st_release_sys(pad_slot(c, threadIdx.x, kBcFlagBank, row, root), flag_base + k + 1);
        } else {
            if (threadIdx.x == 0) wait_flag(pad_slot(c, c.rank, kBcFlagBank, row, root), flag_base + k + 1, root, c.watchdog_ns);
            __syncthreads();
            if (view.ndim != 0) {
                for (int64_t i = c_lo + threadIdx.x; i < c_hi; i += blockDim.x)
                    st_view_v4(dst, view, es, b0 + (i << 4), ld_peer_v4(mine + (i << 4)), dst_vec && (b0 & 15) == 0);
            } else if (dst_vec) {
                for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
                    uint4 v[kU];
#pragma unroll
                    for (int u = 0; u < kU; ++u)   // written through the switch during this kernel: bypass L1
                        if (i0 + u * blockDim.x < c_hi) v[u] = ld_peer_v4(mine + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < c_hi) st_v4(dst + ((i0 + u * blockDim.x) << 4), v[u]);
                }
            } else {
                for (int64_t b = (c_lo << 4) + threadIdx.x; b < (c_hi << 4); b += blockDim.x) dst[b] = ld_peer_scalar(mine + b);
            }
            if (row == 0 && k == 0)
                for (int64_t b = tail_lo + threadIdx.x; b < bytes; b += blockDim.x) dst[view_byte_off(view, es, b0 + b)] = ld_peer_scalar(mine + b);
        }
    }
    // The root never waits inside the loop; this closing barrier keeps it from running ahead and reusing the
    // staging region while a receiver is still copying out of it.
    row_barrier(c, bank, row, end_epoch);
}

// ------------------------------------------------------------------------------------------
// Typed pack (with optional strided view) used by the reducing kernels.
template <typename T>
__device__ __forceinline__ void pack_typed(T* stage, const T* __restrict__ src, const ViewDev& view, int64_t src_off,
                                           int64_t n, int64_t tid, int64_t nth) {
    constexpr int VE = 16 / sizeof(T);
    if (view.ndim == 0) {
        const T* s = src + src_off;
        if (((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(stage)) & 15) == 0) {
            const int64_t nvec = n / VE;
            const uint4* s4 = reinterpret_cast<const uint4*>(s);
            uint4* d4 = reinterpret_cast<uint4*>(stage);
            for (int64_t i0 = tid; i0 < nvec; i0 += nth * kU) {
                uint4 b[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * nth < nvec) b[u] = ld_stream_v4(s4 + i0 + u * nth);
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * nth < nvec) st_v4(d4 + i0 + u * nth, b[u]);
            }
            for (int64_t i = nvec * VE + tid; i < n; i += nth) stage[i] = s[i];
        } else {
            for (int64_t i = tid; i < n; i += nth) stage[i] = s[i];
        }
        return;
    }
    // Strided view.  When the innermost dimension is dense and a whole number of 16 B vectors,
    // gather vector-wise (one index decomposition per vector), else element-wise.
    const int last = view.ndim - 1;
    const bool vec_ok = view.stride[last] == 1 && view.size[last] % VE == 0 && src_off % VE == 0 && n % VE == 0 &&
                        (reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(stage) & 15) == 0;
    bool strides_ok = vec_ok;
    if (vec_ok)
        for (int d = 0; d < last; ++d) strides_ok = strides_ok && (view.stride[d] % VE == 0);
    if (strides_ok) {
        const int64_t nvec = n / VE;
        for (int64_t i = tid; i < nvec; i += nth) {
            const int64_t off = view_offset(view, src_off + i * VE);
            st_v4(reinterpret_cast<uint4*>(stage) + i, ld_stream_v4(src + off));
        }
    } else {
        for (int64_t i = tid; i < n; i += nth) stage[i] = src[view_offset(view, src_off + i)];
    }
}

// Reduce kU vector positions (i0 + u*stride, u < kU) across all ranks, in rank order, so every rank
// computes bit-identical results.  All kU loads of one peer are issued before any is consumed.
template <typename T, int OP>
__device__ __forceinline__ void reduce_batch(const CommDev& c, char* const (&bases)[DLBB_MAX_RANKS], int64_t base_byte_off,
                                             int64_t i0, int64_t stride, int64_t nvec, uint4 (&out)[kU]) {
    constexpr int VE = 16 / sizeof(T);
    using A = typename Acc<T>::type;
    A acc[kU][VE];
    for (int p = 0; p < c.world; ++p) {
        uint4 buf[kU];
        const char* sp = bases[p] + base_byte_off;
#pragma unroll
        for (int u = 0; u < kU; ++u)
            if (i0 + u * stride < nvec) buf[u] = ld_peer_v4(sp + ((i0 + u * stride) << 4));
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            if (i0 + u * stride < nvec) {
                const T* t = reinterpret_cast<const T*>(&buf[u]);
#pragma unroll
                for (int e = 0; e < VE; ++e) {
                    const A x = to_acc<T>(t[e]);
                    acc[u][e] = (p == 0) ? x : red_op<OP, A>(acc[u][e], x);
                }
            }
        }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        T* ot = reinterpret_cast<T*>(&out[u]);
#pragma unroll
        for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(acc[u][e]);
    }
}

template <typename T, int OP>
__device__ __forceinline__ void reduce_batch(const CommDev& c, int64_t base_byte_off, int64_t i0, int64_t stride,
                                             int64_t nvec, uint4 (&out)[kU]) {
    reduce_batch<T, OP>(c, c.stage, base_byte_off, i0, stride, nvec, out);
}

// One vector, the loads from ALL peers in flight before the first one is consumed (reduce_batch keeps kU vectors of ONE
// peer in flight and walks the peers one after the other: right for bandwidth, but a message of at most one vector per
// thread then pays `world` NVLink round trips in a row).  Summed in rank order like reduce_batch: bit-identical results.
template <typename T, int OP>
__device__ __forceinline__ uint4 reduce_wide(const CommDev& c, char* const (&bases)[DLBB_MAX_RANKS], int64_t base_byte_off, int64_t i) {
    constexpr int VE = 16 / sizeof(T);
    constexpr int kW = 8;
    using A = typename Acc<T>::type;
    A acc[VE];
    for (int p0 = 0; p0 < c.world; p0 += kW) {
        uint4 buf[kW];
#pragma unroll
        for (int q = 0; q < kW; ++q)
            if (p0 + q < c.world) buf[q] = ld_peer_v4(bases[p0 + q] + base_byte_off + (i << 4));
#pragma unroll
        for (int q = 0; q < kW; ++q) {
            if (p0 + q < c.world) {
                const T* x = reinterpret_cast<const T*>(&buf[q]);
#pragma unroll
                for (int e = 0; e < VE; ++e) acc[e] = (p0 + q == 0) ? to_acc<T>(x[e]) : red_op<OP, A>(acc[e], to_acc<T>(x[e]));
            }
        }
    }
    uint4 o;
    T* ot = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(acc[e]);
    return o;
}

template <typename T, int OP>
__device__ __forceinline__ typename Acc<T>::type reduce_scalar(const CommDev& c, char* const (&bases)[DLBB_MAX_RANKS],
                                                               int64_t byte_off) {
    using A = typename Acc<T>::type;
    A acc = to_acc<T>(ld_peer_scalar(reinterpret_cast<const T*>(bases[0] + byte_off)));
    for (int p = 1; p < c.world; ++p)
        acc = red_op<OP, A>(acc, to_acc<T>(ld_peer_scalar(reinterpret_cast<const T*>(bases[p] + byte_off))));
    return acc;
}

// Reduce red_n elements at byte offset base_b of every rank's buffer into dst (+ extra), whole grid.
template <typename T, int OP>
__device__ __forceinline__ void reduce_range(const CommDev& c, char* const (&bases)[DLBB_MAX_RANKS], T* __restrict__ dst,
                                             const T* __restrict__ extra, int64_t base_b, int64_t red_n, int64_t tid,
                                             int64_t nth, const ViewDev& ov = ViewDev{}) {
    constexpr int VE = 16 / sizeof(T);
    using A = typename Acc<T>::type;
    if (ov.ndim != 0) {
        // strided destination: element i of the reduced range lands at dst[view_offset(ov, i)] (no `extra` here)
        const bool ovec = view_vec_ok<T>(ov, dst) && (base_b & 15) == 0;
        int64_t done_v = 0;
        if (ovec) {
            const int64_t nvec = red_n / VE;
            for (int64_t i0 = tid; i0 < nvec; i0 += nth * kU) {
                uint4 o[kU];
                reduce_batch<T, OP>(c, bases, base_b, i0, nth, nvec, o);
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * nth < nvec) store_view_vec<T>(dst, ov, i0 + u * nth, o[u], true);
            }
            done_v = nvec * VE;
        }
        for (int64_t i = done_v + tid; i < red_n; i += nth)
            dst[view_offset(ov, i)] = from_acc<T>(reduce_scalar<T, OP>(c, bases, base_b + i * (int64_t)sizeof(T)));
        return;
    }
    const bool vec_ok = (base_b & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0 &&
                        (extra == nullptr || (reinterpret_cast<uintptr_t>(extra) & 15) == 0);
    int64_t done = 0;
    if (vec_ok) {
        const int64_t nvec = red_n / VE;
        if (nvec <= nth && extra == nullptr) {   // at most one vector per thread: latency-bound, all peers at once
            if (tid < nvec) st_v4(reinterpret_cast<uint4*>(dst) + tid, reduce_wide<T, OP>(c, bases, base_b, tid));
        } else for (int64_t i0 = tid; i0 < nvec; i0 += nth * kU) {
            uint4 o[kU];
            reduce_batch<T, OP>(c, bases, base_b, i0, nth, nvec, o);
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t i = i0 + u * nth;
                if (i < nvec) {
                    if (extra) {
                        const uint4 x = ld_stream_v4(reinterpret_cast<const uint4*>(extra) + i);
                        const T* xt = reinterpret_cast<const T*>(&x);
                        T* ot = reinterpret_cast<T*>(&o[u]);
#pragma unroll
                        for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(to_acc<T>(ot[e]) + to_acc<T>(xt[e]));
                    }
                    st_v4(reinterpret_cast<uint4*>(dst) + i, o[u]);
                }
            }
        }
        done = nvec * VE;
    }
    for (int64_t i = done + tid; i < red_n; i += nth) {
        A a = reduce_scalar<T, OP>(c, bases, base_b + i * (int64_t)sizeof(T));
        if (extra) a += to_acc<T>(extra[i]);
        dst[i] = from_acc<T>(a);
    }
}

// ------------------------------------------------------------------------------------------
// K2/K5/K8: single-barrier reductions (one-shot all-reduce, reduce, reduce-scatter).
// Every rank packs `pack` pieces; ranks with red_n > 0 reduce stage elements
// [red_off, red_off + red_n) over all ranks into dst (+ extra, if given).
// Zero-copy form (direct != 0): nothing is packed, the reduction reads elements [red_off, red_off + red_n) of the
// peers' own source tensors (pv), and a closing barrier keeps the sources untouched until everyone is done.
template <typename T, int OP>
__global__ void __launch_bounds__(kCollThreads)
reduce_kernel(CommDev c, const T* __restrict__ src, T* __restrict__ dst, const T* __restrict__ extra, PieceList pack,
              ViewDev view, int64_t red_off, int64_t red_n, int64_t stage_off, int bank, uint32_t epoch,
              const __grid_constant__ PeerView pv, int direct, ViewDev oview) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * blockDim.x;
    if (direct) {
        row_barrier(c, bank, blockIdx.x, epoch);
    } else {
        T* my_stage = reinterpret_cast<T*>(c.stage[c.rank] + stage_off);
        for (int i = 0; i < pack.count; ++i)
            pack_typed<T>(my_stage + pack.p[i].b_off, src, view, pack.p[i].a_off, pack.p[i].n, tid, nth);
        grid_barrier(c, bank, gridDim.x, epoch);
    }
    const int64_t red_b = red_off * (int64_t)sizeof(T);
    if (red_n > 0) {
        if (direct) reduce_range<T, OP>(c, pv.ptr, dst, extra, red_b, red_n, tid, nth);
        else reduce_range<T, OP>(c, c.stage, dst, extra, stage_off + red_b, red_n, tid, nth, oview);
    }
    if (direct) row_barrier(c, bank, blockIdx.x, epoch + 1);
}

// ------------------------------------------------------------------------------------------
// K3 zero-copy: in-place two-shot all-reduce directly on the ranks' own tensors (pv, peer_registry.cpp).
// CTA b owns range b on every rank; inside it rank r reduces its 1/world share out of all ranks' tensors
// (loads from every peer in flight at once, summed in rank order -> bit-identical everywhere) and stores the
// result into all ranks' tensors.  No pack, no unpack: the only HBM traffic is the NVLink traffic itself.
// Rank q's share-r elements are read and written by rank r only, so in-place is race-free between the barriers.
constexpr int kUD = 2;            // vectors per peer in flight per thread ...
constexpr int kDirectMaxWorld = 8;   // ... times at most this many peers (one NVSwitch domain), all in registers

template <typename T, int OP>
__global__ void __launch_bounds__(kCollThreads)
allreduce_direct_kernel(CommDev c, const __grid_constant__ PeerView pv, int64_t n, int bank, uint32_t epoch) {
    constexpr int VE = 16 / sizeof(T);
    using A = typename Acc<T>::type;
    const int64_t nvec = n / VE;   // n is a multiple of VE * world (the host sends the tail through staging)
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t v_lo = std::min<int64_t>((int64_t)blockIdx.x * per, nvec);
    const int64_t v_hi = std::min<int64_t>(v_lo + per, nvec);
    const int64_t share = (v_hi - v_lo + c.world - 1) / c.world;
    const int64_t s_lo = std::min<int64_t>(v_lo + c.rank * share, v_hi);
    const int64_t s_hi = std::min<int64_t>(s_lo + share, v_hi);
    row_barrier(c, bank, blockIdx.x, epoch);
    for (int64_t i0 = s_lo + threadIdx.x; i0 < s_hi; i0 += (int64_t)blockDim.x * kUD) {
        uint4 buf[kDirectMaxWorld][kUD];
#pragma unroll
        for (int p = 0; p < kDirectMaxWorld; ++p) {
            if (p < c.world) {
#pragma unroll
                for (int u = 0; u < kUD; ++u)
                    if (i0 + u * blockDim.x < s_hi) buf[p][u] = ld_peer_v4(pv.ptr[p] + ((i0 + u * blockDim.x) << 4));
            }
        }
        uint4 o[kUD];
#pragma unroll
        for (int u = 0; u < kUD; ++u) {
            A acc[VE];
#pragma unroll
            for (int p = 0; p < kDirectMaxWorld; ++p) {
                if (p < c.world) {
                    const T* t = reinterpret_cast<const T*>(&buf[p][u]);
#pragma unroll
                    for (int e = 0; e < VE; ++e) {
                        const A x = to_acc<T>(t[e]);
                        acc[e] = (p == 0) ? x : red_op<OP, A>(acc[e], x);
                    }
                }
            }
            T* ot = reinterpret_cast<T*>(&o[u]);
#pragma unroll
            for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(acc[e]);
        }
#pragma unroll 1
        for (int p = 0; p < c.world; ++p) {
            char* qp = pv.ptr[(c.rank + p) % c.world];   // spread the store targets over the peers
#pragma unroll
            for (int u = 0; u < kUD; ++u)
                if (i0 + u * blockDim.x < s_hi) st_v4(qp + ((i0 + u * blockDim.x) << 4), o[u]);
        }
    }
    row_barrier(c, bank, blockIdx.x, epoch + 1);
}

// ------------------------------------------------------------------------------------------
// K3/K4: two-shot all-reduce.  CTA b owns the element range [b*per, (b+1)*per) on every rank and
// only ever meets CTA b of the peers (row barriers), so CTAs pipeline independently:
//   pack range -> barrier -> reduce my 1/world of the range from all ranks and store the result
//   into every rank's out-staging -> barrier -> unpack the whole range from my out-staging.
// NVLS variant: the reduce is multimem.ld_reduce on the in-staging multicast address and the
// store is multimem.st on the out-staging multicast address (both performed by the switch).
template <typename T> struct MmReduce { static constexpr bool ok = false; };
template <> struct MmReduce<__nv_bfloat16> {
    static constexpr bool ok = true;
    __device__ static uint4 ld(const void* p) { return mm_ld_reduce_add_bf16(p); }
};
template <> struct MmReduce<__half> {
    static constexpr bool ok = true;
    __device__ static uint4 ld(const void* p) { return mm_ld_reduce_add_f16(p); }
};
template <> struct MmReduce<float> {
    static constexpr bool ok = true;
    __device__ static uint4 ld(const void* p) { return mm_ld_reduce_add_f32(p); }
};

// multimem.ld_reduce answers in ~10 us under load (measured: root-pull bandwidth grows linearly with the CTA count,
// 167 -> 420 GB/s from 32 to 128 CTAs at 8 vectors per thread), so the loops that are nothing but ld_reduce keep
// twice as many requests in flight per thread as the peer-load loops.
constexpr int kUR = 16;

// dst[i] = o (+ extra[i]) for one 16-byte vector of T; scalar stores when dst / extra are not 16 B aligned.
template <typename T>
__device__ __forceinline__ void store_result_vec(T* __restrict__ dst, const T* __restrict__ extra, int64_t i, uint4 o,
                                                 bool dst_vec) {
    constexpr int VE = 16 / sizeof(T);
    if (extra) {
        T* ot = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(to_acc<T>(ot[e]) + to_acc<T>(extra[i * VE + e]));
    }
    if (dst_vec) {
        st_v4(reinterpret_cast<uint4*>(dst) + i, o);
    } else {
        const T* ot = reinterpret_cast<const T*>(&o);
#pragma unroll
        for (int e = 0; e < VE; ++e) dst[i * VE + e] = ot[e];
    }
}

template <typename T, int OP, bool NVLS>
__global__ void __launch_bounds__(kCollThreads)
allreduce_twoshot_kernel(CommDev c, const T* __restrict__ src, T* __restrict__ dst, const T* __restrict__ extra,
                         ViewDev view, int64_t src_off, int64_t n, int64_t in_off, int64_t out_off, int bank,
                         uint32_t epoch, int root, ViewDev oview) {
    constexpr int VE = 16 / sizeof(T);
    using A = typename Acc<T>::type;
    // range of this CTA, in vectors; n is a multiple of VE * world (host pads the tail separately)
    const int64_t nvec = n / VE;
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t v_lo = std::min<int64_t>((int64_t)blockIdx.x * per, nvec);
    const int64_t v_hi = std::min<int64_t>(v_lo + per, nvec);
    const int64_t cnt = v_hi - v_lo;
    T* my_in = reinterpret_cast<T*>(c.stage[c.rank] + in_off);
    pack_typed<T>(my_in + v_lo * VE, src, view, src_off + v_lo * VE, cnt * VE, threadIdx.x, blockDim.x);
    row_barrier(c, bank, blockIdx.x, epoch);
    if constexpr (NVLS && MmReduce<T>::ok) {
        if (root >= 0) {
            // Reduce-to-root through the switch: the root alone issues multimem.ld_reduce over the whole range and
            // writes the user tensor directly - every rank's staging leaves its GPU once, the root receives the
            // message once, and there is neither an out-staging hop nor an unpack.
            if (c.rank == root) {
                const bool dv = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
                for (int64_t i0 = v_lo + threadIdx.x; i0 < v_hi; i0 += (int64_t)blockDim.x * kUR) {
                    uint4 o[kUR];
#pragma unroll
                    for (int u = 0; u < kUR; ++u)
                        if (i0 + u * blockDim.x < v_hi) o[u] = MmReduce<T>::ld(c.stage_mc + in_off + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                    for (int u = 0; u < kUR; ++u) {
                        if (i0 + u * blockDim.x >= v_hi) continue;
                        if (oview.ndim != 0) store_view_vec<T>(dst, oview, i0 + u * blockDim.x, o[u], view_vec_ok<T>(oview, dst));
                        else store_result_vec<T>(dst, extra, i0 + u * blockDim.x, o[u], dv);
                    }
                }
            }
            row_barrier(c, bank, blockIdx.x, epoch + 1);   // nobody repacks this staging region before the root is done
            return;
        }
    }
    // my share of the CTA range
    const int64_t share = (cnt + c.world - 1) / c.world;
    const int64_t s_lo = std::min<int64_t>(v_lo + c.rank * share, v_hi);
    const int64_t s_hi = std::min<int64_t>(s_lo + share, v_hi);
    for (int64_t i0 = s_lo + threadIdx.x; i0 < s_hi; i0 += (int64_t)blockDim.x * kU) {
        uint4 o[kU];
        if constexpr (NVLS && MmReduce<T>::ok) {
#pragma unroll
            for (int u = 0; u < kU; ++u)
                if (i0 + u * blockDim.x < s_hi) o[u] = MmReduce<T>::ld(c.stage_mc + in_off + ((i0 + u * blockDim.x) << 4));
        } else {
            reduce_batch<T, OP>(c, in_off, i0, blockDim.x, s_hi, o);
        }
        if (root >= 0) {
            // reduce-to-root: every rank still reduces its share (the load is spread), only the root receives it
            char* qp = c.stage[root] + out_off;
#pragma unroll
            for (int u = 0; u < kU; ++u)
                if (i0 + u * blockDim.x < s_hi) st_v4(qp + ((i0 + u * blockDim.x) << 4), o[u]);
        } else if constexpr (NVLS && MmReduce<T>::ok) {
#pragma unroll
            for (int u = 0; u < kU; ++u)
                if (i0 + u * blockDim.x < s_hi) mm_st_v4(c.stage_mc + out_off + ((i0 + u * blockDim.x) << 4), o[u]);
        } else {
#pragma unroll 1
            for (int p = 0; p < c.world; ++p) {
                const int q = (c.rank + p) % c.world;  // spread the store targets over the links
                char* qp = c.stage[q] + out_off;
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < s_hi) st_v4(qp + ((i0 + u * blockDim.x) << 4), o[u]);
            }
        }
    }
    row_barrier(c, bank, blockIdx.x, epoch + 1);
    if (root >= 0 && c.rank != root) return;
    const uint4* my_out = reinterpret_cast<const uint4*>(c.stage[c.rank] + out_off);
    const bool dst_vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0 &&
                         (extra == nullptr || (reinterpret_cast<uintptr_t>(extra) & 15) == 0);
    const bool ovec = view_vec_ok<T>(oview, dst);   // strided destination: scatter by the view (no `extra` then)
    for (int64_t i0 = v_lo + threadIdx.x; i0 < v_hi; i0 += (int64_t)blockDim.x * kU) {
        uint4 ob[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u)   // written by peers during this kernel: bypass L1
            if (i0 + u * blockDim.x < v_hi) ob[u] = ld_peer_v4(my_out + i0 + u * blockDim.x);
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const int64_t i = i0 + u * blockDim.x;
            if (i >= v_hi) continue;
            uint4 o = ob[u];
            if (extra) {
                T* ot = reinterpret_cast<T*>(&o);
                if (dst_vec) {
                    const uint4 x = ld_stream_v4(reinterpret_cast<const uint4*>(extra) + i);
                    const T* xt = reinterpret_cast<const T*>(&x);
#pragma unroll
                    for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(to_acc<T>(ot[e]) + to_acc<T>(xt[e]));
                } else {
#pragma unroll
                    for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(to_acc<T>(ot[e]) + to_acc<T>(extra[i * VE + e]));
                }
            }
            if (oview.ndim != 0) {
                store_view_vec<T>(dst, oview, i, o, ovec);
            } else if (dst_vec) {
                st_v4(reinterpret_cast<uint4*>(dst) + i, o);
            } else {
                const T* ot = reinterpret_cast<const T*>(&o);
#pragma unroll
                for (int e = 0; e < VE; ++e) dst[i * VE + e] = ot[e];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// K3/K4 for large messages: the same two-shot data flow, software-pipelined inside the CTA.
// Three warp groups work on different 64 KiB sub-chunks of the CTA's range at the same time:
//   PACK    user tensor (or strided view) -> my in-staging, then "chunk k packed" flag to every rank
//   REDUCE  waits for that flag from all ranks, reduces my 1/world of chunk k out of everyone's in-staging
//           (multimem.ld_reduce or peer loads) and writes it into everyone's out-staging (multimem.st or peer
//           stores), then "chunk k reduced" flag to every rank
//   UNPACK  waits for that flag from all ranks, copies chunk k from my out-staging to the user tensor
// so the two local copies (HBM) run under the NVLink phase instead of before and after it.
// Flags are per (CTA row, source rank), epoch = base + chunk index + 1, on pad banks 4 and 5.
constexpr int kPackT = 128, kRedT = 256, kUnpackT = 128;   // threads per role (the NVLink role gets the most)
constexpr int kPipeThreads = kPackT + kRedT + kUnpackT;    // 512 -> 128 registers per thread, no spills
constexpr int kPipeChunkVecsDefault = 4096;     // 64 KiB sub-chunks (DLBB_PIPE_CHUNK_KB overrides)
constexpr int kPackBank = 4, kRedBank = 5;

template <typename T, int OP, bool NVLS>
__global__ void __launch_bounds__(kPipeThreads, 1)
allreduce_pipelined_kernel(CommDev c, const T* __restrict__ src, T* __restrict__ dst, const T* __restrict__ extra,
                           ViewDev view, int64_t src_off, int64_t n, int64_t in_off, int64_t out_off, uint32_t base_pack,
                           uint32_t base_red, int kPipeChunkVecs, ViewDev oview) {
    constexpr int VE = 16 / sizeof(T);
    const int64_t nvec = n / VE;
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t v_lo = std::min<int64_t>((int64_t)blockIdx.x * per, nvec);
    const int64_t v_hi = std::min<int64_t>(v_lo + per, nvec);
    const int nchunks = static_cast<int>((per + kPipeChunkVecs - 1) / kPipeChunkVecs);   // same on every CTA and rank
    const int role = threadIdx.x < kPackT ? 0 : (threadIdx.x < kPackT + kRedT ? 1 : 2);
    const int t = threadIdx.x - (role == 0 ? 0 : (role == 1 ? kPackT : kPackT + kRedT));
    const int row = blockIdx.x;
    for (int k = 0; k < nchunks; ++k) {
        const int64_t c_lo = std::min<int64_t>(v_lo + (int64_t)k * kPipeChunkVecs, v_hi);
        const int64_t c_hi = std::min<int64_t>(c_lo + kPipeChunkVecs, v_hi);
        const int64_t cnt = c_hi - c_lo;
        if (role == 0) {
            // ---------------- PACK
            T* my_in = reinterpret_cast<T*>(c.stage[c.rank] + in_off);
            if (cnt > 0) pack_typed<T>(my_in + c_lo * VE, src, view, src_off + c_lo * VE, cnt * VE, t, kPackT);
            group_bar_sync(1, kPackT);
            if (t < c.world) st_release_sys(pad_slot(c, t, kPackBank, row, c.rank), base_pack + k + 1);
        } else if (role == 1) {
            // ---------------- REDUCE my share of the chunk
            if (t < c.world) wait_flag(pad_slot(c, c.rank, kPackBank, row, t), base_pack + k + 1, t, c.watchdog_ns);
            group_bar_sync(2, kRedT);
            const int64_t share = (cnt + c.world - 1) / c.world;
            const int64_t s_lo = std::min<int64_t>(c_lo + c.rank * share, c_hi);
            const int64_t s_hi = std::min<int64_t>(s_lo + share, c_hi);
            for (int64_t i0 = s_lo + t; i0 < s_hi; i0 += (int64_t)kRedT * kU) {
                uint4 o[kU];
                if constexpr (NVLS && MmReduce<T>::ok) {
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * kRedT < s_hi) o[u] = MmReduce<T>::ld(c.stage_mc + in_off + ((i0 + u * kRedT) << 4));
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * kRedT < s_hi) mm_st_v4(c.stage_mc + out_off + ((i0 + u * kRedT) << 4), o[u]);
                } else {
                    reduce_batch<T, OP>(c, in_off, i0, kRedT, s_hi, o);
#pragma unroll 1
                    for (int p = 0; p < c.world; ++p) {
                        char* qp = c.stage[(c.rank + p) % c.world] + out_off;
#pragma unroll
                        for (int u = 0; u < kU; ++u)
                            if (i0 + u * kRedT < s_hi) st_v4(qp + ((i0 + u * kRedT) << 4), o[u]);
                    }
                }
            }
            group_bar_sync(2, kRedT);
            if (t < c.world) st_release_sys(pad_slot(c, t, kRedBank, row, c.rank), base_red + k + 1);
        } else {
            // ---------------- UNPACK
            if (t < c.world) wait_flag(pad_slot(c, c.rank, kRedBank, row, t), base_red + k + 1, t, c.watchdog_ns);
            group_bar_sync(3, kUnpackT);
            const uint4* my_out = reinterpret_cast<const uint4*>(c.stage[c.rank] + out_off);
            const bool dst_vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0 &&
                                 (extra == nullptr || (reinterpret_cast<uintptr_t>(extra) & 15) == 0);
            for (int64_t i0 = c_lo + t; i0 < c_hi; i0 += (int64_t)kUnpackT * kU) {
                uint4 ob[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * kUnpackT < c_hi) ob[u] = ld_peer_v4(my_out + i0 + u * kUnpackT);
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const int64_t i = i0 + u * kUnpackT;
                    if (i >= c_hi) continue;
                    uint4 o = ob[u];
                    if (extra) {
                        T* ot = reinterpret_cast<T*>(&o);
#pragma unroll
                        for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(to_acc<T>(ot[e]) + to_acc<T>(extra[i * VE + e]));
                    }
                    if (oview.ndim != 0) {
                        store_view_vec<T>(dst, oview, i, o, view_vec_ok<T>(oview, dst));
                    } else if (dst_vec) {
                        st_v4(reinterpret_cast<uint4*>(dst) + i, o);
                    } else {
                        const T* ot = reinterpret_cast<const T*>(&o);
#pragma unroll
                        for (int e = 0; e < VE; ++e) dst[i * VE + e] = ot[e];
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// K5/K7 for large messages on more than two ranks: pipelined chains.
//
// A single root that pulls from (or is pulled by) world-1 peers moves the message world-1 times over its own
// link, and the multicast forms are limited by what one GPU can issue as multimem traffic (measured at 8 GPUs,
// 512 MiB: broadcast 388 GB/s with multimem.st, reduce 272 GB/s with multimem.ld_reduce, NCCL ~600 for both).  In a
// chain every link carries the message exactly once and all links work at the same time:
//
//   broadcast:  root -> root+1 -> ... -> root-1      each rank pulls chunk k out of its predecessor's staging
//                                                     into its tensor and into its own staging for the successor
//   reduce:     root+1 -> root+2 -> ... -> root       each rank adds its own chunk k to the predecessor's partial
//                                                     sum (fp32 add, partials travel in the tensor's dtype - the
//                                                     same rounding a ring reduction has) and the root writes dst
//
// CTA b owns byte range b on every rank and walks it in 64 KiB chunks; "chunk k is in my staging" travels as a
// one-directional flag on the successor's signal pad (bank 4, epoch = base + k + 1), so a rank works on chunk k
// while its predecessor already produces chunk k+1.  A closing row barrier bounds the skew between ranks (nobody
// re-uses the staging region while a neighbour still reads it).  Everything goes through symmetric staging, so
// any tensor (any allocator, any alignment) can take this path.
constexpr int kChainChunkVecs = 4096;   // 64 KiB
constexpr int kChainBank = 4;

__device__ __forceinline__ uint4 ld_any_v4(const char* p, bool vec) {   // 16 bytes from a possibly unaligned local pointer
    if (vec) return ld_stream_v4(p);
    uint4 v;
    char* vb = reinterpret_cast<char*>(&v);
#pragma unroll
    for (int b = 0; b < 16; ++b) vb[b] = p[b];
    return v;
}
__device__ __forceinline__ void st_any_v4(char* p, const uint4& v, bool vec) {
    if (vec) { st_v4(p, v); return; }
    const char* vb = reinterpret_cast<const char*>(&v);
#pragma unroll
    for (int b = 0; b < 16; ++b) p[b] = vb[b];
}

__global__ void __launch_bounds__(kCollThreads)
chain_broadcast_kernel(CommDev c, const char* __restrict__ src, char* __restrict__ dst, int64_t bytes, int root,
                       int64_t stage_off, uint32_t flag_base, int nchunks, uint32_t end_epoch) {
    const int pos = (c.rank - root + c.world) % c.world;
    const bool first = pos == 0, last = pos == c.world - 1;
    const int pred = (c.rank + c.world - 1) % c.world, succ = (c.rank + 1) % c.world;
    const int row = blockIdx.x;
    const int64_t nvec = bytes >> 4;
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t v_lo = min((int64_t)blockIdx.x * per, nvec), v_hi = min(v_lo + per, nvec);
    const int64_t tail_lo = nvec << 4;   // < 16 trailing bytes ride with chunk 0 of CTA 0
    const bool src_vec = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    char* mine = c.stage[c.rank] + stage_off;
    const char* theirs = c.stage[pred] + stage_off;
    for (int k = 0; k < nchunks; ++k) {
        const int64_t c_lo = min(v_lo + (int64_t)k * kChainChunkVecs, v_hi), c_hi = min(c_lo + kChainChunkVecs, v_hi);
        if (!first) {
            if (threadIdx.x == 0) wait_flag(pad_slot(c, c.rank, kChainBank, row, pred), flag_base + k + 1, pred, c.watchdog_ns);
            __syncthreads();
        }
        for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
            uint4 v[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t i = i0 + u * blockDim.x;
                if (i < c_hi) v[u] = first ? ld_any_v4(src + (i << 4), src_vec) : ld_peer_v4(theirs + (i << 4));
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t i = i0 + u * blockDim.x;
                if (i >= c_hi) continue;
                if (!last) st_v4(mine + (i << 4), v[u]);
                if (!first) st_any_v4(dst + (i << 4), v[u], dst_vec);
            }
        }
        if (row == 0 && k == 0) {
            for (int64_t b = tail_lo + threadIdx.x; b < bytes; b += blockDim.x) {
                const char x = first ? src[b] : ld_peer_scalar(theirs + b);
                if (!last) mine[b] = x;
                if (!first) dst[b] = x;
            }
        }
        if (!last) {
            __syncthreads();
            if (threadIdx.x == 0) st_release_sys(pad_slot(c, succ, kChainBank, row, c.rank), flag_base + k + 1);
        }
    }
    row_barrier(c, 0, row, end_epoch);
}

template <typename T, int OP>
__global__ void __launch_bounds__(kCollThreads)
chain_reduce_kernel(CommDev c, const T* __restrict__ src, T* __restrict__ dst, int64_t n, int root, int64_t stage_off,
                    uint32_t flag_base, int nchunks, uint32_t end_epoch) {
    constexpr int VE = 16 / sizeof(T);
    using A = typename Acc<T>::type;
    const int pos = (c.rank - root - 1 + 2 * c.world) % c.world;   // the chain starts at root+1 and ends at the root
    const bool first = pos == 0, last = pos == c.world - 1;
    const int pred = (c.rank + c.world - 1) % c.world, succ = (c.rank + 1) % c.world;
    const int row = blockIdx.x;
    const int64_t nvec = n / VE;   // n is a multiple of VE (the host reduces a ragged tail separately)
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t v_lo = min((int64_t)blockIdx.x * per, nvec), v_hi = min(v_lo + per, nvec);
    const bool src_vec = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    char* mine = c.stage[c.rank] + stage_off;
    const char* theirs = c.stage[pred] + stage_off;
    const char* srcb = reinterpret_cast<const char*>(src);
    for (int k = 0; k < nchunks; ++k) {
        const int64_t c_lo = min(v_lo + (int64_t)k * kChainChunkVecs, v_hi), c_hi = min(c_lo + kChainChunkVecs, v_hi);
        if (!first) {
            if (threadIdx.x == 0) wait_flag(pad_slot(c, c.rank, kChainBank, row, pred), flag_base + k + 1, pred, c.watchdog_ns);
            __syncthreads();
        }
        for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
            uint4 own[kU], part[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t i = i0 + u * blockDim.x;
                if (i < c_hi) {
                    if (!first) part[u] = ld_peer_v4(theirs + (i << 4));
                    own[u] = ld_any_v4(srcb + (i << 4), src_vec);
                }
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t i = i0 + u * blockDim.x;
                if (i >= c_hi) continue;
                uint4 o = own[u];
                if (!first) {
                    T* ot = reinterpret_cast<T*>(&o);
                    const T* pt = reinterpret_cast<const T*>(&part[u]);
#pragma unroll
                    for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(red_op<OP, A>(to_acc<T>(pt[e]), to_acc<T>(ot[e])));
                }
                if (last) st_any_v4(reinterpret_cast<char*>(dst) + (i << 4), o, dst_vec);
                else st_v4(mine + (i << 4), o);
            }
        }
        if (!last) {
            __syncthreads();
            if (threadIdx.x == 0) st_release_sys(pad_slot(c, succ, kChainBank, row, c.rank), flag_base + k + 1);
        }
    }
    row_barrier(c, 0, row, end_epoch);
}

// ------------------------------------------------------------------------------------------
// K11: point-to-point.  Only the two ranks of a message take part, so it has its own staging region
// (one slot per destination inside the sender's region), its own pad bank and per-pair epochs:
//   sender:   wait until the receiver has consumed the previous message of this pair (ack flag),
//             pack into the slot, raise the data flag on the receiver's pad;
//   receiver: wait for the data flag, pull from the sender's slot, raise the ack flag.
// Each CTA handles a fixed 1/kP2PCtas of the chunk with its own pair of flags, so the CTAs of one
// kernel never wait for each other.  A ring sendrecv runs both roles in one kernel.
constexpr int kP2PCtas = 16;
constexpr int kP2PBank = 3;

__global__ void __launch_bounds__(kCollThreads)
p2p_kernel(CommDev c, const char* __restrict__ send_src, int64_t send_bytes, int send_peer, uint32_t send_epoch,
           char* __restrict__ recv_dst, int64_t recv_bytes, int recv_peer, uint32_t recv_epoch, int64_t region_off,
           int64_t slot_bytes) {
    const int b = blockIdx.x;
    auto range = [&](int64_t bytes, int64_t& lo, int64_t& hi) {
        const int64_t per = ((bytes + gridDim.x - 1) / gridDim.x + 255) / 256 * 256;
        lo = std::min<int64_t>((int64_t)b * per, bytes);
        hi = std::min<int64_t>(lo + per, bytes);
    };
    auto cta_copy = [&](char* dst, const char* src, int64_t n, bool peer) {
        if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
            const int64_t nvec = n >> 4;
            for (int64_t i = threadIdx.x; i < nvec; i += blockDim.x)
                st_v4(dst + (i << 4), peer ? ld_peer_v4(src + (i << 4)) : ld_stream_v4(src + (i << 4)));
            for (int64_t i = (nvec << 4) + threadIdx.x; i < n; i += blockDim.x)
                dst[i] = peer ? ld_peer_scalar(src + i) : src[i];
        } else {
            for (int64_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = peer ? ld_peer_scalar(src + i) : src[i];
        }
    };
    if (send_peer >= 0) {
        int64_t lo, hi;
        range(send_bytes, lo, hi);
        // The slot is free once EVERY CTA of the receiver has copied its part of the previous message: the byte ranges
        // depend on the message length, so my range of this message may overlap other CTAs' ranges of the last one.
        for (int t = threadIdx.x; t < (int)gridDim.x; t += blockDim.x)
            wait_flag(pad_slot(c, c.rank, kP2PBank, kP2PCtas + t, send_peer), send_epoch - 1, send_peer, c.watchdog_ns);
        __syncthreads();
        cta_copy(c.stage[c.rank] + region_off + send_peer * slot_bytes + lo, send_src + lo, hi - lo, false);
        __syncthreads();
        if (threadIdx.x == 0) st_release_sys(pad_slot(c, send_peer, kP2PBank, b, c.rank), send_epoch);
    }
    if (recv_peer >= 0) {
        int64_t lo, hi;
        range(recv_bytes, lo, hi);
        if (threadIdx.x == 0) wait_flag(pad_slot(c, c.rank, kP2PBank, b, recv_peer), recv_epoch, recv_peer, c.watchdog_ns);
        __syncthreads();
        cta_copy(recv_dst + lo, c.stage[recv_peer] + region_off + c.rank * slot_bytes + lo, hi - lo, true);
        __syncthreads();
        if (threadIdx.x == 0) st_release_sys(pad_slot(c, recv_peer, kP2PBank, kP2PCtas + b, c.rank), recv_epoch);
    }
}

// ------------------------------------------------------------------------------------------
// K13: bucket fusion.  Many small tensors, ONE launch: every tensor of the table is packed into its own 16 B aligned
// slot of this rank's staging region, one barrier, then every rank reduces every slot across the ranks straight into
// the tensors (one-shot form: these are small messages, launch and barrier latency is what fusion removes).
// The analogue of oneCCL's CCL_FUSION (collectives/3d/launch_dsccl.sh:49-59); also what the DeepSpeed facade uses for
// the gradients of a ZeRO-0 step (test/ds_mpi_test.py:36-47).
constexpr int kMultiMax = 96;
struct MultiTable {
    int count;
    void* buf[kMultiMax];          // in place: source and destination
    int64_t n[kMultiMax];          // elements
    int64_t off[kMultiMax];        // element offset of the slot inside the staging region (16 B aligned)
};

// The table is walked as ONE index space of 16-byte staging vectors (slot t starts at vector off[t] * sizeof(T) / 16), so
// the threads of the grid spread over all tensors at once - a bucket of many tiny tensors costs one round trip, not one
// per tensor - and every thread has the loads from ALL peers of its vector in flight before it adds them up (in rank
// order: every rank computes bit-identical sums).  A slot's last vector may be partly padding: it is moved whole inside
// the staging area, and only its valid elements are read from / written to the tensor.
template <typename T>
__device__ __forceinline__ int multi_slot_of(const MultiTable& tb, int64_t g) {   // last slot that starts at or before vector g
    int lo = 0, hi = tb.count - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (((tb.off[mid] * (int64_t)sizeof(T)) >> 4) <= g) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

template <typename T, int OP>
__global__ void __launch_bounds__(kCollThreads)
reduce_multi_kernel(CommDev c, const __grid_constant__ MultiTable tb, int64_t total_vecs, int64_t stage_off, int bank,
                    uint32_t epoch) {
    constexpr int VE = 16 / sizeof(T);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * blockDim.x;
    char* my_stage = c.stage[c.rank] + stage_off;
    for (int64_t g = tid; g < total_vecs; g += nth) {
        const int t = multi_slot_of<T>(tb, g);
        const int64_t e0 = (g - ((tb.off[t] * (int64_t)sizeof(T)) >> 4)) * VE;   // first element of this vector in tensor t
        const int64_t left = tb.n[t] - e0;                                        // its valid elements (>= 1)
        const T* src = static_cast<const T*>(tb.buf[t]);
        if (left >= VE && (reinterpret_cast<uintptr_t>(src) & 15) == 0) {
            st_v4(my_stage + (g << 4), ld_stream_v4(src + e0));
        } else {
            T* d = reinterpret_cast<T*>(my_stage + (g << 4));
            for (int e = 0; e < VE && e < left; ++e) d[e] = src[e0 + e];
        }
    }
    grid_barrier(c, bank, gridDim.x, epoch);
    for (int64_t g = tid; g < total_vecs; g += nth) {
        const int t = multi_slot_of<T>(tb, g);
        const int64_t e0 = (g - ((tb.off[t] * (int64_t)sizeof(T)) >> 4)) * VE;
        const int64_t left = tb.n[t] - e0;
        const uint4 o = reduce_wide<T, OP>(c, c.stage, stage_off, g);
        const T* ot = reinterpret_cast<const T*>(&o);
        T* dst = static_cast<T*>(tb.buf[t]);
        if (left >= VE && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            st_v4(dst + e0, o);
        } else {
            for (int e = 0; e < VE && e < left; ++e) dst[e0 + e] = ot[e];
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side
// Kernels with a GRID barrier (staged reduce_kernel / move_kernel, reduce_multi_kernel) need all their CTAs resident at
// once: with 256 threads and up to ~190 registers that is one CTA per SM, so DLBB_CTAS above the SM count would leave CTAs
// waiting for ones that can never start: they are clamped to dlbb_comm::max_resident_ctas.  (Row-barrier kernels have no
// such limit: CTA b only meets CTA b of the peers and, in the two-hop kernels, lower-numbered CTAs - the host emulation
// runs them with fewer CTA slots than CTAs to prove it.)
static int pick_ctas(const dlbb_comm* c, int64_t bytes) {
    int n = c->ctas;
    if (n <= 0) {
        // one CTA per 128 KiB, capped: enough loads in flight to cover ~2 us of NVLink latency
        n = static_cast<int>(std::min<int64_t>(96, std::max<int64_t>(1, bytes / (64 << 10))));
    }
    return std::min(n, c->max_resident_ctas);
}

// Debug aid (DLBB_DEBUG_POISON=1): fill a staging region with 0xFF - NaN in every floating-point
// format, -1 in the integer ones - on the caller's stream right before the kernel that is going to (re)write it, so a
// protocol bug that lets a rank read staging data before its producer has written it shows up as NaN / -1 in the result
// instead of as plausible numbers from an earlier collective.  Safe for the same reason packing into the region is:
// when this call is issued, every peer has finished reading what the call before last left there.
static void poison_region(dlbb_comm* c, int64_t off, cudaStream_t s) {
    if (c->debug_poison) cudaMemsetAsync(c->dev.stage[c->rank] + off, 0xFF, c->stage_bytes, s);
}

static int64_t next_stage_off(dlbb_comm* c, cudaStream_t s, bool may_poison = true) {
    const int64_t off = (c->parity & 1) ? static_cast<int64_t>(c->stage_bytes) : 0;
    c->parity ^= 1;
    if (may_poison) poison_region(c, off, s);
    return off;
}

static ViewDev make_view(const dlbb_view_t* v) {
    ViewDev d{};
    if (v && v->ndim > 0) {
        d.ndim = v->ndim;
        for (int i = 0; i < 4; ++i) { d.size[i] = v->size[i]; d.stride[i] = v->stride[i]; }
    }
    return d;
}

// The strided destination of the running *_v call (comm_internal.h: out_view / out_base), for the kernels' unpack side.
static ViewDev out_view_of(const dlbb_comm* c) {
    ViewDev v = make_view(c->out_view);
    v.base = c->out_view ? c->out_base : 0;
    return v;
}

static size_t dtype_size(int dt) {
    switch (dt) {
        case DLBB_F16: case DLBB_BF16: return 2;
        case DLBB_F32: case DLBB_I32: return 4;
        case DLBB_F64: case DLBB_I64: return 8;
        case DLBB_U8: return 1;
    }
    return 0;
}

// CTAs of a zero-copy kernel: these only run for large messages, so take the full channel count.
static int direct_ctas(const dlbb_comm* c) { return c->ctas > 0 ? std::min(c->ctas, kMaxCtas) : 128; }
// ... scaled down for small messages (symmetric tensors take the direct kernels at every size): one CTA per 32 KiB
static int direct_ctas(const dlbb_comm* c, int64_t bytes) {
    if (c->ctas > 0) return std::min(c->ctas, kMaxCtas);
    return static_cast<int>(std::min<int64_t>(128, std::max<int64_t>(1, bytes / (32 << 10))));
}

template <typename T, int OP>
static int launch_reduce_T(dlbb_comm* c, const void* src, void* dst, const void* extra, const PieceList& pack,
                           const ViewDev& view, int64_t red_off, int64_t red_n, int64_t bytes_hint, cudaStream_t s,
                           const PeerView* direct) {
    if (direct) {
        const uint32_t ep = c->epoch[0] + 1;
        c->epoch[0] += 2;
        reduce_kernel<T, OP><<<direct_ctas(c), kCollThreads, 0, s>>>(
            c->dev, static_cast<const T*>(src), static_cast<T*>(dst), static_cast<const T*>(extra), pack, view, red_off,
            red_n, 0, 0, ep, *direct, 1, ViewDev{});
        DLBB_CHECK_LAUNCH("reduce_kernel (zero-copy)");
        return 0;
    }
    const int ctas = pick_ctas(c, bytes_hint);
    const int64_t so = next_stage_off(c, s);
    const uint32_t ep = ++c->epoch[0];
    reduce_kernel<T, OP><<<ctas, kCollThreads, 0, s>>>(c->dev, static_cast<const T*>(src), static_cast<T*>(dst),
                                                       static_cast<const T*>(extra), pack, view, red_off, red_n, so, 0,
                                                       ep, PeerView{}, 0, out_view_of(c));
    DLBB_CHECK_LAUNCH("reduce_kernel");
    return 0;
}

template <typename T>
static int launch_reduce_op(dlbb_comm* c, int op, const void* src, void* dst, const void* extra, const PieceList& pack,
                            const ViewDev& view, int64_t red_off, int64_t red_n, int64_t bytes_hint, cudaStream_t s,
                            const PeerView* direct) {
    switch (op) {
        case DLBB_SUM: return launch_reduce_T<T, DLBB_SUM>(c, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_MAX: return launch_reduce_T<T, DLBB_MAX>(c, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_MIN: return launch_reduce_T<T, DLBB_MIN>(c, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_PROD: return launch_reduce_T<T, DLBB_PROD>(c, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}

static int launch_reduce(dlbb_comm* c, int dtype, int op, const void* src, void* dst, const void* extra,
                         const PieceList& pack, const ViewDev& view, int64_t red_off, int64_t red_n,
                         int64_t bytes_hint, cudaStream_t s, const PeerView* direct = nullptr) {
    switch (dtype) {
        case DLBB_F16: return launch_reduce_op<__half>(c, op, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_BF16: return launch_reduce_op<__nv_bfloat16>(c, op, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_F32: return launch_reduce_op<float>(c, op, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_F64: return launch_reduce_op<double>(c, op, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_I32: return launch_reduce_op<int32_t>(c, op, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
        case DLBB_I64: return launch_reduce_op<int64_t>(c, op, src, dst, extra, pack, view, red_off, red_n, bytes_hint, s, direct);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

template <typename T, int OP>
static int launch_twoshot_T(dlbb_comm* c, bool nvls, const void* src, void* dst, const void* extra, const ViewDev& view,
                            int64_t src_off, int64_t n, cudaStream_t s, int root) {
    const int64_t bytes = n * (int64_t)sizeof(T);
    const int64_t in_off = next_stage_off(c, s);
    const int64_t out_off = in_off == 0 ? static_cast<int64_t>(c->stage_bytes) : 0;
    // (the out half is NOT poisoned here: a slower peer may still be reading it as the input half of the preceding
    //  single-barrier kernel - the kernel itself first writes there after its first barrier, a host-side fill would not wait)
    const bool use_nvls = nvls && MmReduce<T>::ok && OP == DLBB_SUM;
    static const int64_t pipe_min = getenv("DLBB_PIPELINE_MIN_BYTES") ? atoll(getenv("DLBB_PIPELINE_MIN_BYTES")) : (8ll << 20);
    if (bytes >= pipe_min && use_nvls && root < 0) {
        // large message on the multicast path: pack / reduce / unpack overlapped inside each CTA
        // (measured at 8 GPUs, 256 MiB: 631 GB/s busbw vs 515 for the phase-by-phase kernel and 646 for NCCL;
        // the peer-load variant does not profit - 362 vs 430 - and keeps the phase-by-phase kernel)
        constexpr int VE = 16 / sizeof(T);
        // one 512-thread CTA per SM, and a row's reduce role waits for the pack roles of ALL rows of the peers: every
        // CTA must be resident
        const int ctas = std::min(c->ctas > 0 ? c->ctas : 128, c->max_resident_ctas);
        const int64_t per = (n / VE + ctas - 1) / ctas;
        static const int kPipeChunkVecs = getenv("DLBB_PIPE_CHUNK_KB") ? atoi(getenv("DLBB_PIPE_CHUNK_KB")) * 64 : kPipeChunkVecsDefault;
        const uint32_t nchunks = static_cast<uint32_t>((per + kPipeChunkVecs - 1) / kPipeChunkVecs);
        const uint32_t bp = c->epoch[kPackBank], br = c->epoch[kRedBank];
        c->epoch[kPackBank] += nchunks;
        c->epoch[kRedBank] += nchunks;
        if (use_nvls)
            allreduce_pipelined_kernel<T, OP, true><<<ctas, kPipeThreads, 0, s>>>(
                c->dev, static_cast<const T*>(src), static_cast<T*>(dst), static_cast<const T*>(extra), view, src_off, n,
                in_off, out_off, bp, br, kPipeChunkVecs, out_view_of(c));
        else
            allreduce_pipelined_kernel<T, OP, false><<<ctas, kPipeThreads, 0, s>>>(
                c->dev, static_cast<const T*>(src), static_cast<T*>(dst), static_cast<const T*>(extra), view, src_off, n,
                in_off, out_off, bp, br, kPipeChunkVecs, out_view_of(c));
        DLBB_CHECK_LAUNCH("allreduce_pipelined_kernel");
        return 0;
    }
    const int ctas = c->ctas > 0 ? std::min(c->ctas, kMaxCtas) : static_cast<int>(std::min<int64_t>(
        96, std::max<int64_t>(1, bytes / (128 << 10))));
    const uint32_t ep = c->epoch[0] + 1;
    c->epoch[0] += 2;
    if (use_nvls)
        allreduce_twoshot_kernel<T, OP, true><<<ctas, kCollThreads, 0, s>>>(
            c->dev, static_cast<const T*>(src), static_cast<T*>(dst), static_cast<const T*>(extra), view, src_off, n,
            in_off, out_off, 0, ep, root, out_view_of(c));
    else
        allreduce_twoshot_kernel<T, OP, false><<<ctas, kCollThreads, 0, s>>>(
            c->dev, static_cast<const T*>(src), static_cast<T*>(dst), static_cast<const T*>(extra), view, src_off, n,
            in_off, out_off, 0, ep, root, out_view_of(c));
    DLBB_CHECK_LAUNCH("allreduce_twoshot_kernel");
    return 0;
}

template <typename T>
static int launch_twoshot_op(dlbb_comm* c, int op, bool nvls, const void* src, void* dst, const void* extra,
                             const ViewDev& view, int64_t src_off, int64_t n, cudaStream_t s, int root) {
    switch (op) {
        case DLBB_SUM: return launch_twoshot_T<T, DLBB_SUM>(c, nvls, src, dst, extra, view, src_off, n, s, root);
        case DLBB_MAX: return launch_twoshot_T<T, DLBB_MAX>(c, false, src, dst, extra, view, src_off, n, s, root);
        case DLBB_MIN: return launch_twoshot_T<T, DLBB_MIN>(c, false, src, dst, extra, view, src_off, n, s, root);
        case DLBB_PROD: return launch_twoshot_T<T, DLBB_PROD>(c, false, src, dst, extra, view, src_off, n, s, root);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}

// root < 0: all-reduce; root >= 0: reduce to that rank (same data flow, only the root receives and unpacks)
static int launch_twoshot(dlbb_comm* c, int dtype, int op, bool nvls, const void* src, void* dst, const void* extra,
                          const ViewDev& view, int64_t src_off, int64_t n, cudaStream_t s, int root = -1) {
    switch (dtype) {
        case DLBB_F16: return launch_twoshot_op<__half>(c, op, nvls, src, dst, extra, view, src_off, n, s, root);
        case DLBB_BF16: return launch_twoshot_op<__nv_bfloat16>(c, op, nvls, src, dst, extra, view, src_off, n, s, root);
        case DLBB_F32: return launch_twoshot_op<float>(c, op, nvls, src, dst, extra, view, src_off, n, s, root);
        case DLBB_F64: return launch_twoshot_op<double>(c, op, false, src, dst, extra, view, src_off, n, s, root);
        case DLBB_I32: return launch_twoshot_op<int32_t>(c, op, false, src, dst, extra, view, src_off, n, s, root);
        case DLBB_I64: return launch_twoshot_op<int64_t>(c, op, false, src, dst, extra, view, src_off, n, s, root);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

static int launch_move(dlbb_comm* c, const void* src, void* dst, const PieceList& pack, const PieceList& pull,
                       const ViewDev& view, int elem_size, int64_t bytes_hint, cudaStream_t s,
                       const PeerView* direct = nullptr) {
    if (direct) {
        const uint32_t ep = c->epoch[0] + 1;
        c->epoch[0] += 2;
        move_kernel<<<direct_ctas(c), kCollThreads, 0, s>>>(c->dev, static_cast<const char*>(src), static_cast<char*>(dst),
                                                            pack, pull, view, elem_size, 0, 0, ep, *direct, 1, ViewDev{});
        DLBB_CHECK_LAUNCH("move_kernel (zero-copy)");
        return 0;
    }
    const int ctas = pick_ctas(c, bytes_hint);
    const int64_t so = next_stage_off(c, s);
    const uint32_t ep = ++c->epoch[0];
    move_kernel<<<ctas, kCollThreads, 0, s>>>(c->dev, static_cast<const char*>(src), static_cast<char*>(dst), pack, pull,
                                              view, elem_size, so, 0, ep, PeerView{}, 0, out_view_of(c));
    DLBB_CHECK_LAUNCH("move_kernel");
    return 0;
}

// Zero-copy gate shared by every collective: large enough, enabled, and every rank's buffer exportable.
// Collective and host-blocking when it goes as far as the exchange (see peer_registry.cpp).
// Tensors that live in a symmetric segment (SymmComm.alloc / symm_empty) take this path at every size: the peers have
// them mapped already and, with a multicast mapping, the switch can reduce / replicate them in place.  `raw` is the
// caller's pointer before any "not usable on this rank" substitution, so that all ranks decide alike (SPMD contract:
// either every rank passes a symmetric tensor or none does).
static int try_zero_copy(dlbb_comm* c, const void* buf, int64_t bytes, int64_t gate_bytes, PeerView* pv, bool* ok,
                         const void* raw = nullptr) {
    *ok = false;
    if (c->world == 1 || c->zero_copy_min == 0) return 0;
    // (below 256 KiB the staged single-barrier kernels are as fast or faster than the two-barrier direct ones)
    static const int64_t symm_min = getenv("DLBB_SYMM_MIN_BYTES") ? atoll(getenv("DLBB_SYMM_MIN_BYTES")) : (256 << 10);
    if (gate_bytes < (int64_t)c->zero_copy_min &&
        (gate_bytes < symm_min || symm_segment_of(c, raw ? raw : buf, static_cast<size_t>(bytes)) < 0))
        return 0;
    return peer_registry_resolve(c, c->reg, buf, static_cast<size_t>(bytes), pv, ok);
}

static bool disjoint(const void* a, int64_t an, const void* b, int64_t bn) {
    const char* pa = static_cast<const char*>(a);
    const char* pb = static_cast<const char*>(b);
    return pa + an <= pb || pb + bn <= pa;
}
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T, int OP>
static int launch_allreduce_direct_T(dlbb_comm* c, const PeerView& pv, int64_t n, cudaStream_t s) {
    const uint32_t ep = c->epoch[0] + 1;
    c->epoch[0] += 2;
    allreduce_direct_kernel<T, OP><<<direct_ctas(c), kCollThreads, 0, s>>>(c->dev, pv, n, 0, ep);
    DLBB_CHECK_LAUNCH("allreduce_direct_kernel");
    return 0;
}
template <typename T>
static int launch_allreduce_direct_op(dlbb_comm* c, int op, const PeerView& pv, int64_t n, cudaStream_t s) {
    switch (op) {
        case DLBB_SUM: return launch_allreduce_direct_T<T, DLBB_SUM>(c, pv, n, s);
        case DLBB_MAX: return launch_allreduce_direct_T<T, DLBB_MAX>(c, pv, n, s);
        case DLBB_MIN: return launch_allreduce_direct_T<T, DLBB_MIN>(c, pv, n, s);
        case DLBB_PROD: return launch_allreduce_direct_T<T, DLBB_PROD>(c, pv, n, s);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}
static int launch_allreduce_direct(dlbb_comm* c, int dtype, int op, const PeerView& pv, int64_t n, cudaStream_t s) {
    switch (dtype) {
        case DLBB_F16: return launch_allreduce_direct_op<__half>(c, op, pv, n, s);
        case DLBB_BF16: return launch_allreduce_direct_op<__nv_bfloat16>(c, op, pv, n, s);
        case DLBB_F32: return launch_allreduce_direct_op<float>(c, op, pv, n, s);
        case DLBB_F64: return launch_allreduce_direct_op<double>(c, op, pv, n, s);
        case DLBB_I32: return launch_allreduce_direct_op<int32_t>(c, op, pv, n, s);
        case DLBB_I64: return launch_allreduce_direct_op<int64_t>(c, op, pv, n, s);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

// ---- chain launches (large broadcast / reduce on more than two ranks)
// Opt-in (DLBB_CHAIN_MIN_BYTES=<bytes>): validated on 4 GPUs with the chains forced at every size (97 checks,
// profiles/raw/zero_copy_4gpu_chain_forced.json), but there they only match the multicast forms (broadcast 442 vs
// 532 GB/s, reduce 428 vs 430 at 512 MiB; NCCL 628): one CTA handles wait -> load -> store -> signal of a chunk
// strictly in sequence (~19 us per 64 KiB).  Overlapping the next chunk's loads with the signalling is the open item.
static int64_t chain_min_bytes() {
    static const int64_t v = getenv("DLBB_CHAIN_MIN_BYTES") ? atoll(getenv("DLBB_CHAIN_MIN_BYTES")) : 0;
    return v;   // <= 0: chains off (the default)
}
struct ChainGeom { int ctas; int nchunks; uint32_t flag_base; int64_t stage_off; uint32_t end_epoch; };
static ChainGeom chain_geom(dlbb_comm* c, int64_t nvec, cudaStream_t s) {
    ChainGeom g;
    g.ctas = c->ctas > 0 ? std::min(c->ctas, kMaxCtas) : 256;   // each CTA walks its chunks strictly in sequence: rows feed the pipeline
    const int64_t per = (nvec + g.ctas - 1) / g.ctas;
    g.nchunks = static_cast<int>(std::max<int64_t>(1, (per + kChainChunkVecs - 1) / kChainChunkVecs));
    g.flag_base = c->epoch[kChainBank];
    c->epoch[kChainBank] += g.nchunks;
    g.stage_off = next_stage_off(c, s);
    g.end_epoch = ++c->epoch[0];
    return g;
}

template <typename T, int OP>
static int launch_chain_reduce_T(dlbb_comm* c, const void* src, void* dst, int64_t n, int root, cudaStream_t s) {
    const ChainGeom g = chain_geom(c, n / (16 / (int64_t)sizeof(T)), s);
    chain_reduce_kernel<T, OP><<<g.ctas, kCollThreads, 0, s>>>(c->dev, static_cast<const T*>(src), static_cast<T*>(dst), n,
                                                              root, g.stage_off, g.flag_base, g.nchunks, g.end_epoch);
    DLBB_CHECK_LAUNCH("chain_reduce_kernel");
    return 0;
}
template <typename T>
static int launch_chain_reduce_op(dlbb_comm* c, int op, const void* src, void* dst, int64_t n, int root, cudaStream_t s) {
    switch (op) {
        case DLBB_SUM: return launch_chain_reduce_T<T, DLBB_SUM>(c, src, dst, n, root, s);
        case DLBB_MAX: return launch_chain_reduce_T<T, DLBB_MAX>(c, src, dst, n, root, s);
        case DLBB_MIN: return launch_chain_reduce_T<T, DLBB_MIN>(c, src, dst, n, root, s);
        case DLBB_PROD: return launch_chain_reduce_T<T, DLBB_PROD>(c, src, dst, n, root, s);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}
static int launch_chain_reduce(dlbb_comm* c, int dtype, int op, const void* src, void* dst, int64_t n, int root,
                               cudaStream_t s) {
    switch (dtype) {
        case DLBB_F16: return launch_chain_reduce_op<__half>(c, op, src, dst, n, root, s);
        case DLBB_BF16: return launch_chain_reduce_op<__nv_bfloat16>(c, op, src, dst, n, root, s);
        case DLBB_F32: return launch_chain_reduce_op<float>(c, op, src, dst, n, root, s);
        case DLBB_F64: return launch_chain_reduce_op<double>(c, op, src, dst, n, root, s);
        case DLBB_I32: return launch_chain_reduce_op<int32_t>(c, op, src, dst, n, root, s);
        case DLBB_I64: return launch_chain_reduce_op<int64_t>(c, op, src, dst, n, root, s);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

}  // namespace dlbb
#include "direct_kernels.cuh"
namespace dlbb {

static int allreduce_impl(dlbb_comm* c, const void* in, void* out, const void* extra, int64_t count, int dtype, int op,
                          int algo, const dlbb_view_t* in_view, cudaStream_t s) {
    const size_t es = dtype_size(dtype);
    DLBB_REQUIRE(es != 0 && dtype != DLBB_U8, "allreduce: unsupported dtype %d", dtype);
    DLBB_REQUIRE(count >= 0, "allreduce: negative count");
    if (count == 0) return 0;
    const ViewDev view = make_view(in_view);
    if (algo == DLBB_ALGO_AUTO) algo = c->default_algo;
    const bool force_direct = algo == DLBB_ALGO_DIRECT;
    if (force_direct) algo = DLBB_ALGO_AUTO;   // for whatever cannot go zero-copy
    const int64_t VE = 16 / (int64_t)es;
    const int64_t cap = static_cast<int64_t>(c->stage_bytes / es) / (VE * c->world) * (VE * c->world);
    const bool nvls_ok = dlbb_comm_has_multicast(c) && op == DLBB_SUM && (dtype == DLBB_F16 || dtype == DLBB_BF16 || dtype == DLBB_F32);
    // Measured crossovers (profiles/collectives_busbw.md).  Two ranks: one-shot moves the minimum number of bytes
    // (each rank reads the peer once) and wins at every size.  More ranks with NVLS: the multimem kernel wins from
    // 1 KiB (14 us) to 1 GiB.  Without multicast: one-shot up to 512 KiB, then two-shot.
    int64_t oneshot_max = c->oneshot_max ? (int64_t)c->oneshot_max
                        : (c->world <= 2 ? INT64_MAX : (nvls_ok && algo == DLBB_ALGO_AUTO ? 0 : (512 << 10)));
    int64_t done = 0;
    // Zero-copy: in place on contiguous plain allocations, at most one NVSwitch domain (8 ranks).  The part that is
    // a multiple of 16 B x world goes direct, a ragged tail (if any) through the staged kernels below.
    // (Conditions that may differ between ranks - out of place, misaligned - are folded into the exchange as a
    // null buffer so that every rank still takes part in it and all of them fall back together.)
    if (extra == nullptr && view.ndim == 0 && c->world <= kDirectMaxWorld && algo == DLBB_ALGO_AUTO) {
        PeerView pv;
        bool zc = false;
        const int64_t gate = force_direct ? INT64_MAX : count * (int64_t)es;
        const bool mine_ok = in == out && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
        const int rc = try_zero_copy(c, mine_ok ? out : nullptr, count * (int64_t)es, gate, &pv, &zc, out);
        if (rc) return rc;
        if (zc) {
            const int64_t n = count / (VE * c->world) * (VE * c->world);
            if (n > 0) {
                int rc2;
                if (pv.mc && nvls_ok && c->world > 2 - (int)env_flag("DLBB_MC_WORLD2", false) && env_flag("DLBB_MC_ALLREDUCE", true)) {
                    // symmetric tensor with a multicast mapping: the switch reduces (multimem.ld_reduce) and
                    // replicates (multimem.st) in place - every byte crosses each link about once per direction
                    DirectPlan pl{};
                    pl.mc_src = true;
                    pl.sink = SINK_MC;
                    pl.a.in = pv;
                    pl.a.out_mc = pv.mc;
                    const int64_t nv = n / VE, share = nv / c->world;
                    pl.a.v_lo = c->rank * share;
                    pl.a.v_hi = pl.a.v_lo + share;
                    pl.a.out_shift = 0;
                    rc2 = launch_direct(c, dtype, op, pl, s, n * (int64_t)es);
                } else if (env_flag("DLBB_DIRECT_LEGACY", false)) {
                    rc2 = launch_allreduce_direct(c, dtype, op, pv, n, s);
                } else {
                    // in place on everybody's tensor: rank r reduces its 1/P share and stores it into all of them
                    DirectPlan pl{};
                    pl.mc_src = false;
                    pl.sink = SINK_ALL;
                    pl.a.in = pv;
                    pl.a.out = pv;
                    const int64_t nv = n / VE, share = nv / c->world;
                    pl.a.v_lo = c->rank * share;
                    pl.a.v_hi = pl.a.v_lo + share;
                    pl.a.out_shift = 0;
                    rc2 = launch_direct(c, dtype, op, pl, s, n * (int64_t)es);
                }
                if (rc2) return rc2;
                done = n;
            }
        }
    }
    while (done < count) {
        const int64_t left = count - done;
        const int64_t bytes = left * (int64_t)es;
        bool oneshot = algo == DLBB_ALGO_ONESHOT || (algo == DLBB_ALGO_AUTO && bytes <= oneshot_max) || c->world == 1;
        // two-shot works on multiples of VE*world elements; a ragged tail goes through one-shot
        int64_t n = std::min(left, cap);
        if (!oneshot) {
            n = n / (VE * c->world) * (VE * c->world);
            if (n == 0) { oneshot = true; n = std::min(left, cap); }
        }
        const char* ext = extra ? static_cast<const char*>(extra) + done * es : nullptr;
        char* o = static_cast<char*>(out) + (c->out_view ? 0 : done * es);   // a strided output is addressed through its view
        c->out_base = done;
        int rc;
        if (oneshot) {
            PieceList pack{};
            pack.count = 1;
            pack.p[0] = Piece{0, done, 0, n};
            rc = launch_reduce(c, dtype, op, in, o, ext, pack, view, 0, n, n * es, s);
        } else {
            const bool nvls = nvls_ok && (algo == DLBB_ALGO_NVLS || algo == DLBB_ALGO_AUTO);
            rc = launch_twoshot(c, dtype, op, nvls, in, o, ext, view, done, n, s);
        }
        if (rc) return rc;
        done += n;
    }
    return 0;
}

template <typename T, int OP>
static int launch_multi_T(dlbb_comm* c, const MultiTable& tb, int64_t bytes, cudaStream_t s) {
    // latency-bound: about two staging vectors per thread, so that a bucket of small tensors is one round trip deep
    const int64_t total_vecs = bytes >> 4;
    const int want = c->ctas > 0 ? c->ctas : static_cast<int>(std::min<int64_t>(96, std::max<int64_t>(1, total_vecs / (2 * kCollThreads))));
    const int ctas = std::min(want, c->max_resident_ctas);   // grid barrier inside
    const int64_t so = next_stage_off(c, s);
    const uint32_t ep = ++c->epoch[0];
    reduce_multi_kernel<T, OP><<<ctas, kCollThreads, 0, s>>>(c->dev, tb, total_vecs, so, 0, ep);
    DLBB_CHECK_LAUNCH("reduce_multi_kernel");
    return 0;
}
template <typename T>
static int launch_multi_op(dlbb_comm* c, int op, const MultiTable& tb, int64_t bytes, cudaStream_t s) {
    switch (op) {
        case DLBB_SUM: return launch_multi_T<T, DLBB_SUM>(c, tb, bytes, s);
        case DLBB_MAX: return launch_multi_T<T, DLBB_MAX>(c, tb, bytes, s);
        case DLBB_MIN: return launch_multi_T<T, DLBB_MIN>(c, tb, bytes, s);
        case DLBB_PROD: return launch_multi_T<T, DLBB_PROD>(c, tb, bytes, s);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}
static int launch_multi(dlbb_comm* c, int dtype, int op, const MultiTable& tb, int64_t bytes, cudaStream_t s) {
    switch (dtype) {
        case DLBB_F16: return launch_multi_op<__half>(c, op, tb, bytes, s);
        case DLBB_BF16: return launch_multi_op<__nv_bfloat16>(c, op, tb, bytes, s);
        case DLBB_F32: return launch_multi_op<float>(c, op, tb, bytes, s);
        case DLBB_F64: return launch_multi_op<double>(c, op, tb, bytes, s);
        case DLBB_I32: return launch_multi_op<int32_t>(c, op, tb, bytes, s);
        case DLBB_I64: return launch_multi_op<int64_t>(c, op, tb, bytes, s);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

int refresh_flag_bank_if_due(dlbb_comm* c, int bank, void* stream) {
    static const uint32_t every = getenv("DLBB_FLAG_REFRESH_EVERY") ? static_cast<uint32_t>(strtoul(getenv("DLBB_FLAG_REFRESH_EVERY"), nullptr, 0))
                                                                    : (1u << 28);
    if (c->world == 1 || c->epoch[bank] - c->last_refresh[bank] < every) return 0;
    const uint32_t ep = ++c->epoch[bank];
    barrier_kernel<<<kMaxCtas, 32, 0, static_cast<cudaStream_t>(stream)>>>(c->dev, bank, ep);   // row barriers: any residency
    DLBB_CHECK_LAUNCH("barrier_kernel (flag refresh)");
    c->last_refresh[bank] = ep;
    return 0;
}

// the banks of this file: 0 = barriers of every kernel, 4 / 5 = chunk flags of the pipelined / chain / relay kernels
// (bank 3, point-to-point, counts per pair of ranks and rewrites all of a pair's flags with every message)
static int refresh_collective_flags(dlbb_comm* c, void* stream) {
    for (int bank : {0, 4, 5})
        if (const int rc = refresh_flag_bank_if_due(c, bank, stream)) return rc;
    return 0;
}

}  // namespace dlbb

using namespace dlbb;

extern "C" {

int dlbb_barrier(dlbb_comm_t* c, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const uint32_t ep = ++c->epoch[0];
    barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(c->dev, 0, ep);
    DLBB_CHECK_LAUNCH("barrier_kernel");
    return 0;
}

int dlbb_allreduce(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int algo,
                   const dlbb_view_t* in_view, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    return allreduce_impl(c, in, out, nullptr, count, dtype, op, algo, in_view, static_cast<cudaStream_t>(stream));
}

int dlbb_allreduce_v(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int algo,
                     const dlbb_view_t* in_view, const dlbb_view_t* out_view, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    // a strided destination keeps the call on the staged kernels, whose unpack phase scatters by the view
    c->out_view = (out_view && out_view->ndim > 0) ? out_view : nullptr;
    static const dlbb_view_t kDense1 = {1, {0, 0, 0, 0}, {1, 0, 0, 0}};
    dlbb_view_t dense = kDense1;
    dense.size[0] = count;
    // (zero-copy needs dense tensors on both sides: a 1-D dense descriptor for the input switches it off as well)
    const dlbb_view_t* iv = (in_view && in_view->ndim > 0) ? in_view : (c->out_view ? &dense : nullptr);
    const int rc = allreduce_impl(c, in, out, nullptr, count, dtype, op, algo, iv, static_cast<cudaStream_t>(stream));
    c->out_view = nullptr;
    return rc;
}

int dlbb_allreduce_residual_bf16(dlbb_comm_t* c, const void* in, const void* residual, void* out, int64_t count,
                                 void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    return allreduce_impl(c, in, out, residual, count, DLBB_BF16, DLBB_SUM, DLBB_ALGO_AUTO, nullptr,
                          static_cast<cudaStream_t>(stream));
}

int dlbb_reduce(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int root, void* stream) {
    return dlbb_reduce_v(c, in, out, count, dtype, op, root, nullptr, nullptr, stream);
}

// Strided tensors (K12): the input view is gathered by the pack phase, the root's output view is scattered by the
// store / unpack phase of the same kernel.  Views keep the call on the staged kernels.
static int reduce_views(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int root,
                        const ViewDev& view, cudaStream_t s) {
    const size_t es = dtype_size(dtype);
    const int64_t VE = 16 / (int64_t)es;
    const bool nvls_ok = dlbb_comm_has_multicast(c) && op == DLBB_SUM && (dtype == DLBB_F16 || dtype == DLBB_BF16 || dtype == DLBB_F32);
    int64_t done = 0;
    if (c->world > 2) {
        const int64_t unit = VE * c->world;
        const int64_t cap2 = static_cast<int64_t>(c->stage_bytes / es) / unit * unit;
        while (count - done >= unit) {
            const int64_t n = std::min(count - done, cap2) / unit * unit;
            c->out_base = done;
            const int rc = launch_twoshot(c, dtype, op, nvls_ok, in, static_cast<char*>(out) + (c->out_view ? 0 : done * es), nullptr, view, done, n, s, root);
            if (rc) return rc;
            done += n;
        }
    }
    if (done < count) {
        const int64_t cap = static_cast<int64_t>(c->stage_bytes / es);
        for (; done < count; done += cap) {
            const int64_t n = std::min(cap, count - done);
            c->out_base = done;
            PieceList pack{};
            pack.count = 1;
            pack.p[0] = Piece{0, done, 0, n};
            const int rc = launch_reduce(c, dtype, op, in, static_cast<char*>(out) + (c->out_view ? 0 : done * es), nullptr, pack, view, 0,
                                         c->rank == root ? n : 0, n * es, s);
            if (rc) return rc;
        }
    }
    return 0;
}

int dlbb_reduce_v(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int root,
                  const dlbb_view_t* in_view, const dlbb_view_t* out_view, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const size_t es = dtype_size(dtype);
    if ((in_view && in_view->ndim > 0) || (out_view && out_view->ndim > 0)) {
        DLBB_REQUIRE(es != 0 && dtype != DLBB_U8, "reduce: unsupported dtype %d", dtype);
        DLBB_REQUIRE(root >= 0 && root < c->world, "reduce: bad root %d", root);
        c->out_view = (out_view && out_view->ndim > 0) ? out_view : nullptr;
        const int rc = reduce_views(c, in, out, count, dtype, op, root, make_view(in_view), static_cast<cudaStream_t>(stream));
        c->out_view = nullptr;
        return rc;
    }
    DLBB_REQUIRE(es != 0 && dtype != DLBB_U8, "reduce: unsupported dtype %d", dtype);
    DLBB_REQUIRE(root >= 0 && root < c->world, "reduce: bad root %d", root);
    const int64_t VE = 16 / (int64_t)es;
    const bool nvls_ok = dlbb_comm_has_multicast(c) && op == DLBB_SUM &&
                         (dtype == DLBB_F16 || dtype == DLBB_BF16 || dtype == DLBB_F32);
    if (c->world > 2 && c->world <= kDirectMaxWorld && (count * (int64_t)es) % 16 == 0 && count > 0 && env_flag("DLBB_REDUCE_DIRECT", true)) {
        // Zero-copy, in place (what dist.reduce is): the non-root ranks each reduce 1/(P-1) of the message straight out
        // of all ranks' tensors and store the result into the root's tensor - every link carries the message once in
        // each direction and nothing is staged.  With a multicast mapping all P ranks reduce 1/P in the switch.
        PeerView pv;
        bool zc = false;
        const int64_t bytes = count * (int64_t)es;
        const bool mine_ok = aligned16(in) && in == out;
        const int rc = try_zero_copy(c, mine_ok ? in : nullptr, bytes, bytes, &pv, &zc, in);
        if (rc) return rc;
        if (zc) {
            const int64_t nv = bytes >> 4;
            const bool use_mc = pv.mc && nvls_ok && env_flag("DLBB_MC_REDUCE", true);
            static const int64_t pull_min = getenv("DLBB_REDUCE_PULL_MIN_BYTES") ? atoll(getenv("DLBB_REDUCE_PULL_MIN_BYTES")) : (4ll << 20);
            if (bytes >= pull_min && env_flag("DLBB_REDUCE_PULL", false))
                return launch_reduce_pull(c, dtype, op, pv, out, nv, root, use_mc, static_cast<cudaStream_t>(stream));
            // small messages: one hop - the reducers store their share straight into the root's tensor
            DirectPlan pl{};
            pl.mc_src = use_mc;
            pl.sink = SINK_ONE;
            pl.a.in = pv;
            pl.a.out.ptr[0] = pv.ptr[root];
            pl.a.out_shift = 0;
            if (pl.mc_src) {          // P reducers, one share each (direct_math.h; rows = 1: the whole share)
                int64_t s_lo;
                share_row_range(nv, c->world, 1, c->rank, 0, s_lo, pl.a.v_lo, pl.a.v_hi);
            } else if (c->rank != root) {   // P-1 reducers numbered from root+1
                int64_t s_lo;
                share_row_range(nv, c->world - 1, 1, relay_index(c->rank, root, c->world), 0, s_lo, pl.a.v_lo, pl.a.v_hi);
            }
            return launch_direct(c, dtype, op, pl, static_cast<cudaStream_t>(stream), bytes);
        }
    }
    if (c->world > 2 && chain_min_bytes() > 0 && count * (int64_t)es >= chain_min_bytes()) {
        // Large message, several senders: pipelined chain ending at the root, every link carries the message once.
        const int64_t cap2 = static_cast<int64_t>(c->stage_bytes / es) / VE * VE;
        int64_t done = 0;
        while (count - done >= VE) {
            const int64_t n = std::min(count - done, cap2) / VE * VE;
            const int rc = launch_chain_reduce(c, dtype, op, static_cast<const char*>(in) + done * es,
                                               static_cast<char*>(out) + done * es, n, root, static_cast<cudaStream_t>(stream));
            if (rc) return rc;
            done += n;
        }
        if (done < count) {   // fewer than 16 bytes left: single-barrier form
            const int64_t n = count - done;
            PieceList pack{};
            pack.count = 1;
            pack.p[0] = Piece{0, done, 0, n};
            const int rc = launch_reduce(c, dtype, op, in, static_cast<char*>(out) + done * es, nullptr, pack, ViewDev{}, 0,
                                         c->rank == root ? n : 0, n * es, static_cast<cudaStream_t>(stream));
            if (rc) return rc;
        }
        return 0;
    }
    if (c->world > 2 && (nvls_ok || count * (int64_t)es > (512 << 10))) {
        // More than one sender: a root that pulls everything is limited by its own link ((world-1) x the message;
        // measured at 8 GPUs, 128 MiB: 109 GB/s).  Use the all-reduce data flow instead - every rank reduces
        // 1/world of the message (in the switch when NVLS applies) - and deliver the shares to the root only.
        const int64_t unit = VE * c->world;
        const int64_t cap2 = static_cast<int64_t>(c->stage_bytes / es) / unit * unit;
        int64_t done = 0;
        while (count - done >= unit) {
            const int64_t n = std::min(count - done, cap2) / unit * unit;
            const int rc = launch_twoshot(c, dtype, op, nvls_ok, in, static_cast<char*>(out) + done * es, nullptr, ViewDev{},
                                          done, n, static_cast<cudaStream_t>(stream), root);
            if (rc) return rc;
            done += n;
        }
        if (done < count) {   // ragged tail: single-barrier form
            const int64_t n = count - done;
            PieceList pack{};
            pack.count = 1;
            pack.p[0] = Piece{0, done, 0, n};
            const int rc = launch_reduce(c, dtype, op, in, static_cast<char*>(out) + done * es, nullptr, pack, ViewDev{}, 0,
                                         c->rank == root ? n : 0, n * es, static_cast<cudaStream_t>(stream));
            if (rc) return rc;
        }
        return 0;
    }
    {   // zero-copy (two ranks): the root reduces straight out of the peer's input tensor
        PeerView pv;
        bool zc = false;
        const bool mine_ok = aligned16(in) && (c->rank != root || aligned16(out));
        const int rc = try_zero_copy(c, mine_ok ? in : nullptr, count * (int64_t)es, count * (int64_t)es, &pv, &zc);
        if (rc) return rc;
        if (zc)
            return launch_reduce(c, dtype, op, in, out, nullptr, PieceList{}, ViewDev{}, 0, c->rank == root ? count : 0,
                                 count * es, static_cast<cudaStream_t>(stream), &pv);
    }
    const int64_t cap = static_cast<int64_t>(c->stage_bytes / es);
    for (int64_t done = 0; done < count; done += cap) {
        const int64_t n = std::min(cap, count - done);
        PieceList pack{};
        pack.count = 1;
        pack.p[0] = Piece{0, done, 0, n};
        const int rc = launch_reduce(c, dtype, op, in, static_cast<char*>(out) + done * es, nullptr, pack, ViewDev{}, 0,
                                     c->rank == root ? n : 0, n * es, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int dlbb_reduce_scatter(dlbb_comm_t* c, const void* in, void* out, int64_t cpr, int dtype, int op, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const size_t es = dtype_size(dtype);
    DLBB_REQUIRE(es != 0 && dtype != DLBB_U8, "reduce_scatter: unsupported dtype %d", dtype);
    if ((cpr * (int64_t)es) % 16 == 0) {   // zero-copy: rank r reduces shard r of every rank's input tensor
        PeerView pv;
        bool zc = false;
        const int64_t total = cpr * (int64_t)es * c->world;
        const bool mine_ok = aligned16(in) && aligned16(out);
        const int rc = try_zero_copy(c, mine_ok ? in : nullptr, total, total, &pv, &zc, in);
        if (rc) return rc;
        if (zc && c->world <= kDirectMaxWorld) {
            // all peers' loads of a vector in flight at once (or one multimem.ld_reduce on a symmetric input)
            const bool nvls_ok = dlbb_comm_has_multicast(c) && op == DLBB_SUM &&
                                 (dtype == DLBB_F16 || dtype == DLBB_BF16 || dtype == DLBB_F32);
            DirectPlan pl{};
            // (measured: the switch reduction never beats the peer loads here - 334 vs 648 GB/s on 2 GPUs, 520 vs 656 on 4,
            //  626 vs 634 on 8 - so multimem.ld_reduce is opt-in for reduce-scatter)
            pl.mc_src = pv.mc && nvls_ok && env_flag("DLBB_MC_REDUCE_SCATTER", false);
            pl.sink = SINK_ONE;
            pl.a.in = pv;
            pl.a.out.ptr[0] = static_cast<char*>(out);
            pl.a.v_lo = (c->rank * cpr * (int64_t)es) >> 4;
            pl.a.v_hi = pl.a.v_lo + ((cpr * (int64_t)es) >> 4);
            pl.a.out_shift = pl.a.v_lo;
            return launch_direct(c, dtype, op, pl, static_cast<cudaStream_t>(stream), total);
        }
        if (zc)
            return launch_reduce(c, dtype, op, in, out, nullptr, PieceList{}, ViewDev{}, c->rank * cpr, cpr, total,
                                 static_cast<cudaStream_t>(stream), &pv);
    }
    // per launch every rank stages `chunk` elements for each destination rank
    int64_t cap = static_cast<int64_t>(c->stage_bytes / es) / c->world;
    cap = cap / 16 * 16;
    DLBB_REQUIRE(cap > 0, "reduce_scatter: staging too small");
    for (int64_t done = 0; done < cpr; done += cap) {
        const int64_t n = std::min(cap, cpr - done);
        const int64_t slot = (n + 15) / 16 * 16;  // keep every destination slot 16 B aligned
        PieceList pack{};
        pack.count = c->world;
        for (int p = 0; p < c->world; ++p) pack.p[p] = Piece{0, p * cpr + done, p * slot, n};
        const int rc = launch_reduce(c, dtype, op, in, static_cast<char*>(out) + done * es, nullptr, pack, ViewDev{},
                                     c->rank * slot, n, n * es * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int dlbb_allgather(dlbb_comm_t* c, const void* in, void* out, int64_t cpr, int dtype, const dlbb_view_t* in_view,
                   void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "allgather: unsupported dtype %d", dtype);
    const int64_t total = cpr * es;
    const int64_t cap = static_cast<int64_t>(c->stage_bytes) / 64 * 64;
    const ViewDev view = make_view(in_view);
    if (view.ndim == 0) {   // zero-copy: pull every rank's input tensor straight into my output
        PeerView pv;
        bool zc = false;
        char* my_slot = static_cast<char*>(out) + c->rank * total;
        const bool mine_ok = in == my_slot || disjoint(in, total, out, total * c->world);
        const int rc = try_zero_copy(c, mine_ok ? in : nullptr, total, total * c->world, &pv, &zc);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            pull.count = c->world;
            for (int i = 0; i < c->world; ++i) {
                const int p = (c->rank + i) % c->world;
                pull.p[i] = Piece{p, 0, p * total, total};
            }
            return launch_move(c, in, out, PieceList{}, pull, view, (int)es, total * c->world,
                               static_cast<cudaStream_t>(stream), &pv);
        }
    }
    for (int64_t done = 0; done < total; done += cap) {
        const int64_t n = std::min(cap, total - done);
        PieceList pack{}, pull{};
        pack.count = 1;
        pack.p[0] = Piece{0, done, 0, n};
        pull.count = c->world;
        for (int i = 0; i < c->world; ++i) {
            const int p = (c->rank + i) % c->world;  // start with my own copy, spread peers over time
            pull.p[i] = Piece{p, 0, p * total + done, n};
        }
        const int rc = launch_move(c, in, out, pack, pull, view, (int)es, n * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int dlbb_broadcast_v(dlbb_comm_t* c, void* buf, int64_t count, int dtype, int root, const dlbb_view_t* buf_view, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "broadcast: unsupported dtype %d", dtype);
    DLBB_REQUIRE(root >= 0 && root < c->world, "broadcast: bad root %d", root);
    const ViewDev view = make_view(buf_view);   // strided tensor (same geometry on every rank): packed / unpacked in-kernel
    const int64_t total = count * es, cap = static_cast<int64_t>(c->stage_bytes) / 64 * 64;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (view.ndim == 0 && c->world > 2 && c->world <= kDirectMaxWorld && total % 16 == 0 && total > 0 && env_flag("DLBB_BCAST_RELAY", true)) {
        // Zero-copy relay (direct_kernels.cuh): the root's link carries the message once, nothing is staged.
        PeerView pv;
        bool zc = false;
        const int rc = try_zero_copy(c, aligned16(buf) ? buf : nullptr, total, total, &pv, &zc, buf);
        if (rc) return rc;
        if (zc) return launch_bcast_relay(c, pv, total >> 4, root, s);
    }
    if (view.ndim == 0 && c->world > 2 && chain_min_bytes() > 0 && total >= chain_min_bytes()) {
        // Large message, several receivers: pipelined chain, every link carries the message once.
        for (int64_t done = 0; done < total; done += cap) {
            const int64_t n = std::min(cap, total - done);
            const ChainGeom g = chain_geom(c, n >> 4, s);
            chain_broadcast_kernel<<<g.ctas, kCollThreads, 0, s>>>(
                c->dev, static_cast<const char*>(buf) + done, static_cast<char*>(buf) + done, n, root, g.stage_off,
                g.flag_base, g.nchunks, g.end_epoch);
            DLBB_CHECK_LAUNCH("chain_broadcast_kernel");
        }
        return 0;
    }
    if (c->world > 2 && dlbb_comm_has_multicast(c)) {
        // More than one receiver: one multicast store through the switch instead of world-1 pulls from the root
        // (measured at 8 GPUs, 128 MiB: 109 GB/s for the pull forms - the root's link carries the message 7 times).
        for (int64_t done = 0; done < total; done += cap) {
            const int64_t n = std::min(cap, total - done);
            const int ctas = c->ctas > 0 ? std::min(c->ctas, kMaxCtas)
                                         : static_cast<int>(std::min<int64_t>(128, std::max<int64_t>(1, n / (64 << 10))));
            const int64_t so = next_stage_off(c, s);
            const int64_t per = ((n >> 4) + ctas - 1) / ctas;
            const int nchunks = static_cast<int>(std::max<int64_t>(1, (per + kBcChunkVecs - 1) / kBcChunkVecs));
            const uint32_t fb = c->epoch[kBcFlagBank];
            c->epoch[kBcFlagBank] += nchunks;
            c->epoch[0] += 2;   // opening and closing barrier
            const uint32_t ep = c->epoch[0];
            const int64_t ptr_off = view.ndim ? 0 : done;   // a strided tensor is addressed through the view from its base
            broadcast_nvls_kernel<<<ctas, kCollThreads, 0, s>>>(
                c->dev, static_cast<const char*>(buf) + ptr_off, static_cast<char*>(buf) + ptr_off, n, root, so, fb, nchunks, 0, ep,
                view, (int)es, view.ndim ? done : 0);
            DLBB_CHECK_LAUNCH("broadcast_nvls_kernel");
        }
        return 0;
    }
    if (view.ndim == 0) {   // zero-copy: everyone pulls from the root's tensor
        PeerView pv;
        bool zc = false;
        const int rc = try_zero_copy(c, buf, total, total, &pv, &zc);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            if (c->rank != root) { pull.count = 1; pull.p[0] = Piece{root, 0, 0, total}; }
            return launch_move(c, buf, buf, PieceList{}, pull, ViewDev{}, (int)es, total, s, &pv);
        }
    }
    c->out_view = buf_view;   // receivers scatter into the view inside the kernel
    c->out_base = 0;
    int rc = 0;
    for (int64_t done = 0; done < total && rc == 0; done += cap) {
        const int64_t n = std::min(cap, total - done);
        PieceList pack{}, pull{};
        if (c->rank == root) { pack.count = 1; pack.p[0] = Piece{0, done, 0, n}; }
        else { pull.count = 1; pull.p[0] = Piece{root, 0, done, n}; }
        rc = launch_move(c, buf, buf, pack, pull, view, (int)es, n, s);
    }
    c->out_view = nullptr;
    return rc;
}

int dlbb_broadcast(dlbb_comm_t* c, void* buf, int64_t count, int dtype, int root, void* stream) {
    return dlbb_broadcast_v(c, buf, count, dtype, root, nullptr, stream);
}

int dlbb_gather(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int root, void* stream) {
    return dlbb_gather_v(c, in, out, count, dtype, root, nullptr, stream);
}

int dlbb_gather_v(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int root, const dlbb_view_t* in_view,
                  void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "gather: unsupported dtype %d", dtype);
    DLBB_REQUIRE(root >= 0 && root < c->world, "gather: bad root %d", root);
    const int64_t total = count * es, cap = static_cast<int64_t>(c->stage_bytes) / 64 * 64;
    const ViewDev gview = make_view(in_view);   // strided input: gathered by the pack phase of the kernel
    if (gview.ndim == 0) {   // zero-copy: the root pulls every rank's input tensor
        PeerView pv;
        bool zc = false;
        const bool mine_ok = c->rank != root || in == static_cast<char*>(out) + root * total ||
                             disjoint(in, total, out, total * c->world);
        const int rc = try_zero_copy(c, mine_ok ? in : nullptr, total, total * c->world, &pv, &zc);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            if (c->rank == root) {
                pull.count = c->world;
                for (int p = 0; p < c->world; ++p) pull.p[p] = Piece{p, 0, p * total, total};
            }
            return launch_move(c, in, out, PieceList{}, pull, ViewDev{}, (int)es, total * c->world,
                               static_cast<cudaStream_t>(stream), &pv);
        }
    }
    for (int64_t done = 0; done < total; done += cap) {
        const int64_t n = std::min(cap, total - done);
        PieceList pack{}, pull{};
        pack.count = 1;
        pack.p[0] = Piece{0, done, 0, n};
        if (c->rank == root) {
            pull.count = c->world;
            for (int p = 0; p < c->world; ++p) pull.p[p] = Piece{p, 0, p * total + done, n};
        }
        const int rc = launch_move(c, in, out, pack, pull, gview, (int)es, n * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int dlbb_scatter(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int root, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "scatter: unsupported dtype %d", dtype);
    DLBB_REQUIRE(root >= 0 && root < c->world, "scatter: bad root %d", root);
    const int64_t total = count * es;
    {   // zero-copy: everyone pulls its slice of the root's tensor (non-roots export their output as a stand-in)
        PeerView pv;
        bool zc = false;
        const bool is_root = c->rank == root;
        const bool mine_ok = !is_root || (in != nullptr && disjoint(in, total * c->world, out, total));
        const void* mine = is_root ? in : out;
        const int rc = try_zero_copy(c, mine_ok ? mine : nullptr, is_root ? total * c->world : total, total * c->world,
                                     &pv, &zc);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            pull.count = 1;
            pull.p[0] = Piece{root, c->rank * total, 0, total};
            return launch_move(c, in, out, PieceList{}, pull, ViewDev{}, (int)es, total * c->world,
                               static_cast<cudaStream_t>(stream), &pv);
        }
    }
    const int64_t cap = static_cast<int64_t>(c->stage_bytes) / c->world / 64 * 64;
    DLBB_REQUIRE(cap > 0, "scatter: staging too small");
    for (int64_t done = 0; done < total; done += cap) {
        const int64_t n = std::min(cap, total - done);
        const int64_t slot = (n + 63) / 64 * 64;
        PieceList pack{}, pull{};
        if (c->rank == root) {
            pack.count = c->world;
            for (int p = 0; p < c->world; ++p) pack.p[p] = Piece{0, p * total + done, p * slot, n};
        }
        pull.count = 1;
        pull.p[0] = Piece{root, c->rank * slot, done, n};
        const int rc = launch_move(c, in, out, pack, pull, ViewDev{}, (int)es, n * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int dlbb_alltoall(dlbb_comm_t* c, const void* in, void* out, int64_t cpr, int dtype, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "alltoall: unsupported dtype %d", dtype);
    const int64_t total = cpr * es;
    {   // zero-copy: pull my slice out of every rank's input tensor
        PeerView pv;
        bool zc = false;
        const bool mine_ok = disjoint(in, total * c->world, out, total * c->world);
        const int rc = try_zero_copy(c, mine_ok ? in : nullptr, total * c->world, total * c->world, &pv, &zc);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            pull.count = c->world;
            for (int i = 0; i < c->world; ++i) {
                const int p = (c->rank + i) % c->world;
                pull.p[i] = Piece{p, c->rank * total, p * total, total};
            }
            return launch_move(c, in, out, PieceList{}, pull, ViewDev{}, (int)es, total * c->world,
                               static_cast<cudaStream_t>(stream), &pv);
        }
    }
    const int64_t cap = static_cast<int64_t>(c->stage_bytes) / c->world / 64 * 64;
    DLBB_REQUIRE(cap > 0, "alltoall: staging too small");
    for (int64_t done = 0; done < total; done += cap) {
        const int64_t n = std::min(cap, total - done);
        const int64_t slot = (n + 63) / 64 * 64;
        PieceList pack{}, pull{};
        pack.count = c->world;
        for (int p = 0; p < c->world; ++p) pack.p[p] = Piece{0, p * total + done, p * slot, n};
        pull.count = c->world;
        for (int i = 0; i < c->world; ++i) {
            const int p = (c->rank + i) % c->world;
            pull.p[i] = Piece{p, c->rank * slot, p * total + done, n};
        }
        const int rc = launch_move(c, in, out, pack, pull, ViewDev{}, (int)es, n * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int dlbb_sendrecv(dlbb_comm_t* c, const void* send_buf, int64_t send_count, int send_peer, void* recv_buf,
                  int64_t recv_count, int recv_peer, int dtype, void* stream) {
    // (no flag refresh here: point-to-point calls may run on their own stream beside a collective, and bank 3 needs none)
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "sendrecv: unsupported dtype %d", dtype);
    DLBB_REQUIRE(send_peer < c->world && recv_peer < c->world, "sendrecv: bad peer");
    DLBB_REQUIRE(send_peer != c->rank && recv_peer != c->rank, "sendrecv: a rank cannot message itself");
    const int64_t sb = send_peer >= 0 ? send_count * es : 0, rb = recv_peer >= 0 ? recv_count * es : 0;
    const int64_t slot = static_cast<int64_t>(c->stage_bytes) / c->world / 256 * 256;
    DLBB_REQUIRE(slot > 0, "sendrecv: staging too small");
    const int64_t region = 2 * static_cast<int64_t>(c->stage_bytes);
    // Message k of a directed pair uses epoch k on that pair's flags; long messages are cut into
    // slot-sized chunks, each its own epoch, sender and receiver advancing in lock step.
    const int64_t ns = send_peer >= 0 ? std::max<int64_t>(1, (sb + slot - 1) / slot) : 0;
    const int64_t nr = recv_peer >= 0 ? std::max<int64_t>(1, (rb + slot - 1) / slot) : 0;
    for (int64_t k = 0; k < std::max(ns, nr); ++k) {
        const bool do_s = k < ns, do_r = k < nr;
        const int64_t sn = do_s ? std::max<int64_t>(0, std::min(slot, sb - k * slot)) : 0;
        const int64_t rn = do_r ? std::max<int64_t>(0, std::min(slot, rb - k * slot)) : 0;
        const uint32_t se = do_s ? ++c->send_epoch[send_peer] : 0;
        const uint32_t re = do_r ? ++c->recv_epoch[recv_peer] : 0;
        p2p_kernel<<<kP2PCtas, kCollThreads, 0, static_cast<cudaStream_t>(stream)>>>(
            c->dev, static_cast<const char*>(send_buf) + k * slot, sn, do_s ? send_peer : -1, se,
            static_cast<char*>(recv_buf) + k * slot, rn, do_r ? recv_peer : -1, re, region, slot);
        DLBB_CHECK_LAUNCH("p2p_kernel");
    }
    return 0;
}

int dlbb_allreduce_multi(dlbb_comm_t* c, void* const* bufs, const int64_t* counts, int n_tensors, int dtype, int op,
                         void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0 && dtype != DLBB_U8, "allreduce_multi: unsupported dtype %d", dtype);
    DLBB_REQUIRE(n_tensors >= 0, "allreduce_multi: negative tensor count");
    // DLBB_FUSION=0: one collective per tensor; DLBB_FUSION_BYTES: the most a fused launch carries (default 16 MiB);
    // a tensor above that size travels alone through the size-selected all-reduce.
    static const bool fusion = env_flag("DLBB_FUSION", true);
    static const int64_t fusion_bytes_env = getenv("DLBB_FUSION_BYTES") ? atoll(getenv("DLBB_FUSION_BYTES")) : (16ll << 20);
    const int64_t limit = std::min<int64_t>(fusion_bytes_env, static_cast<int64_t>(c->stage_bytes));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    MultiTable tb{};
    int64_t used = 0;   // bytes of staging taken by the open bucket
    auto flush = [&]() -> int {
        if (tb.count == 0) return 0;
        const int rc = launch_multi(c, dtype, op, tb, used, s);
        tb.count = 0;
        used = 0;
        return rc;
    };
    for (int i = 0; i < n_tensors; ++i) {
        const int64_t bytes = counts[i] * es;
        if (counts[i] == 0) continue;
        if (!fusion || bytes > limit || c->world == 1) {
            const int rc = allreduce_impl(c, bufs[i], bufs[i], nullptr, counts[i], dtype, op, DLBB_ALGO_AUTO, nullptr, s);
            if (rc) return rc;
            continue;
        }
        const int64_t slot = (bytes + 15) / 16 * 16;
        if (tb.count == kMultiMax || used + slot > limit) {
            const int rc = flush();
            if (rc) return rc;
        }
        tb.buf[tb.count] = bufs[i];
        tb.n[tb.count] = counts[i];
        tb.off[tb.count] = used / es;
        ++tb.count;
        used += slot;
    }
    return flush();
}

// All-to-all on lists of tensors (the torch / mpi4py list API the sweeps use, collectives/1d/dsccl.py:134-146): chunk p of
// the message is its own tensor on both sides.  The pieces carry each tensor's offset from the first one, so the kernel
// packs straight out of the list and pulls straight into it - no flattening copy before or after.
int dlbb_alltoall_ptrs(dlbb_comm_t* c, const void* const* in_ptrs, void* const* out_ptrs, int64_t bytes_per_peer, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t total = bytes_per_peer;
    const int64_t cap = static_cast<int64_t>(c->stage_bytes) / c->world / 64 * 64;
    DLBB_REQUIRE(cap > 0, "alltoall: staging too small");
    const char* in0 = static_cast<const char*>(in_ptrs[0]);
    char* out0 = static_cast<char*>(out_ptrs[0]);
    {   // zero-copy when the input list is the chunks of one tensor (what tensor.chunk(world) gives): every rank pulls its
        // slice out of every peer's tensor straight into the output list's tensors.  (A rank whose list is not dense, or
        // whose outputs overlap its input, enters the exchange with a null buffer and all ranks stage together.)
        bool dense = total > 0;
        for (int p = 0; p < c->world && dense; ++p) {
            dense = static_cast<const char*>(in_ptrs[p]) == in0 + p * total;
            dense = dense && disjoint(in0, total * c->world, out_ptrs[p], total);
        }
        PeerView pv;
        bool zc = false;
        const int rc = try_zero_copy(c, dense ? in0 : nullptr, total * c->world, total * c->world, &pv, &zc, in0);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            pull.count = c->world;
            for (int i = 0; i < c->world; ++i) {
                const int p = (c->rank + i) % c->world;
                pull.p[i] = Piece{p, c->rank * total, static_cast<char*>(out_ptrs[p]) - out0, total};
            }
            return launch_move(c, in0, out0, PieceList{}, pull, ViewDev{}, 1, total * c->world, static_cast<cudaStream_t>(stream), &pv);
        }
        // Separate tensors (the reference's loop clones every chunk): one exchange per list position - exchange j tells
        // rank j where every peer keeps the tensor meant for it.  All ranks take part in all of them and see every
        // outcome, so they agree on falling back.  Worth `world` host round trips (~0.2 ms on 8 ranks) only for large
        // messages: the gate is the size of ONE tensor (DLBB_A2A_LIST_ZC_MIN_BYTES, default 64 MiB - at 16 MiB per peer
        // the staged form measured 1.27x NCCL on 8 GPUs and stays; at 128 MiB per peer it measured 0.74x).
        static const int64_t list_zc_min = getenv("DLBB_A2A_LIST_ZC_MIN_BYTES") ? atoll(getenv("DLBB_A2A_LIST_ZC_MIN_BYTES")) : (64ll << 20);
        if (!dense && c->zero_copy_min > 0 && total >= list_zc_min && total >= (int64_t)c->zero_copy_min) {
            PeerView mine{};
            bool all = true;
            for (int j = 0; j < c->world; ++j) {
                bool ok_j = false;
                bool usable = true;
                for (int p = 0; p < c->world; ++p) usable = usable && disjoint(in_ptrs[j], total, out_ptrs[p], total);
                PeerView pj;
                const int rc2 = try_zero_copy(c, usable ? in_ptrs[j] : nullptr, total, total, &pj, &ok_j, in_ptrs[j]);
                if (rc2) return rc2;
                all = all && ok_j;
                if (j == c->rank) mine = pj;
            }
            if (all) {
                PieceList pull{};
                pull.count = c->world;
                for (int i = 0; i < c->world; ++i) {
                    const int p = (c->rank + i) % c->world;
                    pull.p[i] = Piece{p, 0, static_cast<char*>(out_ptrs[p]) - out0, total};
                }
                mine.mc = nullptr;
                return launch_move(c, in0, out0, PieceList{}, pull, ViewDev{}, 1, total * c->world, static_cast<cudaStream_t>(stream), &mine);
            }
        }
    }
    for (int64_t done = 0; done < total || (total == 0 && done == 0); done += cap) {
        const int64_t n = std::min(cap, total - done);
        if (n <= 0) break;
        const int64_t slot = (n + 63) / 64 * 64;
        PieceList pack{}, pull{};
        pack.count = c->world;
        for (int p = 0; p < c->world; ++p)
            pack.p[p] = Piece{0, (static_cast<const char*>(in_ptrs[p]) - in0) + done, p * slot, n};
        pull.count = c->world;
        for (int i = 0; i < c->world; ++i) {
            const int p = (c->rank + i) % c->world;
            pull.p[i] = Piece{p, c->rank * slot, (static_cast<char*>(out_ptrs[p]) - out0) + done, n};
        }
        const int rc = launch_move(c, in0, out0, pack, pull, ViewDev{}, 1, n * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

// All-gather / gather into a LIST of output tensors (torch's list API, collectives/1d/dsccl.py:76,102): rank p's
// contribution is pulled straight into out_ptrs[p].  root < 0: all-gather; otherwise only the root receives.
int dlbb_allgather_ptrs(dlbb_comm_t* c, const void* in, void* const* out_ptrs, int64_t count, int dtype, int root,
                        const dlbb_view_t* in_view, void* stream) {
    if (const int rg = refresh_collective_flags(c, stream)) return rg;
    const int64_t es = (int64_t)dtype_size(dtype);
    DLBB_REQUIRE(es != 0, "allgather: unsupported dtype %d", dtype);
    DLBB_REQUIRE(root < c->world, "gather: bad root %d", root);
    const int64_t total = count * es, cap = static_cast<int64_t>(c->stage_bytes) / 64 * 64;
    const ViewDev view = make_view(in_view);
    const bool receiver = root < 0 || c->rank == root;
    char* out0 = receiver ? static_cast<char*>(out_ptrs[0]) : nullptr;
    if (view.ndim == 0) {   // zero-copy: the receivers pull every rank's input tensor straight into their list
        PeerView pv;
        bool zc = false;
        const int rc = try_zero_copy(c, in, total, total * c->world, &pv, &zc, in);
        if (rc) return rc;
        if (zc) {
            PieceList pull{};
            if (receiver) {
                pull.count = c->world;
                for (int i = 0; i < c->world; ++i) {
                    const int p = (c->rank + i) % c->world;
                    pull.p[i] = Piece{p, 0, static_cast<char*>(out_ptrs[p]) - out0, total};
                }
            }
            return launch_move(c, in, out0, PieceList{}, pull, view, (int)es, total * c->world, static_cast<cudaStream_t>(stream), &pv);
        }
    }
    for (int64_t done = 0; done < total; done += cap) {
        const int64_t n = std::min(cap, total - done);
        PieceList pack{}, pull{};
        pack.count = 1;
        pack.p[0] = Piece{0, done, 0, n};
        if (receiver) {
            pull.count = c->world;
            for (int i = 0; i < c->world; ++i) {
                const int p = (c->rank + i) % c->world;
                pull.p[i] = Piece{p, 0, (static_cast<char*>(out_ptrs[p]) - out0) + done, n};
            }
        }
        const int rc = launch_move(c, in, out0, pack, pull, view, (int)es, n * c->world, static_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

}  // extern "C"
