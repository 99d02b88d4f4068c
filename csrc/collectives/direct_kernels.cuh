// Zero-copy reductions and the relay broadcast: kernels that work on the ranks' OWN tensors (PeerView: CUDA-IPC
// mappings of plain cudaMalloc blocks, or the peers' mappings of a symmetric segment - peer_registry.cpp) with no
// staging copy on either side.  Included by collectives.cu (one translation unit: they share its helpers).
//
//   direct_reduce_kernel   every reducer rank owns a vector range of the logical buffer, reduces it across all
//                          ranks and delivers it to a sink.  Sources: plain peer loads (all peers' loads in flight
//                          at once, summed in rank order) or ONE multimem.ld_reduce when the buffer has a multicast
//                          mapping (symmetric tensors).  Sinks: this rank's output (reduce-scatter), one rank's
//                          tensor (reduce), every rank's tensor by peer stores or by ONE multimem.st (all-reduce).
//   bcast_relay_kernel     broadcast in two hops so that the root's link carries the message once: the non-root
//                          ranks each pull 1/(P-1) of the message from the root and pass it on - by multimem.st
//                          into everybody's tensor, or by letting the others pull it behind per-chunk flags.
//
// ref (what these replace): dist.reduce / dist.broadcast / dist.all_reduce / reduce-scatter of the sweeps,
// collectives/1d/dsccl.py:59-163, collectives/3d/dsccl.py:46-118; reduce-scatter is the ZeRO-2 gradient exchange of
// test/ccl.py:86-88.
#pragma once
#include "direct_math.h"

namespace dlbb {

constexpr int kDirThreads = 256;
// Peer loads in flight per thread: U vectors from each of up to MAXW peers, all in registers.  The product stays at 16
// (64 registers of payload): 2 ranks x 8, 4 x 4, 8 x 2 - a two-rank reduce-scatter with 2 vectors per peer reached only
// half the link rate (397 GB/s), it is latency-bound until ~4 MiB per GPU are in flight.
constexpr int kDirUR = 8;     // multimem.ld_reduce requests in flight per thread
constexpr int kRelayBank = 4;           // one-directional chunk flags (pad bank shared with the other pipelined kernels)
constexpr int kRelayChunkVecsDefault = 8192;   // 128 KiB chunks of the two-hop kernels (DLBB_RELAY_CHUNK_KB)
enum { SINK_ONE = 0, SINK_ALL = 1, SINK_MC = 2 };

struct DirectArgs {
    PeerView in;          // rank p's copy of the logical input
    PeerView out;         // SINK_ALL: rank p's copy of the logical output; SINK_ONE: out.ptr[0] is the sink
    char* out_mc;         // SINK_MC: multicast address of the logical output
    int64_t v_lo, v_hi;   // the 16-byte vectors of the input this rank reduces (empty on ranks that only synchronise)
    int64_t out_shift;    // output vector index = input vector index - out_shift
};

// Reduce the vectors [lo, hi) of the logical buffer across all ranks with the calling CTA and hand each result vector to
// `sink(i, value)`.  MC_SRC: one multimem.ld_reduce per vector (kDirUR in flight per thread); otherwise 16 / MAXW vectors
// from each of the (up to MAXW) ranks in flight per thread, summed in rank order -> bit-identical on every rank.
template <typename T, int OP, bool MC_SRC, int MAXW, typename Sink>
__device__ __forceinline__ void reduce_span(const CommDev& c, const PeerView& in, int64_t lo, int64_t hi, Sink&& sink) {
    constexpr int VE = 16 / sizeof(T);
    using A = typename Acc<T>::type;
    if constexpr (MC_SRC) {
        for (int64_t i0 = lo + threadIdx.x; i0 < hi; i0 += (int64_t)blockDim.x * kDirUR) {
            uint4 o[kDirUR];
#pragma unroll
            for (int u = 0; u < kDirUR; ++u)
                if (i0 + u * blockDim.x < hi) o[u] = MmReduce<T>::ld(in.mc + ((i0 + u * blockDim.x) << 4));
#pragma unroll
            for (int u = 0; u < kDirUR; ++u)
                if (i0 + u * blockDim.x < hi) sink(i0 + u * blockDim.x, o[u]);
        }
    } else {
        constexpr int kDirU = 16 / MAXW;
        for (int64_t i0 = lo + threadIdx.x; i0 < hi; i0 += (int64_t)blockDim.x * kDirU) {
            uint4 buf[MAXW][kDirU];
#pragma unroll
            for (int p = 0; p < MAXW; ++p) {
                if (p < c.world) {
#pragma unroll
                    for (int u = 0; u < kDirU; ++u)
                        if (i0 + u * blockDim.x < hi) buf[p][u] = ld_peer_v4(in.ptr[p] + ((i0 + u * blockDim.x) << 4));
                }
            }
#pragma unroll
            for (int u = 0; u < kDirU; ++u) {
                if (i0 + u * blockDim.x >= hi) continue;
                A acc[VE];
#pragma unroll
                for (int p = 0; p < MAXW; ++p) {
                    if (p < c.world) {
                        const T* t = reinterpret_cast<const T*>(&buf[p][u]);
#pragma unroll
                        for (int e = 0; e < VE; ++e) {
                            const A x = to_acc<T>(t[e]);
                            acc[e] = (p == 0) ? x : red_op<OP, A>(acc[e], x);
                        }
                    }
                }
                uint4 o;
                T* ot = reinterpret_cast<T*>(&o);
#pragma unroll
                for (int e = 0; e < VE; ++e) ot[e] = from_acc<T>(acc[e]);
                sink(i0 + u * blockDim.x, o);
            }
        }
    }
}

template <typename T, int OP, bool MC_SRC, int SINK, int MAXW>
__global__ void __launch_bounds__(kDirThreads)
direct_reduce_kernel(CommDev c, const __grid_constant__ DirectArgs a, int bank, uint32_t epoch) {
    const int64_t per = (a.v_hi - a.v_lo + gridDim.x - 1) / gridDim.x;
    const int64_t lo = min(a.v_lo + (int64_t)blockIdx.x * per, a.v_hi), hi = min(lo + per, a.v_hi);
    // entry: the peers' kernels run, so their inputs are final (CTA b only meets CTA b of the peers)
    row_barrier(c, bank, blockIdx.x, epoch);
    reduce_span<T, OP, MC_SRC, MAXW>(c, a.in, lo, hi, [&](int64_t i, const uint4& o) {
        const int64_t ob = (i - a.out_shift) << 4;
        if (SINK == SINK_ONE) {
            st_v4(a.out.ptr[0] + ob, o);
        } else if (SINK == SINK_MC) {
            mm_st_v4(a.out_mc + ob, o);
        } else {
#pragma unroll 1
            for (int p = 0; p < c.world; ++p) st_v4(a.out.ptr[(c.rank + p) % c.world] + ob, o);   // spread over the links
        }
    });
    // exit: nobody's stream may touch its tensor again before every peer has read / written what it had to
    row_barrier(c, bank, blockIdx.x, epoch + 1);
}

// ------------------------------------------------------------------------------------------
// Reduce to a root in two pulled hops.  Measured on 4 GPUs (1 GiB): reducers that load their share from the peers AND
// push the result into the root's tensor reach 430 GB/s although the loads alone run at 604 and the pushes alone at
// 629 - the two streams do not overlap on the reducers' links.  So nothing is pushed: reducer q reduces its share chunk
// by chunk into its OWN staging region and raises a per-chunk flag on the root; CTAs of the root wait for the flags and
// pull the reduced chunks into the root's tensor.  Every byte still crosses each link once per direction.
//   reducers: the non-root ranks (peer loads) or all ranks (multimem.ld_reduce; the root writes its share directly);
//   grid = groups x rows; group 0 reduces, the root's other groups pull from one reducer each.
// In place on the root (its tensor is input and output): a chunk is overwritten only after its reducer has read it.
template <typename T, int OP, bool MC_SRC, int MAXW>
__global__ void __launch_bounds__(kDirThreads)
reduce_pull_kernel(CommDev c, const __grid_constant__ PeerView in, char* __restrict__ out, int64_t nvec, int root, int rows,
                   int64_t stage_off, uint32_t flag_base, int chunk_vecs, int bank, uint32_t epoch) {
    const int P = c.world;
    const int nred = MC_SRC ? P : P - 1;
    const int my_j = MC_SRC ? c.rank : relay_index(c.rank, root, P);    // my reducer index (P2P: the root gets P-1 = none)
    const bool is_root = c.rank == root;
    const int group = blockIdx.x / rows, row = blockIdx.x - group * rows;
    const int nchunks = share_nchunks(nvec, nred, rows, chunk_vecs);   // same everywhere
    row_barrier(c, bank, blockIdx.x, epoch);
    auto row_range = [&](int j, int64_t& s_lo, int64_t& r_lo, int64_t& r_hi) { share_row_range(nvec, nred, rows, j, row, s_lo, r_lo, r_hi); };
    if (group == 0 && my_j < nred) {
        // ---- reduce my share; the root (multimem mode only) writes its share straight into its tensor
        int64_t s_lo, r_lo, r_hi;
        row_range(my_j, s_lo, r_lo, r_hi);
        char* stage = c.stage[c.rank] + stage_off;
        for (int k = 0; k < nchunks; ++k) {
            int64_t c_lo, c_hi;
            chunk_range(r_lo, r_hi, k, chunk_vecs, c_lo, c_hi);
            if (is_root) {
                reduce_span<T, OP, MC_SRC, MAXW>(c, in, c_lo, c_hi, [&](int64_t i, const uint4& o) { st_v4(out + (i << 4), o); });
            } else {
                reduce_span<T, OP, MC_SRC, MAXW>(c, in, c_lo, c_hi,
                                                 [&](int64_t i, const uint4& o) { st_v4(stage + ((i - s_lo) << 4), o); });
                __syncthreads();
                if (threadIdx.x == 0) st_release_sys(pad_slot(c, root, kRelayBank, row, c.rank), flag_base + k + 1);
            }
        }
    } else if (is_root && group >= (MC_SRC ? 1 : 0)) {
        // ---- the root pulls one reducer's share out of that reducer's staging
        const int j = MC_SRC ? (root + group) % P : group;              // reducer index served by this group
        const int q = MC_SRC ? j : relay_rank(j, root, P);              // its rank
        int64_t s_lo, r_lo, r_hi;
        row_range(j, s_lo, r_lo, r_hi);
        const char* from = c.stage[q] + stage_off;
        for (int k = 0; k < nchunks; ++k) {
            int64_t c_lo, c_hi;
            chunk_range(r_lo, r_hi, k, chunk_vecs, c_lo, c_hi);
            if (threadIdx.x == 0) wait_flag(pad_slot(c, root, kRelayBank, row, q), flag_base + k + 1, q, c.watchdog_ns);
            __syncthreads();
            for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
                uint4 v[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < c_hi) v[u] = ld_peer_v4(from + ((i0 + u * blockDim.x - s_lo) << 4));
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < c_hi) st_v4(out + ((i0 + u * blockDim.x) << 4), v[u]);
            }
        }
    }
    row_barrier(c, bank, blockIdx.x, epoch + 1);
}

// ------------------------------------------------------------------------------------------
// Broadcast relay.  The message is cut into P-1 shares, one per non-root rank ("relay" j = (rank-root-1) mod P).
// A share is cut into `rows` row ranges; a row range is walked in chunks.
//   MC:   CTA (row) of relay q pulls its row of share j(q) from the root's tensor (plain peer loads: the root's link
//         sends every byte once, in full-size packets) and multimem.st's it into EVERY rank's tensor.  No flags.
//   P2P:  grid = (P-1) x rows CTAs per rank.  CTAs [0, rows) pull the own share from the root into the own tensor and
//         raise "chunk k of my row is here" on every other relay's pad; CTAs of group g >= 1 wait for those flags
//         from the g-th other relay and pull that relay's share from its tensor.
// The root takes part in the barriers only.  In place on every rank (buf is both source and destination).

template <bool MC>
__global__ void __launch_bounds__(kDirThreads)
bcast_relay_kernel(CommDev c, const __grid_constant__ PeerView pv, int64_t nvec, int root, int rows, uint32_t flag_base,
                   int bank, uint32_t epoch, int kRelayChunkVecs) {
    const int P = c.world;
    const int me = relay_index(c.rank, root, P);         // relay index; P-1 = the root
    const bool is_root = c.rank == root;
    const int group = blockIdx.x / rows, row = blockIdx.x - group * rows;
    row_barrier(c, bank, blockIdx.x, epoch);
    if (!is_root) {
        // which relay's share this CTA moves: group 0 = mine (from the root), group g = the g-th other relay
        const int j = group == 0 ? me : other_relay(me, group, P);
        const int owner = relay_rank(j, root, P);         // rank that relays share j
        int64_t s_lo, r_lo, r_hi;
        share_row_range(nvec, P - 1, rows, j, row, s_lo, r_lo, r_hi);
        const int nchunks = share_nchunks(nvec, P - 1, rows, kRelayChunkVecs);   // same everywhere
        const char* from = group == 0 ? pv.ptr[root] : pv.ptr[owner];
        char* mine = pv.ptr[c.rank];
        if (MC) {
            if (group == 0) {
                for (int64_t i0 = r_lo + threadIdx.x; i0 < r_hi; i0 += (int64_t)blockDim.x * kU) {
                    uint4 v[kU];
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < r_hi) v[u] = ld_peer_v4(from + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < r_hi) mm_st_v4(pv.mc + ((i0 + u * blockDim.x) << 4), v[u]);
                }
            }
        } else {
            for (int k = 0; k < nchunks; ++k) {
                int64_t c_lo, c_hi;
                chunk_range(r_lo, r_hi, k, kRelayChunkVecs, c_lo, c_hi);
                if (group != 0) {
                    if (threadIdx.x == 0) wait_flag(pad_slot(c, c.rank, kRelayBank, row, owner), flag_base + k + 1, owner, c.watchdog_ns);
                    __syncthreads();
                }
                for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
                    uint4 v[kU];
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < c_hi) v[u] = ld_peer_v4(from + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                    for (int u = 0; u < kU; ++u)
                        if (i0 + u * blockDim.x < c_hi) st_v4(mine + ((i0 + u * blockDim.x) << 4), v[u]);
                }
                if (group == 0) {
                    __syncthreads();   // the whole chunk is in my tensor (st.release below orders it for the readers)
                    if (threadIdx.x < P && threadIdx.x != root && threadIdx.x != c.rank)
                        st_release_sys(pad_slot(c, threadIdx.x, kRelayBank, row, c.rank), flag_base + k + 1);
                }
            }
        }
    }
    row_barrier(c, bank, blockIdx.x, epoch + 1);
}

// ------------------------------------------------------------------------------------------
// Broadcast relay, push form (opt-in: DLBB_BCAST_PUSH=1; verified on the host emulation, not yet measured on a box).
// Same two hops and the same shares as the pull relay above, but every hop is a STORE: the root writes share j into relay
// j's tensor chunk by chunk and raises "chunk k is there" on that relay's pad; relay j, one group of CTAs per destination,
// waits for the flag, reads the chunk out of its OWN tensor (coherent loads: it was written during this kernel) and
// writes it into the tensor of one of the other P-2 relays.  No read request crosses NVLink, the only flags go
// root -> relay, and the closing row barrier tells every rank that all stores into its tensor have landed.
// Grid: (P-1) groups x rows on every rank.  Root: group g pushes share g.  Relay: group g < P-2 serves its g-th other
// relay, the last group has nothing to do (it only takes part in the barriers).
__global__ void __launch_bounds__(kDirThreads)
bcast_push_kernel(CommDev c, const __grid_constant__ PeerView pv, int64_t nvec, int root, int rows, uint32_t flag_base,
                  int bank, uint32_t epoch, int chunk_vecs) {
    const int P = c.world;
    const bool is_root = c.rank == root;
    const int me = relay_index(c.rank, root, P);
    // CTAs are numbered row-major (the P-1 groups of a row are neighbours): a relay's CTA waits for the root's CTA of the
    // same row, which is at most P-2 blocks away - with in-order block scheduling the kernel makes progress on any device
    // that holds P-1 CTAs at a time, however many rows there are (group-major numbering stalled in the emulation with
    // fewer slots than CTAs: the root's early CTAs sat in the closing barrier while its late ones could not start).
    const int row = blockIdx.x / (P - 1), group = blockIdx.x - row * (P - 1);
    const int nchunks = share_nchunks(nvec, P - 1, rows, chunk_vecs);
    row_barrier(c, bank, blockIdx.x, epoch);
    const int j = is_root ? group : me;                       // the share this CTA moves
    int64_t s_lo, r_lo, r_hi;
    share_row_range(nvec, P - 1, rows, j, row, s_lo, r_lo, r_hi);
    char* mine = pv.ptr[c.rank];
    if (is_root) {
        const int relay = relay_rank(group, root, P);
        char* to = pv.ptr[relay];
        for (int k = 0; k < nchunks; ++k) {
            int64_t c_lo, c_hi;
            chunk_range(r_lo, r_hi, k, chunk_vecs, c_lo, c_hi);
            for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
                uint4 v[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < c_hi) v[u] = ld_stream_v4(mine + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < c_hi) st_v4(to + ((i0 + u * blockDim.x) << 4), v[u]);
            }
            __syncthreads();   // the whole chunk has been issued by this CTA; the release below orders it for the relay
            if (threadIdx.x == 0) st_release_sys(pad_slot(c, relay, kRelayBank, row, root), flag_base + k + 1);
        }
    } else if (group < P - 2) {
        char* to = pv.ptr[relay_rank(other_relay(me, group + 1, P), root, P)];
        for (int k = 0; k < nchunks; ++k) {
            int64_t c_lo, c_hi;
            chunk_range(r_lo, r_hi, k, chunk_vecs, c_lo, c_hi);
            if (threadIdx.x == 0) wait_flag(pad_slot(c, c.rank, kRelayBank, row, root), flag_base + k + 1, root, c.watchdog_ns);
            __syncthreads();
            for (int64_t i0 = c_lo + threadIdx.x; i0 < c_hi; i0 += (int64_t)blockDim.x * kU) {
                uint4 v[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < c_hi) v[u] = ld_peer_v4(mine + ((i0 + u * blockDim.x) << 4));
#pragma unroll
                for (int u = 0; u < kU; ++u)
                    if (i0 + u * blockDim.x < c_hi) st_v4(to + ((i0 + u * blockDim.x) << 4), v[u]);
            }
        }
    }
    row_barrier(c, bank, blockIdx.x, epoch + 1);
}

// ------------------------------------------------------------------------------------------ host side
// DLBB_MC_* / DLBB_BCAST_MC: 1 = use the multimem form whenever the buffers have a multicast mapping
// (the default), 0 = plain peer loads / stores even then (for the tuning study).
static bool env_flag(const char* name, bool dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) != 0 : dflt;
}

struct DirectPlan {
    bool mc_src;
    int sink;
    DirectArgs a;
};

template <typename T, int OP>
static int launch_direct_T(dlbb_comm* c, const DirectPlan& pl, cudaStream_t s, int64_t bytes) {
    const uint32_t ep = c->epoch[0] + 1;
    c->epoch[0] += 2;
    const int ctas = direct_ctas(c, bytes);
    const bool mc = pl.mc_src && MmReduce<T>::ok && OP == DLBB_SUM;
    // (instantiations are limited to what the entry points use: a multimem source ignores MAXW; peer sources never feed a
    //  multimem sink)
#define DLBB_DIRECT_P2P(SNK)                                                                                        \
    do {                                                                                                            \
        if (c->world > 4) direct_reduce_kernel<T, OP, false, SNK, 8><<<ctas, kDirThreads, 0, s>>>(c->dev, pl.a, 0, ep);      \
        else if (c->world > 2) direct_reduce_kernel<T, OP, false, SNK, 4><<<ctas, kDirThreads, 0, s>>>(c->dev, pl.a, 0, ep); \
        else direct_reduce_kernel<T, OP, false, SNK, 2><<<ctas, kDirThreads, 0, s>>>(c->dev, pl.a, 0, ep);                   \
    } while (0)
    if constexpr (MmReduce<T>::ok && OP == DLBB_SUM) {
        if (mc) {
            if (pl.sink == SINK_ONE) direct_reduce_kernel<T, OP, true, SINK_ONE, 8><<<ctas, kDirThreads, 0, s>>>(c->dev, pl.a, 0, ep);
            else if (pl.sink == SINK_MC) direct_reduce_kernel<T, OP, true, SINK_MC, 8><<<ctas, kDirThreads, 0, s>>>(c->dev, pl.a, 0, ep);
            else direct_reduce_kernel<T, OP, true, SINK_ALL, 8><<<ctas, kDirThreads, 0, s>>>(c->dev, pl.a, 0, ep);
            DLBB_CHECK_LAUNCH("direct_reduce_kernel (multimem)");
            return 0;
        }
    }
    if (pl.sink == SINK_MC) { set_error("direct reduce: a multimem sink needs a multimem source"); return -2; }
    if (pl.sink == SINK_ONE) DLBB_DIRECT_P2P(SINK_ONE);
    else DLBB_DIRECT_P2P(SINK_ALL);
#undef DLBB_DIRECT_P2P
    DLBB_CHECK_LAUNCH("direct_reduce_kernel");
    return 0;
}

template <typename T>
static int launch_direct_op(dlbb_comm* c, int op, const DirectPlan& pl, cudaStream_t s, int64_t bytes) {
    switch (op) {
        case DLBB_SUM: return launch_direct_T<T, DLBB_SUM>(c, pl, s, bytes);
        case DLBB_MAX: return launch_direct_T<T, DLBB_MAX>(c, pl, s, bytes);
        case DLBB_MIN: return launch_direct_T<T, DLBB_MIN>(c, pl, s, bytes);
        case DLBB_PROD: return launch_direct_T<T, DLBB_PROD>(c, pl, s, bytes);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}

static int launch_direct(dlbb_comm* c, int dtype, int op, const DirectPlan& pl, cudaStream_t s, int64_t bytes) {
    switch (dtype) {
        case DLBB_F16: return launch_direct_op<__half>(c, op, pl, s, bytes);
        case DLBB_BF16: return launch_direct_op<__nv_bfloat16>(c, op, pl, s, bytes);
        case DLBB_F32: return launch_direct_op<float>(c, op, pl, s, bytes);
        case DLBB_F64: return launch_direct_op<double>(c, op, pl, s, bytes);
        case DLBB_I32: return launch_direct_op<int32_t>(c, op, pl, s, bytes);
        case DLBB_I64: return launch_direct_op<int64_t>(c, op, pl, s, bytes);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

// Reduce to `root` (in place on the root) through reduce_pull_kernel, in slices that fit the reducers' staging region.
template <typename T, int OP>
static int launch_reduce_pull_T(dlbb_comm* c, const PeerView& pv, void* out, int64_t nvec, int root, bool mc, cudaStream_t s) {
    const int P = c->world;
    const int nred = mc ? P : P - 1;
    const int groups = mc ? P : P - 1;
    static const int chunk_vecs = getenv("DLBB_RELAY_CHUNK_KB") ? std::max(1, atoi(getenv("DLBB_RELAY_CHUNK_KB"))) * 64
                                                                : kRelayChunkVecsDefault;
    const int rows = std::max(1, (c->ctas > 0 ? std::min(c->ctas, kMaxCtas) : 296) / groups);
    const int64_t cap = static_cast<int64_t>(c->stage_bytes >> 4) * nred;   // vectors per launch
    for (int64_t done = 0; done < nvec; done += cap) {
        const int64_t n = std::min(cap, nvec - done);
        PeerView in = pv;
        for (int p = 0; p < P; ++p) in.ptr[p] = pv.ptr[p] + (done << 4);
        in.mc = pv.mc ? pv.mc + (done << 4) : nullptr;
        const uint32_t nchunks = static_cast<uint32_t>(share_nchunks(n, nred, rows, chunk_vecs));
        const uint32_t fb = c->epoch[kRelayBank];
        c->epoch[kRelayBank] += nchunks;
        const int64_t so = next_stage_off(c, s, false);
        const uint32_t ep = c->epoch[0] + 1;
        c->epoch[0] += 2;
        char* o = static_cast<char*>(out) + (done << 4);
        const int ctas = groups * rows;
        if constexpr (MmReduce<T>::ok && OP == DLBB_SUM) {
            if (mc) {
                reduce_pull_kernel<T, OP, true, 8><<<ctas, kDirThreads, 0, s>>>(c->dev, in, o, n, root, rows, so, fb, chunk_vecs, 0, ep);
                DLBB_CHECK_LAUNCH("reduce_pull_kernel (multimem)");
                continue;
            }
        }
        if (P > 4) reduce_pull_kernel<T, OP, false, 8><<<ctas, kDirThreads, 0, s>>>(c->dev, in, o, n, root, rows, so, fb, chunk_vecs, 0, ep);
        else reduce_pull_kernel<T, OP, false, 4><<<ctas, kDirThreads, 0, s>>>(c->dev, in, o, n, root, rows, so, fb, chunk_vecs, 0, ep);
        DLBB_CHECK_LAUNCH("reduce_pull_kernel");
    }
    return 0;
}
template <typename T>
static int launch_reduce_pull_op(dlbb_comm* c, int op, const PeerView& pv, void* out, int64_t nvec, int root, bool mc, cudaStream_t s) {
    switch (op) {
        case DLBB_SUM: return launch_reduce_pull_T<T, DLBB_SUM>(c, pv, out, nvec, root, mc, s);
        case DLBB_MAX: return launch_reduce_pull_T<T, DLBB_MAX>(c, pv, out, nvec, root, false, s);
        case DLBB_MIN: return launch_reduce_pull_T<T, DLBB_MIN>(c, pv, out, nvec, root, false, s);
        case DLBB_PROD: return launch_reduce_pull_T<T, DLBB_PROD>(c, pv, out, nvec, root, false, s);
    }
    set_error("unknown reduce op %d", op);
    return -2;
}
static int launch_reduce_pull(dlbb_comm* c, int dtype, int op, const PeerView& pv, void* out, int64_t nvec, int root, bool mc,
                              cudaStream_t s) {
    switch (dtype) {
        case DLBB_F16: return launch_reduce_pull_op<__half>(c, op, pv, out, nvec, root, mc, s);
        case DLBB_BF16: return launch_reduce_pull_op<__nv_bfloat16>(c, op, pv, out, nvec, root, mc, s);
        case DLBB_F32: return launch_reduce_pull_op<float>(c, op, pv, out, nvec, root, mc, s);
        case DLBB_F64: return launch_reduce_pull_op<double>(c, op, pv, out, nvec, root, false, s);
        case DLBB_I32: return launch_reduce_pull_op<int32_t>(c, op, pv, out, nvec, root, false, s);
        case DLBB_I64: return launch_reduce_pull_op<int64_t>(c, op, pv, out, nvec, root, false, s);
    }
    set_error("reduction on unsupported dtype %d", dtype);
    return -2;
}

static int launch_bcast_relay(dlbb_comm* c, const PeerView& pv, int64_t nvec, int root, cudaStream_t s) {
    const int P = c->world;
    // multimem relay for messages below 8 MiB (no flags, one hop less: 158 vs 101 GB/s at 4 MiB on 8 GPUs), the flag-
    // driven peer relay above (multicast delivery tops out near 400-420 GB/s per GPU, the peer pulls reach 540-600)
    static const int64_t mc_max = getenv("DLBB_BCAST_MC_MAX_BYTES") ? atoll(getenv("DLBB_BCAST_MC_MAX_BYTES")) : (8ll << 20);
    const bool mc = pv.mc != nullptr && env_flag("DLBB_BCAST_MC", true) && (nvec << 4) < mc_max;
    // P2P: every CTA walks its row chunk by chunk (wait -> load -> store -> signal, ~3 NVLink round trips per chunk), so
    // the pipeline is fed by the number of rows: two CTAs per SM unless DLBB_CTAS says otherwise.
    static const int kRelayChunkVecs = getenv("DLBB_RELAY_CHUNK_KB") ? std::max(1, atoi(getenv("DLBB_RELAY_CHUNK_KB"))) * 64
                                                                     : kRelayChunkVecsDefault;
    int rows;
    if (mc) rows = direct_ctas(c);
    else rows = std::max(1, (c->ctas > 0 ? std::min(c->ctas, kMaxCtas) : 294) / (P - 1));
    const int ctas = mc ? rows : rows * (P - 1);
    const uint32_t nchunks = static_cast<uint32_t>(share_nchunks(nvec, P - 1, rows, kRelayChunkVecs));
    const uint32_t fb = c->epoch[kRelayBank];
    if (!mc) c->epoch[kRelayBank] += nchunks;
    const uint32_t ep = c->epoch[0] + 1;
    c->epoch[0] += 2;
    if (!mc && env_flag("DLBB_BCAST_PUSH", false)) {   // same shares, rows, chunks and flag epochs; stores instead of loads
        bcast_push_kernel<<<ctas, kDirThreads, 0, s>>>(c->dev, pv, nvec, root, rows, fb, 0, ep, kRelayChunkVecs);
        DLBB_CHECK_LAUNCH("bcast_push_kernel");
        return 0;
    }
    if (mc) bcast_relay_kernel<true><<<ctas, kDirThreads, 0, s>>>(c->dev, pv, nvec, root, rows, fb, 0, ep, kRelayChunkVecs);
    else bcast_relay_kernel<false><<<ctas, kDirThreads, 0, s>>>(c->dev, pv, nvec, root, rows, fb, 0, ep, kRelayChunkVecs);
    DLBB_CHECK_LAUNCH("bcast_relay_kernel");
    return 0;
}

}  // namespace dlbb
