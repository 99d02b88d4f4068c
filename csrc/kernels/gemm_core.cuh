// Device body of the persistent warp-specialised tcgen05 GEMM plus the host-side tensor-map helpers,
// shared by the plain GEMM (gemm_tcgen05.cu) and the GEMM fused with its tensor-parallel all-reduce
// (../collectives/gemm_allreduce.cu).  See gemm_tcgen05.cu for the design notes.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdlib.h>

#include <algorithm>
#include <mutex>
#include <unordered_map>

#include "../../include/dlbb.h"
#include "common.cuh"
#include "gemm_tiles.h"
#include "ptx.cuh"

namespace dlbb {

constexpr int kBK = 64;          // 64 bf16 = 128 bytes = one swizzle row
constexpr int kEpiCols = 32;     // epilogue chunk width (64 B rows, 64B swizzle)
constexpr int kEpiBufBytes = kBM * kEpiCols * 2;  // 8 KiB
constexpr int kEpiThreads = 128;

// CG = CTAs per MMA (1, or 2 = a CTA pair sharing one 256 x BN tile; each CTA then stages only half of B).
template <int BN, int CG = 1>
struct GemmCfg {
    static constexpr int kBRows = BN / CG;   // B rows (N extent) staged by one CTA
    static constexpr int kStages = (kBRows == 256) ? 4 : 6;
    static constexpr int kABytes = kBM * kBK * 2;
    static constexpr int kBBytes = kBRows * kBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kTileBytes = kStages * kStageBytes;
    static constexpr int kEpiBytes = 4 * kEpiBufBytes;  // out[2] + aux[2]
    static constexpr int kBarBytes = 256;
    static constexpr int kSmemBytes = kTileBytes + kEpiBytes + kBarBytes + 1024;  // + align slack
    static constexpr int kTmemCols = 2 * BN;
};

struct GemmParams {
    int M, N, K;
    int epi;
    int aux_out;   // EPI_GELU: also store the pre-activation
    int group_m;   // tile rasterisation: m-blocks per group
    uint64_t a_policy, b_policy;  // L2 eviction priority of the operand panels
};

// Hook points for kernels that fuse more work into the GEMM (see csrc/collectives/gemm_allreduce.cu):
//   tile_stored(t, m0, n0)  epilogue leader thread, after the last TMA store of tile t was committed
//   extra_role(warp, lane)  warps >= 8 (present only when kExtraWarps > 0), concurrent with the GEMM roles
//   finish()                every thread, after all roles are done and before TMEM is released
struct NoFusion {
    static constexpr int kExtraWarps = 0;
    __device__ __forceinline__ void tile_stored(int, int, int) {}
    __device__ __forceinline__ void extra_role(int, int) {}
    __device__ __forceinline__ void finish() {}
};

template <int BN, bool A_MN, bool B_MN, class Fusion, int CG = 1>
__device__ __forceinline__ void gemm_cta_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                              const CUtensorMap& tmAux, const GemmParams& p, Fusion& fusion) {
    using Cfg = GemmCfg<BN, CG>;
    static_assert(CG == 1 || (CG == 2 && Fusion::kExtraWarps == 0), "pairs: plain GEMM only");
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled operand tiles need 1024 B alignment.
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* epi_smem = smem + Cfg::kTileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + Cfg::kEpiBytes);
    // barrier layout: full[S], empty[S], tmem_full[2], tmem_empty[2], aux_full[2], then tmem ptr
    const uint32_t bar_full = smem_u32(bars);
    const uint32_t bar_empty = bar_full + 8 * Cfg::kStages;
    const uint32_t bar_tfull = bar_empty + 8 * Cfg::kStages;
    const uint32_t bar_tempty = bar_tfull + 16;
    const uint32_t bar_aux = bar_tempty + 16;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 6);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // CTA pair: rank within the pair, pair index and pair count replace block index and grid size.
    const int cta_rank = (CG == 2) ? static_cast<int>(cluster_ctarank()) : 0;
    const int worker = blockIdx.x / CG, num_workers = gridDim.x / CG;
    constexpr int kTileM = kBM * CG;
    const int num_m = (p.M + kTileM - 1) / kTileM;
    const int num_n = (p.N + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int num_k = (p.K + kBK - 1) / kBK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmC);
        if (p.epi != DLBB_EPI_NONE) tma_prefetch_desc(&tmAux);
    } else if (warp == 1 && lane == 0) {
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_tfull + 8 * s, 1);
            mbar_init(bar_tempty + 8 * s, kEpiThreads * CG);
            mbar_init(bar_aux + 8 * s, 1);
        }
        fence_mbar_init();
    } else if (warp == 2) {
        if constexpr (CG == 2) {
            tmem_alloc_2sm(smem_u32(tmem_ptr_smem), Cfg::kTmemCols);
            tmem_relinquish_2sm();
        } else {
            tmem_alloc(smem_u32(tmem_ptr_smem), Cfg::kTmemCols);
            tmem_relinquish();
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if constexpr (CG == 2) cluster_sync_all();  // the peer's barriers exist before anything signals them
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr_smem;

    auto tma_ld = [](uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, uint64_t pol) {
        if constexpr (CG == 2) tma_load_2d_hint_2sm(dst, m, bar, c0, c1, pol);
        else tma_load_2d_hint(dst, m, bar, c0, c1, pol);
    };
    if (warp == 0) {
        // ============================== TMA producer ==============================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = worker; t < num_tiles; t += num_workers) {
                int m_blk, n_blk;
                tile_coords(t, num_m, num_n, p.group_m, m_blk, n_blk);
                const int m0 = m_blk * kTileM + cta_rank * kBM, n0 = n_blk * BN + cta_rank * Cfg::kBRows;
                for (int kb = 0; kb < num_k; ++kb) {
                    const uint32_t full = bar_full + 8 * stage;
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    // pair: both CTAs' loads count on the leader's barrier, which expects both halves
                    if (CG == 1 || cta_rank == 0) mbar_arrive_expect_tx(full, Cfg::kStageBytes * CG);
                    const uint32_t a_dst = smem_u32(smem + stage * Cfg::kStageBytes);
                    const uint32_t b_dst = a_dst + Cfg::kABytes;
                    const int k0 = kb * kBK;
                    if constexpr (!A_MN) {
                        tma_ld(a_dst, &tmA, full, k0, m0, p.a_policy);  // box {64 k, 128 m}
                    } else {
#pragma unroll
                        for (int i = 0; i < kBM / 64; ++i)      // box {64 m, 64 k} per atom
                            tma_ld(a_dst + i * (kBK * 128), &tmA, full, m0 + 64 * i, k0, p.a_policy);
                    }
                    if constexpr (!B_MN) {
                        tma_ld(b_dst, &tmB, full, k0, n0, p.b_policy);  // box {64 k, BN / CG n}
                    } else {
#pragma unroll
                        for (int i = 0; i < Cfg::kBRows / 64; ++i)
                            tma_ld(b_dst + i * (kBK * 128), &tmB, full, n0 + 64 * i, k0, p.b_policy);
                    }
                    if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ============================== MMA issuer ==============================
        if (lane == 0 && cta_rank == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(kTileM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            // K-major: 8-row groups are 1024 B apart (SBO), LBO unused inside one 128 B atom.
            // MN-major: 8 k-rows are 1024 B apart (SBO), 64-element MN atoms kBK*128 B apart (LBO).
            constexpr uint32_t a_lbo = A_MN ? kBK * 128 : 0, b_lbo = B_MN ? kBK * 128 : 0;
            constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32, b_kstep = B_MN ? 16 * 128 : 32;
            int stage = 0;
            uint32_t phase = 0;
            int as = 0;
            uint32_t aphase = 0;
            for (int t = worker; t < num_tiles; t += num_workers) {
                mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after_sync();
                    const uint32_t a_addr = smem_u32(smem + stage * Cfg::kStageBytes);
                    const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k) {
                        const uint64_t da = make_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, 1024);
                        const uint64_t db = make_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, 1024);
                        if constexpr (CG == 2) umma_f16_ss_2sm(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                        else umma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    // smem slot free once these MMAs retire (pair: in both CTAs)
                    if constexpr (CG == 2) umma_commit_2sm(bar_empty + 8 * stage, 3);
                    else umma_commit(bar_empty + 8 * stage);
                    if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
                }
                // accumulator complete
                if constexpr (CG == 2) umma_commit_2sm(bar_tfull + 8 * as, 3);
                else umma_commit(bar_tfull + 8 * as);
                if (++as == 2) { as = 0; aphase ^= 1; }
            }
        }
    } else if (warp >= 8) {
        fusion.extra_role(warp, lane);
    } else if (warp >= 4) {
        // ============================== epilogue ==============================
        const int tid_e = threadIdx.x - 128;       // 0..127 == tile row == TMEM lane
        const int warp_e = warp - 4;                // TMEM lane quarter (== warp % 4)
        const bool leader = tid_e == 0;
        const bool aux_in = (p.epi == DLBB_EPI_ADD) || (p.epi == DLBB_EPI_MUL_DGELU);
        const bool aux_out = (p.epi == DLBB_EPI_GELU) && p.aux_out;
        // 64B swizzle: 16 B chunk index ^= (row >> 1) & 3
        const uint32_t row_off = tid_e * 64;
        const uint32_t sw = (tid_e >> 1) & 3;
        int as = 0;
        uint32_t aphase = 0;
        uint32_t chunk_it = 0;
        // The MMA issuer (pair: the leader's) learns that an accumulator stage was drained.
        auto release_acc = [&](int stage_idx) {
            tc_fence_before_sync();
            if constexpr (CG == 2) mbar_arrive_cluster(mapa_shared(bar_tempty + 8 * stage_idx, 0));
            else mbar_arrive(bar_tempty + 8 * stage_idx);
        };
        for (int t = worker; t < num_tiles; t += num_workers) {
            int m_blk, n_blk;
            tile_coords(t, num_m, num_n, p.group_m, m_blk, n_blk);
            const int m0 = m_blk * kTileM + cta_rank * kBM, n0 = n_blk * BN;
            mbar_wait(bar_tfull + 8 * as, aphase);
            tc_fence_after_sync();
#pragma unroll 1
            for (int c = 0; c < BN / kEpiCols; ++c) {
                const int nc = n0 + c * kEpiCols;
                const bool last = (c == BN / kEpiCols - 1);
                if (nc >= p.N) {
                    // chunk entirely right of the matrix edge: nothing to store
                    if (last) release_acc(as);
                    continue;
                }
                const int buf = chunk_it & 1;
                const uint32_t aux_phase = (chunk_it >> 1) & 1;
                ++chunk_it;
                uint8_t* out_buf = epi_smem + buf * kEpiBufBytes;
                uint8_t* aux_buf = epi_smem + (2 + buf) * kEpiBufBytes;
                if (leader) {
                    tma_wait_group_read<1>();  // the store that last used this buffer has drained
                    if (aux_in) {
                        mbar_arrive_expect_tx(bar_aux + 8 * buf, kEpiBufBytes);
                        tma_load_2d(smem_u32(aux_buf), &tmAux, bar_aux + 8 * buf, nc, m0);
                    }
                }
                named_bar_sync(1, kEpiThreads);

                uint32_t acc[32];
                tmem_ld_32x32b_x32(tmem_base + as * BN + c * kEpiCols + ((warp_e * 32u) << 16), acc);
                tmem_ld_wait();
                if (last) release_acc(as);  // accumulator fully drained into registers
                uint32_t auxw[16];
                if (aux_in) {
                    mbar_wait(bar_aux + 8 * buf, aux_phase);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint4 v =
                            *reinterpret_cast<const uint4*>(aux_buf + row_off + ((j ^ sw) << 4));
                        auxw[4 * j + 0] = v.x; auxw[4 * j + 1] = v.y;
                        auxw[4 * j + 2] = v.z; auxw[4 * j + 3] = v.w;
                    }
                }
                uint32_t outw[16];
                if (p.epi == DLBB_EPI_NONE) {
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        outw[j] = pack_bf16x2(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1]));
                } else if (p.epi == DLBB_EPI_GELU) {
                    if (aux_out) {
                        // pre-activation goes out through the aux buffer (same swizzle)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            uint4 v;
                            v.x = pack_bf16x2(__uint_as_float(acc[8 * j + 0]), __uint_as_float(acc[8 * j + 1]));
                            v.y = pack_bf16x2(__uint_as_float(acc[8 * j + 2]), __uint_as_float(acc[8 * j + 3]));
                            v.z = pack_bf16x2(__uint_as_float(acc[8 * j + 4]), __uint_as_float(acc[8 * j + 5]));
                            v.w = pack_bf16x2(__uint_as_float(acc[8 * j + 6]), __uint_as_float(acc[8 * j + 7]));
                            *reinterpret_cast<uint4*>(aux_buf + row_off + ((j ^ sw) << 4)) = v;
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        outw[j] = pack_bf16x2(gelu_exact(__uint_as_float(acc[2 * j])),
                                              gelu_exact(__uint_as_float(acc[2 * j + 1])));
                } else if (p.epi == DLBB_EPI_ADD) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float2 a = unpack_bf16x2(auxw[j]);
                        outw[j] = pack_bf16x2(__uint_as_float(acc[2 * j]) + a.x,
                                              __uint_as_float(acc[2 * j + 1]) + a.y);
                    }
                } else {  // DLBB_EPI_MUL_DGELU
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float2 a = unpack_bf16x2(auxw[j]);
                        outw[j] = pack_bf16x2(__uint_as_float(acc[2 * j]) * dgelu_exact(a.x),
                                              __uint_as_float(acc[2 * j + 1]) * dgelu_exact(a.y));
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<uint4*>(out_buf + row_off + ((j ^ sw) << 4)) =
                        make_uint4(outw[4 * j], outw[4 * j + 1], outw[4 * j + 2], outw[4 * j + 3]);
                fence_proxy_async_smem();  // generic-proxy smem writes -> visible to TMA
                named_bar_sync(1, kEpiThreads);
                if (leader) {
                    tma_store_2d(&tmC, smem_u32(out_buf), nc, m0);
                    if (aux_out) tma_store_2d(&tmAux, smem_u32(aux_buf), nc, m0);
                    tma_commit_group();
                }
            }
            if (leader) fusion.tile_stored(t, m0, n0);
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
        if (leader) tma_wait_group<0>();
    }

    tc_fence_before_sync();
    __syncthreads();
    if constexpr (CG == 2) cluster_sync_all();  // no CTA leaves while its peer may still signal it
    tc_fence_after_sync();
    fusion.finish();
    if (warp == 2) {
        if constexpr (CG == 2) tmem_dealloc_2sm(tmem_base, Cfg::kTmemCols);
        else tmem_dealloc(tmem_base, Cfg::kTmemCols);
    }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kNumThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA,
                         const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmC,
                         const __grid_constant__ CUtensorMap tmAux, const GemmParams p) {
    NoFusion nf;
    gemm_cta_body<BN, A_MN, B_MN, NoFusion>(tmA, tmB, tmC, tmAux, p, nf);
}

// CTA-pair variant: cluster (2,1,1), one 256 x 256 tile per pair, tcgen05.mma.cta_group::2.
template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kNumThreads, 1)
gemm_bf16_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tmA,
                              const __grid_constant__ CUtensorMap tmB,
                              const __grid_constant__ CUtensorMap tmC,
                              const __grid_constant__ CUtensorMap tmAux, const GemmParams p) {
    NoFusion nf;
    gemm_cta_body<256, A_MN, B_MN, NoFusion, 2>(tmA, tmB, tmC, tmAux, p, nf);
}

// The same kernel capped at 168 registers per thread, for steps that run a second kernel next to the GEMM on the same
// SMs (the data-parallel bucket kernel): the launch of the uncapped cluster kernel allocates 230 registers per thread
// (ncu: launch__registers_per_thread) and leaves no room for a co-resident CTA.
template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(168)
gemm_bf16_tcgen05_pair_lowreg_kernel(const __grid_constant__ CUtensorMap tmA,
                              const __grid_constant__ CUtensorMap tmB,
                              const __grid_constant__ CUtensorMap tmC,
                              const __grid_constant__ CUtensorMap tmAux, const GemmParams p) {
    NoFusion nf;
    gemm_cta_body<256, A_MN, B_MN, NoFusion, 2>(tmA, tmB, tmC, tmAux, p, nf);
}

// ------------------------------------------------------------------------------------------
// Host side: cached TMA tensor maps
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
                cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    });
    return fn;
}

struct MapKey {
    const void* ptr;
    uint64_t d0, d1, ld;
    uint32_t b0, b1, swz;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && ld == o.ld && b0 == o.b0 && b1 == o.b1 &&
               swz == o.swz;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        auto mix = [&h](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
        mix(k.d0); mix(k.d1); mix(k.ld); mix(k.b0); mix(k.b1); mix(k.swz);
        return h;
    }
};

// bf16 2-D tensor map: dims {d0 (contiguous), d1}, row pitch ld elements, box {b0, b1}.
// Encoding costs a few microseconds on the host, so maps are cached per (pointer, shape, box).
inline int make_map_2d(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t ld,
                       uint32_t b0, uint32_t b1, CUtensorMapSwizzle swz) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    MapKey key{ptr, d0, d1, ld, b0, b1, static_cast<uint32_t>(swz)};
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return 0; }
    }
    EncodeTiledFn enc = get_encode_fn();
    DLBB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[2] = {d0, d1};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {b0, b1};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DLBB_REQUIRE(r == CUDA_SUCCESS,
                 "cuTensorMapEncodeTiled failed (%d) ptr=%p dims=%llu,%llu ld=%llu box=%u,%u", (int)r,
                 ptr, (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)ld, b0, b1);
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 8192) cache.clear();
    cache.emplace(key, *out);
    return 0;
}


}  // namespace dlbb
