/* dlbb communication layer — C ABI.
 *
 * One process per GPU on one NVLink/NVSwitch node.  A communicator owns
 *   - a bootstrap channel (unix-domain sockets, star through rank 0) used for rendezvous,
 *     small all-gathers, host barriers and file-descriptor passing;
 *   - symmetric memory segments: CUDA VMM allocations exported as POSIX file descriptors,
 *     imported and mapped by every peer, optionally also bound to an NVLS multicast object;
 *   - signal pads for device-side barriers and a staging segment for collectives on
 *     ordinary (non-symmetric) tensors.
 * Collective kernels read and write the peers' memory directly from inside the kernel
 * (P2P loads/stores, multimem.ld_reduce / multimem.st when multicast is available).
 *
 * Replaces, on B200, what libmpi / Gloo / oneCCL do for the reference:
 * collectives/1d/dsccl.py:59-163, collectives/3d/dsccl.py:46-118, models.py:95.
 */
#ifndef DLBB_COMM_H_
#define DLBB_COMM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLBB_MAX_RANKS 16

typedef struct dlbb_comm dlbb_comm_t;

enum { DLBB_SUM = 0, DLBB_MAX = 1, DLBB_MIN = 2, DLBB_PROD = 3 };
/* all-reduce algorithms (the analogue of oneCCL's CCL_ALLREDUCE, collectives/3d/launch_dsccl.sh:46) */
enum { DLBB_ALGO_AUTO = 0, DLBB_ALGO_ONESHOT = 1, DLBB_ALGO_TWOSHOT = 2, DLBB_ALGO_NVLS = 3,
       DLBB_ALGO_DIRECT = 4 /* zero-copy on the ranks' own tensors at any size (in place only; else falls back) */ };

/* Create a communicator.  `job` names the rendezvous (all ranks pass the same string, e.g.
 * "<MASTER_PORT>-<sequence>"); `device` is the CUDA ordinal this rank drives.  Collective call:
 * every rank must call it.  staging_bytes: size of each of the three staging regions (0 = default
 * 256 MiB, or DLBB_STAGING_MB). */
int dlbb_comm_create(int rank, int world, int device, const char* job, size_t staging_bytes,
                     dlbb_comm_t** out);
int dlbb_comm_destroy(dlbb_comm_t* c);
int dlbb_comm_rank(const dlbb_comm_t* c);
int dlbb_comm_world(const dlbb_comm_t* c);
/* 1 when every segment of this communicator is also mapped through an NVLS multicast object. */
int dlbb_comm_has_multicast(const dlbb_comm_t* c);

/* Host-side helpers over the bootstrap channel. */
int dlbb_comm_host_barrier(dlbb_comm_t* c);
int dlbb_comm_host_allgather(dlbb_comm_t* c, const void* in, void* out, size_t bytes_per_rank);

/* Symmetric allocation (collective: same size on every rank, same call order).  Returns a segment
 * id; the local mapping, a peer's mapping and the multicast mapping (NULL when unavailable) are
 * queried per segment.  Memory is zero-initialised. */
int dlbb_symm_alloc(dlbb_comm_t* c, size_t bytes, int* seg_id);
void* dlbb_symm_ptr(dlbb_comm_t* c, int seg_id, int peer);
void* dlbb_symm_mc_ptr(dlbb_comm_t* c, int seg_id);
size_t dlbb_symm_size(dlbb_comm_t* c, int seg_id);
int dlbb_symm_free(dlbb_comm_t* c, int seg_id);

/* Knobs (the analogue of CCL_WORKER_COUNT etc., collectives/3d/launch_dsccl.sh:37-59).
 * ctas: number of CTAs ("channels") a collective kernel uses, 0 = auto. */
void dlbb_comm_set_ctas(dlbb_comm_t* c, int ctas);
void dlbb_comm_set_oneshot_max_bytes(dlbb_comm_t* c, size_t bytes);
/* Zero-copy collectives: messages of at least `bytes` bytes whose tensors are plain cudaMalloc memory (e.g. the
 * PyTorch caching allocator) are read from / written to the peers' tensors directly instead of going through the
 * symmetric staging area (CUDA IPC mappings, cached; all-reduce in place, all-gather, broadcast, gather, scatter,
 * all-to-all, reduce, reduce-scatter).  Such a call blocks the calling host thread until every rank has issued the
 * same call (a shared-memory mailbox exchange, ~1 us when the peers are there).  0 disables; default 16 MiB
 * (env DLBB_ZEROCOPY_MIN_BYTES).  No counterpart in the reference. */
void dlbb_comm_set_zero_copy_min_bytes(dlbb_comm_t* c, size_t bytes);
size_t dlbb_comm_zero_copy_min_bytes(const dlbb_comm_t* c);

/* Device barrier across all ranks, enqueued on `stream` (K1; ref comm.Barrier run_mpi.py:177). */
int dlbb_barrier(dlbb_comm_t* c, void* stream);

/* Strided view descriptor for in-kernel packing (K12): up to 4 dims, element strides.
 * ndim == 0 means contiguous. */
typedef struct {
    int ndim;
    int64_t size[4];
    int64_t stride[4];
} dlbb_view_t;

/* All collectives: `count` elements of `dtype` (DLBB_F16/BF16/F32/F64 from dlbb.h; DLBB_U8 for
 * data-movement ops), buffers are ordinary device pointers (staged through symmetric memory
 * inside the kernel) unless stated otherwise.  in/out may alias where the op allows in-place. */
int dlbb_allreduce(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op,
                   int algo, const dlbb_view_t* in_view, void* stream);
/* Bucket fusion (K13; the analogue of CCL_FUSION, collectives/3d/launch_dsccl.sh:49-59): all-reduce `n_tensors` tensors
 * of one dtype in place with as few launches as possible - up to 96 tensors / DLBB_FUSION_BYTES (default 16 MiB) per
 * launch are packed, reduced and unpacked by ONE kernel; larger tensors go through dlbb_allreduce one by one.
 * DLBB_FUSION=0 switches fusion off (one collective per tensor). */
int dlbb_allreduce_multi(dlbb_comm_t* c, void* const* bufs, const int64_t* counts, int n_tensors, int dtype,
                         int op, void* stream);
/* out = sum over ranks(in) + residual (bf16, SUM): the TP row-parallel reduction fused with the
 * residual add (K16 without the GEMM; ref models.py:95 + :173/:188). */
int dlbb_allreduce_residual_bf16(dlbb_comm_t* c, const void* in, const void* residual, void* out,
                                 int64_t count, void* stream);
/* Strided tensors (K12): *_v variants take view descriptors for the source and / or the destination; the kernels
 * gather the source view while packing and scatter into the destination view while unpacking - there is no
 * .contiguous() copy before or after (no counterpart in the reference, whose tensors are dense: SURVEY.md N4).
 * For broadcast the same view describes the tensor on every rank. */
int dlbb_allreduce_v(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int algo,
                     const dlbb_view_t* in_view, const dlbb_view_t* out_view, void* stream);
int dlbb_broadcast_v(dlbb_comm_t* c, void* buf, int64_t count, int dtype, int root, const dlbb_view_t* buf_view,
                     void* stream);
int dlbb_reduce_v(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op, int root,
                  const dlbb_view_t* in_view, const dlbb_view_t* out_view, void* stream);
int dlbb_gather_v(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int root,
                  const dlbb_view_t* in_view, void* stream);
int dlbb_reduce_scatter(dlbb_comm_t* c, const void* in, void* out, int64_t count_per_rank,
                        int dtype, int op, void* stream);
int dlbb_allgather(dlbb_comm_t* c, const void* in, void* out, int64_t count_per_rank, int dtype,
                   const dlbb_view_t* in_view, void* stream);
int dlbb_broadcast(dlbb_comm_t* c, void* buf, int64_t count, int dtype, int root, void* stream);
int dlbb_reduce(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int op,
                int root, void* stream);
/* gather: root's out receives world*count; scatter: root's in holds world*count. */
int dlbb_gather(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int root,
                void* stream);
int dlbb_scatter(dlbb_comm_t* c, const void* in, void* out, int64_t count, int dtype, int root,
                 void* stream);
/* all-to-all: chunk p of `in` goes to rank p; chunk r of `out` comes from rank r. */
int dlbb_alltoall(dlbb_comm_t* c, const void* in, void* out, int64_t count_per_rank, int dtype,
                  void* stream);
/* All-gather (root < 0) or gather (to `root`) into a list of output tensors: rank p's `count` elements land in
 * out_ptrs[p]; the input may be a strided view.  out_ptrs is only read on ranks that receive. */
int dlbb_allgather_ptrs(dlbb_comm_t* c, const void* in, void* const* out_ptrs, int64_t count, int dtype, int root,
                        const dlbb_view_t* in_view, void* stream);
/* The same on lists of tensors: in_ptrs[p] goes to rank p, out_ptrs[r] comes from rank r (bytes_per_peer bytes each);
 * packed and unpacked inside the kernel, no flattening copies. */
int dlbb_alltoall_ptrs(dlbb_comm_t* c, const void* const* in_ptrs, void* const* out_ptrs, int64_t bytes_per_peer,
                       void* stream);
/* Point-to-point: send `send_count` elements to send_peer and / or receive `recv_count` from
 * recv_peer in one kernel (ring sendrecv, ref collectives/1d/openmpi.py:173-198; plain send / recv
 * with the other peer = -1, ref test/test_deepseed.py:132-153).  Only the ranks involved call it.
 * Messages between a pair are matched in issue order (no tags). */
int dlbb_sendrecv(dlbb_comm_t* c, const void* send_buf, int64_t send_count, int send_peer,
                  void* recv_buf, int64_t recv_count, int recv_peer, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused gradient reduce + Adam + parameter broadcast (K20).
 * grads / params are *symmetric* buffers (segment ids + byte offsets identical on all ranks).
 * For the bucket [elem_off, elem_off + n) rank r owns the r-th of `world` equal shards (the
 * bucket is logically padded to a multiple of world*8 elements): it sums that shard of the bf16
 * gradients of all ranks (peer loads, or multimem.ld_reduce), multiplies by grad_scale, applies
 * Adam with its local fp32 m / v / master shard and stores the new bf16 parameters into every
 * rank's parameter buffer (peer stores, or multimem.st).
 * ref: DeepSpeed ZeRO engine.backward()/step(), test/ccl.py:86-88,114-115.
 */
int dlbb_fused_reduce_adam_bf16(dlbb_comm_t* c, int grads_seg, int params_seg, int64_t elem_off,
                                int64_t n, float* m, float* v, float* master, float lr, float beta1,
                                float beta2, float eps, float weight_decay, float grad_scale,
                                int step, int use_multicast, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused row-parallel linear (K16): out = all_reduce_sum(A . B) + residual in ONE kernel - the tcgen05
 * GEMM whose epilogue publishes partial tiles in symmetric memory while reducer warps of the same
 * kernel sum finished tiles across ranks (multimem.ld_reduce or peer loads) and write the result to
 * every rank (multimem.st or peer stores).  ref models.py:73-100 + :173/:188.
 * A/B as in dlbb_gemm_bf16 (this rank's K-shard).  The partial buffer (M*N bf16) and the output
 * buffer (M*N bf16, row-major, ld = N) live in symmetric segments at the given element offsets;
 * residual (may be NULL) is a local [M, ld_res] bf16 matrix, identical on every rank.
 */
int dlbb_gemm_allreduce_bf16(dlbb_comm_t* c, const void* A, const void* B, int M, int N, int K,
                             int64_t lda, int64_t ldb, int a_mn_major, int b_mn_major, int part_seg,
                             int64_t part_elem_off, int out_seg, int64_t out_elem_off,
                             const void* residual, int64_t ld_res, int use_multicast, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLBB_COMM_H_ */
