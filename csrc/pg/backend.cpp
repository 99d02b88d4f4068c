// torch.distributed process-group backend "dlbb" (also reachable as "ccl" through the
// oneccl_bindings_for_pytorch shim): a c10d::Backend whose collectives are the hand-written
// sm_100a kernels of libdlbb.so running over NVLink peer / multicast memory.
//
// Takes the place of oneCCL's c10d backend and of Gloo in the reference's DeepSpeed / torch paths
// (collectives/1d/dsccl.py:47-163, collectives/3d/dsccl.py:34-118, test/test_deepseed.py:24-181,
// test/ccl.py:24-57) and sits under the mpi4py shim (models.py:95, collectives/1d/openmpi.py).
//
//  * CUDA tensors: the kernel is enqueued on the caller's current stream; the returned Work
//    carries a CUDA event (wait() = stream-side wait, like ProcessGroupNCCL).
//  * CPU tensors (what the unchanged reference scripts create, e.g. collectives/1d/dsgloo.py:267):
//    copied to the rank's GPU, reduced / moved there, copied back, complete on return - so the
//    reference's host-clock timing brackets the whole operation.
//  * point-to-point ops run on two private streams (one for sends, one for receives) so that a
//    ring of isend / irecv cannot deadlock on issue order.
#include <ATen/cuda/CUDAContext.h>
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/csrc/distributed/c10d/Backend.hpp>
#include <torch/csrc/distributed/c10d/Store.hpp>
#include <torch/csrc/distributed/c10d/Types.hpp>
#include <torch/csrc/distributed/c10d/Work.hpp>
#include <torch/extension.h>
#include <pybind11/chrono.h>
#include <unistd.h>

#include <chrono>
#include <functional>
#include <string>
#include <vector>

#include "../../include/dlbb.h"
#include "../../include/dlbb_comm.h"

namespace dlbb_pg {

using c10d::Work;

#define DLBB_CALL(expr)                                                                       \
    do {                                                                                      \
        int _rc = (expr);                                                                     \
        TORCH_CHECK(_rc == 0, "dlbb backend: ", #expr, " failed (", _rc, "): ", dlbb_last_error()); \
    } while (0)

class WorkDlbb : public Work {
  public:
    // host_blocking: wait() blocks the calling thread (CPU-tensor semantics) instead of only ordering
    // the current stream; on_wait runs once after completion (device -> host copy of a received tensor).
    WorkDlbb(c10d::OpType op, std::vector<at::Tensor> outputs, c10::optional<at::cuda::CUDAEvent> ev,
             c10::DeviceIndex dev, bool host_blocking = false, std::function<void()> on_wait = nullptr,
             std::vector<at::Tensor> keep_alive = {})
        : Work(-1, op), outputs_(std::move(outputs)), event_(std::move(ev)), dev_(dev), host_blocking_(host_blocking),
          on_wait_(std::move(on_wait)), keep_alive_(std::move(keep_alive)) {
        future_ = c10::make_intrusive<at::ivalue::Future>(c10::ListType::create(c10::TensorType::get()));
        future_->markCompleted(at::IValue(outputs_));
    }
    bool isCompleted() override { return !event_.has_value() || event_->query(); }
    bool isSuccess() const override { return true; }
    bool wait(std::chrono::milliseconds /*timeout*/ = kNoTimeout) override {
        if (event_.has_value()) {
            if (host_blocking_) event_->synchronize();
            else event_->block(c10::cuda::getCurrentCUDAStream(dev_));
        }
        if (on_wait_) {
            auto fn = std::move(on_wait_);
            on_wait_ = nullptr;
            fn();
        }
        keep_alive_.clear();
        return true;
    }
    void synchronize() override { wait(); }
    std::vector<at::Tensor> result() override { return outputs_; }
    c10::intrusive_ptr<c10::ivalue::Future> getFuture() override { return future_; }

  private:
    std::vector<at::Tensor> outputs_;
    c10::optional<at::cuda::CUDAEvent> event_;
    c10::DeviceIndex dev_;
    bool host_blocking_;
    std::function<void()> on_wait_;
    std::vector<at::Tensor> keep_alive_;
    c10::intrusive_ptr<at::ivalue::Future> future_;
};

static int to_dlbb_dtype(at::ScalarType t, bool reduce) {
    switch (t) {
        case at::kHalf: return DLBB_F16;
        case at::kBFloat16: return DLBB_BF16;
        case at::kFloat: return DLBB_F32;
        case at::kDouble: return DLBB_F64;
        case at::kInt: return DLBB_I32;
        case at::kLong: return DLBB_I64;
        default: break;
    }
    TORCH_CHECK(!reduce, "dlbb backend: reductions support fp16/bf16/fp32/fp64/int32/int64, got ", t);
    return DLBB_U8;
}

static int to_dlbb_op(const c10d::ReduceOp& op) {
    switch (op) {
        case c10d::ReduceOp::SUM: case c10d::ReduceOp::AVG: return DLBB_SUM;
        case c10d::ReduceOp::MAX: return DLBB_MAX;
        case c10d::ReduceOp::MIN: return DLBB_MIN;
        case c10d::ReduceOp::PRODUCT: return DLBB_PROD;
        default: TORCH_CHECK(false, "dlbb backend: unsupported reduce op");
    }
    return DLBB_SUM;
}

// Data-movement ops only need the element size: pick any dtype of that size.
static int move_dtype(int64_t elem_size) {
    switch (elem_size) {
        case 1: return DLBB_U8;
        case 2: return DLBB_F16;
        case 4: return DLBB_F32;
        case 8: return DLBB_F64;
        default: return -1;
    }
}

// Descriptor of a non-contiguous tensor for in-kernel packing / unpacking (K12).  False when the tensor has more than
// four dimensions, zero or negative strides or an element size the kernels do not move as a unit: the caller then makes
// a dense copy.
static bool strided_view(const at::Tensor& t, dlbb_view_t* v) {
    if (t.dim() == 0 || t.dim() > 4 || move_dtype(t.element_size()) < 0) return false;
    v->ndim = static_cast<int>(t.dim());
    for (int i = 0; i < 4; ++i) { v->size[i] = 0; v->stride[i] = 0; }
    for (int i = 0; i < v->ndim; ++i) {
        if (t.stride(i) <= 0 && t.size(i) > 1) return false;
        v->size[i] = t.size(i);
        v->stride[i] = t.stride(i);
    }
    return true;
}

class BackendDlbb : public c10d::Backend {
  public:
    BackendDlbb(const c10::intrusive_ptr<c10d::Store>& store, int rank, int size) : Backend(rank, size) {
        dev_ = at::cuda::current_device();
        // all ranks must name the same rendezvous: rank 0 publishes a unique id through the store
        std::string job;
        if (rank == 0) {
            job = std::to_string(getpid()) + "-" + std::to_string(std::chrono::steady_clock::now().time_since_epoch().count());
            store->set("dlbb_job", std::vector<uint8_t>(job.begin(), job.end()));
        } else {
            auto v = store->get("dlbb_job");
            job.assign(v.begin(), v.end());
        }
        DLBB_CALL(dlbb_comm_create(rank, size, dev_, job.c_str(), 0, &comm_));
    }
    ~BackendDlbb() override { shutdown(); }
    void shutdown() override {
        if (comm_) {
            cudaDeviceSynchronize();
            dlbb_comm_destroy(comm_);
            comm_ = nullptr;
        }
    }
    const std::string getBackendName() const override { return "dlbb"; }
    bool has_multicast() const { return dlbb_comm_has_multicast(comm_) != 0; }
    void set_ctas(int n) { dlbb_comm_set_ctas(comm_, n); }
    void set_oneshot_max_bytes(int64_t n) { dlbb_comm_set_oneshot_max_bytes(comm_, static_cast<size_t>(n)); }
    void set_algo(int a) { algo_ = a; }

    // A tensor in symmetric memory (collective call: same sizes, same order on every rank).  Collectives on such
    // tensors run in place on the peers' mappings at every size - multimem.ld_reduce / multimem.st where the box has
    // NVLS - with no staging copy.  The memory belongs to the communicator and lives until it is shut down.
    at::Tensor symm_empty(std::vector<int64_t> sizes, at::ScalarType dtype) {
        int64_t numel = 1;
        for (auto v : sizes) numel *= v;
        const size_t bytes = static_cast<size_t>(std::max<int64_t>(numel, 1)) * c10::elementSize(dtype);
        int seg = -1;
        DLBB_CALL(dlbb_symm_alloc(comm_, bytes, &seg));
        void* ptr = dlbb_symm_ptr(comm_, seg, getRank());
        TORCH_CHECK(ptr != nullptr, "dlbb backend: symmetric allocation failed");
        auto opts = at::TensorOptions().dtype(dtype).device(c10::Device(c10::kCUDA, dev_));
        return at::from_blob(ptr, sizes, [](void*) {}, opts);
    }

    // ------------------------------------------------------------------ collectives
    c10::intrusive_ptr<Work> allreduce(std::vector<at::Tensor>& tensors, const c10d::AllreduceOptions& opts) override {
        TORCH_CHECK(tensors.size() == 1, "dlbb backend: one tensor per rank");
        at::Tensor& t = tensors[0];
        check_avg(opts.reduceOp, tensors);
        Staged s = stage_in(t, /*need_copy=*/true);
        const int dt = to_dlbb_dtype(s.dev.scalar_type(), true);
        if (s.dev.is_contiguous()) {
            DLBB_CALL(dlbb_allreduce(comm_, s.dev.data_ptr(), s.dev.data_ptr(), s.dev.numel(), dt,
                                     to_dlbb_op(opts.reduceOp), algo_, nullptr, stream()));
        } else {
            // strided view: gathered by the kernel itself (no .contiguous() pre-pass), result scattered back
            TORCH_CHECK(s.dev.dim() <= 4, "dlbb backend: strided all_reduce supports up to 4 dims");
            dlbb_view_t v{};
            v.ndim = static_cast<int>(s.dev.dim());
            for (int i = 0; i < v.ndim; ++i) { v.size[i] = s.dev.size(i); v.stride[i] = s.dev.stride(i); }
            // in place on the view: packed by the kernel's first phase, scattered back by its last one
            DLBB_CALL(dlbb_allreduce_v(comm_, s.dev.data_ptr(), s.dev.data_ptr(), s.dev.numel(), dt,
                                       to_dlbb_op(opts.reduceOp), algo_, &v, &v, stream()));
        }
        if (opts.reduceOp == c10d::ReduceOp::AVG) s.dev.div_(getSize());
        return finish(c10d::OpType::ALLREDUCE, s, {t});
    }

    // Bucket fusion (the analogue of CCL_FUSION, collectives/3d/launch_dsccl.sh:49-59): the tensors of one dtype travel
    // as a pointer table to ONE multi-tensor kernel (pack, barrier, reduce, unpack - dlbb_allreduce_multi); nothing is
    // concatenated or copied around it.  DLBB_FUSION=0 issues one collective per tensor, DLBB_FUSION_BYTES bounds a
    // fused launch.  CPU tensors and strided views are brought to a dense device tensor first and copied back after.
    c10::intrusive_ptr<Work> allreduce_coalesced(std::vector<at::Tensor>& tensors,
                                                 const c10d::AllreduceCoalescedOptions& opts) override {
        TORCH_CHECK(!tensors.empty(), "dlbb backend: empty tensor list");
        c10::cuda::CUDAGuard g(dev_);
        check_avg(opts.reduceOp, tensors);
        std::vector<at::Tensor> work(tensors.size());
        bool any_cpu = false;
        for (size_t i = 0; i < tensors.size(); ++i) {
            const at::Tensor& t = tensors[i];
            any_cpu = any_cpu || !t.is_cuda();
            work[i] = (t.is_cuda() && t.is_contiguous()) ? t : t.to(c10::Device(c10::kCUDA, dev_)).contiguous();
        }
        std::vector<bool> taken(tensors.size(), false);
        for (size_t i = 0; i < tensors.size(); ++i) {
            if (taken[i]) continue;
            const auto st = work[i].scalar_type();
            std::vector<void*> ptrs;
            std::vector<int64_t> counts;
            for (size_t j = i; j < tensors.size(); ++j) {
                if (taken[j] || work[j].scalar_type() != st) continue;
                taken[j] = true;
                ptrs.push_back(work[j].data_ptr());
                counts.push_back(work[j].numel());
            }
            DLBB_CALL(dlbb_allreduce_multi(comm_, ptrs.data(), counts.data(), static_cast<int>(ptrs.size()),
                                           to_dlbb_dtype(st, true), to_dlbb_op(opts.reduceOp), stream()));
        }
        if (opts.reduceOp == c10d::ReduceOp::AVG) at::_foreach_div_(work, static_cast<double>(getSize()));
        for (size_t i = 0; i < tensors.size(); ++i)
            if (!work[i].is_same(tensors[i])) tensors[i].copy_(work[i].view_as(tensors[i]));
        return done(c10d::OpType::ALLREDUCE_COALESCED, tensors, !any_cpu);
    }

    c10::intrusive_ptr<Work> broadcast(std::vector<at::Tensor>& tensors, const c10d::BroadcastOptions& opts) override {
        TORCH_CHECK(tensors.size() == 1, "dlbb backend: one tensor per rank");
        at::Tensor& t = tensors[0];
        const int root = static_cast<int>(opts.rootRank);
        dlbb_view_t v{};
        const bool view = t.is_cuda() && !t.is_contiguous() && strided_view(t, &v);
        Staged s = stage_in(t, getRank() == root, /*force_contig=*/!view);
        if (view) DLBB_CALL(dlbb_broadcast_v(comm_, s.dev.data_ptr(), s.dev.numel(), move_dtype(s.dev.element_size()), root, &v, stream()));
        else DLBB_CALL(dlbb_broadcast(comm_, s.dev.data_ptr(), nbytes(s.dev), DLBB_U8, root, stream()));
        return finish(c10d::OpType::BROADCAST, s, {t});
    }

    c10::intrusive_ptr<Work> reduce(std::vector<at::Tensor>& tensors, const c10d::ReduceOptions& opts) override {
        TORCH_CHECK(tensors.size() == 1, "dlbb backend: one tensor per rank");
        at::Tensor& t = tensors[0];
        const int root = static_cast<int>(opts.rootRank);
        check_avg(opts.reduceOp, tensors);
        dlbb_view_t v{};
        const bool view = t.is_cuda() && !t.is_contiguous() && strided_view(t, &v);
        Staged s = stage_in(t, true, /*force_contig=*/!view);
        if (view)
            DLBB_CALL(dlbb_reduce_v(comm_, s.dev.data_ptr(), s.dev.data_ptr(), s.dev.numel(), to_dlbb_dtype(s.dev.scalar_type(), true),
                                    to_dlbb_op(opts.reduceOp), root, &v, &v, stream()));
        else
            DLBB_CALL(dlbb_reduce(comm_, s.dev.data_ptr(), s.dev.data_ptr(), s.dev.numel(),
                                  to_dlbb_dtype(s.dev.scalar_type(), true), to_dlbb_op(opts.reduceOp), root, stream()));
        if (opts.reduceOp == c10d::ReduceOp::AVG && getRank() == root) s.dev.div_(getSize());
        s.copy_back = s.copy_back && getRank() == root;
        return finish(c10d::OpType::REDUCE, s, {t});
    }

    c10::intrusive_ptr<Work> _allgather_base(at::Tensor& out, at::Tensor& in, const c10d::AllgatherOptions&) override {
        TORCH_CHECK(out.numel() == in.numel() * getSize(), "dlbb backend: all_gather_into_tensor size mismatch");
        dlbb_view_t v{};
        const bool view = in.is_cuda() && !in.is_contiguous() && strided_view(in, &v);
        Staged si = stage_in(in, true, /*force_contig=*/!view);
        Staged so = stage_in(out, false, true);
        if (view) DLBB_CALL(dlbb_allgather(comm_, si.dev.data_ptr(), so.dev.data_ptr(), si.dev.numel(), move_dtype(si.dev.element_size()), &v, stream()));
        else DLBB_CALL(dlbb_allgather(comm_, si.dev.data_ptr(), so.dev.data_ptr(), nbytes(si.dev), DLBB_U8, nullptr, stream()));
        return finish(c10d::OpType::_ALLGATHER_BASE, so, {out});
    }

    c10::intrusive_ptr<Work> allgather(std::vector<std::vector<at::Tensor>>& outs, std::vector<at::Tensor>& ins,
                                       const c10d::AllgatherOptions& opts) override {
        TORCH_CHECK(ins.size() == 1 && outs.size() == 1 && (int)outs[0].size() == getSize(),
                    "dlbb backend: all_gather expects one input and world_size outputs");
        at::Tensor& in = ins[0];
        dlbb_view_t v{};
        const bool view = in.is_cuda() && !in.is_contiguous() && strided_view(in, &v);
        Staged si = stage_in(in, true, /*force_contig=*/!view);
        if (in.is_cuda() && move_dtype(in.element_size()) >= 0) {
            // device list: every rank's contribution is pulled straight into its output tensor (dense temporaries stand in
            // for outputs that are not dense, so that every rank runs the same kernel)
            const bool direct = list_is_dense_cuda(outs[0], in);
            std::vector<void*> op(getSize());
            std::vector<at::Tensor> tmp;
            for (int p = 0; p < getSize(); ++p) {
                if (!direct) tmp.push_back(at::empty(in.sizes(), in.options().memory_format(at::MemoryFormat::Contiguous)));
                op[p] = direct ? outs[0][p].data_ptr() : tmp[p].data_ptr();
            }
            DLBB_CALL(dlbb_allgather_ptrs(comm_, si.dev.data_ptr(), op.data(), si.dev.numel(), move_dtype(si.dev.element_size()), -1,
                                          view ? &v : nullptr, stream()));
            for (size_t p = 0; p < tmp.size(); ++p) outs[0][p].copy_(tmp[p].view_as(outs[0][p]));
            return done(c10d::OpType::ALLGATHER, outs[0], true);
        }
        at::Tensor flat = at::empty({getSize(), in.numel()}, si.dev.options());
        if (view) DLBB_CALL(dlbb_allgather(comm_, si.dev.data_ptr(), flat.data_ptr(), si.dev.numel(), move_dtype(si.dev.element_size()), &v, stream()));
        else DLBB_CALL(dlbb_allgather(comm_, si.dev.data_ptr(), flat.data_ptr(), nbytes(si.dev), DLBB_U8, nullptr, stream()));
        for (int p = 0; p < getSize(); ++p) outs[0][p].copy_(flat[p].view_as(in), /*non_blocking=*/false);
        return done(c10d::OpType::ALLGATHER, outs[0], in.is_cuda());
    }

    c10::intrusive_ptr<Work> gather(std::vector<std::vector<at::Tensor>>& outs, std::vector<at::Tensor>& ins,
                                    const c10d::GatherOptions& opts) override {
        TORCH_CHECK(ins.size() == 1, "dlbb backend: gather expects one input");
        const int root = static_cast<int>(opts.rootRank);
        at::Tensor& in = ins[0];
        dlbb_view_t v{};
        const bool view = in.is_cuda() && !in.is_contiguous() && strided_view(in, &v);
        Staged si = stage_in(in, true, /*force_contig=*/!view);
        at::Tensor flat;
        if (getRank() == root)
            TORCH_CHECK(outs.size() == 1 && (int)outs[0].size() == getSize(), "dlbb backend: gather needs world_size outputs on the root");
        if (in.is_cuda() && move_dtype(in.element_size()) >= 0) {
            // CUDA tensors (on every rank: the SPMD contract): the kernel writes the root's list directly; a list that is
            // not dense is gathered into dense temporaries first - the choice of kernel must not depend on the rank
            std::vector<void*> op(getSize(), nullptr);
            std::vector<at::Tensor> tmp;
            if (getRank() == root) {
                const bool direct = list_is_dense_cuda(outs[0], in);
                for (int p = 0; p < getSize(); ++p) {
                    if (!direct) tmp.push_back(at::empty(in.sizes(), in.options().memory_format(at::MemoryFormat::Contiguous)));
                    op[p] = direct ? outs[0][p].data_ptr() : tmp[p].data_ptr();
                }
            }
            DLBB_CALL(dlbb_allgather_ptrs(comm_, si.dev.data_ptr(), op.data(), si.dev.numel(), move_dtype(si.dev.element_size()), root,
                                          view ? &v : nullptr, stream()));
            for (size_t p = 0; p < tmp.size(); ++p) outs[0][p].copy_(tmp[p].view_as(outs[0][p]));
            return done(c10d::OpType::GATHER, getRank() == root ? outs[0] : std::vector<at::Tensor>{}, true);
        }
        if (getRank() == root) flat = at::empty({getSize(), in.numel()}, si.dev.options());
        if (view)
            DLBB_CALL(dlbb_gather_v(comm_, si.dev.data_ptr(), flat.defined() ? flat.data_ptr() : nullptr, si.dev.numel(),
                                    move_dtype(si.dev.element_size()), root, &v, stream()));
        else
            DLBB_CALL(dlbb_gather(comm_, si.dev.data_ptr(), flat.defined() ? flat.data_ptr() : nullptr, nbytes(si.dev), DLBB_U8,
                                  root, stream()));
        std::vector<at::Tensor> res;
        if (getRank() == root) {
            for (int p = 0; p < getSize(); ++p) outs[0][p].copy_(flat[p].view_as(in));
            res = outs[0];
        }
        return done(c10d::OpType::GATHER, res, in.is_cuda());
    }

    c10::intrusive_ptr<Work> scatter(std::vector<at::Tensor>& outs, std::vector<std::vector<at::Tensor>>& ins,
                                     const c10d::ScatterOptions& opts) override {
        TORCH_CHECK(outs.size() == 1, "dlbb backend: scatter expects one output");
        const int root = static_cast<int>(opts.rootRank);
        at::Tensor& out = outs[0];
        Staged so = stage_in(out, false, true);
        at::Tensor flat;
        if (getRank() == root) {
            TORCH_CHECK(ins.size() == 1 && (int)ins[0].size() == getSize(), "dlbb backend: scatter needs world_size inputs on the root");
            flat = at::empty({getSize(), out.numel()}, so.dev.options());
            for (int p = 0; p < getSize(); ++p) flat[p].copy_(ins[0][p].reshape({-1}));
        }
        DLBB_CALL(dlbb_scatter(comm_, flat.defined() ? flat.data_ptr() : nullptr, so.dev.data_ptr(), nbytes(so.dev), DLBB_U8,
                               root, stream()));
        return finish(c10d::OpType::SCATTER, so, {out});
    }

    c10::intrusive_ptr<Work> _reduce_scatter_base(at::Tensor& out, at::Tensor& in, const c10d::ReduceScatterOptions& opts) override {
        TORCH_CHECK(in.numel() == out.numel() * getSize(), "dlbb backend: reduce_scatter_tensor size mismatch");
        check_avg(opts.reduceOp, {out});
        Staged si = stage_in(in, true, true);
        Staged so = stage_in(out, false, true);
        DLBB_CALL(dlbb_reduce_scatter(comm_, si.dev.data_ptr(), so.dev.data_ptr(), so.dev.numel(),
                                      to_dlbb_dtype(si.dev.scalar_type(), true), to_dlbb_op(opts.reduceOp), stream()));
        if (opts.reduceOp == c10d::ReduceOp::AVG) so.dev.div_(getSize());
        return finish(c10d::OpType::_REDUCE_SCATTER_BASE, so, {out});
    }

    c10::intrusive_ptr<Work> reduce_scatter(std::vector<at::Tensor>& outs, std::vector<std::vector<at::Tensor>>& ins,
                                            const c10d::ReduceScatterOptions& opts) override {
        TORCH_CHECK(outs.size() == 1 && ins.size() == 1 && (int)ins[0].size() == getSize(),
                    "dlbb backend: reduce_scatter expects one output and world_size inputs");
        at::Tensor& out = outs[0];
        check_avg(opts.reduceOp, outs);
        Staged so = stage_in(out, false, true);
        at::Tensor flat = at::empty({getSize(), out.numel()}, so.dev.options());
        for (int p = 0; p < getSize(); ++p) flat[p].copy_(ins[0][p].reshape({-1}));
        DLBB_CALL(dlbb_reduce_scatter(comm_, flat.data_ptr(), so.dev.data_ptr(), so.dev.numel(),
                                      to_dlbb_dtype(so.dev.scalar_type(), true), to_dlbb_op(opts.reduceOp), stream()));
        if (opts.reduceOp == c10d::ReduceOp::AVG) so.dev.div_(getSize());
        return finish(c10d::OpType::REDUCE_SCATTER, so, {out});
    }

    c10::intrusive_ptr<Work> alltoall_base(at::Tensor& out, at::Tensor& in, std::vector<int64_t>& osplit,
                                           std::vector<int64_t>& isplit, const c10d::AllToAllOptions&) override {
        TORCH_CHECK(osplit.empty() && isplit.empty(), "dlbb backend: all_to_all_single supports equal splits only");
        TORCH_CHECK(in.numel() == out.numel() && in.numel() % getSize() == 0, "dlbb backend: all_to_all size mismatch");
        Staged si = stage_in(in, true, true);
        Staged so = stage_in(out, false, true);
        DLBB_CALL(dlbb_alltoall(comm_, si.dev.data_ptr(), so.dev.data_ptr(), nbytes(si.dev) / getSize(), DLBB_U8, stream()));
        return finish(c10d::OpType::ALLTOALL_BASE, so, {out});
    }

    c10::intrusive_ptr<Work> alltoall(std::vector<at::Tensor>& outs, std::vector<at::Tensor>& ins,
                                      const c10d::AllToAllOptions&) override {
        TORCH_CHECK((int)outs.size() == getSize() && (int)ins.size() == getSize(), "dlbb backend: all_to_all needs world_size tensors");
        const int64_t n = ins[0].numel();
        for (int p = 0; p < getSize(); ++p)
            TORCH_CHECK(ins[p].numel() == n && outs[p].numel() == n, "dlbb backend: all_to_all supports equal chunk sizes only");
        const bool cuda = ins[0].is_cuda();
        bool dense = cuda;
        for (int p = 0; p < getSize() && dense; ++p)
            dense = ins[p].is_cuda() && outs[p].is_cuda() && ins[p].is_contiguous() && outs[p].is_contiguous();
        if (dense) {   // device lists: the kernel reads and writes the list's tensors directly
            c10::cuda::CUDAGuard g(dev_);
            std::vector<const void*> ip(getSize());
            std::vector<void*> op(getSize());
            for (int p = 0; p < getSize(); ++p) { ip[p] = ins[p].data_ptr(); op[p] = outs[p].data_ptr(); }
            DLBB_CALL(dlbb_alltoall_ptrs(comm_, ip.data(), op.data(), n * ins[0].element_size(), stream()));
            return done(c10d::OpType::ALLTOALL, outs, true);
        }
        auto options = ins[0].options().device(c10::Device(c10::kCUDA, dev_));
        at::Tensor fin = at::empty({getSize(), n}, options), fout = at::empty({getSize(), n}, options);
        for (int p = 0; p < getSize(); ++p) fin[p].copy_(ins[p].reshape({-1}));
        DLBB_CALL(dlbb_alltoall(comm_, fin.data_ptr(), fout.data_ptr(), n * ins[0].element_size(), DLBB_U8, stream()));
        for (int p = 0; p < getSize(); ++p) outs[p].copy_(fout[p].view_as(outs[p]));
        return done(c10d::OpType::ALLTOALL, outs, cuda);
    }

    // ------------------------------------------------------------------ point to point
    c10::intrusive_ptr<Work> send(std::vector<at::Tensor>& tensors, int dst, int /*tag*/) override {
        TORCH_CHECK(tensors.size() == 1, "dlbb backend: one tensor per send");
        return p2p(tensors[0], dst, /*is_send=*/true);
    }
    c10::intrusive_ptr<Work> recv(std::vector<at::Tensor>& tensors, int src, int /*tag*/) override {
        TORCH_CHECK(tensors.size() == 1, "dlbb backend: one tensor per recv");
        return p2p(tensors[0], src, /*is_send=*/false);
    }

    c10::intrusive_ptr<Work> barrier(const c10d::BarrierOptions&) override {
        c10::cuda::CUDAGuard g(dev_);
        DLBB_CALL(dlbb_barrier(comm_, stream()));
        // host-blocking, like an MPI barrier: the reference brackets its timed regions with it
        C10_CUDA_CHECK(cudaStreamSynchronize(c10::cuda::getCurrentCUDAStream(dev_)));
        return c10::make_intrusive<WorkDlbb>(c10d::OpType::BARRIER, std::vector<at::Tensor>{}, c10::nullopt, dev_);
    }

  private:
    // every tensor of an output list is a dense CUDA tensor of the input's size on this device
    bool list_is_dense_cuda(const std::vector<at::Tensor>& list, const at::Tensor& like) const {
        if (!like.is_cuda()) return false;
        for (const auto& t : list)
            if (!t.is_cuda() || t.get_device() != dev_ || !t.is_contiguous() || t.numel() != like.numel() ||
                t.element_size() != like.element_size())
                return false;
        return true;
    }

    // ReduceOp::AVG is SUM followed by a division: only meaningful on floating-point tensors
    static void check_avg(const c10d::ReduceOp& op, const std::vector<at::Tensor>& ts) {
        if (op != c10d::ReduceOp::AVG) return;
        for (const auto& t : ts)
            TORCH_CHECK(at::isFloatingType(t.scalar_type()), "dlbb backend: ReduceOp.AVG needs a floating-point tensor, got ",
                        t.scalar_type());
    }

    struct Staged {
        at::Tensor dev;        // tensor the kernel works on (the original if it already lives on the GPU)
        at::Tensor host;       // original CPU tensor (undefined for CUDA inputs)
        at::Tensor orig;       // original CUDA tensor when a contiguous copy had to be made
        bool copy_back = false;
    };

    void* stream() { return c10::cuda::getCurrentCUDAStream(dev_).stream(); }
    static int64_t nbytes(const at::Tensor& t) { return t.numel() * t.element_size(); }

    // Bring a tensor to the GPU (CPU callers) and, when the op needs it, make it dense.
    Staged stage_in(at::Tensor& t, bool need_copy, bool force_contig = false) {
        Staged s;
        c10::cuda::CUDAGuard g(dev_);
        if (t.is_cuda()) {
            TORCH_CHECK(t.get_device() == dev_, "dlbb backend: tensor is on cuda:", t.get_device(), " but this rank drives cuda:", (int)dev_);
            if (force_contig && !t.is_contiguous()) {
                s.orig = t;
                s.dev = need_copy ? t.contiguous() : at::empty(t.sizes(), t.options());
                s.copy_back = true;
            } else {
                s.dev = t;
            }
        } else {
            s.host = t;
            auto opt = t.options().device(c10::Device(c10::kCUDA, dev_));
            s.dev = need_copy ? t.contiguous().to(opt.device(), /*non_blocking=*/false) : at::empty(t.sizes(), opt);
            s.copy_back = true;
        }
        return s;
    }

    c10::intrusive_ptr<Work> finish(c10d::OpType op, Staged& s, std::vector<at::Tensor> outs) {
        if (s.host.defined()) {
            if (s.copy_back) s.host.copy_(s.dev);  // synchronous D2H: the op is complete on return
            else C10_CUDA_CHECK(cudaStreamSynchronize(c10::cuda::getCurrentCUDAStream(dev_)));
            return c10::make_intrusive<WorkDlbb>(op, std::move(outs), c10::nullopt, dev_);
        }
        if (s.orig.defined() && s.copy_back) s.orig.copy_(s.dev);
        return done(op, std::move(outs), true);
    }

    c10::intrusive_ptr<Work> done(c10d::OpType op, std::vector<at::Tensor> outs, bool cuda) {
        if (!cuda) {
            C10_CUDA_CHECK(cudaStreamSynchronize(c10::cuda::getCurrentCUDAStream(dev_)));
            return c10::make_intrusive<WorkDlbb>(op, std::move(outs), c10::nullopt, dev_);
        }
        at::cuda::CUDAEvent ev;
        ev.record(c10::cuda::getCurrentCUDAStream(dev_));
        return c10::make_intrusive<WorkDlbb>(op, std::move(outs), std::move(ev), dev_);
    }

    c10::intrusive_ptr<Work> p2p(at::Tensor& t, int peer, bool is_send) {
        TORCH_CHECK(peer >= 0 && peer < getSize() && peer != getRank(), "dlbb backend: bad p2p peer ", peer);
        c10::cuda::CUDAGuard g(dev_);
        if (!p2p_streams_) {
            send_stream_ = c10::cuda::getStreamFromPool(false, dev_);
            recv_stream_ = c10::cuda::getStreamFromPool(false, dev_);
            p2p_streams_ = true;
        }
        auto& side = is_send ? *send_stream_ : *recv_stream_;
        const bool cpu = !t.is_cuda();
        at::Tensor devt;
        if (cpu) devt = is_send ? t.contiguous().to(c10::Device(c10::kCUDA, dev_)) : at::empty(t.sizes(), t.options().device(c10::Device(c10::kCUDA, dev_)));
        else {
            TORCH_CHECK(t.is_contiguous(), "dlbb backend: p2p tensors must be contiguous");
            devt = t;
        }
        // the side stream starts after everything already queued on the caller's stream
        at::cuda::CUDAEvent before;
        before.record(c10::cuda::getCurrentCUDAStream(dev_));
        before.block(side);
        const int64_t nb = nbytes(devt);
        if (is_send) DLBB_CALL(dlbb_sendrecv(comm_, devt.data_ptr(), nb, peer, nullptr, 0, -1, DLBB_U8, side.stream()));
        else DLBB_CALL(dlbb_sendrecv(comm_, nullptr, 0, -1, devt.data_ptr(), nb, peer, DLBB_U8, side.stream()));
        c10::cuda::CUDACachingAllocator::recordStream(devt.storage().data_ptr(), side);
        at::cuda::CUDAEvent ev;
        ev.record(side);
        const auto op = is_send ? c10d::OpType::SEND : c10d::OpType::RECV;
        if (cpu) {
            // asynchronous for host tensors too: the kernel may have to wait for the peer (a second
            // message to the same rank needs the first to be consumed), and blocking here would
            // deadlock two ranks that both isend before they irecv.  wait() blocks and, for a
            // receive, copies the payload back to the host tensor.
            std::function<void()> fin = nullptr;
            if (!is_send) {
                at::Tensor host = t, dev_copy = devt;
                fin = [host, dev_copy]() mutable { host.copy_(dev_copy); };
            }
            return c10::make_intrusive<WorkDlbb>(op, std::vector<at::Tensor>{t}, std::move(ev), dev_, /*host_blocking=*/true,
                                                 std::move(fin), std::vector<at::Tensor>{devt});
        }
        return c10::make_intrusive<WorkDlbb>(op, std::vector<at::Tensor>{t}, std::move(ev), dev_);
    }

    dlbb_comm_t* comm_ = nullptr;
    c10::DeviceIndex dev_ = 0;
    int algo_ = DLBB_ALGO_AUTO;
    bool p2p_streams_ = false;
    c10::optional<c10::cuda::CUDAStream> send_stream_, recv_stream_;
};

}  // namespace dlbb_pg

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    namespace py = pybind11;
    using dlbb_pg::BackendDlbb;
    m.doc() = "dlbb process-group backend (c10d::Backend over libdlbb.so)";
    py::class_<BackendDlbb, c10d::Backend, c10::intrusive_ptr<BackendDlbb>>(m, "BackendDlbb")
        .def("has_multicast", &BackendDlbb::has_multicast)
        .def("set_ctas", &BackendDlbb::set_ctas)
        .def("set_oneshot_max_bytes", &BackendDlbb::set_oneshot_max_bytes)
        .def("set_algo", &BackendDlbb::set_algo)
        .def("symm_empty", [](BackendDlbb& b, std::vector<int64_t> sizes, py::object dtype) {
            return b.symm_empty(std::move(sizes), torch::python::detail::py_object_to_dtype(dtype));
        }, py::arg("sizes"), py::arg("dtype"));
    m.def(
        "create_backend",
        [](const c10::intrusive_ptr<c10d::Store>& store, int rank, int size, const std::chrono::milliseconds&) {
            return c10::make_intrusive<BackendDlbb>(store, rank, size);
        },
        py::arg("store"), py::arg("rank"), py::arg("size"), py::arg("timeout") = std::chrono::milliseconds(0),
        py::call_guard<py::gil_scoped_release>());
}
