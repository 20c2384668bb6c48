// C++ autograd node for the training step over the C ABI of libpe_rotation.so (include/pe_rotation.h).
//
// Why: at the reference's batch sizes (32-128 matrices per step, Comparison/main.py:100, Iterative/main.py:216) a call is
// bound by the host.  The Python route (ops._FusedStep: a torch.autograd.Function whose forward and backward are Python,
// two ctypes calls) costs ~99 us per training step against PyTorch's own ~73 us floor for one CUDA op with autograd; the
// same node in C++ - no Python frames in forward or backward, no GIL hand-over on the autograd engine's device thread,
// direct calls instead of ctypes - takes most of that difference out (tools/latency.py).
//
// This file contains NO kernels and no arithmetic: checks, allocations through PyTorch's allocator, and the very same
// entry points the ctypes binding calls.  Built by poseestimation_b200/build.py (g++, linked against libpe_rotation.so
// next to it and libtorch); loaded by ops.py.
#include <torch/extension.h>

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <torch/custom_class.h>

#include <atomic>
#include <limits>
#include <mutex>

#include "../../include/pe_rotation.h"

namespace {

void check_rc(int rc, const char* what) {
    if (rc == 0) return;
    const char* msg = pe_last_error();
    TORCH_CHECK(false, what, rc > 0 ? ": CUDA error " : ": bad call ", rc, ": ", msg ? msg : "");
}

// pe_init once per device (the library's launch slots are allocated here, never by a launch)
void ensure_init(int dev) {
    static std::atomic<bool> done[64];
    if (dev >= 0 && dev < 64 && !done[dev].load(std::memory_order_acquire)) {
        check_rc(pe_init(dev), "pe_init");
        done[dev].store(true, std::memory_order_release);
    }
}

// (n, 9) float32 records with unit inner stride: a view when the rows are dense or wider (out[:, :9]), else a copy
at::Tensor records9(const at::Tensor& x, int64_t* stride) {
    TORCH_CHECK(x.is_cuda(), "pe_b200 ops compute on CUDA tensors; there is no CPU path");
    TORCH_CHECK(x.scalar_type() == at::kFloat, "pe_b200 ops compute in float32; got ", x.scalar_type());
    TORCH_CHECK(x.dim() == 2 && x.size(1) == 9, "expected a (n, 9) tensor, got ", x.sizes());
    *stride = 9;
    if (x.size(0) <= 1 || x.stride(1) != 1 || x.stride(0) < 9) return x.contiguous();
    *stride = x.stride(0);
    return x;
}

at::Tensor matrices9(const at::Tensor& r, int64_t n, const at::Tensor& like) {
    TORCH_CHECK(r.is_cuda() && r.scalar_type() == at::kFloat, "expected float32 CUDA rotation targets");
    TORCH_CHECK(r.device() == like.device(), "tensors on different devices: ", like.device(), " and ", r.device());
    TORCH_CHECK((r.dim() == 3 && r.size(1) == 3 && r.size(2) == 3) || (r.dim() == 2 && r.size(1) == 9),
                "expected (n, 3, 3) matrices, got ", r.sizes());
    TORCH_CHECK(r.size(0) == n, "batch mismatch: ", n, " records but ", r.size(0), " target rotations");
    return r.contiguous();
}

struct FusedStep : public torch::autograd::Function<FusedStep> {
    // -> {loss (0-d f32)} or {loss, mean angle (0-d f64), flag (1,) i32}
    static torch::autograd::variable_list forward(torch::autograd::AutogradContext* ctx, const at::Tensor& x_in,
                                                  const at::Tensor& r_true, bool with_angle) {
        int64_t stride = 9;
        at::Tensor x = records9(x_in, &stride);
        const int64_t n = x.size(0);
        at::Tensor rt = matrices9(r_true, n, x);
        const c10::cuda::CUDAGuard guard(x.device());
        ensure_init(x.device().index());
        auto f32 = x.options().requires_grad(false);
        at::Tensor dx = at::empty({n, 9}, f32);
        at::Tensor loss, angle, flag;
        if (n == 0) {
            const double nan = std::numeric_limits<double>::quiet_NaN();
            loss = at::full({}, nan, f32);
            if (with_angle) {
                angle = at::full({}, nan, f32.dtype(at::kDouble));
                flag = at::zeros({1}, f32.dtype(at::kInt));
            }
        } else {
            void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
            loss = at::empty({}, f32);
            if (with_angle) {
                at::Tensor sums = at::empty({3}, f32.dtype(at::kDouble));
                angle = at::empty({}, f32.dtype(at::kDouble));
                flag = at::empty({1}, f32.dtype(at::kInt));
                check_rc(pe_svd_orth_frobenius_step_metrics(x.data_ptr<float>(), stride, rt.data_ptr<float>(),
                                                            (float)(1.0 / (double)n), dx.data_ptr<float>(), sums.data_ptr<double>(),
                                                            loss.data_ptr<float>(), angle.data_ptr<double>(),
                                                            flag.data_ptr<int>(), nullptr, n, st),
                         "pe_svd_orth_frobenius_step_metrics");
            } else {
                at::Tensor sum = at::empty({1}, f32.dtype(at::kDouble));
                check_rc(pe_svd_orth_frobenius_step_mean(x.data_ptr<float>(), stride, rt.data_ptr<float>(), (float)(1.0 / (double)n),
                                                         dx.data_ptr<float>(), sum.data_ptr<double>(),
                                                         loss.data_ptr<float>(), n, st),
                         "pe_svd_orth_frobenius_step_mean");
            }
        }
        // the gradient for a unit upstream gradient stays private to the node; x and the targets are kept (no copies)
        // for a second traversal of the graph (retain_graph), which recomputes instead
        ctx->saved_data["dx"] = dx;
        ctx->saved_data["stride"] = stride;
        ctx->save_for_backward({x, rt});
        if (with_angle) {
            ctx->mark_non_differentiable({angle, flag});
            return {loss, angle, flag};
        }
        return {loss};
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads);
};

torch::autograd::variable_list FusedStep::backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads) {
    at::Tensor g = grads[0];
    TORCH_CHECK(g.defined(), "svd_frobenius_loss: undefined upstream gradient");
    if (g.scalar_type() != at::kFloat || g.numel() != 1 || !g.is_contiguous()) g = g.to(at::kFloat).reshape({1}).contiguous();
    auto saved = ctx->get_saved_variables();
    const at::Tensor& x = saved[0];
    const at::Tensor& rt = saved[1];
    const int64_t n = x.size(0);
    const c10::cuda::CUDAGuard guard(x.device());
    void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
    at::Tensor dx;
    auto it = ctx->saved_data.find("dx");
    if (it != ctx->saved_data.end() && it->second.isTensor() && it->second.toTensor().defined()) {
        dx = it->second.toTensor();
        ctx->saved_data["dx"] = at::Tensor();  // sole owner from here on: AccumulateGrad can take the buffer
        if (n) check_rc(pe_scale_by_scalar_inplace(dx.data_ptr<float>(), g.data_ptr<float>(), 9 * n, st),
                        "pe_scale_by_scalar_inplace");  // nothing on the device when the upstream gradient is 1
    } else {  // second traversal (retain_graph): the recompute kernel, 108 B/matrix
        dx = at::empty({n, 9}, x.options().requires_grad(false));
        const int64_t stride = ctx->saved_data["stride"].toInt();
        if (n) check_rc(pe_svd_orth_frobenius_bwd(x.data_ptr<float>(), stride, rt.data_ptr<float>(), g.data_ptr<float>(),
                                                  (float)(1.0 / (double)n), dx.data_ptr<float>(), n, st),
                        "pe_svd_orth_frobenius_bwd");
    }
    return {dx, at::Tensor(), at::Tensor()};
}

// ---------------------------------------------------------------------------------------------------------------------
// The UNCHANGED call surface in C++:  R = symmetric_orthogonalization(out);  loss = loss_frobenius(R, R_true)
// (Comparison/main.py:59-62, UPNA/main.py:54-57, 3D-Pose/main.py:60,85-91).  Same scheme as ops.frobenius_of_svd_output:
// loss_frobenius recognises an untouched output of the SvdOrth node, runs the single-pass step on that node's own input
// and parks the gradient (for a unit upstream gradient) where the SvdOrth node's backward finds it; its own backward
// hands R a zero-stride all-zero gradient, so gradients of other consumers of R still add up in autograd's buffer.

// what the two nodes share: gradients parked by the loss node's backward for the SvdOrth node's backward
struct Park : torch::CustomClassHolder {
    std::vector<std::pair<at::Tensor, at::Tensor>> pending;  // (d loss / d x for upstream 1, upstream gradient)
};

at::Tensor zero_scalar(const at::Device& dev) {
    static std::mutex mu;
    static at::Tensor z[64];
    std::lock_guard<std::mutex> lk(mu);
    const int i = dev.index();
    TORCH_CHECK(i >= 0 && i < 64, "device index out of range");
    if (!z[i].defined()) z[i] = at::zeros({}, at::TensorOptions().dtype(at::kFloat).device(dev));
    return z[i];
}
bool is_zero_marker(const at::Tensor& g) {
    if (!g.defined() || !g.is_cuda() || g.scalar_type() != at::kFloat) return false;
    for (auto s : g.strides())
        if (s != 0) return false;
    return g.data_ptr() == zero_scalar(g.device()).data_ptr();
}
at::Tensor scalar32(at::Tensor g) {
    if (g.scalar_type() != at::kFloat || g.numel() != 1 || !g.is_contiguous()) g = g.to(at::kFloat).reshape({1}).contiguous();
    return g;
}

struct SvdOrth : public torch::autograd::Function<SvdOrth> {
    static at::Tensor forward(torch::autograd::AutogradContext* ctx, const at::Tensor& x_in) {
        int64_t stride = 9;
        at::Tensor x = records9(x_in, &stride);
        const int64_t n = x.size(0);
        const c10::cuda::CUDAGuard guard(x.device());
        ensure_init(x.device().index());
        at::Tensor r = at::empty({n, 3, 3}, x.options().requires_grad(false));
        void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
        check_rc(pe_svd_orth_fwd(x.data_ptr<float>(), stride, r.data_ptr<float>(), nullptr, n, st), "pe_svd_orth_fwd");
        ctx->saved_data["stride"] = stride;
        ctx->saved_data["park"] = c10::IValue::make_capsule(c10::make_intrusive<Park>());
        ctx->save_for_backward({x});
        return r;
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads) {
        auto saved = ctx->get_saved_variables();
        const at::Tensor& x = saved[0];
        const int64_t n = x.size(0);
        const int64_t stride = ctx->saved_data["stride"].toInt();
        const c10::cuda::CUDAGuard guard(x.device());
        void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
        auto park = c10::static_intrusive_pointer_cast<Park>(ctx->saved_data["park"].toCapsule());
        at::Tensor total;
        const at::Tensor& g = grads[0];
        // `g` is the zero marker unless R had other consumers: their gradients go through the recompute pullback
        if (g.defined() && !(is_zero_marker(g) && !park->pending.empty())) {
            TORCH_CHECK(g.numel() == 9 * n, "batch mismatch: ", n, " records but ", g.numel() / 9, " upstream gradients");
            at::Tensor gc = g.scalar_type() == at::kFloat ? g.contiguous() : g.to(at::kFloat).contiguous();
            total = at::empty({n, 9}, x.options().requires_grad(false));
            if (n) check_rc(pe_svd_orth_bwd(x.data_ptr<float>(), stride, nullptr, gc.data_ptr<float>(), total.data_ptr<float>(), n, st),
                            "pe_svd_orth_bwd");
        }
        for (auto& p : park->pending) {
            at::Tensor gl = scalar32(p.second);
            if (n) check_rc(pe_scale_by_scalar_inplace(p.first.data_ptr<float>(), gl.data_ptr<float>(), 9 * n, st),
                            "pe_scale_by_scalar_inplace");  // nothing on the device when the upstream gradient is 1
            if (total.defined()) total.add_(p.first); else total = p.first;
        }
        park->pending.clear();
        if (!total.defined()) total = at::zeros({n, 9}, x.options().requires_grad(false));
        return {total};
    }
};

struct FusedFromSvdOutput : public torch::autograd::Function<FusedFromSvdOutput> {
    static at::Tensor forward(torch::autograd::AutogradContext* ctx, const at::Tensor& r, const at::Tensor& r_true,
                              const at::Tensor& x, int64_t stride, c10::IValue park) {
        const int64_t n = x.size(0);
        at::Tensor rt = matrices9(r_true, n, x);
        const c10::cuda::CUDAGuard guard(x.device());
        auto f32 = x.options().requires_grad(false);
        at::Tensor dx = at::empty({n, 9}, f32), loss;
        if (n == 0) {
            loss = at::full({}, std::numeric_limits<double>::quiet_NaN(), f32);
        } else {
            void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
            loss = at::empty({}, f32);
            at::Tensor sum = at::empty({1}, f32.dtype(at::kDouble));
            check_rc(pe_svd_orth_frobenius_step_mean(x.data_ptr<float>(), stride, rt.data_ptr<float>(), (float)(1.0 / (double)n),
                                                     dx.data_ptr<float>(), sum.data_ptr<double>(), loss.data_ptr<float>(), n, st),
                     "pe_svd_orth_frobenius_step_mean");
        }
        ctx->saved_data["dx"] = dx;
        ctx->saved_data["park"] = park;
        ctx->save_for_backward({r, rt});
        return loss;
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads) {
        TORCH_CHECK(grads[0].defined(), "loss_frobenius: undefined upstream gradient");
        auto saved = ctx->get_saved_variables();
        const at::Tensor& r = saved[0];
        const at::Tensor& rt = saved[1];
        const int64_t n = rt.size(0);
        at::Tensor dx = ctx->saved_data["dx"].toTensor();
        if (dx.defined()) {
            ctx->saved_data["dx"] = at::Tensor();
            auto park = c10::static_intrusive_pointer_cast<Park>(ctx->saved_data["park"].toCapsule());
            park->pending.emplace_back(dx, grads[0]);
            return {zero_scalar(r.device()).expand(r.sizes()), at::Tensor(), at::Tensor(), at::Tensor(), at::Tensor()};
        }
        // second traversal (retain_graph): the unfused route, a dense gradient for R
        const c10::cuda::CUDAGuard guard(r.device());
        void* st = c10::cuda::getCurrentCUDAStream(r.device().index()).stream();
        at::Tensor gl = scalar32(grads[0]);
        at::Tensor rc = r.contiguous();
        at::Tensor da = at::empty(r.sizes(), r.options().requires_grad(false));
        if (n) check_rc(pe_frobenius_bwd(rc.data_ptr<float>(), rt.data_ptr<float>(), gl.data_ptr<float>(), (float)(1.0 / (double)n),
                                         da.data_ptr<float>(), nullptr, n, st), "pe_frobenius_bwd");
        return {da, at::Tensor(), at::Tensor(), at::Tensor(), at::Tensor()};
    }
};

at::Tensor svd_orth(const at::Tensor& x) { return SvdOrth::apply(x); }

// loss_frobenius(a, b) through the single-pass step if one argument is an untouched output of the SvdOrth node and the
// other a plain target of the same batch; nullopt otherwise (the caller then takes the generic op)
c10::optional<at::Tensor> frobenius_of_svd_output(const at::Tensor& a, const at::Tensor& b) {
    if (!at::GradMode::is_enabled()) return c10::nullopt;
    const at::Tensor* pair[2][2] = {{&a, &b}, {&b, &a}};
    for (auto& p : pair) {
        const at::Tensor& r = *p[0];
        const at::Tensor& other = *p[1];
        if (!r.defined() || !r.requires_grad()) continue;
        auto fn = r.grad_fn();
        auto* node = dynamic_cast<torch::autograd::CppNode<SvdOrth>*>(fn.get());
        if (!node || r.output_nr() != 0 || r._version() != 0 || r.retains_grad() || r.is_view() ||
            !torch::autograd::impl::hooks(r).empty())
            continue;
        if (!other.defined() || other.requires_grad() || !other.is_cuda() || other.scalar_type() != at::kFloat ||
            other.numel() != r.numel() || other.device() != r.device())
            return c10::nullopt;
        at::Tensor x;
        try {
            x = node->ctx_.get_saved_variables()[0];  // throws if x was modified in place or the graph was freed
        } catch (const std::exception&) {
            return c10::nullopt;
        }
        // x enters detached: the step only reads its storage; its gradient reaches it through the SvdOrth node
        return FusedFromSvdOutput::apply(r, other, x.detach(), node->ctx_.saved_data["stride"].toInt(),
                                         node->ctx_.saved_data["park"]);
    }
    return c10::nullopt;
}

// ---------------------------------------------------------------------------------------------------------------------
// The other representations and the generic losses (Comparison/main.py at batch 64: map -> loss_frobenius -> backward
// took ~186 us through the registered Python ops; these nodes call the same entry points without Python frames).

// (n, w) float32 records with unit inner stride: a view when the rows are dense or wider, else a copy
at::Tensor records_w(const at::Tensor& x, int64_t w, int64_t* stride) {
    TORCH_CHECK(x.is_cuda(), "pe_b200 ops compute on CUDA tensors; there is no CPU path");
    TORCH_CHECK(x.scalar_type() == at::kFloat, "pe_b200 ops compute in float32; got ", x.scalar_type());
    TORCH_CHECK(x.dim() == 2 && x.size(1) == w, "expected a (n, ", w, ") tensor, got ", x.sizes());
    *stride = w;
    if (x.size(0) <= 1 || x.stride(1) != 1 || x.stride(0) < w) return x.contiguous();
    *stride = x.stride(0);
    return x;
}

using MapFwdFn = int (*)(const float*, int64_t, float*, int64_t, void*);
using MapBwdFn = int (*)(const float*, int64_t, const float*, float*, int64_t, void*);
struct MapEntry {
    const char* name;
    int64_t width;
    MapFwdFn fwd;
    MapBwdFn bwd;
};
// ids as ops.MAP_IDS / PE_MAP_* of include/pe_rotation.h (0 = SVD has its own nodes above)
const MapEntry* map_entry(int64_t id) {
    static const MapEntry table[] = {
        {"svd", 9, nullptr, nullptr},
        {"pe_ortho6d", 6, pe_ortho6d_fwd, pe_ortho6d_bwd},
        {"pe_quat", 4, pe_quat_fwd, pe_quat_bwd},
        {"pe_euler", 3, pe_euler_fwd, pe_euler_bwd},
        {"pe_ortho5d", 5, pe_ortho5d_fwd, pe_ortho5d_bwd},
        {"pe_expmap", 3, pe_expmap_fwd, pe_expmap_bwd},
    };
    TORCH_CHECK(id >= 1 && id <= 5, "unknown map id ", id);
    return &table[id];
}

// x (n, w) -> R (n, 3, 3); eps > 0 selects so3_exp_map's clamp bound for the exponential map (0: the default 1e-4)
struct MapNode : public torch::autograd::Function<MapNode> {
    static at::Tensor forward(torch::autograd::AutogradContext* ctx, const at::Tensor& x_in, int64_t id, double eps) {
        const MapEntry* e = map_entry(id);
        TORCH_CHECK(eps == 0.0 || id == PE_MAP_EXPMAP, "eps belongs to the exponential map");
        int64_t stride = e->width;
        at::Tensor x = records_w(x_in, e->width, &stride);
        const int64_t n = x.size(0);
        const c10::cuda::CUDAGuard guard(x.device());
        ensure_init(x.device().index());
        at::Tensor r = at::empty({n, 3, 3}, x.options().requires_grad(false));
        void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
        if (eps != 0.0)
            check_rc(pe_expmap_eps_fwd(x.data_ptr<float>(), stride, (float)eps, r.data_ptr<float>(), n, st), "pe_expmap_eps_fwd");
        else
            check_rc(e->fwd(x.data_ptr<float>(), stride, r.data_ptr<float>(), n, st), e->name);
        ctx->saved_data["id"] = id;
        ctx->saved_data["eps"] = eps;
        ctx->saved_data["stride"] = stride;
        ctx->save_for_backward({x});
        return r;
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads) {
        auto saved = ctx->get_saved_variables();
        const at::Tensor& x = saved[0];
        const int64_t id = ctx->saved_data["id"].toInt();
        const double eps = ctx->saved_data["eps"].toDouble();
        const int64_t stride = ctx->saved_data["stride"].toInt();
        const MapEntry* e = map_entry(id);
        const int64_t n = x.size(0);
        const at::Tensor& g = grads[0];
        if (!g.defined()) return {at::Tensor(), at::Tensor(), at::Tensor()};
        TORCH_CHECK(g.numel() == 9 * n, "batch mismatch: ", n, " records but ", g.numel() / 9, " upstream gradients");
        TORCH_CHECK(g.device() == x.device(), "tensors on different devices: ", x.device(), " and ", g.device());
        const c10::cuda::CUDAGuard guard(x.device());
        at::Tensor gc = g.scalar_type() == at::kFloat ? g.contiguous() : g.to(at::kFloat).contiguous();
        at::Tensor dx = at::empty({n, e->width}, x.options().requires_grad(false));
        void* st = c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
        if (eps != 0.0)
            check_rc(pe_expmap_eps_bwd(x.data_ptr<float>(), stride, gc.data_ptr<float>(), (float)eps, dx.data_ptr<float>(), n, st),
                     "pe_expmap_eps_bwd");
        else
            check_rc(e->bwd(x.data_ptr<float>(), stride, gc.data_ptr<float>(), dx.data_ptr<float>(), n, st), e->name);
        return {dx, at::Tensor(), at::Tensor()};
    }
};

at::Tensor map_fwd(const at::Tensor& x, int64_t id, double eps) { return MapNode::apply(x, id, eps); }

// two dense (n, 9) batches of the SAME length on one device (the kernels read n records of both)
at::Tensor dense9(const at::Tensor& x) {
    TORCH_CHECK(x.is_cuda(), "pe_b200 ops compute on CUDA tensors; there is no CPU path");
    TORCH_CHECK(x.scalar_type() == at::kFloat, "pe_b200 ops compute in float32; got ", x.scalar_type());
    TORCH_CHECK((x.dim() == 3 && x.size(1) == 3 && x.size(2) == 3) || (x.dim() == 2 && x.size(1) == 9),
                "expected (n, 3, 3) matrices, got ", x.sizes());
    return x.contiguous();
}
void same_batch(const at::Tensor& a, const at::Tensor& b) {
    TORCH_CHECK(a.size(0) == b.size(0), "batch mismatch: ", a.size(0), " and ", b.size(0), " matrices");
    TORCH_CHECK(a.device() == b.device(), "tensors on different devices: ", a.device(), " and ", b.device());
}

// loss_frobenius on two materialised batches (Comparison/main.py:12-16): mean_b |a_b - b_b|_F, 0-d float32
struct FrobeniusMean : public torch::autograd::Function<FrobeniusMean> {
    static at::Tensor forward(torch::autograd::AutogradContext* ctx, const at::Tensor& a_in, const at::Tensor& b_in) {
        at::Tensor a = dense9(a_in), b = dense9(b_in);
        same_batch(a, b);
        const int64_t n = a.size(0);
        const c10::cuda::CUDAGuard guard(a.device());
        ensure_init(a.device().index());
        at::Tensor loss = at::empty({}, a.options().requires_grad(false));
        if (n == 0) {
            loss.fill_(std::numeric_limits<float>::quiet_NaN());
        } else {
            at::Tensor sum = at::empty({1}, a.options().dtype(at::kDouble).requires_grad(false));
            void* st = c10::cuda::getCurrentCUDAStream(a.device().index()).stream();
            check_rc(pe_frobenius_mean_fwd(a.data_ptr<float>(), b.data_ptr<float>(), sum.data_ptr<double>(),
                                           loss.data_ptr<float>(), n, st), "pe_frobenius_mean_fwd");
        }
        ctx->save_for_backward({a, b});
        ctx->saved_data["a_sizes"] = a_in.sizes().vec();
        ctx->saved_data["b_sizes"] = b_in.sizes().vec();
        return loss;
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads) {
        auto saved = ctx->get_saved_variables();
        const at::Tensor &a = saved[0], &b = saved[1];
        const int64_t n = a.size(0);
        const bool na = ctx->needs_input_grad(0), nb = ctx->needs_input_grad(1);
        if (!grads[0].defined() || !(na || nb)) return {at::Tensor(), at::Tensor()};
        const c10::cuda::CUDAGuard guard(a.device());
        at::Tensor g = scalar32(grads[0]);
        at::Tensor da, db;
        if (na) da = at::empty(ctx->saved_data["a_sizes"].toIntVector(), a.options().requires_grad(false));
        if (nb) db = at::empty(ctx->saved_data["b_sizes"].toIntVector(), a.options().requires_grad(false));
        if (n) {
            void* st = c10::cuda::getCurrentCUDAStream(a.device().index()).stream();
            check_rc(pe_frobenius_bwd(a.data_ptr<float>(), b.data_ptr<float>(), g.data_ptr<float>(), (float)(1.0 / (double)n),
                                      na ? da.data_ptr<float>() : nullptr, nb ? db.data_ptr<float>() : nullptr, n, st),
                     "pe_frobenius_bwd");
        }
        return {da, db};
    }
};

at::Tensor frobenius_mean(const at::Tensor& a, const at::Tensor& b) { return FrobeniusMean::apply(a, b); }

// geodesic(R1, R2, reduction) of point_cloud/main.py:61-73 and compute_geodesic_distance_from_two_matrices
// (rotation_representation.py:209-227).  mode 0: per-record theta (n,), 1: mean, 2: sum (0-d float32)
struct Geodesic : public torch::autograd::Function<Geodesic> {
    static at::Tensor forward(torch::autograd::AutogradContext* ctx, const at::Tensor& a_in, const at::Tensor& b_in,
                              double lo, double hi, int64_t mode) {
        at::Tensor a = dense9(a_in), b = dense9(b_in);
        same_batch(a, b);
        TORCH_CHECK(mode >= 0 && mode <= 2, "mode: 0 none, 1 mean, 2 sum");
        const int64_t n = a.size(0);
        const c10::cuda::CUDAGuard guard(a.device());
        ensure_init(a.device().index());
        void* st = c10::cuda::getCurrentCUDAStream(a.device().index()).stream();
        at::Tensor out;
        if (mode == 0) {
            out = at::empty({n}, a.options().requires_grad(false));
            if (n) check_rc(pe_geodesic_fwd(a.data_ptr<float>(), b.data_ptr<float>(), (float)lo, (float)hi, out.data_ptr<float>(),
                                            nullptr, n, st), "pe_geodesic_fwd");
        } else {
            out = at::empty({}, a.options().requires_grad(false));
            if (n == 0) {
                out.fill_(mode == 1 ? std::numeric_limits<float>::quiet_NaN() : 0.0f);
            } else {
                at::Tensor sum = at::empty({1}, a.options().dtype(at::kDouble).requires_grad(false));
                check_rc(pe_geodesic_reduce_fwd(a.data_ptr<float>(), b.data_ptr<float>(), (float)lo, (float)hi,
                                                sum.data_ptr<double>(), out.data_ptr<float>(), mode == 1 ? 1.0 / (double)n : 1.0, n, st),
                         "pe_geodesic_reduce_fwd");
            }
        }
        ctx->save_for_backward({a, b});
        ctx->saved_data["lo"] = lo;
        ctx->saved_data["hi"] = hi;
        ctx->saved_data["mode"] = mode;
        ctx->saved_data["a_sizes"] = a_in.sizes().vec();
        ctx->saved_data["b_sizes"] = b_in.sizes().vec();
        return out;
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grads) {
        auto saved = ctx->get_saved_variables();
        const at::Tensor &a = saved[0], &b = saved[1];
        const int64_t n = a.size(0);
        const int64_t mode = ctx->saved_data["mode"].toInt();
        const bool na = ctx->needs_input_grad(0), nb = ctx->needs_input_grad(1);
        torch::autograd::variable_list none = {at::Tensor(), at::Tensor(), at::Tensor(), at::Tensor(), at::Tensor()};
        if (!grads[0].defined() || !(na || nb)) return none;
        const c10::cuda::CUDAGuard guard(a.device());
        at::Tensor g;
        if (mode == 0) {
            TORCH_CHECK(grads[0].numel() == n, grads[0].numel(), " upstream gradients for ", n, " matrices");
            g = grads[0].to(at::kFloat).reshape({-1}).contiguous();
        } else {
            g = scalar32(grads[0]);
        }
        at::Tensor da, db;
        if (na) da = at::empty(ctx->saved_data["a_sizes"].toIntVector(), a.options().requires_grad(false));
        if (nb) db = at::empty(ctx->saved_data["b_sizes"].toIntVector(), a.options().requires_grad(false));
        if (n) {
            void* st = c10::cuda::getCurrentCUDAStream(a.device().index()).stream();
            check_rc(pe_geodesic_bwd(a.data_ptr<float>(), b.data_ptr<float>(), (float)ctx->saved_data["lo"].toDouble(),
                                     (float)ctx->saved_data["hi"].toDouble(), g.data_ptr<float>(), mode == 0 ? 0 : 1,
                                     mode == 1 ? (float)(1.0 / (double)n) : 1.0f, na ? da.data_ptr<float>() : nullptr,
                                     nb ? db.data_ptr<float>() : nullptr, n, st), "pe_geodesic_bwd");
        }
        none[0] = da;
        none[1] = db;
        return none;
    }
};

at::Tensor geodesic(const at::Tensor& a, const at::Tensor& b, double lo, double hi, int64_t mode) {
    return Geodesic::apply(a, b, lo, hi, mode);
}

std::vector<at::Tensor> fused_step_loss(const at::Tensor& x, const at::Tensor& r_true, bool with_angle) {
    return FusedStep::apply(x, r_true, with_angle);
}
}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.doc() = "C++ autograd node of the single-pass training step over libpe_rotation.so (host-latency path)";
    m.def("fused_step_loss", &fused_step_loss, "loss (and mean angle error, range flag) of the step; differentiable in x",
          py::arg("x"), py::arg("r_true"), py::arg("with_angle"));
    m.def("svd_orth", &svd_orth, "symmetric_orthogonalization of (n, 9) records as a C++ autograd node", py::arg("x"));
    m.def("map_fwd", &map_fwd, "representation -> rotation as a C++ autograd node (map id as PE_MAP_*; eps: exp-map clamp, 0 = default)",
          py::arg("x"), py::arg("map_id"), py::arg("eps") = 0.0);
    m.def("frobenius_mean", &frobenius_mean, "loss_frobenius on two materialised batches as a C++ autograd node", py::arg("a"),
          py::arg("b"));
    m.def("geodesic", &geodesic, "geodesic distance, mode 0 none / 1 mean / 2 sum, as a C++ autograd node", py::arg("a"),
          py::arg("b"), py::arg("lo"), py::arg("hi"), py::arg("mode"));
    m.def("frobenius_of_svd_output", &frobenius_of_svd_output,
          "loss_frobenius(a, b) through the single-pass step when one argument is an untouched svd_orth output, else None",
          py::arg("a"), py::arg("b"));
}
