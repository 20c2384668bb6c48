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

#include <atomic>
#include <limits>

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

at::Tensor matrices9(const at::Tensor& r, int64_t n) {
    TORCH_CHECK(r.is_cuda() && r.scalar_type() == at::kFloat, "expected float32 CUDA rotation targets");
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
        at::Tensor rt = matrices9(r_true, n);
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
                                                            1.0f / (float)n, dx.data_ptr<float>(), sums.data_ptr<double>(),
                                                            loss.data_ptr<float>(), angle.data_ptr<double>(),
                                                            flag.data_ptr<int>(), nullptr, n, st),
                         "pe_svd_orth_frobenius_step_metrics");
            } else {
                at::Tensor sum = at::empty({1}, f32.dtype(at::kDouble));
                check_rc(pe_svd_orth_frobenius_step_mean(x.data_ptr<float>(), stride, rt.data_ptr<float>(), 1.0f / (float)n,
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
                                                  1.0f / (float)n, dx.data_ptr<float>(), n, st),
                        "pe_svd_orth_frobenius_bwd");
    }
    return {dx, at::Tensor(), at::Tensor()};
}

std::vector<at::Tensor> fused_step_loss(const at::Tensor& x, const at::Tensor& r_true, bool with_angle) {
    return FusedStep::apply(x, r_true, with_angle);
}
}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.doc() = "C++ autograd node of the single-pass training step over libpe_rotation.so (host-latency path)";
    m.def("fused_step_loss", &fused_step_loss, "loss (and mean angle error, range flag) of the step; differentiable in x",
          py::arg("x"), py::arg("r_true"), py::arg("with_angle"));
}
