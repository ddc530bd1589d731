// agemm_ext.cpp -- the reference's pybind11 module `agemm` (kernels/src/bindings.cpp:547-582) over the C ABI of
// include/arcq.h.  Same function names, keyword names (bindings.cpp:555-575), shapes, dtypes and allocation sizes
// (bindings.cpp:83-95,109-110,133-134,181-182,228-229); what differs from the reference is deliberate:
//   * kernels run on at::cuda::getCurrentCUDAStream() (the reference launches on the legacy default stream);
//   * arguments are validated (device, dtype, contiguity, shape) and every non-zero status of the C ABI becomes a
//     RuntimeError (the reference only throws for an unsupported KQ, bindings.cpp:157-160);
//   * KQ is generic; `scale` of matmul may be a Python float OR a 1-element CUDA float tensor, which is then read by the
//     kernel (the reference converts a 0-dim CUDA tensor to `const float`, an implicit D2H sync per call, :104);
//   * extra entry points beyond the reference's: matmul(..., col_scale, epilogue, out), reorder_quantize_x_scaled,
//     absmax, set_option, sf_bytes, flavour_for_kq, launch_count.
// This file is glue only: no arithmetic happens here, and there is no CPU path.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGraphsC10Utils.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <mutex>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "arcq.h"

namespace py = pybind11;

namespace {

void fail(int rc, const char *what) {
  if (rc != ARCQ_OK) throw std::runtime_error(std::string(what) + " failed (status " + std::to_string(rc) + "): " + arcq_last_error());
}

void need(const torch::Tensor &t, const char *name, c10::ScalarType dtype, int ndim = -1) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (arcquant_b200 has no CPU path)");
  TORCH_CHECK(t.scalar_type() == dtype, name, " must have dtype ", dtype, ", got ", t.scalar_type());
  TORCH_CHECK(ndim < 0 || t.dim() == ndim, name, " must be ", ndim, "-dimensional");
  TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
}

int64_t sf_bytes(int64_t rows, int64_t K) { return arcq_sf_bytes(rows, K); }

void *stream_of(const torch::Tensor &t) { return (void *)at::cuda::getCurrentCUDAStream(t.get_device()).stream(); }

// ---- placement plans (include/arcq.h: arcq_plan_*): one per reorder_index, built on first use, kept on the device ----
// Keyed by (device, data pointer, KQ).  A plan built for an index whose storage was later reused by another index of the
// same length is still a valid placement (the kernel needs a bijection, nothing more), so a stale hit costs speed only.
struct PlanKey {
  const void *ptr; int64_t kq; int dev;
  bool operator==(const PlanKey &o) const { return ptr == o.ptr && kq == o.kq && dev == o.dev; }
};
struct PlanKeyHash { size_t operator()(const PlanKey &k) const { return std::hash<const void *>()(k.ptr) ^ (size_t)(k.kq * 1315423911u) ^ (size_t)k.dev; } };
std::mutex g_plan_mu;
std::unordered_map<PlanKey, torch::Tensor, PlanKeyHash> g_plans;
bool g_plans_enabled = true;

const uint16_t *plan_for(const torch::Tensor &idx, int64_t KQ) {
  if (!g_plans_enabled || KQ % 128 || KQ > 32768) return nullptr;   // no tile height has a usable section
  const PlanKey key{idx.data_ptr(), KQ, (int)idx.get_device()};
  std::lock_guard<std::mutex> lock(g_plan_mu);
  auto it = g_plans.find(key);
  if (it != g_plans.end()) return it->second.defined() ? reinterpret_cast<const uint16_t *>(it->second.data_ptr<int16_t>()) : nullptr;
  if (c10::cuda::currentStreamCaptureStatusMayInitCtx() != c10::cuda::CaptureStatus::None) return nullptr;  // building needs a D2H copy
  torch::Tensor host = idx.to(torch::kCPU).contiguous();
  std::vector<uint16_t> plan((size_t)arcq_plan_entries((int)KQ));
  torch::Tensor dev_plan;   // stays undefined if the index is not a permutation: remembered, never retried
  if (arcq_plan_build(host.data_ptr<int16_t>(), (int)KQ, plan.data()) == ARCQ_OK) {
    dev_plan = torch::from_blob(plan.data(), {(int64_t)plan.size()}, torch::kInt16).clone().to(idx.device());
  }
  if (g_plans.size() > 4096) g_plans.clear();
  g_plans.emplace(key, dev_plan);
  return dev_plan.defined() ? reinterpret_cast<const uint16_t *>(dev_plan.data_ptr<int16_t>()) : nullptr;
}

enum QuantKind { KIND_X, KIND_W };

std::tuple<torch::Tensor, torch::Tensor> quant(QuantKind kind, const torch::Tensor &X, const torch::Tensor &reorder_index, int KE,
                                               int flavour) {
  need(X, kind == KIND_W ? "W" : "X", torch::kBFloat16, 2);
  need(reorder_index, "reorder_index", torch::kInt16, 1);
  const int64_t M = X.size(0), KQ = X.size(1), K = KQ + KE;
  TORCH_CHECK(reorder_index.numel() == KQ && reorder_index.device() == X.device(), "reorder_index must hold ", KQ,
              " int16 entries on ", X.device());
  const c10::cuda::CUDAGuard guard(X.device());
  auto opt = torch::dtype(torch::kUInt8).device(X.device());
  auto Q = torch::empty({M, K / 2}, opt);
  auto SF = torch::empty({sf_bytes(M, K)}, opt);
  const int rc = kind == KIND_W
                     ? arcq_quantize_w(X.data_ptr(), M, (int)KQ, reorder_index.data_ptr<int16_t>(), KE, flavour, Q.data_ptr(), SF.data_ptr(), stream_of(X))
                     : arcq_quantize_x_ex(X.data_ptr(), M, (int)KQ, reorder_index.data_ptr<int16_t>(), KE, flavour, plan_for(reorder_index, KQ), nullptr, 0,
                                          Q.data_ptr(), SF.data_ptr(), nullptr, stream_of(X));
  fail(rc, kind == KIND_W ? "arcq_quantize_w" : "arcq_quantize_x");
  return std::make_tuple(Q, SF);
}

std::tuple<torch::Tensor, torch::Tensor> reorder_quantize_x(const torch::Tensor &X, const torch::Tensor &reorder_index, int KE, int flavour) {
  return quant(KIND_X, X, reorder_index, KE, flavour);
}
std::tuple<torch::Tensor, torch::Tensor> reorder_quantize_w(const torch::Tensor &W, const torch::Tensor &reorder_index, int KE, int flavour) {
  return quant(KIND_W, W, reorder_index, KE, flavour);
}

std::tuple<torch::Tensor, torch::Tensor> rmsnorm_quantize_x(const torch::Tensor &X, const torch::Tensor &W, float eps,
                                                            const torch::Tensor &reorder_index, int KE) {
  need(X, "X", torch::kBFloat16, 2);
  need(W, "W", torch::kBFloat16, 1);
  need(reorder_index, "reorder_index", torch::kInt16, 1);
  const int64_t M = X.size(0), H = X.size(1), K = H + KE;
  TORCH_CHECK(W.numel() == H && reorder_index.numel() == H, "W and reorder_index must have ", H, " entries");
  TORCH_CHECK(W.device() == X.device() && reorder_index.device() == X.device(), "W and reorder_index must be on ", X.device());
  const c10::cuda::CUDAGuard guard(X.device());
  auto opt = torch::dtype(torch::kUInt8).device(X.device());
  auto Q = torch::empty({M, K / 2}, opt);
  auto SF = torch::empty({sf_bytes(M, K)}, opt);
  fail(arcq_rmsnorm_quantize_x(X.data_ptr(), W.data_ptr(), eps, M, (int)H, reorder_index.data_ptr<int16_t>(), KE, Q.data_ptr(),
                               SF.data_ptr(), stream_of(X)),
       "arcq_rmsnorm_quantize_x");
  return std::make_tuple(Q, SF);
}

// residual + hidden_states, then the norm + quantise of the next block, in one launch: returns (QX, SFX, sum)
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> add_rmsnorm_quantize_x(const torch::Tensor &X, const torch::Tensor &residual, const torch::Tensor &W,
                                                                               float eps, const torch::Tensor &reorder_index, int KE) {
  need(X, "X", torch::kBFloat16, 2);
  need(residual, "residual", torch::kBFloat16, 2);
  need(W, "W", torch::kBFloat16, 1);
  need(reorder_index, "reorder_index", torch::kInt16, 1);
  const int64_t M = X.size(0), H = X.size(1), K = H + KE;
  TORCH_CHECK(residual.size(0) == M && residual.size(1) == H, "residual must have the shape of X");
  TORCH_CHECK(W.numel() == H && reorder_index.numel() == H, "W and reorder_index must have ", H, " entries");
  TORCH_CHECK(W.device() == X.device() && reorder_index.device() == X.device() && residual.device() == X.device(), "all inputs must be on ", X.device());
  const c10::cuda::CUDAGuard guard(X.device());
  auto opt = torch::dtype(torch::kUInt8).device(X.device());
  auto Q = torch::empty({M, K / 2}, opt);
  auto SF = torch::empty({sf_bytes(M, K)}, opt);
  auto S = torch::empty({M, H}, torch::dtype(torch::kBFloat16).device(X.device()));
  fail(arcq_add_rmsnorm_quantize_x(X.data_ptr(), residual.data_ptr(), W.data_ptr(), eps, M, (int)H, reorder_index.data_ptr<int16_t>(), KE, S.data_ptr(),
                                   Q.data_ptr(), SF.data_ptr(), stream_of(X)),
       "arcq_add_rmsnorm_quantize_x");
  return std::make_tuple(Q, SF, S);
}

// max|x| of a bf16 tensor as a 1-element fp32 CUDA tensor (no host sync)
torch::Tensor absmax(const torch::Tensor &X) {
  need(X, "X", torch::kBFloat16);
  const c10::cuda::CUDAGuard guard(X.device());
  auto out = torch::empty({1}, torch::dtype(torch::kFloat32).device(X.device()));
  fail(arcq_absmax(X.data_ptr(), X.numel(), out.data_ptr<float>(), stream_of(X)), "arcq_absmax");
  return out;
}

// NVFP4_reorder_quantize_x of the reference callers (qLlamaLayer.py:73-77) in two launches: returns (QX, SFX, scale)
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> reorder_quantize_x_scaled(const torch::Tensor &X, const torch::Tensor &reorder_index,
                                                                                  int KE, int flavour, const c10::optional<torch::Tensor> &amax, int mode) {
  need(X, "X", torch::kBFloat16, 2);
  need(reorder_index, "reorder_index", torch::kInt16, 1);
  const int64_t M = X.size(0), KQ = X.size(1), K = KQ + KE;
  TORCH_CHECK(reorder_index.numel() == KQ && reorder_index.device() == X.device(), "reorder_index must hold ", KQ,
              " int16 entries on ", X.device());
  const c10::cuda::CUDAGuard guard(X.device());
  torch::Tensor am;
  if (amax.has_value()) {
    am = *amax;
    need(am, "amax", torch::kFloat32);
    TORCH_CHECK(am.numel() == 1 && am.device() == X.device(), "amax must be one float32 on ", X.device());
  } else {
    am = absmax(X);
  }
  auto opt = torch::dtype(torch::kUInt8).device(X.device());
  auto Q = torch::empty({M, K / 2}, opt);
  auto SF = torch::empty({sf_bytes(M, K)}, opt);
  auto scale = torch::empty({1}, torch::dtype(torch::kFloat32).device(X.device()));
  fail(arcq_quantize_x_ex(X.data_ptr(), M, (int)KQ, reorder_index.data_ptr<int16_t>(), KE, flavour, plan_for(reorder_index, KQ), am.data_ptr<float>(),
                          mode, Q.data_ptr(), SF.data_ptr(), scale.data_ptr<float>(), stream_of(X)),
       "arcq_quantize_x_scaled");
  return std::make_tuple(Q, SF, scale);
}

torch::Tensor matmul(const torch::Tensor &A, const torch::Tensor &B, const torch::Tensor &SFA, const torch::Tensor &SFB, const py::object &scale,
                     const c10::optional<torch::Tensor> &col_scale, int epilogue, const c10::optional<torch::Tensor> &out,
                     const c10::optional<torch::Tensor> &bias) {
  need(A, "A", torch::kUInt8, 2);
  need(B, "B", torch::kUInt8, 2);
  need(SFA, "SFA", torch::kUInt8);
  need(SFB, "SFB", torch::kUInt8);
  const int64_t M = A.size(0), N = B.size(0), K = A.size(1) * 2;  // 4-bit packing is on the columns (bindings.cpp:109)
  TORCH_CHECK(B.size(1) == A.size(1), "A and B must share the packed K dimension (", A.size(1), " vs ", B.size(1), ")");
  TORCH_CHECK(B.device() == A.device() && SFA.device() == A.device() && SFB.device() == A.device(), "B, SFA and SFB must be on ", A.device());
  TORCH_CHECK(SFA.numel() >= (M + 127) / 128 * 128 * K / 16 && SFB.numel() >= (N + 127) / 128 * 128 * K / 16,
              "scale-factor buffers too small for M=", M, " N=", N, " K=", K);
  float alpha = 0.0f;
  const float *alpha_dev = nullptr;
  torch::Tensor scale_t;
  // THPVariable_Check, not py::isinstance<torch::Tensor>: at::Tensor is not a pybind11-registered class, so the latter is
  // always false and a tensor `scale` would silently go through __float__ (.item(): a host sync, and an error under capture)
  const bool scale_is_tensor = THPVariable_Check(scale.ptr());
  if (scale_is_tensor) scale_t = scale.cast<torch::Tensor>();
  if (scale_is_tensor && scale_t.is_cuda() && scale_t.scalar_type() == torch::kFloat32 && scale_t.numel() == 1 && scale_t.device() == A.device()) {
    alpha_dev = scale_t.data_ptr<float>();  // read by the kernel: no host sync
  } else if (scale_is_tensor) {
    TORCH_CHECK(scale_t.numel() == 1, "`scale` must hold one value");
    alpha = scale_t.item<float>();  // any other tensor: .item(), a D2H sync like the reference's `const float scale` (:104)
  } else {
    alpha = scale.cast<float>();
  }
  const float *cs = nullptr;
  if (col_scale.has_value()) {
    need(*col_scale, "col_scale", torch::kFloat32, 1);
    TORCH_CHECK(col_scale->numel() == N && col_scale->device() == A.device(), "col_scale must hold ", N, " float32 on ", A.device());
    cs = col_scale->data_ptr<float>();
  }
  const void *bias_ptr = nullptr;
  if (bias.has_value()) {   // y + self.bias of the reference's linear (qLinearLayer.py:74-75), in the epilogue
    need(*bias, "bias", torch::kBFloat16, 1);
    TORCH_CHECK(bias->numel() == N && bias->device() == A.device() && epilogue == ARCQ_EPILOGUE_NONE,
                "bias must hold ", N, " bfloat16 on ", A.device(), " and needs epilogue == EPILOGUE_NONE");
    bias_ptr = bias->data_ptr();
  }
  const int64_t out_cols = epilogue == ARCQ_EPILOGUE_SWIGLU ? N / 2 : N;
  const c10::cuda::CUDAGuard guard(A.device());
  torch::Tensor C;
  int64_t ldd = out_cols;
  if (out.has_value()) {
    C = *out;
    TORCH_CHECK(C.is_cuda() && C.device() == A.device() && C.scalar_type() == torch::kBFloat16 && C.dim() == 2, "out must be a 2-D bf16 tensor on ", A.device());
    TORCH_CHECK(C.size(0) == M && C.size(1) == out_cols && C.stride(1) == 1 && C.stride(0) >= out_cols,
                "out must be [", M, ", ", out_cols, "] with unit column stride, got ", C.sizes(), " strides ", C.strides());
    ldd = C.stride(0);
  } else {
    C = torch::empty({M, out_cols}, torch::dtype(torch::kBFloat16).device(A.device()));
  }
  fail(arcq_gemm_nvfp4_ex2(A.data_ptr(), B.data_ptr(), SFA.data_ptr(), SFB.data_ptr(), M, N, K, alpha, alpha_dev, cs, bias_ptr, epilogue, C.data_ptr(),
                           ldd, stream_of(A)),
       "arcq_gemm_nvfp4");
  return C;
}

void set_option(const std::string &name, const std::string &value) { fail(arcq_set_option(name.c_str(), value.c_str()), "arcq_set_option"); }

[[noreturn]] void not_on_this_path(const char *name) {
  PyErr_SetString(PyExc_NotImplementedError,
                  (std::string("agemm.") + name + " is the reference's int4 KV-cache path (bindings.cpp:258-543), outside the NVFP4+ARC hot path").c_str());
  throw py::error_already_set();
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "B200-native agemm: NVFP4 + ARC quantise and block-scaled GEMM (drop-in for kernels/src/bindings.cpp)";
  m.def("matmul", &matmul, "nvfp4 matmul, return bfloat16 tensor", py::arg("A"), py::arg("B"), py::arg("SFA"), py::arg("SFB"), py::arg("scale"),
        py::arg("col_scale") = py::none(), py::arg("epilogue") = 0, py::arg("out") = py::none(), py::arg("bias") = py::none());
  m.def("reorder_quantize_x", &reorder_quantize_x, "Reorder and quantize activation, return K + KE channels", py::arg("X"),
        py::arg("reorder_index"), py::arg("KE"), py::arg("flavour") = -1);
  m.def("reorder_quantize_w", &reorder_quantize_w, "Reorder and quantize weight, return K + KE channels", py::arg("W"),
        py::arg("reorder_index"), py::arg("KE"), py::arg("flavour") = -1);
  m.def("rmsnorm_quantize_x", &rmsnorm_quantize_x, "Rmsnorm and quantize activation, return K + KE channels", py::arg("X"), py::arg("W"),
        py::arg("eps"), py::arg("reorder_index"), py::arg("KE"));
  m.def("add_rmsnorm_quantize_x", &add_rmsnorm_quantize_x,
        "residual add fused in front of rmsnorm_quantize_x (modeling_arc.py:294-301): (QX, SFX, X + residual)", py::arg("X"), py::arg("residual"),
        py::arg("W"), py::arg("eps"), py::arg("reorder_index"), py::arg("KE"));
  m.def("reorder_quantize_x_scaled", &reorder_quantize_x_scaled,
        "max|x| / 2688 pre-scale fused into reorder_quantize_x (NVFP4_reorder_quantize_x of the reference callers): (QX, SFX, scale)",
        py::arg("X"), py::arg("reorder_index"), py::arg("KE"), py::arg("flavour") = -1, py::arg("amax") = py::none(), py::arg("mode") = 0);
  m.def("absmax", &absmax, "max|x| of a bf16 CUDA tensor as a 1-element float32 CUDA tensor", py::arg("X"));
  m.def("sf_bytes", &sf_bytes, "scale-factor allocation of the reference (bindings.cpp:83-95)", py::arg("rows"), py::arg("K"));
  m.def("flavour_for_kq", [](int KQ) { return arcq_flavour_for_kq(KQ); }, py::arg("KQ"));
  m.def("launch_count", []() { return arcq_launch_count(); });
  m.def("set_option", &set_option, py::arg("name"), py::arg("value") = "");
  m.def("use_plans", [](bool on) { std::lock_guard<std::mutex> lock(g_plan_mu); g_plans_enabled = on; },
        "enable / disable the conflict-free placement plans of reorder_quantize_x (A/B; same bytes either way)", py::arg("on"));
  m.def("plan_build", [](const torch::Tensor &idx) {
          TORCH_CHECK(idx.scalar_type() == torch::kInt16 && idx.dim() == 1, "reorder_index must be a 1-D int16 tensor");
          torch::Tensor host = idx.to(torch::kCPU).contiguous();
          const int KQ = (int)host.numel();
          torch::Tensor plan = torch::empty({arcq_plan_entries(KQ)}, torch::kInt16);
          fail(arcq_plan_build(host.data_ptr<int16_t>(), KQ, reinterpret_cast<uint16_t *>(plan.data_ptr<int16_t>())), "arcq_plan_build");
          return plan;
        }, "host plan (int16 CPU tensor, 6 * KQ entries) of a reorder_index", py::arg("reorder_index"));
  m.def("plan_max_conflict", [](const torch::Tensor &idx, int R, const torch::Tensor &plan) {
          torch::Tensor host = idx.to(torch::kCPU).contiguous(), hp = plan.to(torch::kCPU).contiguous();
          return arcq_plan_max_conflict(host.data_ptr<int16_t>(), (int)host.numel(), R, reinterpret_cast<const uint16_t *>(hp.data_ptr<int16_t>()));
        }, py::arg("reorder_index"), py::arg("R"), py::arg("plan"));
  m.def("reorder_quantize_x_planned", [](const torch::Tensor &X, const torch::Tensor &reorder_index, int KE, const torch::Tensor &plan, int flavour) {
          need(X, "X", torch::kBFloat16, 2);
          need(reorder_index, "reorder_index", torch::kInt16, 1);
          need(plan, "plan", torch::kInt16, 1);
          const int64_t M = X.size(0), KQ = X.size(1), K = KQ + KE;
          TORCH_CHECK(reorder_index.numel() == KQ && plan.numel() == 6 * KQ && plan.device() == X.device() && reorder_index.device() == X.device(),
                      "reorder_index / plan do not match X");
          const c10::cuda::CUDAGuard guard(X.device());
          auto opt = torch::dtype(torch::kUInt8).device(X.device());
          auto Q = torch::empty({M, K / 2}, opt);
          auto SF = torch::empty({sf_bytes(M, K)}, opt);
          fail(arcq_quantize_x_ex(X.data_ptr(), M, (int)KQ, reorder_index.data_ptr<int16_t>(), KE, flavour,
                                  reinterpret_cast<const uint16_t *>(plan.data_ptr<int16_t>()), nullptr, 0, Q.data_ptr(), SF.data_ptr(), nullptr, stream_of(X)),
               "arcq_quantize_x_ex");
          return std::make_tuple(Q, SF);
        }, "reorder_quantize_x with an explicit device plan (tests: any bijection gives the same bytes)", py::arg("X"), py::arg("reorder_index"),
        py::arg("KE"), py::arg("plan"), py::arg("flavour") = -1);
  m.def("version", []() { return std::string(arcq_version()); });
  for (const char *n : {"batch_decode_i4", "init_kv_i4", "append_kv_i4", "batch_decode_f16", "init_kv_f16", "append_kv_f16"})
    m.def(n, [n](const py::args &, const py::kwargs &) { not_on_this_path(n); });
  m.attr("EPILOGUE_NONE") = ARCQ_EPILOGUE_NONE;
  m.attr("EPILOGUE_SWIGLU") = ARCQ_EPILOGUE_SWIGLU;
  m.attr("FLAVOUR_16") = ARCQ_FLAVOUR_16;
  m.attr("FLAVOUR_32") = ARCQ_FLAVOUR_32;
}
