// wkv6_torch.cpp -- custom-op shim: registers the reference's dispatcher namespace `wkv6` with the exact
// schemas of wenet/rwkv_v6/cuda/wkv6_op.cpp:9-41 (six void ops writing into caller-allocated outputs) on
// top of the C ABI in include/wkv6_b200.h.  The unmodified reference model.py (WKV_6 / WKV_6_FP32,
// model.py:108-214) resolves torch.ops.wkv6.* to these -- also from inside its TorchScript ScriptModule.
//
// Differences from the reference shim: arguments are validated (TORCH_CHECK instead of C assert), the launch
// goes to PyTorch's current stream (the reference uses the legacy default stream), T is unlimited in backward.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>
#include <torch/types.h>

#include "../../include/wkv6_b200.h"

namespace {

void check_btc(const at::Tensor& t, const char* name, at::ScalarType dt, int64_t B, int64_t T, int64_t C) {
    TORCH_CHECK(t.is_cuda(), "wkv6: ", name, " must be a CUDA tensor (this build has no CPU path)");
    TORCH_CHECK(t.scalar_type() == dt, "wkv6: ", name, " has dtype ", t.scalar_type(), ", expected ", dt);
    TORCH_CHECK(t.is_contiguous(), "wkv6: ", name, " must be contiguous");
    TORCH_CHECK(t.numel() == B * T * C, "wkv6: ", name, " has ", t.numel(), " elements, expected B*T*C = ", B * T * C);
}

void check_rc(int rc) { TORCH_CHECK(rc == 0, "wkv6_b200: ", wkv6_b200_last_error(), " (code ", rc, ")"); }

void fwd_impl(int dtype, at::ScalarType dt, int64_t B, int64_t T, int64_t C, int64_t H, const at::Tensor& r,
              const at::Tensor& k, const at::Tensor& v, const at::Tensor& w, const at::Tensor& u, at::Tensor& y) {
    check_btc(r, "r", dt, B, T, C); check_btc(k, "k", dt, B, T, C); check_btc(v, "v", dt, B, T, C);
    check_btc(w, "w", dt, B, T, C); check_btc(y, "y", dt, B, T, C);
    TORCH_CHECK(u.is_cuda() && u.scalar_type() == dt && u.is_contiguous() && u.numel() == C, "wkv6: bad u");
    const c10::cuda::CUDAGuard guard(r.device());
    // few, long sequences run time-segmented on a scratch buffer from the caching allocator (stream-ordered)
    const size_t sc_bytes = wkv6_b200_forward_scratch_bytes(dtype, (int)B, (int)T, (int)C, (int)H, 1);
    at::Tensor sc = at::empty({(int64_t)sc_bytes}, r.options().dtype(at::kByte));
    check_rc(wkv6_b200_forward(dtype, (int)B, (int)T, (int)C, (int)H, r.data_ptr(), k.data_ptr(), v.data_ptr(),
                               w.data_ptr(), u.data_ptr(), nullptr, y.data_ptr(), nullptr, nullptr, 0,
                               sc_bytes ? sc.data_ptr() : nullptr, sc_bytes, at::cuda::getCurrentCUDAStream().stream()));
}

void bwd_impl(int dtype, at::ScalarType dt, int64_t B, int64_t T, int64_t C, int64_t H, const at::Tensor& r,
              const at::Tensor& k, const at::Tensor& v, const at::Tensor& w, const at::Tensor& u, const at::Tensor& gy,
              at::Tensor& gr, at::Tensor& gk, at::Tensor& gv, at::Tensor& gw, at::Tensor& gu) {
    check_btc(r, "r", dt, B, T, C); check_btc(k, "k", dt, B, T, C); check_btc(v, "v", dt, B, T, C);
    check_btc(w, "w", dt, B, T, C); check_btc(gy, "gy", dt, B, T, C); check_btc(gr, "gr", dt, B, T, C);
    check_btc(gk, "gk", dt, B, T, C); check_btc(gv, "gv", dt, B, T, C); check_btc(gw, "gw", dt, B, T, C);
    TORCH_CHECK(u.is_cuda() && u.scalar_type() == dt && u.is_contiguous() && u.numel() == C, "wkv6: bad u");
    TORCH_CHECK(gu.is_cuda() && gu.scalar_type() == dt && gu.is_contiguous() && gu.numel() == B * C, "wkv6: bad gu");
    const c10::cuda::CUDAGuard guard(r.device());
    const size_t ws_bytes = wkv6_b200_backward_workspace_bytes(dtype, (int)B, (int)T, (int)C, (int)H, 1);
    at::Tensor ws = at::empty({(int64_t)ws_bytes}, r.options().dtype(at::kByte));   // caching allocator, stream-ordered
    check_rc(wkv6_b200_backward(dtype, (int)B, (int)T, (int)C, (int)H, r.data_ptr(), k.data_ptr(), v.data_ptr(),
                                w.data_ptr(), u.data_ptr(), nullptr, gy.data_ptr(), nullptr, nullptr, gr.data_ptr(),
                                gk.data_ptr(), gv.data_ptr(), gw.data_ptr(), gu.data_ptr(), nullptr, 0, ws.data_ptr(),
                                ws_bytes, at::cuda::getCurrentCUDAStream().stream()));
}

// --- the six reference ops (wkv6_op.cpp:9-41) ---
void forward(int64_t B, int64_t T, int64_t C, int64_t H, at::Tensor& r, at::Tensor& k, at::Tensor& v, at::Tensor& w,
             at::Tensor& u, at::Tensor& y) {
    fwd_impl(WKV6_DTYPE_BF16, at::kBFloat16, B, T, C, H, r, k, v, w, u, y);
}
void backward(int64_t B, int64_t T, int64_t C, int64_t H, at::Tensor& r, at::Tensor& k, at::Tensor& v, at::Tensor& w,
              at::Tensor& u, at::Tensor& gy, at::Tensor& gr, at::Tensor& gk, at::Tensor& gv, at::Tensor& gw, at::Tensor& gu) {
    bwd_impl(WKV6_DTYPE_BF16, at::kBFloat16, B, T, C, H, r, k, v, w, u, gy, gr, gk, gv, gw, gu);
}
void forward_fp32(int64_t B, int64_t T, int64_t C, int64_t H, at::Tensor& r, at::Tensor& k, at::Tensor& v, at::Tensor& w,
                  at::Tensor& u, at::Tensor& y) {
    fwd_impl(WKV6_DTYPE_F32, at::kFloat, B, T, C, H, r, k, v, w, u, y);
}
void backward_fp32(int64_t B, int64_t T, int64_t C, int64_t H, at::Tensor& r, at::Tensor& k, at::Tensor& v, at::Tensor& w,
                   at::Tensor& u, at::Tensor& gy, at::Tensor& gr, at::Tensor& gk, at::Tensor& gv, at::Tensor& gw,
                   at::Tensor& gu) {
    bwd_impl(WKV6_DTYPE_F32, at::kFloat, B, T, C, H, r, k, v, w, u, gy, gr, gk, gv, gw, gu);
}
// The reference's "fp64" instantiation keeps fp32 state and __expf (wkv6_cuda.cu:18-19,28): fp64 is an I/O
// format only.  Same here: convert, run the fp32 kernels, convert back.
void forward_fp64(int64_t B, int64_t T, int64_t C, int64_t H, at::Tensor& r, at::Tensor& k, at::Tensor& v, at::Tensor& w,
                  at::Tensor& u, at::Tensor& y) {
    TORCH_CHECK(r.scalar_type() == at::kDouble && y.scalar_type() == at::kDouble, "wkv6::forward_fp64 expects float64");
    at::Tensor y32 = at::empty(y.sizes(), y.options().dtype(at::kFloat));
    fwd_impl(WKV6_DTYPE_F32, at::kFloat, B, T, C, H, r.to(at::kFloat), k.to(at::kFloat), v.to(at::kFloat),
             w.to(at::kFloat), u.to(at::kFloat), y32);
    y.copy_(y32);
}
void backward_fp64(int64_t B, int64_t T, int64_t C, int64_t H, at::Tensor& r, at::Tensor& k, at::Tensor& v, at::Tensor& w,
                   at::Tensor& u, at::Tensor& gy, at::Tensor& gr, at::Tensor& gk, at::Tensor& gv, at::Tensor& gw,
                   at::Tensor& gu) {
    TORCH_CHECK(r.scalar_type() == at::kDouble && gy.scalar_type() == at::kDouble, "wkv6::backward_fp64 expects float64");
    auto o = r.options().dtype(at::kFloat);
    at::Tensor a = at::empty(gr.sizes(), o), b = at::empty(gk.sizes(), o), c = at::empty(gv.sizes(), o),
               d = at::empty(gw.sizes(), o), e = at::empty(gu.sizes(), o);
    bwd_impl(WKV6_DTYPE_F32, at::kFloat, B, T, C, H, r.to(at::kFloat), k.to(at::kFloat), v.to(at::kFloat),
             w.to(at::kFloat), u.to(at::kFloat), gy.to(at::kFloat), a, b, c, d, e);
    gr.copy_(a); gk.copy_(b); gv.copy_(c); gw.copy_(d); gu.copy_(e);
}

}  // namespace

TORCH_LIBRARY(wkv6, m) {
    m.def("forward", forward);
    m.def("backward", backward);
    m.def("forward_fp32", forward_fp32);
    m.def("backward_fp32", backward_fp32);
    m.def("forward_fp64", forward_fp64);
    m.def("backward_fp64", backward_fp64);
}
