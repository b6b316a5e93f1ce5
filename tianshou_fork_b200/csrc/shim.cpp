// Compiled host shim over the C-ABI of include/ts_b200.h for the latency-critical entry points.
//
// The kernels behind prev / next, the n-step target and the fused PPO loss run 2-6 us; called
// through ctypes from Python, marshalling ~15 arguments, allocating outputs one torch call at a
// time and (for the loss) crossing a Python autograd.Function twice cost 7-180 us per call
// (profiles/r01k_rows.md).  This file does the same calls from C++: tensors in, tensors out, the
// PPO loss as a torch::autograd::Function.  It contains no kernel and no arithmetic: every
// function forwards to the extern "C" entry point of libts_b200.so it names.  Built in-tree by
// __graft_entry__.build() (torch.utils.cpp_extension); when it is absent the Python layer calls
// the same C-ABI through ctypes.
#include <torch/extension.h>

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime.h>

#include "../../include/ts_b200.h"

namespace {

inline void* stream_of(const at::Tensor& t) {
    return (void*)c10::cuda::getCurrentCUDAStream(t.get_device()).stream();
}

inline void check(int32_t code, const char* what) {
    TORCH_CHECK(code == TS_OK, what, ": ", ts_error_string(code), " (code ", code, ")");
}

inline int32_t dtype_code(const at::Tensor& t) {
    if (t.scalar_type() == at::kFloat) return TS_F32;
    TORCH_CHECK(t.scalar_type() == at::kDouble, "expected float32 / float64");
    return TS_F64;
}

// ---- prev / next (ts_prev_index / ts_next_index) ------------------------------------------------
// `state` = the buffer's packed bookkeeping [offset(B+1) | last_index(B) | lengths(B)] (int64).
at::Tensor neighbour(const at::Tensor& index, const at::Tensor& state, int64_t B,
                     int64_t sub_size, const at::Tensor& done, bool next) {
    TORCH_CHECK(index.is_cuda() && index.scalar_type() == at::kLong && index.is_contiguous());
    at::Tensor out = at::empty_like(index);
    const int64_t n = index.numel();
    if (n == 0) return out;
    const int64_t* st = state.data_ptr<int64_t>();
    auto fn = next ? ts_next_index : ts_prev_index;
    check(fn(index.data_ptr<int64_t>(), n, st, B, sub_size, (const uint8_t*)done.data_ptr(),
             st + B + 1, st + 2 * B + 1, out.data_ptr<int64_t>(), stream_of(index)),
          next ? "ts_next_index" : "ts_prev_index");
    return out;
}

// ---- n-step (ts_nstep_chain, ts_nstep_combine) ---------------------------------------------------
std::vector<at::Tensor> nstep_chain(const at::Tensor& idx, const at::Tensor& rew,
                                    const at::Tensor& done, const at::Tensor& term,
                                    const at::Tensor& state, int64_t B, int64_t sub_size,
                                    double gamma, int64_t n_step) {
    TORCH_CHECK(idx.is_cuda() && idx.scalar_type() == at::kLong && idx.is_contiguous());
    const int64_t bsz = idx.numel();
    at::Tensor terminal = at::empty_like(idx);
    at::Tensor fl = at::empty({2, bsz}, idx.options().dtype(at::kDouble));
    const int64_t* st = state.data_ptr<int64_t>();
    check(ts_nstep_chain(idx.data_ptr<int64_t>(), bsz, rew.data_ptr(), dtype_code(rew),
                         (const uint8_t*)done.data_ptr(), (const uint8_t*)term.data_ptr(), st, B,
                         sub_size, st + B + 1, st + 2 * B + 1, gamma, (int32_t)n_step,
                         terminal.data_ptr<int64_t>(), fl.data_ptr<double>(),
                         fl.data_ptr<double>() + bsz, stream_of(idx)),
          "ts_nstep_chain");
    return {terminal, fl};
}

at::Tensor nstep_combine(const at::Tensor& tq, const at::Tensor& fl) {
    TORCH_CHECK(tq.is_cuda() && tq.is_contiguous() && tq.dim() == 2);
    const int64_t bsz = tq.size(0);
    at::Tensor out = at::empty_like(tq);
    check(ts_nstep_combine(tq.data_ptr(), dtype_code(tq), bsz, tq.size(1), fl.data_ptr<double>(),
                           fl.data_ptr<double>() + bsz, out.data_ptr(), stream_of(tq)),
          "ts_nstep_combine");
    return out;
}

// ---- fused PPO loss (ts_ppo_loss) as an autograd function ---------------------------------------
// forward: (clip_loss, vf_loss, [adv_mean, adv_std]); backward: the kernel already produced
// d clip / d logp and d vf / d value, so it is two scalings (ppo.py:109-139).
struct PpoLossFn : public torch::autograd::Function<PpoLossFn> {
    static torch::autograd::variable_list forward(
        torch::autograd::AutogradContext* ctx, const at::Tensor& logp, const at::Tensor& value,
        const at::Tensor& logp_old, const at::Tensor& adv, const at::Tensor& returns,
        const c10::optional<at::Tensor>& v_s, const c10::optional<at::Tensor>& mb,
        const at::Tensor& ws, double eps_clip, double dual_clip, bool value_clip, bool norm_adv,
        double eps) {
        at::Tensor lp = logp.detach().reshape({-1}).to(at::kFloat).contiguous();
        at::Tensor val = value.detach().reshape({-1}).to(at::kFloat).contiguous();
        const int64_t m = lp.numel();
        TORCH_CHECK(val.numel() == m, "logp and value must have one entry per sample");
        // one allocation: [clip, vf, mean, std | dlogp(m) | dvalue(m)]
        at::Tensor buf = at::empty({4 + 2 * m}, lp.options());
        float* b = buf.data_ptr<float>();
        check(ts_ppo_loss(lp.data_ptr<float>(), val.data_ptr<float>(), logp_old.data_ptr<float>(),
                          adv.data_ptr<float>(), returns.data_ptr<float>(),
                          v_s.has_value() ? v_s->data_ptr<float>() : nullptr,
                          mb.has_value() ? mb->data_ptr<int64_t>() : nullptr, m, (float)eps_clip,
                          (float)dual_clip, value_clip ? 1 : 0, norm_adv ? 1 : 0, (float)eps, b,
                          b + 4, b + 4 + m, ws.data_ptr(), (size_t)ws.numel(), stream_of(lp)),
              "ts_ppo_loss");
        ctx->save_for_backward({buf});
        ctx->saved_data["m"] = m;
        ctx->saved_data["lshape"] = logp.sizes().vec();
        ctx->saved_data["vshape"] = value.sizes().vec();
        ctx->saved_data["ldtype"] = (int64_t)logp.scalar_type();
        ctx->saved_data["vdtype"] = (int64_t)value.scalar_type();
        return {buf[0], buf[1], buf.slice(0, 2, 4)};
    }

    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                   torch::autograd::variable_list grad) {
        const at::Tensor buf = ctx->get_saved_variables()[0];
        const int64_t m = ctx->saved_data["m"].toInt();
        at::Tensor g_logp = (buf.slice(0, 4, 4 + m) * grad[0])
                                .reshape(ctx->saved_data["lshape"].toIntVector())
                                .to((at::ScalarType)ctx->saved_data["ldtype"].toInt());
        at::Tensor g_value = (buf.slice(0, 4 + m, 4 + 2 * m) * grad[1])
                                 .reshape(ctx->saved_data["vshape"].toIntVector())
                                 .to((at::ScalarType)ctx->saved_data["vdtype"].toInt());
        at::Tensor none;
        return {g_logp, g_value, none, none, none, none, none, none, none, none, none, none, none};
    }
};

std::vector<at::Tensor> ppo_loss(const at::Tensor& logp, const at::Tensor& value,
                                 const at::Tensor& logp_old, const at::Tensor& adv,
                                 const at::Tensor& returns, const c10::optional<at::Tensor>& v_s,
                                 const c10::optional<at::Tensor>& mb, const at::Tensor& ws,
                                 double eps_clip, double dual_clip, bool value_clip, bool norm_adv,
                                 double eps) {
    return PpoLossFn::apply(logp, value, logp_old, adv, returns, v_s, mb, ws, eps_clip, dual_clip,
                            value_clip, norm_adv, eps);
}

// ---- prioritized replay step with HOST arrays (prio.py:59-64, 78-102) ----------------------------
// The learner of the reference holds indices, importance weights and td-errors as NumPy arrays.
// These two calls are the copies and the launch of data/prio.py's host path issued from C++
// (one call instead of ~8 torch calls): pinned staging in, pinned staging out.
inline void cuda_ok(cudaError_t e, const char* what) {
    TORCH_CHECK(e == cudaSuccess, what, ": ", cudaGetErrorString(e));
}

// stage_u: pinned f64[m] holding the np.random.rand draws; dev_u f64[m]; dev_out / host_out:
// i64[2m] (device / pinned) = indices | importance weights (f64 bits).  Returns when host_out is valid.
void per_sample_host(const at::Tensor& heap, int64_t bound, const at::Tensor& prio,
                     const at::Tensor& err, const at::Tensor& stage_u, const at::Tensor& dev_u,
                     const at::Tensor& dev_out, const at::Tensor& host_out, int64_t m, double beta,
                     bool weight_norm) {
    TORCH_CHECK(stage_u.is_pinned() && host_out.is_pinned(), "staging buffers must be pinned");
    TORCH_CHECK(dev_u.numel() >= m && dev_out.numel() >= 2 * m && host_out.numel() >= 2 * m);
    c10::cuda::CUDAGuard guard(heap.device());
    cudaStream_t st = c10::cuda::getCurrentCUDAStream(heap.get_device()).stream();
    cuda_ok(cudaMemcpyAsync(dev_u.data_ptr(), stage_u.data_ptr(), 8 * (size_t)m,
                            cudaMemcpyHostToDevice, st), "upload of the draws");
    int64_t* out = dev_out.data_ptr<int64_t>();
    check(ts_per_sample_lazy(heap.data_ptr<double>(), bound, dev_u.data_ptr<double>(), m,
                             prio.data_ptr<double>(), beta, weight_norm ? 1 : 0, out,
                             reinterpret_cast<double*>(out + m), err.data_ptr<int32_t>(), (void*)st),
          "ts_per_sample_lazy");
    cuda_ok(cudaMemcpyAsync(host_out.data_ptr(), out, 16 * (size_t)m, cudaMemcpyDeviceToHost, st),
            "download of indices and weights");
    pybind11::gil_scoped_release nogil;
    cuda_ok(cudaStreamSynchronize(st), "cudaStreamSynchronize");
}

// stage: pinned bytes = idx (i64[m]) | td (f32[m] or f64[m]); dev_buf: device bytes of the same size
void per_update_host(const at::Tensor& heap, int64_t size, int64_t bound, const at::Tensor& prio,
                     const at::Tensor& stamp, const at::Tensor& counter, const at::Tensor& err,
                     const at::Tensor& stage, const at::Tensor& dev_buf, int64_t m, bool td_f32,
                     double alpha) {
    const size_t nb = (size_t)m * (8 + (td_f32 ? 4 : 8));
    TORCH_CHECK(stage.is_pinned() && (size_t)stage.numel() >= nb && (size_t)dev_buf.numel() >= nb);
    c10::cuda::CUDAGuard guard(heap.device());
    cudaStream_t st = c10::cuda::getCurrentCUDAStream(heap.get_device()).stream();
    uint8_t* d = static_cast<uint8_t*>(dev_buf.data_ptr());
    cuda_ok(cudaMemcpyAsync(d, stage.data_ptr(), nb, cudaMemcpyHostToDevice, st),
            "upload of indices and td-errors");
    check(ts_per_update_lazy(heap.data_ptr<double>(), size, bound, reinterpret_cast<int64_t*>(d),
                             d + 8 * (size_t)m, td_f32 ? TS_F32 : TS_F64, m, alpha,
                             prio.data_ptr<double>(), reinterpret_cast<uint32_t*>(stamp.data_ptr()),
                             reinterpret_cast<uint32_t*>(counter.data_ptr()),
                             err.data_ptr<int32_t>(), (void*)st),
          "ts_per_update_lazy");
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.def("per_sample_host", &per_sample_host, "draws up, ts_per_sample_lazy, indices + weights down");
    m.def("per_update_host", &per_update_host, "indices + td up, ts_per_update_lazy");
    m.def("neighbour", &neighbour, "ts_prev_index / ts_next_index");
    m.def("nstep_chain", &nstep_chain, "ts_nstep_chain");
    m.def("nstep_combine", &nstep_combine, "ts_nstep_combine");
    m.def("ppo_loss", &ppo_loss, "ts_ppo_loss with autograd");
}
