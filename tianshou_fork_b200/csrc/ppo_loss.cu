// PPO clipped-surrogate loss fused with per-minibatch advantage normalisation.
//
// Replaces the loss arithmetic of PPOPolicy.learn (tianshou/policy/modelfree/ppo.py:109-139):
// ~25 eager launches + 4 host syncs per minibatch become one launch (two above 16 K
// elements) that gathers the minibatch through its index vector, normalises the advantage
// with the UNBIASED std (torch.Tensor.std, ppo.py:110), evaluates ratio / clip / dual-clip /
// value-clip in f64, reduces both losses deterministically and emits the gradients
// d clip_loss / d logp and d vf_loss / d value for the autograd backward.
#include "ts_common.cuh"

#include <math.h>

namespace ts {
namespace {

constexpr int PPO_NT = 256;
constexpr int64_t kPpoSmall = 16384;  // up to here a single CTA does everything in one launch
constexpr int PPO_MAX_GRID = kNumSMs * 4;

struct PpoArgs {
    const float *logp, *value, *logp_old, *adv, *returns, *v_s;
    const int64_t* mb;
    int64_t m;
    double eps_clip, dual_clip, eps;
    int value_clip, norm_adv;
    float *out, *dlogp, *dvalue;
    double* ws;          // [0] mean  [1] 1/(std+eps)  [2] std  [3] pad, then partials
    unsigned int* cnt;   // [0] stats arrivals  [1] loss arrivals (self-resetting)
};

__device__ __forceinline__ double block_sum(double v, double* s_red) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    const int tid = threadIdx.x;
    __syncthreads();  // s_red may still be read from a previous call
    if ((tid & 31) == 0) s_red[tid >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < PPO_NT / 32; ++w) t += s_red[w];  // fixed order: deterministic
    return t;
}

__device__ __forceinline__ int64_t mb_at(const PpoArgs& a, int64_t i) {
    return a.mb ? a.mb[i] : i;
}

// sum and sum of squares of the gathered advantages over [begin, end) with stride
__device__ __forceinline__ void stats_partial(const PpoArgs& a, int64_t first, int64_t stride,
                                              double& s1, double& s2) {
    s1 = 0.0; s2 = 0.0;
    for (int64_t i = first; i < a.m; i += stride) {
        const double x = (double)a.adv[mb_at(a, i)];
        s1 += x;
        s2 = fma(x, x, s2);
    }
}

__device__ __forceinline__ void stats_finish(const PpoArgs& a, double s1, double s2) {
    const double n = (double)a.m;
    const double mean = s1 / n;
    // unbiased variance (torch.std default, ppo.py:110); clamp tiny negative round-off
    double var = (s2 - s1 * mean) / (n - 1.0);
    if (var < 0.0) var = 0.0;
    const double sd = sqrt(var);
    a.ws[0] = mean;
    a.ws[1] = 1.0 / (sd + a.eps);
    a.ws[2] = sd;
}

// per-element forward + gradient; returns the two loss contributions
__device__ __forceinline__ void element(const PpoArgs& a, int64_t i, double mean, double inv,
                                        double inv_m, double& clip_term, double& vf_term) {
    const int64_t g = mb_at(a, i);
    double A = (double)a.adv[g];
    if (a.norm_adv) A = (A - mean) * inv;                              // ppo.py:109-113
    const double ratio = exp((double)a.logp[i] - (double)a.logp_old[g]);  // :114-116
    const double lo = 1.0 - a.eps_clip, hi = 1.0 + a.eps_clip;
    const double rc = fmin(fmax(ratio, lo), hi);
    const double surr1 = ratio * A, surr2 = rc * A;                    // :118-122
    const bool in_range = ratio >= lo && ratio <= hi;
    double obj = fmin(surr1, surr2);
    double dobj = surr1 <= surr2 ? A : (in_range ? A : 0.0);           // d obj / d ratio
    if (a.dual_clip > 0.0 && A < 0.0) {                                // :123-126
        const double floor_ = a.dual_clip * A;
        if (obj < floor_) { obj = floor_; dobj = 0.0; }
    }
    clip_term = -obj;                                                  // :126 / :128
    a.dlogp[i] = (float)(-dobj * ratio * inv_m);

    const double ret = (double)a.returns[g], v = (double)a.value[i];
    double d_v;
    if (a.value_clip) {                                                // :131-137
        const double vs = (double)a.v_s[g];
        const double dv = v - vs;
        const double dvc = fmin(fmax(dv, -a.eps_clip), a.eps_clip);
        const double v_clip = vs + dvc;
        const double e1 = ret - v, e2 = ret - v_clip;
        const double vf1 = e1 * e1, vf2 = e2 * e2;
        if (vf1 >= vf2) { vf_term = vf1; d_v = -2.0 * e1; }
        else { vf_term = vf2; d_v = (dv >= -a.eps_clip && dv <= a.eps_clip) ? -2.0 * e2 : 0.0; }
    } else {                                                           // :139
        const double e1 = ret - v;
        vf_term = e1 * e1;
        d_v = -2.0 * e1;
    }
    a.dvalue[i] = (float)(d_v * inv_m);
}

__device__ __forceinline__ void write_out(const PpoArgs& a, double clip_sum, double vf_sum) {
    const double inv_m = 1.0 / (double)a.m;
    a.out[0] = (float)(clip_sum * inv_m);
    a.out[1] = (float)(vf_sum * inv_m);
    a.out[2] = (float)a.ws[0];
    a.out[3] = (float)a.ws[2];
}

// ---- one CTA, one launch ---------------------------------------------------------------------
__global__ void __launch_bounds__(PPO_NT) ppo_loss_small_kernel(PpoArgs a) {
    __shared__ double s_red[PPO_NT / 32];
    const int tid = threadIdx.x;
    if (a.norm_adv) {
        double s1, s2;
        stats_partial(a, tid, PPO_NT, s1, s2);
        s1 = block_sum(s1, s_red);
        s2 = block_sum(s2, s_red);
        if (tid == 0) stats_finish(a, s1, s2);
        __syncthreads();
    } else if (tid == 0) {
        a.ws[0] = 0.0; a.ws[1] = 1.0; a.ws[2] = 0.0;
    }
    __syncthreads();
    const double mean = a.ws[0], inv = a.ws[1], inv_m = 1.0 / (double)a.m;
    double c = 0.0, v = 0.0;
    for (int64_t i = tid; i < a.m; i += PPO_NT) {
        double ct, vt;
        element(a, i, mean, inv, inv_m, ct, vt);
        c += ct; v += vt;
    }
    c = block_sum(c, s_red);
    v = block_sum(v, s_red);
    if (tid == 0) write_out(a, c, v);
}

// ---- large minibatches: stats launch + loss launch, last-arriving CTA finishes each ----------
__global__ void __launch_bounds__(PPO_NT) ppo_stats_kernel(PpoArgs a) {
    __shared__ double s_red[PPO_NT / 32];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    const int64_t gtid = blockIdx.x * (int64_t)PPO_NT + tid;
    double s1, s2;
    stats_partial(a, gtid, (int64_t)gridDim.x * PPO_NT, s1, s2);
    s1 = block_sum(s1, s_red);
    s2 = block_sum(s2, s_red);
    double* part = a.ws + 4;
    if (tid == 0) {
        part[2 * blockIdx.x] = s1;
        part[2 * blockIdx.x + 1] = s2;
        __threadfence();
        s_last = atomicAdd(a.cnt, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && tid == 0) {
        __threadfence();
        double t1 = 0.0, t2 = 0.0;
        for (unsigned b = 0; b < gridDim.x; ++b) {
            t1 += __ldcg(part + 2 * b);
            t2 += __ldcg(part + 2 * b + 1);
        }
        stats_finish(a, t1, t2);
        a.cnt[0] = 0u;
    }
}

__global__ void __launch_bounds__(PPO_NT) ppo_loss_kernel(PpoArgs a) {
    __shared__ double s_red[PPO_NT / 32];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    const double mean = a.norm_adv ? a.ws[0] : 0.0, inv = a.norm_adv ? a.ws[1] : 1.0;
    const double inv_m = 1.0 / (double)a.m;
    double c = 0.0, v = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)PPO_NT + tid; i < a.m;
         i += (int64_t)gridDim.x * PPO_NT) {
        double ct, vt;
        element(a, i, mean, inv, inv_m, ct, vt);
        c += ct; v += vt;
    }
    c = block_sum(c, s_red);
    v = block_sum(v, s_red);
    double* part = a.ws + 4 + 2 * PPO_MAX_GRID;
    if (tid == 0) {
        part[2 * blockIdx.x] = c;
        part[2 * blockIdx.x + 1] = v;
        __threadfence();
        s_last = atomicAdd(a.cnt + 1, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && tid == 0) {
        __threadfence();
        double t1 = 0.0, t2 = 0.0;
        for (unsigned b = 0; b < gridDim.x; ++b) {
            t1 += __ldcg(part + 2 * b);
            t2 += __ldcg(part + 2 * b + 1);
        }
        if (!a.norm_adv) { a.ws[0] = 0.0; a.ws[2] = 0.0; }
        write_out(a, t1, t2);
        a.cnt[1] = 0u;
    }
}

}  // namespace
}  // namespace ts

using namespace ts;

extern "C" size_t ts_ppo_loss_workspace_bytes(int64_t m) {
    (void)m;
    // 4 doubles of statistics + 2 partial arrays of 2*PPO_MAX_GRID doubles + 4 counters
    return (size_t)(4 + 4 * PPO_MAX_GRID) * sizeof(double) + 16;
}

extern "C" int32_t ts_ppo_loss(const float* logp, const float* value, const float* logp_old,
                               const float* adv, const float* returns, const float* v_s,
                               const int64_t* mb_index, int64_t m, float eps_clip,
                               float dual_clip, int32_t value_clip, int32_t norm_adv, float eps,
                               float* out, float* dlogp, float* dvalue, void* workspace,
                               size_t workspace_bytes, ts_stream_t stream) {
    if (m <= 0) return TS_ERR_BAD_ARG;
    if (!logp || !value || !logp_old || !adv || !returns || !out || !dlogp || !dvalue ||
        !workspace)
        return TS_ERR_BAD_ARG;
    if (value_clip && !v_s) return TS_ERR_BAD_ARG;
    if (workspace_bytes < ts_ppo_loss_workspace_bytes(m)) return TS_ERR_WORKSPACE;
    if (!aligned(workspace, 8)) return TS_ERR_ALIGN;
    PpoArgs a;
    a.logp = logp; a.value = value; a.logp_old = logp_old; a.adv = adv; a.returns = returns;
    a.v_s = v_s; a.mb = mb_index; a.m = m;
    a.eps_clip = (double)eps_clip; a.dual_clip = (double)dual_clip; a.eps = (double)eps;
    a.value_clip = value_clip; a.norm_adv = norm_adv;
    a.out = out; a.dlogp = dlogp; a.dvalue = dvalue;
    a.ws = (double*)workspace;
    a.cnt = (unsigned int*)((double*)workspace + 4 + 4 * PPO_MAX_GRID);
    cudaStream_t st = (cudaStream_t)stream;
    if (m <= kPpoSmall) {
        ppo_loss_small_kernel<<<1, PPO_NT, 0, st>>>(a);
        return check_launch();
    }
    int64_t g = (m + PPO_NT * 8 - 1) / (PPO_NT * 8);
    if (g > PPO_MAX_GRID) g = PPO_MAX_GRID;
    if (norm_adv) ppo_stats_kernel<<<(unsigned)g, PPO_NT, 0, st>>>(a);
    ppo_loss_kernel<<<(unsigned)g, PPO_NT, 0, st>>>(a);
    return check_launch();
}
