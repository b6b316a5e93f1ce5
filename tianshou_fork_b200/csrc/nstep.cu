// n-step TD target: the buffer.next chain fused with the discounted reward sum.
//
// Replaces BasePolicy.compute_nstep_return (tianshou/policy/base.py:386-440) and numba
// _nstep_return (:500-524).  The reference materialises an [n_step, bsz] index matrix with
// n_step-1 calls of buffer.next (:422-425, each O(buffer_num * bsz)), copies the whole `done`
// array to add the unfinished tails (:433-434, O(capacity) per update) and then loops
// n = n_step-1 .. 0 over it.  Here one thread walks the chain of its sample in registers.
//
// The reference's reverse loop
//     gammas[end[now]] = n + 1; returns[end[now]] = 0; returns = rew[now] + gamma * returns
// evaluated for n = n_step-1 .. 0 keeps only the steps up to the FIRST end flag met walking
// forward (every later contribution is zeroed when that flag is processed), as a Horner sum.
// Walking forward we therefore collect rewards until the first end flag (m steps), then run
// the same Horner recurrence backwards over those m rewards: identical operations in the
// identical order, so `ret_sum` is bit-identical to the reference's f64 `returns`.
// Once a step has end_flag set, next() returns the same index (manager.py:295-296), which is
// why the chain can stop there: `terminal` (= indices[-1], :428) is that index.
#include "ts_common.cuh"

namespace ts {
namespace {

constexpr int kMaxRegSteps = 16;  // chains up to this length keep their rewards in registers

template <typename RewT>
__global__ void __launch_bounds__(128)
nstep_chain_kernel(const int64_t* __restrict__ indice, int64_t bsz,
                   const RewT* __restrict__ rew, const uint8_t* __restrict__ done,
                   const uint8_t* __restrict__ terminated, const int64_t* __restrict__ offset,
                   int64_t B, int64_t sub_size, const int64_t* __restrict__ last_index,
                   const int64_t* __restrict__ lengths, double gamma, int n_step,
                   int64_t* __restrict__ terminal, double* __restrict__ ret_sum,
                   double* __restrict__ coef) {
    const int64_t total = __ldg(offset + B);
    for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < bsz;
         b += (int64_t)gridDim.x * blockDim.x) {
        int64_t now = indice[b];
        // step 0 is read at the caller's index as given (policy/base.py:422, :519);
        // negative indices address from the end like numpy fancy indexing does.
        if (now < 0) now += total;
        const int64_t k = sub_of(now, offset, B, sub_size);
        const int64_t start = __ldg(offset + k);
        int64_t cur_len = __ldg(lengths + k);
        cur_len = cur_len > 1 ? cur_len : 1;
        const int64_t last = __ldg(last_index + k);
        // unfinished tail of this sub-buffer (base.py:148-151): end_flag[unfinished] = True
        const bool unfinished_exists = __ldg(lengths + k) > 0 && !done[last];

        double r[kMaxRegSteps];
        int m = 0;         // number of rewards collected = gammas (:520)
        int64_t first = now;
        if (n_step <= kMaxRegSteps) {
            // A chain that has not ended moves one slot forward per step (a step at `last`
            // always ends it: either done[last] or the unfinished tail), so the n_step
            // candidate slots are known up front: fetch their flags and rewards in ONE round
            // of independent loads and resolve the chain in registers, instead of one
            // dependent L2 round trip per step.
            int64_t pos[kMaxRegSteps];
            uint8_t dn[kMaxRegSteps];
            int64_t rel = pymod(now - start + 1, cur_len);
#pragma unroll
            for (int n = 0; n < kMaxRegSteps; ++n) {
                if (n < n_step) {
                    pos[n] = n == 0 ? now : rel + start;
                    if (n > 0 && ++rel == cur_len) rel = 0;
                    dn[n] = done[pos[n]];
                    r[n] = (double)rew[pos[n]];
                }
            }
#pragma unroll
            for (int n = 0; n < kMaxRegSteps; ++n) {
                if (n < n_step && m == n) {  // the chain is still alive at step n
                    now = pos[n];
                    m = n + 1;
                    if (dn[n] != 0 || (unfinished_exists && now == last)) m = -m;  // ended
                }
            }
            if (m < 0) m = -m;
        } else {
#pragma unroll 1
            for (int n = 0; n < n_step; ++n) {
                const bool d = done[now] != 0;
                const bool end = d || (unfinished_exists && now == last);
                m = n + 1;
                if (end) break;
                if (n + 1 < n_step) {
                    // _next_index (manager.py:295-296); `now == last` with done[last] set was
                    // caught above, so here end_flag of next() is zero unless now == last.
                    const int64_t ef = (now == last);
                    now = pymod(now - start + 1 - ef, cur_len) + start;
                }
            }
        }
        // Horner, last collected reward first (:518-522 restricted to the surviving steps)
        double acc = 0.0;
        if (m <= kMaxRegSteps) {
#pragma unroll
            for (int n = kMaxRegSteps - 1; n >= 0; --n)
                if (n < m) acc = __dadd_rn(r[n], __dmul_rn(gamma, acc));
        } else {  // long chains: re-walk from the start for each step (rare: n_step > 16)
            for (int n = m - 1; n >= 0; --n) {
                int64_t p = first;
                for (int q = 0; q < n; ++q) {
                    const int64_t ef = (p == last) | (done[p] != 0);
                    p = pymod(p - start + 1 - ef, cur_len) + start;
                }
                acc = __dadd_rn((double)rew[p], __dmul_rn(gamma, acc));
            }
        }
        // gamma_buffer[gammas] (:509-511, :523): repeated multiplication, not pow()
        double g = 1.0;
        for (int q = 0; q < m; ++q) g = __dmul_rn(g, gamma);
        // the chain's terminal index: next() keeps returning `now` after an end flag
        terminal[b] = now;
        ret_sum[b] = acc;
        // value_mask(buffer, terminal) = ~terminated[terminal]  (:335, :432)
        coef[b] = terminated[now] ? 0.0 : g;
    }
}

// out[b, d] = target_q[b, d] * coef[b] + ret_sum[b]   (policy/base.py:432, :523)
// The reference multiplies target_q by the 0/1 mask first and by gamma^k second; with
// coef = mask * gamma^k the product differs only in the sign of a zero.
template <typename QT>
__global__ void __launch_bounds__(256)
nstep_combine_kernel(const QT* __restrict__ target_q, int64_t bsz, int64_t D,
                     const double* __restrict__ ret_sum, const double* __restrict__ coef,
                     QT* __restrict__ out) {
    const int64_t n = bsz * D;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = i / D;
        const double v = __dadd_rn(__dmul_rn((double)target_q[i], coef[b]), ret_sum[b]);
        out[i] = (QT)v;
    }
}

inline unsigned grid_for(int64_t n, int threads) {
    int64_t g = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * 16;
    return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace ts

using namespace ts;

extern "C" int32_t ts_nstep_chain(const int64_t* indice, int64_t bsz, const void* rew,
                                  int32_t rew_dtype, const uint8_t* done,
                                  const uint8_t* terminated, const int64_t* offset, int64_t B,
                                  int64_t sub_size, const int64_t* last_index,
                                  const int64_t* lengths, double gamma, int32_t n_step,
                                  int64_t* terminal, double* ret_sum, double* coef,
                                  ts_stream_t stream) {
    if (bsz < 0 || B <= 0 || n_step < 1) return TS_ERR_BAD_ARG;
    if (bsz == 0) return TS_OK;
    if (!indice || !rew || !done || !terminated || !offset || !last_index || !lengths ||
        !terminal || !ret_sum || !coef)
        return TS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (rew_dtype == TS_F32)
        nstep_chain_kernel<float><<<grid_for(bsz, 128), 128, 0, st>>>(
            indice, bsz, (const float*)rew, done, terminated, offset, B, sub_size, last_index,
            lengths, gamma, n_step, terminal, ret_sum, coef);
    else if (rew_dtype == TS_F64)
        nstep_chain_kernel<double><<<grid_for(bsz, 128), 128, 0, st>>>(
            indice, bsz, (const double*)rew, done, terminated, offset, B, sub_size, last_index,
            lengths, gamma, n_step, terminal, ret_sum, coef);
    else
        return TS_ERR_BAD_ARG;
    return check_launch();
}

extern "C" int32_t ts_nstep_combine(const void* target_q, int32_t q_dtype, int64_t bsz,
                                    int64_t D, const double* ret_sum, const double* coef,
                                    void* out, ts_stream_t stream) {
    if (bsz < 0 || D < 0) return TS_ERR_BAD_ARG;
    if (bsz == 0 || D == 0) return TS_OK;
    if (!target_q || !ret_sum || !coef || !out) return TS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (q_dtype == TS_F32)
        nstep_combine_kernel<float><<<grid_for(bsz * D, 256), 256, 0, st>>>(
            (const float*)target_q, bsz, D, ret_sum, coef, (float*)out);
    else if (q_dtype == TS_F64)
        nstep_combine_kernel<double><<<grid_for(bsz * D, 256), 256, 0, st>>>(
            (const double*)target_q, bsz, D, ret_sum, coef, (double*)out);
    else
        return TS_ERR_BAD_ARG;
    return check_launch();
}
