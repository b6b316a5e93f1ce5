/*
 * ts_oracle.c -- CPU restatement (plain C) of the reference's return / replay hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity oracle for the CUDA kernels in
 * tianshou_fork_b200/csrc.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call it.  The product path
 * never links or calls anything in oracle/ and has no CPU fallback.
 *
 * Parity status: PINNED.  oracle/oracle.py wraps these functions; tests/test_oracle.py
 * checks them against (a) the reference's own known-answer vectors re-typed into the
 * tests (test/base/test_returns.py, test/base/test_buffer.py) and (b) fixtures under
 * tests/golden/ produced by running the imported reference (numba 0.65 / numpy 2.3.5)
 * through tests/golden/make_golden.py.
 *
 * Every function cites the reference lines it restates (paths relative to the
 * reference checkout).  The arithmetic order is the reference's: IEEE f64, one
 * operation per source operation, compiled with -ffp-contract=off so that no FMA
 * is formed (numba/LLVM does not contract by default either).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* Python-style modulo for a positive modulus (numpy `%` on int64). */
static inline int64_t pymod(int64_t a, int64_t m) {
    int64_t r = a % m;
    return r < 0 ? r + m : r;
}

/* ------------------------------------------------------------------------------------
 * GAE advantages: tianshou/policy/base.py:453-497 (_gae_return).
 *   delta    = rew + v_s_ * gamma - v_s
 *   discount = (1 - end_flag) * (gamma * gae_lambda)
 *   gae      = delta[i] + discount[i] * gae      for i = n-1 .. 0
 * One serial reverse loop over the whole flattened batch.
 * ---------------------------------------------------------------------------------- */
void orc_gae_return(const double* v_s, const double* v_s_, const double* rew,
                    const uint8_t* end_flag, double gamma, double gae_lambda,
                    int64_t n, double* out) {
    const double gl = gamma * gae_lambda;
    double gae = 0.0;
    for (int64_t i = n - 1; i >= 0; --i) {
        double delta = rew[i] + v_s_[i] * gamma - v_s[i];
        double discount = (1.0 - (double)(end_flag[i] != 0)) * gl;
        gae = delta + discount * gae;
        out[i] = gae;
    }
}

/* The (f32, f32, f64, bool) signature that A2C/PPO actually hit
 * (tianshou/policy/base.py:449; a2c.py:91-93 hands f32 critic outputs). */
void orc_gae_return_f32(const float* v_s, const float* v_s_, const double* rew,
                        const uint8_t* end_flag, double gamma, double gae_lambda,
                        int64_t n, double* out) {
    const double gl = gamma * gae_lambda;
    double gae = 0.0;
    for (int64_t i = n - 1; i >= 0; --i) {
        double delta = rew[i] + (double)v_s_[i] * gamma - (double)v_s[i];
        double discount = (1.0 - (double)(end_flag[i] != 0)) * gl;
        gae = delta + discount * gae;
        out[i] = gae;
    }
}

/* ------------------------------------------------------------------------------------
 * compute_episodic_return on an [E, T] row-major rollout whose rows are independent
 * sub-buffer streams (tianshou/policy/base.py:337-384 applied to buffer.sample(0) of an
 * un-wrapped VectorReplayBuffer, manager.py:187-192):
 *   v_s_   <- v_s_ * ~terminated                         (:375-376, value_mask :335)
 *   end    <- terminated | truncated; last step of each row forced True (:379-380;
 *             the row tail is either done or the unfinished index, manager.py:68-74)
 *   adv    <- _gae_return                                (:381)
 *   ret    <- adv + v_s                                  (:382)
 * f32 inputs, f64 arithmetic, f64 outputs.  Rows are split over `threads` OpenMP
 * threads: the reference itself is serial; this is its best case on a many-core host
 * (one reference process per disjoint slice of env streams).
 * ---------------------------------------------------------------------------------- */
void orc_episodic_return_rows(const float* rew, const float* v_s, const float* v_s_,
                              const uint8_t* terminated, const uint8_t* truncated,
                              int64_t E, int64_t T, double gamma, double gae_lambda,
                              double* ret, double* adv, int threads) {
    const double gl = gamma * gae_lambda;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
#endif
    for (int64_t e = 0; e < E; ++e) {
        const int64_t base = e * T;
        double gae = 0.0;
        for (int64_t t = T - 1; t >= 0; --t) {
            const int64_t i = base + t;
            const int term = terminated[i] != 0;
            const int end = term | (truncated[i] != 0) | (t == T - 1);
            double vnext = (double)v_s_[i] * (double)(!term);
            double delta = (double)rew[i] + vnext * gamma - (double)v_s[i];
            double discount = (1.0 - (double)end) * gl;
            gae = delta + discount * gae;
            adv[i] = gae;
            ret[i] = gae + (double)v_s[i];
        }
    }
    (void)threads;
}

/* ------------------------------------------------------------------------------------
 * n-step return: tianshou/policy/base.py:500-524 (_nstep_return).
 *   indices is [n_step, bsz] row-major; target_q is [bsz, D]; returns [bsz, D] f64.
 * ---------------------------------------------------------------------------------- */
void orc_nstep_return(const double* rew, const uint8_t* end_flag, const double* target_q,
                      const int64_t* indices, double gamma, int64_t n_step, int64_t bsz,
                      int64_t D, double* out) {
    double* gamma_buffer = (double*)malloc(sizeof(double) * (size_t)(n_step + 1));
    gamma_buffer[0] = 1.0;
    for (int64_t i = 1; i <= n_step; ++i) gamma_buffer[i] = gamma_buffer[i - 1] * gamma;
    for (int64_t b = 0; b < bsz; ++b) {
        int64_t gammas = n_step;
        double returns = 0.0; /* identical across the D columns (rew broadcast :522) */
        for (int64_t n = n_step - 1; n >= 0; --n) {
            int64_t now = indices[n * bsz + b];
            if (end_flag[now]) {
                gammas = n + 1;
                returns = 0.0;
            }
            returns = rew[now] + gamma * returns;
        }
        for (int64_t d = 0; d < D; ++d)
            out[b * D + d] = target_q[b * D + d] * gamma_buffer[gammas] + returns;
    }
    free(gamma_buffer);
}

/* ------------------------------------------------------------------------------------
 * Manager prev / next: tianshou/data/buffer/manager.py:259-297.
 *   offset has B+1 entries (_extend_offset); index is reduced mod offset[B] first.
 * The reference loops over sub-buffers with a mask; per element the result is the same
 * as locating the element's sub-buffer, done here by linear search.
 * ---------------------------------------------------------------------------------- */
static inline int64_t find_sub(const int64_t* offset, int64_t B, int64_t idx) {
    for (int64_t k = 0; k < B; ++k)
        if (offset[k] <= idx && idx < offset[k + 1]) return k;
    return -1;
}

void orc_prev_index(const int64_t* index, int64_t m, const int64_t* offset, int64_t B,
                    const uint8_t* done, const int64_t* last_index, const int64_t* lengths,
                    int64_t* out) {
    for (int64_t j = 0; j < m; ++j) {
        int64_t idx = pymod(index[j], offset[B]);
        int64_t k = find_sub(offset, B, idx);
        if (k < 0) { out[j] = 0; continue; } /* zeros_like default, manager.py:268 */
        int64_t start = offset[k];
        int64_t cur_len = lengths[k] > 1 ? lengths[k] : 1;
        int64_t sub = pymod(idx - start - 1, cur_len);
        int64_t end_flag = (done[sub + start] != 0) | (sub + start == last_index[k]);
        out[j] = pymod(sub + end_flag, cur_len) + start;
    }
}

void orc_next_index(const int64_t* index, int64_t m, const int64_t* offset, int64_t B,
                    const uint8_t* done, const int64_t* last_index, const int64_t* lengths,
                    int64_t* out) {
    for (int64_t j = 0; j < m; ++j) {
        int64_t idx = pymod(index[j], offset[B]);
        int64_t k = find_sub(offset, B, idx);
        if (k < 0) { out[j] = 0; continue; }
        int64_t start = offset[k];
        int64_t cur_len = lengths[k] > 1 ? lengths[k] : 1;
        int64_t end_flag = (done[idx] != 0) | (idx == last_index[k]);
        out[j] = pymod(idx - start + 1 - end_flag, cur_len) + start;
    }
}

/* Single-ring prev / next: tianshou/data/buffer/base.py:153-168 (modulo by current size). */
void orc_prev_single(const int64_t* index, int64_t m, const uint8_t* done, int64_t last_index,
                     int64_t size, int64_t* out) {
    for (int64_t j = 0; j < m; ++j) {
        int64_t i = pymod(index[j] - 1, size);
        int64_t end_flag = (done[i] != 0) | (i == last_index);
        out[j] = pymod(i + end_flag, size);
    }
}

void orc_next_single(const int64_t* index, int64_t m, const uint8_t* done, int64_t last_index,
                     int64_t size, int64_t* out) {
    for (int64_t j = 0; j < m; ++j) {
        int64_t i = index[j];
        int64_t end_flag = (done[i] != 0) | (i == last_index);
        out[j] = pymod(i + (1 - end_flag), size);
    }
}

/* ------------------------------------------------------------------------------------
 * Segment tree: tianshou/data/utils/segtree.py.
 *   tree has 2*bound f64 nodes, leaf i at bound+i, node i = node 2i + node 2i+1.
 * ---------------------------------------------------------------------------------- */

/* _setitem (:98-104).  `index` already carries +bound and is CONSUMED (halved in place)
 * exactly as the reference mutates its temporary.  Sequential assignment => the last
 * occurrence of a duplicated index wins (:40-49). */
void orc_segtree_setitem(double* tree, int64_t* index, const double* value, int64_t m) {
    if (m <= 0) return;
    for (int64_t j = 0; j < m; ++j) tree[index[j]] = value[j];
    while (index[0] > 1) {
        for (int64_t j = 0; j < m; ++j) index[j] /= 2;
        for (int64_t j = 0; j < m; ++j)
            tree[index[j]] = tree[index[j] * 2] + tree[index[j] * 2 + 1];
    }
}

/* _reduce (:107-119): sum of nodes strictly inside (start, end), heap coordinates. */
double orc_segtree_reduce(const double* tree, int64_t start, int64_t end) {
    double result = 0.0;
    while (end - start > 1) {
        if (start % 2 == 0) result += tree[start + 1];
        start /= 2;
        if (end % 2 == 1) result += tree[end - 1];
        end /= 2;
    }
    return result;
}

/* _get_prefix_sum_idx (:122-137).  `value` is mutated in place like the reference
 * (:134).  Strict `<`; `value -= lsons * direct` keeps the multiply-by-{0,1} form. */
void orc_segtree_prefix_idx(double* value, int64_t m, int64_t bound, const double* sums,
                            int64_t* out) {
    for (int64_t j = 0; j < m; ++j) {
        int64_t index = 1;
        double v = value[j];
        while (index < bound) {
            index *= 2;
            double lsons = sums[index];
            int direct = lsons < v;
            v -= lsons * (double)direct;
            index += direct;
        }
        value[j] = v;
        out[j] = index - bound;
    }
}

/* f32-valued variant (segtree.py:95): compare/subtract happen in f64 after promotion of
 * the f32 value, and the running value is stored back as f32 each level. */
void orc_segtree_prefix_idx_f32(float* value, int64_t m, int64_t bound, const double* sums,
                                int64_t* out) {
    for (int64_t j = 0; j < m; ++j) {
        int64_t index = 1;
        float v = value[j];
        while (index < bound) {
            index *= 2;
            double lsons = sums[index];
            int direct = lsons < (double)v;
            v = (float)((double)v - lsons * (double)direct);
            index += direct;
        }
        value[j] = v;
        out[j] = index - bound;
    }
}

/* ------------------------------------------------------------------------------------
 * One PER index step on one tree, as BasePolicy.update drives it
 * (tianshou/policy/base.py:307-311 -> prio.py:59-64, 66-76, 78-89, 91-102):
 *   scalar = u * tree[1]; idx = get_prefix_sum_idx(scalar)
 *   w      = (tree[idx] / min_prio) ** (-beta);  w /= max(w) if weight_norm
 *   p      = |td| + eps (f32);  tree[idx] = p ** alpha (f32 pow);  max/min prio update
 * Used as the timed CPU baseline for PER samples/s.  `scratch_idx` is bsz int64.
 * ---------------------------------------------------------------------------------- */
void orc_per_step(double* tree, int64_t bound, const double* u, const float* td, int64_t bsz,
                  double alpha, double beta, int weight_norm, double* max_prio,
                  double* min_prio, int64_t* idx_out, double* weight_out, double* scratch_val,
                  int64_t* scratch_idx) {
    const float eps = 1.1920928955078125e-07f; /* np.finfo(np.float32).eps, prio.py:38 */
    const double total = tree[1];
    for (int64_t j = 0; j < bsz; ++j) scratch_val[j] = u[j] * total;
    orc_segtree_prefix_idx(scratch_val, bsz, bound, tree, idx_out);
    double wmax = -INFINITY;
    for (int64_t j = 0; j < bsz; ++j) {
        double w = pow(tree[bound + idx_out[j]] / *min_prio, -beta);
        weight_out[j] = w;
        if (w > wmax) wmax = w;
    }
    if (weight_norm)
        for (int64_t j = 0; j < bsz; ++j) weight_out[j] /= wmax;
    float pmax = -INFINITY, pmin = INFINITY;
    for (int64_t j = 0; j < bsz; ++j) {
        float p = fabsf(td[j]) + eps;
        if (p > pmax) pmax = p;
        if (p < pmin) pmin = p;
        scratch_val[j] = (double)powf(p, (float)alpha);
        scratch_idx[j] = idx_out[j] + bound;
    }
    orc_segtree_setitem(tree, scratch_idx, scratch_val, bsz);
    if ((double)pmax > *max_prio) *max_prio = (double)pmax;
    if ((double)pmin < *min_prio) *min_prio = (double)pmin;
}

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
