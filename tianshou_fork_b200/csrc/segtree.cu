// Segment tree (sum tree) for prioritized replay: batched bottom-up update, interval reduce,
// batched prefix-sum descent, and the fused PER sample / update steps.
//
// Replaces numba _setitem / _reduce / _get_prefix_sum_idx
// (tianshou/data/utils/segtree.py:98-137) and the NumPy arithmetic of
// PrioritizedReplayBuffer (tianshou/data/buffer/prio.py:59-102).
//
// Bit-exactness.  The reference recomputes every touched ancestor as tree[2i] + tree[2i+1]
// in IEEE f64, so at all times node == fl(left + right).  Any schedule that ends with every
// touched ancestor recomputed from its FINAL children reproduces the same bits; untouched
// nodes may be recomputed too (same inputs, same result).  All adds below are __dadd_rn and
// the descent compares/subtracts without FMA, so node values and sampled indices are
// bit-identical to the reference given identical leaves and identical query scalars.
//
// Update schedule (M <= kSmallM, the replay-step case).  The leaves are split into S = 128
// aligned subtrees; CTA s owns subtree s: it scans the whole index batch (L2-resident, 32 KB
// for 4096 indices), keeps the items that fall into its subtree in shared memory, resolves
// duplicated indices ("last occurrence wins", segtree.py:40-49) with an atomicMax stamp per
// leaf, and climbs its subtree level by level with __syncthreads only -- no CTA ever reads a
// node another CTA writes.  The last CTA to finish (arrival counter) rebuilds the top
// log2(S) levels from the S subtree roots in shared memory.  Large batches (bulk loads) take
// the stamp -> leaf-write -> dense level-by-level rebuild path instead.
#include "ts_common.cuh"

#include <math.h>

namespace ts {
namespace {

// ---- where leaf values come from ------------------------------------------------------------
enum ValueMode : int {
    VM_F64 = 0,       // value[j * stride] f64                      (SegmentTree.__setitem__)
    VM_F32 = 1,       // value[j * stride] f32, cast on store       (segtree.py:91-92)
    VM_PER_F32 = 2,   // td f32: p = |td| + eps (f32), leaf = f32(p^alpha)   (prio.py:86-87)
    VM_PER_F64 = 3,   // td f64: p = |td| + eps,        leaf = p^alpha
    VM_INIT_F64 = 4,  // leaf = max_prio^alpha                      (prio.py:43)
    VM_INIT_F32 = 5,  // same, rounded to f32
};

struct ValueSrc {
    const void* ptr;
    int64_t stride;
    double alpha;
    double* prio;  // {max_prio, min_prio} on device
    int mode;
};

constexpr float kEpsF32 = 1.1920928955078125e-07f;  // np.finfo(np.float32).eps (prio.py:38)

// priority before the exponent, promoted to f64 (what _max_prio / _min_prio track, :88-89)
__device__ __forceinline__ double per_priority(const ValueSrc& v, int64_t j) {
    if (v.mode == VM_PER_F32) return (double)(fabsf(((const float*)v.ptr)[j]) + kEpsF32);
    return fabs(((const double*)v.ptr)[j]) + (double)kEpsF32;
}

__device__ __forceinline__ double leaf_value(const ValueSrc& v, int64_t j) {
    switch (v.mode) {
        case VM_F64: return ((const double*)v.ptr)[j * v.stride];
        case VM_F32: return (double)((const float*)v.ptr)[j * v.stride];
        case VM_PER_F32: return (double)(float)pow(per_priority(v, j), v.alpha);
        case VM_PER_F64: return pow(per_priority(v, j), v.alpha);
        case VM_INIT_F64: return pow(v.prio[0], v.alpha);
        default: return (double)(float)pow(v.prio[0], v.alpha);
    }
}

__device__ __forceinline__ bool is_per(const ValueSrc& v) {
    return v.mode == VM_PER_F32 || v.mode == VM_PER_F64;
}

// positive finite doubles order like their bit patterns
__device__ __forceinline__ void atomic_max_pos(double* addr, double v) {
    atomicMax((unsigned long long*)addr, (unsigned long long)__double_as_longlong(v));
}
__device__ __forceinline__ void atomic_min_pos(double* addr, double v) {
    atomicMin((unsigned long long*)addr, (unsigned long long)__double_as_longlong(v));
}

constexpr int UPD_NT = 128;
constexpr int UPD_S = 128;             // subtrees = CTAs of the small-batch update
constexpr int64_t kSmallM = 8192;      // item list of one CTA: M * 8 bytes of shared memory
constexpr int64_t kDenseBound = 4096;  // trees up to this many leaves are rebuilt densely

// ============================================================================================
// Small-batch update: one CTA per aligned subtree.
// ============================================================================================
__global__ void __launch_bounds__(UPD_NT)
segtree_update_small(double* __restrict__ tree, int64_t size, int64_t bound,
                     const int64_t* __restrict__ index, ValueSrc vs, int64_t M,
                     uint32_t* __restrict__ stamp, uint32_t* __restrict__ counter,
                     int32_t* __restrict__ err_flag, int shift, int S) {
    extern __shared__ uint2 items[];  // (heap node, batch position)
    __shared__ int n_items;
    __shared__ bool s_last;
    __shared__ double s_top[2 * UPD_S];
    __shared__ double s_red[2 * (UPD_NT / 32)];

    const int tid = threadIdx.x;
    const int64_t sub = blockIdx.x;
    if (tid == 0) n_items = 0;
    __syncthreads();

    // ---- scan the batch: claim items of this subtree, stamp their leaves ----------------------
    double pmax = 0.0, pmin = INFINITY;
    const bool reduce_prio = is_per(vs) && sub == 0;
    for (int64_t j0 = 0; j0 < M; j0 += UPD_NT * 4) {
        int64_t idx[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int64_t j = j0 + q * UPD_NT + tid;
            idx[q] = j < M ? index[j] : -1;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int64_t j = j0 + q * UPD_NT + tid;
            if (j >= M) continue;
            if (reduce_prio) {
                const double p = per_priority(vs, j);
                pmax = fmax(pmax, p);
                pmin = fmin(pmin, p);
            }
            if (idx[q] < 0 || idx[q] >= size) {
                if (sub == 0) atomicOr(err_flag, TS_FLAG_INDEX_RANGE);
                continue;
            }
            if ((idx[q] >> shift) == sub) {
                const int pos = atomicAdd(&n_items, 1);
                items[pos] = make_uint2((uint32_t)(idx[q] + bound), (uint32_t)j);
                atomicMax(stamp + idx[q], (uint32_t)j + 1u);
            }
        }
    }
    __threadfence();
    __syncthreads();
    const int n = n_items;

    // ---- leaves: only the last occurrence of an index writes ---------------------------------
    for (int i = tid; i < n; i += UPD_NT) {
        const uint2 it = items[i];
        const int64_t leaf = (int64_t)it.x - bound;
        if (__ldcg(stamp + leaf) == it.y + 1u) {
            tree[it.x] = leaf_value(vs, it.y);
            stamp[leaf] = 0u;  // losers compare against their own position: 0 never matches
        }
    }
    __syncthreads();

    // ---- climb the subtree ---------------------------------------------------------------------
    for (int lvl = 0; lvl < shift; ++lvl) {
        for (int i = tid; i < n; i += UPD_NT) {
            const uint32_t node = items[i].x >> 1;
            items[i].x = node;
            tree[node] = __dadd_rn(__ldcg(tree + 2 * (size_t)node),
                                   __ldcg(tree + 2 * (size_t)node + 1));  // segtree.py:104
        }
        __syncthreads();
    }

    // ---- priority extrema (prio.py:88-89), CTA 0 only -------------------------------------------
    if (reduce_prio) {
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            pmax = fmax(pmax, __shfl_xor_sync(0xffffffffu, pmax, d));
            pmin = fmin(pmin, __shfl_xor_sync(0xffffffffu, pmin, d));
        }
        if ((tid & 31) == 0) { s_red[tid >> 5] = pmax; s_red[UPD_NT / 32 + (tid >> 5)] = pmin; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < UPD_NT / 32; ++w) {
                pmax = fmax(pmax, s_red[w]);
                pmin = fmin(pmin, s_red[UPD_NT / 32 + w]);
            }
            vs.prio[0] = fmax(vs.prio[0], pmax);
            vs.prio[1] = fmin(vs.prio[1], pmin);
        }
    }

    // ---- top log2(S) levels: last CTA to arrive ------------------------------------------------
    if (S > 1) {
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = (atomicAdd(counter, 1u) == (uint32_t)(S - 1));
        __syncthreads();
        if (s_last) {
            __threadfence();
            for (int i = tid; i < S; i += UPD_NT) s_top[S + i] = __ldcg(tree + S + i);
            __syncthreads();
            for (int w = S >> 1; w >= 1; w >>= 1) {
                for (int i = tid; i < w; i += UPD_NT) {
                    const int node = w + i;
                    const double v = __dadd_rn(s_top[2 * node], s_top[2 * node + 1]);
                    s_top[node] = v;
                    tree[node] = v;
                }
                __syncthreads();
            }
            if (tid == 0) *counter = 0u;
        }
    }
}

// ============================================================================================
// Bulk path: stamp, leaf write, dense rebuild.
// ============================================================================================
__global__ void __launch_bounds__(256)
segtree_stamp_kernel(const int64_t* __restrict__ index, int64_t M, int64_t size,
                     uint32_t* __restrict__ stamp, int32_t* __restrict__ err_flag) {
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M;
         j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t idx = index[j];
        if (idx < 0 || idx >= size) { atomicOr(err_flag, TS_FLAG_INDEX_RANGE); continue; }
        atomicMax(stamp + idx, (uint32_t)j + 1u);
    }
}

__global__ void __launch_bounds__(256)
segtree_leaf_kernel(double* __restrict__ tree, int64_t size, int64_t bound,
                    const int64_t* __restrict__ index, ValueSrc vs, int64_t M,
                    uint32_t* __restrict__ stamp) {
    double pmax = 0.0, pmin = INFINITY;
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M;
         j += (int64_t)gridDim.x * blockDim.x) {
        if (is_per(vs)) {
            const double p = per_priority(vs, j);
            pmax = fmax(pmax, p);
            pmin = fmin(pmin, p);
        }
        const int64_t idx = index[j];
        if (idx < 0 || idx >= size) continue;
        if (stamp[idx] == (uint32_t)j + 1u) {
            tree[bound + idx] = leaf_value(vs, j);
            stamp[idx] = 0u;
        }
    }
    if (is_per(vs)) {
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            pmax = fmax(pmax, __shfl_xor_sync(0xffffffffu, pmax, d));
            pmin = fmin(pmin, __shfl_xor_sync(0xffffffffu, pmin, d));
        }
        if ((threadIdx.x & 31) == 0 && pmax > 0.0) {
            atomic_max_pos(vs.prio, pmax);
            atomic_min_pos(vs.prio + 1, pmin);
        }
    }
}

// Rebuilds, for every aligned group of `chunk` nodes of the level that starts at heap index
// `level_base`, all ancestors up to the group's root.  chunk is a power of two <= 2048.
constexpr int REB_NT = 256;
constexpr int REB_CHUNK = 2048;
__global__ void __launch_bounds__(REB_NT)
segtree_rebuild_kernel(double* __restrict__ tree, int64_t level_base, int chunk) {
    __shared__ double s[2 * REB_CHUNK];  // heap layout: s[chunk + i] = node i of the group
    const int tid = threadIdx.x;
    const int64_t first = level_base + (int64_t)blockIdx.x * chunk;
    for (int i = tid; i < chunk; i += REB_NT) s[chunk + i] = __ldcg(tree + first + i);
    __syncthreads();
    int64_t base = first;
    for (int w = chunk >> 1; w >= 1; w >>= 1) {
        base >>= 1;  // heap index of the first node of this group's next level up
        for (int i = tid; i < w; i += REB_NT) {
            const double v = __dadd_rn(s[2 * (w + i)], s[2 * (w + i) + 1]);
            s[w + i] = v;
            tree[base + i] = v;
        }
        __syncthreads();
    }
}

inline unsigned grid_for(int64_t n, int threads) {
    int64_t g = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * 16;
    return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

int32_t dense_rebuild(double* tree, int64_t bound, cudaStream_t st) {
    int64_t cur = bound;  // number of nodes in the current "leaf" level (heap base == cur)
    while (cur > 1) {
        const int chunk = (int)(cur < REB_CHUNK ? cur : REB_CHUNK);
        const int64_t groups = cur / chunk;
        segtree_rebuild_kernel<<<(unsigned)groups, REB_NT, 0, st>>>(tree, cur, chunk);
        cur = groups;
    }
    return check_launch();
}

int32_t update_impl(double* tree, int64_t size, int64_t bound, const int64_t* index,
                    ValueSrc vs, int64_t M, uint32_t* stamp, uint32_t* counter,
                    int32_t* err_flag, cudaStream_t st) {
    if (M < 0 || size <= 0 || bound < size || (bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (M >= 0x7fffffffLL || bound > (1LL << 30)) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!tree || !index || !stamp || !counter || !err_flag) return TS_ERR_BAD_ARG;

    if (M <= kSmallM && bound > kDenseBound) {
        int lg = 0;
        while ((1LL << lg) < bound) ++lg;
        int lgS = 0;
        while ((1 << lgS) < UPD_S) ++lgS;
        const int shift = lg - lgS;
        static bool attr_set = false;
        if (!attr_set) {
            cudaFuncSetAttribute(segtree_update_small,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(kSmallM * sizeof(uint2)));
            attr_set = true;
        }
        segtree_update_small<<<UPD_S, UPD_NT, (size_t)M * sizeof(uint2), st>>>(
            tree, size, bound, index, vs, M, stamp, counter, err_flag, shift, UPD_S);
        return check_launch();
    }
    // bulk or tiny-tree path
    segtree_stamp_kernel<<<grid_for(M, 256), 256, 0, st>>>(index, M, size, stamp, err_flag);
    segtree_leaf_kernel<<<grid_for(M, 256), 256, 0, st>>>(tree, size, bound, index, vs, M, stamp);
    if (bound == 1) {  // single leaf: heap node 1 is the leaf itself, node 0 unused
        return check_launch();
    }
    return dense_rebuild(tree, bound, st);
}

// ============================================================================================
// reduce(start, end): the reference's climb (segtree.py:107-119), one thread.
// ============================================================================================
__global__ void segtree_reduce_kernel(const double* __restrict__ tree, int64_t start,
                                      int64_t end, double* __restrict__ out) {
    double result = 0.0;
    while (end - start > 1) {
        if ((start & 1) == 0) result = __dadd_rn(result, tree[start + 1]);
        start >>= 1;  // start >= 0 here: floor division == shift
        if ((end & 1) == 1) result = __dadd_rn(result, tree[end - 1]);
        end >>= 1;
    }
    *out = result;
}

// ============================================================================================
// Prefix-sum descent (segtree.py:122-137).  One query per thread; the top ten levels
// of the heap are staged in shared memory once per CTA (8 KB), the rest comes from L2 (a
// 2^20-leaf tree is 16 MiB, resident in the 126 MB L2).
// ============================================================================================
constexpr int DESC_NT = 64;
constexpr int kTopCached = 1024;  // heap indices [0, 1024) = levels 0..9, 8 KB per CTA

__device__ __forceinline__ int stage_top(const double* __restrict__ tree, int64_t bound,
                                         double* s_top, int tid) {
    const int cached = (int)(2 * bound < kTopCached ? 2 * bound : kTopCached);
    for (int i = tid; i < cached; i += DESC_NT) s_top[i] = __ldcg(tree + i);
    __syncthreads();
    return cached;  // heap indices below this value are in shared memory
}

// Descent for one query; `v` is updated like the reference mutates `value` (:134).
template <typename VT>
__device__ __forceinline__ int64_t descend(const double* __restrict__ tree, int64_t bound,
                                           const double* s_top, int cached, VT& v) {
    int64_t index = 1;
    while (index < bound) {
        index <<= 1;                                            // :131
        const double lsons = index < cached ? s_top[index] : __ldg(tree + index);  // :132
        const bool direct = lsons < (double)v;                  // :133 (strict)
        if (direct) v = (VT)__dsub_rn((double)v, lsons);        // :134 (v - lsons*1; else v - 0)
        index += direct;                                        // :135
    }
    return index - bound;                                       // :136
}

template <typename VT>
__global__ void __launch_bounds__(DESC_NT)
segtree_prefix_kernel(const double* __restrict__ tree, int64_t bound, VT* __restrict__ value,
                      int64_t M, int scale_by_total, int64_t* __restrict__ out,
                      int32_t* __restrict__ err_flag) {
    __shared__ double s_top[kTopCached];
    const int tid = threadIdx.x;
    const int cached = stage_top(tree, bound, s_top, tid);
    const double total = s_top[1];
    const int64_t j = blockIdx.x * (int64_t)DESC_NT + tid;
    if (j >= M) return;
    VT v = value[j];
    if (scale_by_total) v = (VT)__dmul_rn((double)v, total);     // prio.py:61
    if (!((double)v >= 0.0 && (double)v < total)) atomicOr(err_flag, TS_FLAG_PREFIX_RANGE);  // :79
    out[j] = descend<VT>(tree, bound, s_top, cached, v);
    value[j] = v;
}

// ============================================================================================
// Fused PER sample: scalar = u * total, descent, importance weight, batch-max normalisation.
// prio.py:59-64 (sample_indices), :66-76 (get_weight), :101 (weight / max(weight)).
// ============================================================================================
struct PerScratch {  // 16 bytes, all-zero between calls
    unsigned long long wmax_bits;
    unsigned int arrived;
    unsigned int pad;
};

__device__ __forceinline__ void finish_weights(double w, bool active, double* __restrict__ weight,
                                               int64_t M, int weight_norm, PerScratch* sc) {
    // CTA-level max -> one atomic per CTA -> last CTA normalises all M weights
    __shared__ double s_w[DESC_NT / 32];
    __shared__ bool s_last;
    if (!weight_norm) return;
    double m = active ? w : 0.0;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, d));
    const int tid = threadIdx.x;
    if ((tid & 31) == 0) s_w[tid >> 5] = m;
    __syncthreads();
    if (tid == 0) {
        for (int q = 1; q < DESC_NT / 32; ++q) m = fmax(m, s_w[q]);
        atomicMax(&sc->wmax_bits, (unsigned long long)__double_as_longlong(m));
        __threadfence();
        s_last = (atomicAdd(&sc->arrived, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        const double wmax = __longlong_as_double((long long)__ldcg(&sc->wmax_bits));
        for (int64_t j = tid; j < M; j += DESC_NT)
            weight[j] = __ddiv_rn(__ldcg(weight + j), wmax);     // prio.py:101
        __syncthreads();
        if (tid == 0) { sc->wmax_bits = 0ull; sc->arrived = 0u; }
    }
}

__device__ __forceinline__ double is_weight(double leaf, double min_prio, double beta) {
    return pow(__ddiv_rn(leaf, min_prio), -beta);                // prio.py:76
}

__global__ void __launch_bounds__(DESC_NT)
per_sample_kernel(const double* __restrict__ tree, int64_t bound, const double* __restrict__ u,
                  int64_t M, const double* __restrict__ prio, double beta, int weight_norm,
                  int64_t* __restrict__ idx, double* __restrict__ weight, PerScratch* sc,
                  int32_t* __restrict__ err_flag) {
    __shared__ double s_top[kTopCached];
    const int tid = threadIdx.x;
    const int cached = stage_top(tree, bound, s_top, tid);
    const double total = s_top[1];
    const int64_t j = blockIdx.x * (int64_t)DESC_NT + tid;
    const bool active = j < M;
    double w = 0.0;
    if (active) {
        double v = __dmul_rn(u[j], total);                       // prio.py:61
        if (!(v >= 0.0 && v < total)) atomicOr(err_flag, TS_FLAG_PREFIX_RANGE);
        const int64_t leaf = descend<double>(tree, bound, s_top, cached, v);
        idx[j] = leaf;
        w = is_weight(__ldg(tree + bound + leaf), prio[1], beta);
        weight[j] = w;
    }
    if (weight_norm) {
        __threadfence();
        finish_weights(w, active, weight, M, weight_norm, sc);
    }
}

__global__ void __launch_bounds__(DESC_NT)
per_weight_kernel(const double* __restrict__ tree, int64_t bound, const int64_t* __restrict__ idx,
                  int64_t M, const double* __restrict__ prio, double beta, int weight_norm,
                  double* __restrict__ weight, PerScratch* sc) {
    const int64_t j = blockIdx.x * (int64_t)DESC_NT + threadIdx.x;
    const bool active = j < M;
    double w = 0.0;
    if (active) {
        w = is_weight(__ldg(tree + bound + idx[j]), prio[1], beta);
        weight[j] = w;
    }
    if (weight_norm) {
        __threadfence();
        finish_weights(w, active, weight, M, weight_norm, sc);
    }
}

}  // namespace
}  // namespace ts

using namespace ts;

extern "C" int32_t ts_segtree_update(double* tree, int64_t size, int64_t bound,
                                     const int64_t* index, const void* value,
                                     int32_t value_dtype, int64_t value_stride, int64_t M,
                                     uint32_t* stamp, uint32_t* counter, int32_t* err_flag,
                                     ts_stream_t stream) {
    if (!value || value_stride < 0) return TS_ERR_BAD_ARG;
    if (value_dtype != TS_F32 && value_dtype != TS_F64) return TS_ERR_BAD_ARG;
    ValueSrc vs{value, value_stride, 0.0, nullptr, value_dtype == TS_F64 ? VM_F64 : VM_F32};
    return update_impl(tree, size, bound, index, vs, M, stamp, counter, err_flag,
                       (cudaStream_t)stream);
}

extern "C" int32_t ts_segtree_reduce(const double* tree, int64_t bound, int64_t start,
                                     int64_t end, double* out, ts_stream_t stream) {
    if (!tree || !out || bound <= 0 || start < 0 || end < start) return TS_ERR_BAD_ARG;
    // heap coordinates of the open interval (segtree.py:64)
    segtree_reduce_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(tree, start + bound - 1, end + bound,
                                                             out);
    return check_launch();
}

extern "C" int32_t ts_segtree_prefix_idx(const double* tree, int64_t bound, void* value,
                                         int32_t value_dtype, int64_t M, int32_t scale_by_total,
                                         int64_t* out, int32_t* err_flag, ts_stream_t stream) {
    if (M < 0 || bound <= 0) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!tree || !value || !out || !err_flag) return TS_ERR_BAD_ARG;
    const unsigned grid = (unsigned)((M + DESC_NT - 1) / DESC_NT);
    cudaStream_t st = (cudaStream_t)stream;
    if (value_dtype == TS_F64)
        segtree_prefix_kernel<double><<<grid, DESC_NT, 0, st>>>(tree, bound, (double*)value, M,
                                                               scale_by_total, out, err_flag);
    else if (value_dtype == TS_F32 && !scale_by_total)
        segtree_prefix_kernel<float><<<grid, DESC_NT, 0, st>>>(tree, bound, (float*)value, M, 0,
                                                              out, err_flag);
    else
        return TS_ERR_BAD_ARG;
    return check_launch();
}

extern "C" int32_t ts_per_sample(const double* tree, int64_t bound, const double* u, int64_t M,
                                 const double* prio, double beta, int32_t weight_norm,
                                 int64_t* idx, double* weight, void* scratch16,
                                 int32_t* err_flag, ts_stream_t stream) {
    if (M < 0 || bound <= 0) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!tree || !u || !prio || !idx || !weight || !scratch16 || !err_flag) return TS_ERR_BAD_ARG;
    const unsigned grid = (unsigned)((M + DESC_NT - 1) / DESC_NT);
    per_sample_kernel<<<grid, DESC_NT, 0, (cudaStream_t)stream>>>(
        tree, bound, u, M, prio, beta, weight_norm, idx, weight, (PerScratch*)scratch16, err_flag);
    return check_launch();
}

extern "C" int32_t ts_per_weight(const double* tree, int64_t bound, const int64_t* idx, int64_t M,
                                 const double* prio, double beta, int32_t weight_norm,
                                 double* weight, void* scratch16, ts_stream_t stream) {
    if (M < 0 || bound <= 0) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!tree || !idx || !prio || !weight || !scratch16) return TS_ERR_BAD_ARG;
    const unsigned grid = (unsigned)((M + DESC_NT - 1) / DESC_NT);
    per_weight_kernel<<<grid, DESC_NT, 0, (cudaStream_t)stream>>>(
        tree, bound, idx, M, prio, beta, weight_norm, weight, (PerScratch*)scratch16);
    return check_launch();
}

extern "C" int32_t ts_per_update(double* tree, int64_t size, int64_t bound, const int64_t* idx,
                                 const void* td, int32_t td_dtype, int64_t M, double alpha,
                                 double* prio, uint32_t* stamp, uint32_t* counter,
                                 int32_t* err_flag, ts_stream_t stream) {
    if (!td || !prio) return TS_ERR_BAD_ARG;
    if (td_dtype != TS_F32 && td_dtype != TS_F64) return TS_ERR_BAD_ARG;
    ValueSrc vs{td, 1, alpha, prio, td_dtype == TS_F64 ? VM_PER_F64 : VM_PER_F32};
    return update_impl(tree, size, bound, idx, vs, M, stamp, counter, err_flag,
                       (cudaStream_t)stream);
}

extern "C" int32_t ts_per_init_weight(double* tree, int64_t size, int64_t bound,
                                      const int64_t* idx, int64_t M, double alpha,
                                      const double* prio, int32_t leaf_dtype, uint32_t* stamp,
                                      uint32_t* counter, int32_t* err_flag, ts_stream_t stream) {
    if (!prio) return TS_ERR_BAD_ARG;
    ValueSrc vs{nullptr, 0, alpha, const_cast<double*>(prio),
                leaf_dtype == TS_F32 ? VM_INIT_F32 : VM_INIT_F64};
    return update_impl(tree, size, bound, idx, vs, M, stamp, counter, err_flag,
                       (cudaStream_t)stream);
}
