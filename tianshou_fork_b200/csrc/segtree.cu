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

// x^a for x >= 0 as exp2(a * log2(x)): relative error ~ |a log2 x| * 2^-52 (a few 1e-15), about
// a third of the instructions of the correctly-rounded pow().  Used where the result is rounded
// to f32 anyway (the reference's f32 power, prio.py:87) and for the importance weights
// (parity tolerance 1e-12); f64 leaves keep pow().
__device__ __forceinline__ double pow_fast(double x, double a) { return exp2(a * log2(x)); }

__device__ __forceinline__ double leaf_value(const ValueSrc& v, int64_t j) {
    switch (v.mode) {
        case VM_F64: return ((const double*)v.ptr)[j * v.stride];
        case VM_F32: return (double)((const float*)v.ptr)[j * v.stride];
        case VM_PER_F32: return (double)(float)pow_fast(per_priority(v, j), v.alpha);
        case VM_PER_F64: return pow(per_priority(v, j), v.alpha);
        case VM_INIT_F64: return pow(v.prio[0], v.alpha);
        default: return (double)(float)pow_fast(v.prio[0], v.alpha);
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

// ---- thread-block cluster helpers ---------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;"
                 ::: "memory");
}
// split form: arrive as soon as this CTA is done with its peers' shared memory, wait before exit
__device__ __forceinline__ void cluster_arrive() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// read a double at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ double dsmem_ld_f64(const double* local, uint32_t rank) {
    uint32_t addr = smem_u32(local), remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(addr), "r"(rank));
    double v;
    asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(remote) : "memory");
    return v;
}

// ---- sharded PER: peer-memory exchange helpers --------------------------------------------
constexpr int kMaxShards = 64;
// Slot buffer, 16 * G 64-bit words: records [2][G][6] (root, max, min as three doubles), weight
// maxima [2][G][2] (one double), both indexed by the parity of their sequence number.  Every
// 64-bit word is SELF-VALIDATING: low half = 32 bits of payload, high half = the low 32 bits of
// the sequence number it belongs to (the "LL" scheme of NCCL: an aligned 8-byte store arrives
// whole).  A record therefore needs no flag behind it and no ordering between its stores: one
// NVLink hop instead of payload -> release -> flag (6.7 us vs 5.0 us per round trip,
// scripts/nvlink_pingpong.py).  Zero-initialised words carry sequence 0; sequences start at 1
// and a parity slot is reused two sequence numbers later, so a stale word never matches.
__host__ __device__ inline int slot_record(int G, unsigned long long e, int r) {
    return ((int)(e & 1ull) * G + r) * 6;
}
__host__ __device__ inline int slot_wmax(int G, unsigned long long e, int r) {
    return 12 * G + ((int)(e & 1ull) * G + r) * 2;
}
constexpr int kPeerSpin = 1 << 22;  // bounded wait (~ seconds), then TS_FLAG_PEER_TIMEOUT

__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const void* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u64(void* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_f64(double* p, double v) {
    asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
// one double as two self-validating words at dst[0..1], tagged with sequence number `seq`
__device__ __forceinline__ void ll_store_f64(double* dst, double v, unsigned long long seq) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
    const unsigned long long tag = (seq & 0xffffffffull) << 32;
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(tag | (bits & 0xffffffffull))
                 : "memory");
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst + 1), "l"(tag | (bits >> 32))
                 : "memory");
}
// spin until both words at src[0..1] carry sequence number `seq`; false on timeout
__device__ __forceinline__ bool ll_load_f64(const double* src, unsigned long long seq, double& v) {
    const unsigned long long want = seq & 0xffffffffull;
    unsigned long long w[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        int it = 0;
        for (;;) {
            asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w[h]) : "l"(src + h) : "memory");
            if ((w[h] >> 32) == want) break;
            if (++it >= kPeerSpin) return false;
            __nanosleep(64);
        }
    }
    v = __longlong_as_double((long long)((w[0] & 0xffffffffull) | (w[1] << 32)));
    return true;
}
// spin until the epoch word at `flag` equals `want`; false on timeout (ts_debug_pingpong only)
__device__ __forceinline__ bool wait_epoch(const double* flag, unsigned long long want) {
    for (int it = 0; it < kPeerSpin; ++it) {
        if (ld_acquire_sys_u64(flag) == want) return true;
        __nanosleep(64);
    }
    return false;
}

// Where the update kernel publishes this rank's record when it is the producer side of the
// exchange (state == nullptr: nothing to publish).
struct PublishInfo {
    unsigned long long* state;
    double* const* peer_slots;
    int G, rank;
    // Optional hand-over from ts_per_sample_p2p_owned (nullptr: none): int32[4 + capacity] =
    // {accumulator of even draws, of odd draws, count of the latest draw, -, batch positions this
    // rank owns ...}.  With it the update scans count ~ M / G positions instead of all M.
    const int32_t* owned = nullptr;
};
// G threads: store {root, max_prio, min_prio, epoch} into every peer's slot buffer
__device__ __forceinline__ void publish_record(const double* tree, const double* prio,
                                               const PublishInfo& pub, int tid) {
    const unsigned long long e = __ldcg(pub.state) + 1;
    if (tid < pub.G) {
        double* dst = pub.peer_slots[tid] + slot_record(pub.G, e, pub.rank);
        ll_store_f64(dst, __ldcg(tree + 1), e);
        ll_store_f64(dst + 2, __ldcg(prio), e);
        ll_store_f64(dst + 4, __ldcg(prio + 1), e);
    }
    __syncthreads();
    if (tid == 0) pub.state[0] = e;
}

constexpr int64_t kSmallM = 16383;     // tiered path: batch position + 1 fits the stamp's 14 bits
constexpr int64_t kDenseBound = 4096;  // trees up to this many leaves are rebuilt densely
constexpr int kTagShift = 14;
constexpr int64_t kSkipIndex = -1;  // batch entries to ignore silently (foreign shard)
constexpr uint32_t kTagMax = (1u << (32 - kTagShift)) - 1u;

// ============================================================================================
// Tiered small-batch update (the replay-step case): ONE cooperative launch.
//
//   phase A   every item stamps its leaf: atomicMax(stamp[idx], tag << 14 | (j + 1)).  The tag
//             is a per-call epoch kept on the device, so stale stamps of earlier calls are
//             ignored and nothing has to be reset (CUDA-graph replays advance it themselves).
//   barrier
//   tier 0    one warp per item rebuilds the aligned group of G = 2^LG leaves that holds the
//             item: loads the G leaves and their G stamps, replaces every stamped leaf by its
//             new value (the stamp names the LAST batch position that wrote it,
//             segtree.py:40-49), then forms the LG levels above with in-register adds and
//             shuffles -- node = left + right in IEEE f64, the reference's order
//             (segtree.py:104) -- and writes leaves and nodes back.  Items that share a group
//             do identical work and write identical bits (benign).
//   barrier   (+ further tiers of the same shape for trees deeper than LG + 13 levels)
//   top       CTA 0 loads the remaining level (<= 8192 nodes) into shared memory and rebuilds
//             every level above it densely; the other CTAs have already exited.
//
// Dependent memory round trips: index, stamp RED, group load, group store, top load -- about
// five plus two grid barriers, against ~40 for a level-by-level climb.
// ============================================================================================
// Phase timestamps of the most recent tiered update (globaltimer ns, CTA 0 / thread 0):
// start, after phase A barrier, after tier 0, after the last barrier, after top load, end.
__device__ unsigned long long g_update_phase_ns[8];
// start / batch scanned / groups done / ticket taken, per CTA of the most recent ranged update
__device__ unsigned long long g_update_cta_ns[128][4];

__device__ __forceinline__ void cta_mark(int slot) {
    if (threadIdx.x == 0 && blockIdx.x < 128) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_update_cta_ns[blockIdx.x][slot] = t;
    }
}

__device__ __forceinline__ void phase_mark(int slot) {
    // CTA 1 when there is one: CTA 0 carries the batch-wide extras and is not representative
    if (blockIdx.x == (gridDim.x > 1 ? 1u : 0u) && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_update_phase_ns[slot] = t;
    }
}

constexpr int TIER_NT = 1024;
constexpr int kTopMax = 8192;  // nodes of the level handed to the top phase (128 KB heap)

struct GridBarrier {
    unsigned int* count;  // monotonically increasing within one launch; reset by the last CTA
    unsigned int ctas;
    unsigned int phase = 0;
    // every CTA arrives (its earlier writes become visible to whoever waits) ...
    __device__ __forceinline__ void arrive() {
        ++phase;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(count, 1u);
        }
    }
    // ... and the CTAs that go on to read other CTAs' writes wait for all arrivals
    __device__ __forceinline__ void wait() {
        if (threadIdx.x == 0) {
            const unsigned int target = phase * ctas;
            unsigned int seen;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(count)
                             : "memory");
            } while (seen < target);
        }
        __syncthreads();
    }
    // final rendezvous: nobody waits; returns true in the CTA that arrived last (CTA-uniform)
    __device__ __forceinline__ bool arrive_last(bool* s_flag) {
        ++phase;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            *s_flag = atomicAdd(count, 1u) == phase * ctas - 1u;
            __threadfence();
        }
        __syncthreads();
        return *s_flag;
    }
};

template <int K>
__device__ __forceinline__ double group_levels(double* __restrict__ tree, int64_t base,
                                               double (&v)[K], int lane);

// One warp rebuilds the aligned group of G = 32 * K nodes that starts at heap index `first`,
// LG = log2(G) levels up, and returns the group's root (valid in lane 0).  With STAMPED the
// group is a leaf group: stamped leaves are replaced first.
template <int K, bool STAMPED>
__device__ __forceinline__ double rebuild_group(double* __restrict__ tree, int64_t first,
                                                int64_t leaf0, int64_t size, const ValueSrc& vs,
                                                const uint32_t* __restrict__ stamp, uint32_t tag,
                                                int lane) {
    static_assert(K == 4 || K == 8, "lane strip of 4 or 8 nodes");
    double v[K];
    const int64_t base = first + (int64_t)lane * K;
#pragma unroll
    for (int q = 0; q < K / 2; ++q) {
        const double2 t = __ldcg(reinterpret_cast<const double2*>(tree + base) + q);
        v[2 * q] = t.x; v[2 * q + 1] = t.y;
    }
    if (STAMPED) {
        const int64_t p0 = leaf0 + (int64_t)lane * K;  // leaf number of v[0]
        uint32_t st[K];
        if (p0 + K <= size) {
#pragma unroll
            for (int q = 0; q < K / 4; ++q) {
                const uint4 t = __ldcg(reinterpret_cast<const uint4*>(stamp + p0) + q);
                st[4 * q] = t.x; st[4 * q + 1] = t.y; st[4 * q + 2] = t.z; st[4 * q + 3] = t.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < K; ++k) st[k] = p0 + k < size ? __ldcg(stamp + p0 + k) : 0u;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if ((st[k] >> kTagShift) == tag) {  // written by this call; low bits = position + 1
                v[k] = leaf_value(vs, (int64_t)(st[k] & ((1u << kTagShift) - 1u)) - 1);
                tree[base + k] = v[k];
            }
        }
    }
    return group_levels<K>(tree, base, v, lane);
}

// The LG levels above a group whose 32 * K nodes are already in registers (K per lane).
template <int K>
__device__ __forceinline__ double group_levels(double* __restrict__ tree, int64_t base,
                                               double (&v)[K], int lane) {
    // levels inside the lane's strip
    int64_t node = base;
#pragma unroll
    for (int w = K; w > 1; w >>= 1) {
        node >>= 1;
#pragma unroll
        for (int k = 0; k < w / 2; ++k) v[k] = __dadd_rn(v[2 * k], v[2 * k + 1]);
        if (w == 8) {
            *reinterpret_cast<double2*>(tree + node) = make_double2(v[0], v[1]);
            *reinterpret_cast<double2*>(tree + node + 2) = make_double2(v[2], v[3]);
        } else if (w == 4) {
            *reinterpret_cast<double2*>(tree + node) = make_double2(v[0], v[1]);
        } else {
            tree[node] = v[0];
        }
    }
    // levels across lanes: lane l (multiple of 2d) holds left, lane l + d holds right
    double x = v[0];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const double r = __shfl_down_sync(0xffffffffu, x, d);
        x = __dadd_rn(x, r);
        node >>= 1;
        if ((lane & (2 * d - 1)) == 0) tree[node] = x;
    }
    return x;
}

// One warp rebuilds everything above the level of n <= 128 nodes (heap indices [n, 2n)).
// The level is read from shared memory when `s_level` is given (n <= 32), else from L2.
__device__ __forceinline__ void warp_finish(double* __restrict__ tree, int n, int lane,
                                            const double* s_level) {
    const int c = n >= 32 ? n / 32 : 1;  // nodes per lane: 1, 2 or 4
    double v[4] = {0.0, 0.0, 0.0, 0.0};
    const bool mine = lane * c < n;
    if (s_level != nullptr) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k < c && mine) v[k] = s_level[lane * c + k];
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k < c && mine) v[k] = __ldcg(tree + n + lane * c + k);
    }
    int node = n + lane * c;
#pragma unroll
    for (int w = 4; w > 1; w >>= 1) {
        if (w > c) continue;
        node >>= 1;
#pragma unroll
        for (int k = 0; k < w / 2; ++k) {
            v[k] = __dadd_rn(v[2 * k], v[2 * k + 1]);
            if (mine) tree[node + k] = v[k];
        }
    }
    const int live = n < 32 ? n : 32;
    double x = v[0];
    for (int d = 1; d < live; d <<= 1) {
        const double r = __shfl_down_sync(0xffffffffu, x, d);
        x = __dadd_rn(x, r);
        node >>= 1;
        if ((lane & (2 * d - 1)) == 0 && lane < live) tree[node] = x;
    }
}

// One CTA of 32 warps rebuilds levels lvl-1 .. 0 from level lvl (2^lvl <= 8192 nodes): groups of
// 256 per warp (8 levels, no CTA barrier inside), whose roots go through shared memory to one
// warp that finishes the <= 128 nodes that remain.
__device__ __forceinline__ void rebuild_top(double* __restrict__ tree, int lvl, double* s_root,
                                            int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    const double* s_level = nullptr;
    if (lvl >= 8) {
        const int groups = 1 << (lvl - 8);  // <= 32 = warps of the CTA: a single pass
        if (warp < groups) {
            double v[8];
            const int64_t base = ((int64_t)1 << lvl) + ((int64_t)warp << 8) + (int64_t)lane * 8;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 t = __ldcg(reinterpret_cast<const double2*>(tree + base) + q);
                v[2 * q] = t.x; v[2 * q + 1] = t.y;
            }
            const double root = group_levels<8>(tree, base, v, lane);
            if (lane == 0) s_root[warp] = root;
        }
        lvl -= 8;
        s_level = s_root;
        __syncthreads();
    }
    if (tid < 32) warp_finish(tree, 1 << lvl, lane, s_level);
}

constexpr int kIdxRounds = 8;  // rounds of 32 items per CTA whose indices stay in shared memory

template <int K>
__global__ void __launch_bounds__(TIER_NT)
segtree_update_tiered(double* __restrict__ tree, int64_t size, int64_t bound, int lg_bound,
                      const int64_t* __restrict__ index, ValueSrc vs, int64_t M,
                      uint32_t* __restrict__ stamp, uint32_t* __restrict__ counter,
                      int32_t* __restrict__ err_flag, PublishInfo pub, const double* prio) {
    constexpr int LG = K == 4 ? 7 : 8;
    __shared__ int64_t s_idx[kIdxRounds][32];
    __shared__ double s_root[64];
    __shared__ bool s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t gtid = blockIdx.x * (int64_t)TIER_NT + tid;
    const int64_t gthreads = gridDim.x * (int64_t)TIER_NT;
    GridBarrier bar{counter, gridDim.x};

    phase_mark(0);
    // counter[2] = epoch of the previous call, counter[3] = "stamps must be cleared first"
    const uint32_t tag = __ldcg(counter + 2) + 1u;
    if (__ldcg(counter + 3) != 0u) {  // epoch wrapped: forget every old stamp (rare)
        for (int64_t i = gtid; i < size; i += gthreads) stamp[i] = 0u;
        bar.arrive();
        bar.wait();
    }

    // Item j belongs to CTA (j / 32) % gridDim, round j / (32 * gridDim), slot j % 32: the
    // CTA's first warp stamps slot = lane, and in the tiers warp `slot` rebuilds its group.
    const int64_t per_round = 32 * (int64_t)gridDim.x;
    const int64_t rounds = (M + per_round - 1) / per_round;

    // ---- phase A: stamps (+ priority extrema, prio.py:88-89), one warp per CTA ----------------
    if (warp == 0) {
        double pmax = 0.0, pmin = INFINITY;
        for (int64_t r = 0; r < rounds; ++r) {
            const int64_t j = r * per_round + blockIdx.x * 32 + lane;
            int64_t idx = -1;
            if (j < M) idx = index[j];
            if (idx != kSkipIndex) {  // -1: "not in this shard" (ts_per_sample_sharded), no error
                if (is_per(vs)) {
                    const double p = per_priority(vs, j);
                    pmax = fmax(pmax, p);
                    pmin = fmin(pmin, p);
                }
                if (idx < 0 || idx >= size) {
                    atomicOr(err_flag, TS_FLAG_INDEX_RANGE);
                    idx = -1;
                } else {
                    atomicMax(stamp + idx, (tag << kTagShift) | (uint32_t)(j + 1));
                }
            }
            if (r < kIdxRounds) s_idx[r][lane] = idx;
        }
        if (is_per(vs)) {
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                pmax = fmax(pmax, __shfl_xor_sync(0xffffffffu, pmax, d));
                pmin = fmin(pmin, __shfl_xor_sync(0xffffffffu, pmin, d));
            }
            if (lane == 0 && pmax > 0.0) {
                atomic_max_pos(vs.prio, pmax);
                atomic_min_pos(vs.prio + 1, pmin);
            }
        }
    }
    bar.arrive();
    bar.wait();
    phase_mark(1);

    // ---- tiers: warp per item, LG levels per tier ---------------------------------------------
    int lvl = lg_bound;  // depth of the level being consumed (leaves first)
    bool first_tier = true;
    while (true) {
        const int shift = lg_bound - lvl;  // leaf number -> node number at this level
        for (int64_t r = 0; r < rounds; ++r) {
            int64_t idx;
            if (r < kIdxRounds) {
                idx = s_idx[r][warp];
            } else {
                const int64_t j = r * per_round + blockIdx.x * 32 + warp;
                idx = j < M ? index[j] : -1;
                if (idx >= size) idx = -1;
            }
            if (idx < 0) continue;  // warp-uniform
            const int64_t grp = (idx >> shift) >> LG;        // group of G nodes at this level
            const int64_t first = ((int64_t)1 << lvl) + (grp << LG);
            if (first_tier) (void)rebuild_group<K, true>(tree, first, grp << LG, size, vs, stamp, tag, lane);
            else (void)rebuild_group<K, false>(tree, first, 0, size, vs, stamp, tag, lane);
        }
        lvl -= LG;
        first_tier = false;
        if (((int64_t)1 << lvl) <= kTopMax) break;
        bar.arrive();
        bar.wait();
    }
    // ---- top: the CTA that arrives last rebuilds levels lvl-1 .. 0 from level lvl -------------
    phase_mark(2);
    if (!bar.arrive_last(&s_last)) return;
    rebuild_top(tree, lvl, s_root, tid);
    if (pub.state != nullptr) {
        // fused producer side of the shard exchange: the CTA that just wrote the new root stores
        // this rank's record into every peer's buffer over NVLink
        __syncthreads();
        publish_record(tree, prio, pub, tid);
    }
    if (tid == 0) {
        counter[0] = 0u;
        if (tag >= kTagMax) { counter[2] = 0u; counter[3] = 1u; }
        else { counter[2] = tag; counter[3] = 0u; }
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_update_phase_ns[5] = t;
    }
}

// ============================================================================================
// Ranged small-batch update (trees up to 2^20 leaves): ONE ordinary launch, no grid barrier,
// no global stamps.
//
// CTA c owns the aligned range of L = bound / gridDim leaves (<= 8192).  Every CTA reads the
// WHOLE batch (M indices, coalesced, 8 B each -- the same 32 KB for all CTAs, served by L2) and
// keeps the items that fall in its range in a shared-memory table, one 64-bit word per leaf:
// atomicMax((j + 1) << 32 | f32 bits of the new leaf).  The maximum is the LAST batch position
// that wrote the leaf (segtree.py:40-49) and carries its value, so after one __syncthreads the
// CTA knows every leaf it has to write without another trip to memory.  Then one warp per touched
// group of 128 leaves: load the group, overlay the table, write leaves and the 7 levels above.
// The CTA that finishes last rebuilds the 13 levels that remain (rebuild_top).
//
// Dependent memory round trips: batch read, group read, last-CTA ticket, top read -- four, and
// the only inter-CTA dependency is "last one out", which needs no co-residency.
// ============================================================================================
constexpr int RANGE_NT = 1024;
constexpr int kRangeLgMax = 13;       // 8192 leaves per CTA: 64 KB table
constexpr int kRangeLgMin = 10;       // at least 8 groups per CTA
constexpr int kRangeUnroll = 4;       // items per thread whose loads are in flight together
constexpr int64_t kRangedMaxM = 1 << 18;  // every CTA scans the whole batch: beyond this the bulk path wins
constexpr int kRangeList = 2048;      // compacted in-range items per pass (more: handled in place)

__device__ __forceinline__ bool carries_f32(int mode) {
    return mode == VM_PER_F32 || mode == VM_F32 || mode == VM_INIT_F32;
}

// f32 bits of the new leaf of batch item j (modes whose leaves are f32 values)
__device__ __forceinline__ uint32_t carried_bits(const ValueSrc& vs, int64_t j, float td) {
    if (vs.mode == VM_PER_F32)
        return __float_as_uint((float)pow_fast((double)(fabsf(td) + kEpsF32), vs.alpha));
    return __float_as_uint((float)leaf_value(vs, j));
}

// one warp: store {root, max_prio, min_prio, epoch} into every peer's slot buffer
__device__ __forceinline__ void publish_record_warp(const double* tree, const double* prio,
                                                    const PublishInfo& pub, int lane) {
    const unsigned long long e = __ldcg(pub.state) + 1;
    for (int t = lane; t < pub.G; t += 32) {
        double* dst = pub.peer_slots[t] + slot_record(pub.G, e, pub.rank);
        ll_store_f64(dst, __ldcg(tree + 1), e);
        ll_store_f64(dst + 2, __ldcg(prio), e);
        ll_store_f64(dst + 4, __ldcg(prio + 1), e);
    }
    __syncwarp();
    if (lane == 0) pub.state[0] = e;
}

// Out-of-line copies for the paths that are rarely taken: the kernel's instruction footprint is
// what its latency is made of (every phase runs once per launch, from a cold instruction cache).
__device__ __noinline__ uint32_t carried_bits_cold(const ValueSrc& vs, int64_t j, float td) {
    return carried_bits(vs, j, td);
}
__device__ __noinline__ double leaf_value_cold(const ValueSrc& vs, int64_t j) {
    return leaf_value(vs, j);
}
__device__ __noinline__ double per_priority_cold(const ValueSrc& vs, int64_t j) {
    return per_priority(vs, j);
}

// PERF32: vs.mode == VM_PER_F32, the replay step's update_weight (prio.py:78-89 on a f32 td_error);
// every mode test folds away.  The other modes share the generic instantiation.
// DEFER: stop at the range roots.  The levels above them (heap indices < gridDim.x) are left
// stale for the consumer to rebuild (per_weight_cluster_kernel<.., REBUILD> in shared memory,
// or segtree_finish_top_kernel): no last-CTA ticket, no fence, no serial tail.
template <bool PERF32, bool DEFER = false>
__global__ void __launch_bounds__(RANGE_NT)
segtree_update_ranged(double* __restrict__ tree, int64_t size, int lg_bound, int lg_range,
                      const int64_t* __restrict__ index, ValueSrc vs, int64_t M,
                      uint32_t* __restrict__ counter, int32_t* __restrict__ err_flag,
                      PublishInfo pub, const double* prio) {
    extern __shared__ __align__(16) unsigned long long s_tab[];  // [L]: (j + 1) << 32 | f32 bits
    __shared__ uint2 s_list[kRangeList];                         // {j + 1, leaf - lo}
    __shared__ float s_list_td[kRangeList];
    __shared__ unsigned int s_n[2];         // list length, alternating between passes
    __shared__ unsigned int s_touched[64];  // per group of 128 leaves
    __shared__ double s_root[64];
    __shared__ double s_pmax[32], s_pmin[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int L = 1 << lg_range;
    const int groups = L >> 7;  // 8 .. 64: one or two per warp
    const bool audits = blockIdx.x == 0;  // batch-wide checks and extrema are CTA 0's job
    const int64_t lo = (int64_t)blockIdx.x << lg_range;
    const bool carry = PERF32 || carries_f32(vs.mode);
    const bool tracks_prio = PERF32 || vs.mode == VM_PER_F64;
    constexpr int64_t kPass = (int64_t)kRangeUnroll * RANGE_NT;

    pdl_launch_dependents();
    for (int i = tid; i < L / 2; i += RANGE_NT)
        reinterpret_cast<ulonglong2*>(s_tab)[i] = make_ulonglong2(0ull, 0ull);
    if (tid < 64) s_touched[tid] = 0u;
    if (tid < 2) s_n[tid] = 0u;
    pdl_wait();  // nothing above touches global memory
    phase_mark(0);
    cta_mark(0);
    // warp 0: this range's group roots as they are now; the untouched ones are needed at the end
    const int gper = groups > 32 ? 2 : 1;
    const int64_t groot0 = ((int64_t)1 << (lg_bound - 7)) + (lo >> 7);
    double old_root0 = 0.0, old_root1 = 0.0;
    if (warp == 0 && lane * gper < groups) {
        old_root0 = __ldcg(tree + groot0 + lane * gper);
        if (gper == 2) old_root1 = __ldcg(tree + groot0 + lane * gper + 1);
    }
    __syncthreads();
    phase_mark(6);

    // ---- the batch: every CTA scans all of it, keeps what falls in its range ---------------------
    double pmax = 0.0, pmin = INFINITY;                 // VM_PER_F64
    uint32_t fmax_bits = 0u, fmin_bits = 0x7f800000u;   // VM_PER_F32: positive floats order like bits
    int pass = 0;
    // A batch of several passes whose owner list came with it (sharded sampling: most entries are
    // "not in this shard"): scan the listed positions only.  The first pass's list entries are
    // fetched together with the count (the list's capacity is >= M > kPass: always in bounds).
    const bool listed = pub.owned != nullptr && M > kPass;
    int64_t Mscan = M;
    int first_j[kRangeUnroll];
    if (listed) {
#pragma unroll
        for (int q = 0; q < kRangeUnroll; ++q) first_j[q] = __ldcg(pub.owned + 4 + q * RANGE_NT + tid);
        Mscan = (int64_t)__ldcg(pub.owned + 2);
        if (Mscan > M) Mscan = M;
    }
    for (int64_t j0 = 0; j0 < Mscan; j0 += kPass, pass ^= 1) {
        int64_t id[kRangeUnroll];
        float tdf[kRangeUnroll];
        int jj[kRangeUnroll];  // batch position (M < 2^31 checked by the host)
#pragma unroll
        for (int q = 0; q < kRangeUnroll; ++q) {
            const int64_t i = j0 + (int64_t)q * RANGE_NT + tid;
            int j = i < Mscan ? (int)i : -1;
            if (listed && j >= 0) j = j0 == 0 ? first_j[q] : __ldcg(pub.owned + 4 + i);
            if (listed && (j < 0 || j >= M)) j = -1;  // a corrupt list must not read out of bounds
            jj[q] = j;
            id[q] = j >= 0 ? __ldg(index + j) : kSkipIndex;
            tdf[q] = PERF32 && j >= 0 ? __ldg((const float*)vs.ptr + j) : 0.f;
        }
        if (audits) {
#pragma unroll
            for (int q = 0; q < kRangeUnroll; ++q) {
                const int64_t idx = id[q];
                if (idx == kSkipIndex) continue;  // "not in this shard" (ts_per_sample_sharded)
                if (PERF32) {  // priority extrema over the batch (prio.py:88-89)
                    const uint32_t pb = __float_as_uint(fabsf(tdf[q]) + kEpsF32);
                    fmax_bits = max(fmax_bits, pb);
                    fmin_bits = min(fmin_bits, pb);
                } else if (tracks_prio) {
                    const double p = per_priority_cold(vs, (int64_t)jj[q]);
                    pmax = fmax(pmax, p);
                    pmin = fmin(pmin, p);
                }
                if (idx < 0 || idx >= size) atomicOr(err_flag, TS_FLAG_INDEX_RANGE);
            }
        }
#pragma unroll
        for (int q = 0; q < kRangeUnroll; ++q) {
            const int64_t j = jj[q];
            const int64_t rel = id[q] - lo;
            if (id[q] == kSkipIndex || rel < 0 || rel >= L || id[q] >= size) continue;
            s_touched[rel >> 7] = 1u;
            const unsigned int pos = atomicAdd(&s_n[pass], 1u);
            if (pos < (unsigned)kRangeList) {
                s_list[pos] = make_uint2((uint32_t)(j + 1), (uint32_t)rel);
                s_list_td[pos] = tdf[q];
            } else {  // a batch concentrated on this range: finish the item here
                const uint32_t bits = carry ? carried_bits_cold(vs, j, tdf[q]) : 0u;
                atomicMax(s_tab + rel, ((unsigned long long)(j + 1) << 32) | bits);
            }
        }
        __syncthreads();
        phase_mark(7);
        // the kept items, densely: the exponent is a few hundred dependent f64 instructions
        const unsigned int n = min(s_n[pass], (unsigned)kRangeList);
        if (tid == 0) s_n[pass ^ 1] = 0u;  // the next pass's counter; nobody is using it now
        for (unsigned int i = tid; i < n; i += RANGE_NT) {
            const uint2 it = s_list[i];
            uint32_t bits = 0u;
            if (PERF32)
                bits = __float_as_uint(
                    (float)pow_fast((double)(fabsf(s_list_td[i]) + kEpsF32), vs.alpha));
            else if (carry)
                bits = carried_bits_cold(vs, (int64_t)it.x - 1, 0.f);
            atomicMax(s_tab + it.y, ((unsigned long long)it.x << 32) | bits);
        }
        __syncthreads();
    }
    if (audits && tracks_prio) {
        if (PERF32) {
            fmax_bits = __reduce_max_sync(0xffffffffu, fmax_bits);
            fmin_bits = __reduce_min_sync(0xffffffffu, fmin_bits);
            pmax = fmax_bits != 0u ? (double)__uint_as_float(fmax_bits) : 0.0;
            pmin = (double)__uint_as_float(fmin_bits);
        } else {
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                pmax = fmax(pmax, __shfl_xor_sync(0xffffffffu, pmax, d));
                pmin = fmin(pmin, __shfl_xor_sync(0xffffffffu, pmin, d));
            }
        }
        if (lane == 0) { s_pmax[warp] = pmax; s_pmin[warp] = pmin; }
    }
    phase_mark(1);
    cta_mark(1);

    // ---- touched groups: warp per group of 128 leaves (its two groups fetched together) ---------
    {
        const int64_t leaves = ((int64_t)1 << lg_bound) + lo + lane * 4;
        const int g0 = warp, g1 = warp + 32;
        const bool t0 = g0 < groups && s_touched[g0] != 0u;  // warp-uniform
        const bool t1 = g1 < groups && s_touched[g1] != 0u;
        double v[2][4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 0 ? t0 : t1) {
                const int64_t base = leaves + ((int64_t)(h == 0 ? g0 : g1) << 7);
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const double2 t = __ldcg(reinterpret_cast<const double2*>(tree + base) + q);
                    v[h][2 * q] = t.x; v[h][2 * q + 1] = t.y;
                }
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 0 ? t0 : t1) {
                const int gl = h == 0 ? g0 : g1;
                const int64_t base = leaves + ((int64_t)gl << 7);
                const ulonglong2* e2 =
                    reinterpret_cast<const ulonglong2*>(s_tab + (gl << 7) + lane * 4);
                const ulonglong2 ea = e2[0], eb = e2[1];
                const unsigned long long e[4] = {ea.x, ea.y, eb.x, eb.y};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (e[k] != 0ull) {
                        v[h][k] = carry ? (double)__uint_as_float((uint32_t)e[k])
                                        : leaf_value_cold(vs, (int64_t)(e[k] >> 32) - 1);
                        tree[base + k] = v[h][k];
                    }
                }
                const double root = group_levels<4>(tree, base, v[h], lane);
                if (lane == 0) s_root[gl] = root;
            }
        }
    }
    __syncthreads();
    phase_mark(2);
    cta_mark(2);
    if (warp != 0) return;

    if (audits && tracks_prio) {
        pmax = s_pmax[lane];
        pmin = s_pmin[lane];
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            pmax = fmax(pmax, __shfl_xor_sync(0xffffffffu, pmax, d));
            pmin = fmin(pmin, __shfl_xor_sync(0xffffffffu, pmin, d));
        }
        if (lane == 0 && pmax > 0.0) {
            atomic_max_pos(vs.prio, pmax);
            atomic_min_pos(vs.prio + 1, pmin);
        }
    }
    // ---- warp 0: this range's root from its group roots (untouched ones from the tree) ------------
    {
        const int g0 = lane * gper;
        const bool in = g0 < groups;
        const bool m0 = in && s_touched[g0] != 0u;
        const bool m1 = in && gper == 2 && s_touched[g0 + 1] != 0u;
        if (__any_sync(0xffffffffu, m0 || m1)) {  // else nothing in this range changed
            double x = 0.0;
            int64_t node = groot0 + g0;
            if (in) x = m0 ? s_root[g0] : old_root0;
            if (gper == 2) {
                const double y = m1 ? s_root[g0 + 1] : old_root1;
                x = __dadd_rn(x, y);
                node >>= 1;
                tree[node] = x;
            }
            const int live = groups / gper;  // 8 .. 32 lanes hold one node each
            for (int d = 1; d < live; d <<= 1) {
                const double r = __shfl_down_sync(0xffffffffu, x, d);
                x = __dadd_rn(x, r);
                node >>= 1;
                if ((lane & (2 * d - 1)) == 0 && lane < live) tree[node] = x;
            }
        }
        __syncwarp();
    }
    if (DEFER) return;  // kernel completion publishes the range roots; the top is the reader's job

    // ---- warp 0 of every CTA: the one that arrives last rebuilds the levels above the ranges -------
    int last = 0;
    if (lane == 0) {
        __threadfence();
        last = atomicAdd(counter, 1u) == gridDim.x - 1u;
    }
    cta_mark(3);
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();  // acquire side: the other CTAs' nodes
    if (lane == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_update_phase_ns[3] = t;
    }
    if (gridDim.x > 1) warp_finish(tree, (int)gridDim.x, lane, nullptr);
    if (pub.state != nullptr) {
        __syncwarp();
        publish_record_warp(tree, prio, pub, lane);
    }
    if (lane == 0) {
        counter[0] = 0u;
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_update_phase_ns[5] = t;
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
        if (idx == kSkipIndex) continue;
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
        const int64_t idx = index[j];
        if (idx == kSkipIndex) continue;
        if (is_per(vs)) {
            const double p = per_priority(vs, j);
            pmax = fmax(pmax, p);
            pmin = fmin(pmin, p);
        }
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

inline int ilog2(int64_t pow2) {
    int lg = 0;
    while ((1LL << lg) < pow2) ++lg;
    return lg;
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

int32_t publish_launch(const double* tree, const double* prio, const PublishInfo& pub,
                       cudaStream_t st);

int32_t update_impl(double* tree, int64_t size, int64_t bound, const int64_t* index,
                    ValueSrc vs, int64_t M, uint32_t* stamp, uint32_t* counter,
                    int32_t* err_flag, cudaStream_t st,
                    PublishInfo pub = PublishInfo{nullptr, nullptr, 1, 0}, bool defer_top = false) {
    if (M < 0 || size <= 0 || bound < size || (bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (M >= 0x7fffffffLL || bound > (1LL << 30)) return TS_ERR_BAD_ARG;
    if (M == 0) return pub.state != nullptr && tree ? publish_launch(tree, vs.prio, pub, st) : TS_OK;
    if (!tree || !index || !stamp || !counter || !err_flag) return TS_ERR_BAD_ARG;

    if (!aligned(tree, 16) || !aligned(stamp, 16)) return TS_ERR_ALIGN;
    // (the ranged kernel's table carries the batch position in 32 bits: no 14-bit limit)
    if (M <= kRangedMaxM && bound > kDenseBound && bound <= (1LL << (kRangeLgMax + 7)) &&
        g_tune_segtree_ranged != 0) {
        const int lg = ilog2(bound);
        int lg_range = lg - 7;  // 128 CTAs when the range stays within [2^10, 2^13] leaves
        if (lg_range > kRangeLgMax) lg_range = kRangeLgMax;
        if (lg_range < kRangeLgMin) lg_range = kRangeLgMin;
        const unsigned ctas = (unsigned)(bound >> lg_range);
        const size_t smem = sizeof(unsigned long long) << lg_range;
        static bool attr_done = false;
        if (!attr_done) {
            const int cap = (int)(sizeof(unsigned long long) << kRangeLgMax);
            cudaError_t e = cudaFuncSetAttribute(segtree_update_ranged<true>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(segtree_update_ranged<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(segtree_update_ranged<true, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(segtree_update_ranged<false, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
            if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
            attr_done = true;
        }
        const double* prio_c = vs.prio;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(ctas);
        cfg.blockDim = dim3(RANGE_NT);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = g_tune_pdl ? 1 : 0;
        auto kern = vs.mode == VM_PER_F32 ? segtree_update_ranged<true> : segtree_update_ranged<false>;
        if (defer_top && pub.state == nullptr)
            kern = vs.mode == VM_PER_F32 ? segtree_update_ranged<true, true>
                                         : segtree_update_ranged<false, true>;
        cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tree, size, lg, lg_range, index, vs, M,
                                           counter, err_flag, pub, prio_c);
        if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; (void)cudaGetLastError(); return TS_ERR_CUDA; }
        return check_launch();
    }
    if (M <= kSmallM && bound > kDenseBound) {
        const int lg = ilog2(bound);
        // level handed to the top phase: strip LG levels per tier until <= kTopMax nodes
        const int K = g_tune_segtree_group == 256 ? 8 : 4;
        auto kern = K == 8 ? segtree_update_tiered<8> : segtree_update_tiered<4>;
        int64_t ctas = (M * 32 + TIER_NT - 1) / TIER_NT;  // one warp per item
        if (ctas > g_tune_segtree_max_ctas) ctas = g_tune_segtree_max_ctas;  // co-resident
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)ctas);
        cfg.blockDim = dim3(TIER_NT);
        cfg.dynamicSmemBytes = 0;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;  // the CTAs wait on each other
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        const double* prio_c = vs.prio;
        cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tree, size, bound, lg, index, vs, M, stamp,
                                           counter, err_flag, pub, prio_c);
        if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; (void)cudaGetLastError(); return TS_ERR_CUDA; }
        return check_launch();
    }
    // bulk or tiny-tree path: plain stamps (position + 1), cleared before and after use
    {
        cudaError_t e = cudaMemsetAsync(stamp, 0, (size_t)size * sizeof(uint32_t), st);
        if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
    }
    segtree_stamp_kernel<<<grid_for(M, 256), 256, 0, st>>>(index, M, size, stamp, err_flag);
    segtree_leaf_kernel<<<grid_for(M, 256), 256, 0, st>>>(tree, size, bound, index, vs, M, stamp);
    // single leaf: heap node 1 is the leaf itself, node 0 unused
    int32_t rc = bound == 1 ? check_launch() : dense_rebuild(tree, bound, st);
    if (rc == TS_OK && pub.state != nullptr) rc = publish_launch(tree, vs.prio, pub, st);
    return rc;
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
// Prefix-sum descent (segtree.py:122-137).  One query per thread.  The top twelve levels of
// the heap are staged in shared memory once per CTA (32 KB); below them the tree is read from
// L2 (a 2^20-leaf tree is 16 MiB, resident in the 126 MB L2) THREE LEVELS PER ROUND TRIP: the
// eight great-grandchildren of the current node are one aligned 64-byte run, and because
// every node equals fl(left + right) the two levels above them are recomputed in registers
// bit-exactly instead of being fetched -- 20 dependent loads become 11 shared + 3 global.
// ============================================================================================
constexpr int kTopCached = 4096;  // heap indices [0, 4096) = depths 0..11, 32 KB per CTA

// warp 0: the levels above the n <= 128 nodes s_top[n .. 2n), rebuilt in shared memory in the
// tree's own pairwise order (node = left + right, segtree.py:104) and, with `tree_out`, written
// back to the heap in global memory.
__device__ __forceinline__ void warp_top_smem(double* s_top, int n, int lane, double* tree_out) {
    const int c = n >= 32 ? n / 32 : 1;  // nodes per lane: 1, 2 or 4
    double v[4];
    const bool mine = lane * c < n;
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (k < c && mine) ? s_top[n + lane * c + k] : 0.0;
    int node = n + lane * c;
#pragma unroll
    for (int w = 4; w > 1; w >>= 1) {
        if (w > c) continue;
        node >>= 1;
#pragma unroll
        for (int k = 0; k < w / 2; ++k) {
            v[k] = __dadd_rn(v[2 * k], v[2 * k + 1]);
            if (mine) {
                s_top[node + k] = v[k];
                if (tree_out != nullptr) tree_out[node + k] = v[k];
            }
        }
    }
    const int live = n < 32 ? n : 32;
    double x = v[0];
    for (int d = 1; d < live; d <<= 1) {
        const double r = __shfl_down_sync(0xffffffffu, x, d);
        x = __dadd_rn(x, r);
        node >>= 1;
        if ((lane & (2 * d - 1)) == 0 && lane < live) {
            s_top[node] = x;
            if (tree_out != nullptr) tree_out[node] = x;
        }
    }
}

// Levels above the `n` range roots, for readers of a lazily updated tree other than the sampling
// kernel (reduce, pickling, tests): one warp.
__global__ void __launch_bounds__(32)
segtree_finish_top_kernel(double* __restrict__ tree, int n) {
    __shared__ double s_top[256];
    const int lane = threadIdx.x;
    for (int i = lane; i < n; i += 32) s_top[n + i] = __ldcg(tree + n + i);
    __syncwarp();
    warp_top_smem(s_top, n, lane, tree);
}

// number of range roots of the ranged update for a tree of 2^lg leaves (0: not in that regime)
inline int lazy_top_nodes(int64_t bound) {
    if (bound <= kDenseBound || bound > (1LL << (kRangeLgMax + 7))) return 0;
    int lg = 0;
    while ((1LL << lg) < bound) ++lg;
    int lg_range = lg - 7;
    if (lg_range > kRangeLgMax) lg_range = kRangeLgMax;
    if (lg_range < kRangeLgMin) lg_range = kRangeLgMin;
    return (int)(bound >> lg_range);
}

template <int NT>
__device__ __forceinline__ int stage_top(const double* __restrict__ tree, int64_t bound,
                                         double* s_top, int tid) {
    const int cached = (int)(2 * bound < kTopCached ? 2 * bound : kTopCached);
    for (int i = tid * 2; i < cached; i += NT * 2) {
        const double2 t = __ldcg(reinterpret_cast<const double2*>(tree + i));
        s_top[i] = t.x;
        s_top[i + 1] = t.y;
    }
    __syncthreads();
    return cached;  // heap indices below this value are in shared memory
}

// one level: left son `lsons`, go right iff lsons < v (strict), subtracting it (:132-135)
template <typename VT>
__device__ __forceinline__ int step_down(double lsons, VT& v) {
    const bool direct = lsons < (double)v;
    if (direct) v = (VT)__dsub_rn((double)v, lsons);
    return direct ? 1 : 0;
}

// Descent for one query; `v` is updated like the reference mutates `value` (:134).
// `leaf_out` (optional) receives tree[bound + result]: the last round trip has already fetched it.
template <typename VT>
__device__ __forceinline__ int64_t descend(const double* __restrict__ tree, int64_t bound,
                                           int lg_bound, const double* s_top, int cached,
                                           VT& v, double* leaf_out = nullptr) {
    int64_t index = 1;
    int depth = 0;
    double leaf = 0.0;
    bool have_leaf = false;
    while (depth < lg_bound && 2 * index < cached) {  // sons still in shared memory
        index <<= 1;
        index += step_down(s_top[index], v);
        ++depth;
    }
    for (int r = (lg_bound - depth) % 3; r > 0; --r) {  // leave a multiple of three levels
        index <<= 1;
        index += step_down(__ldg(tree + index), v);
        ++depth;
    }
    while (depth < lg_bound) {
        const double2* q = reinterpret_cast<const double2*>(tree + (index << 3));
        const double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
        const double n0 = __dadd_rn(a.x, a.y), n1 = __dadd_rn(b.x, b.y);  // tree[4i], tree[4i+1]
        const double n2 = __dadd_rn(c.x, c.y);                            // tree[4i+2]
        const double m0 = __dadd_rn(n0, n1);                              // tree[2i]
        const int r1 = step_down(m0, v);
        const int r2 = step_down(r1 ? n2 : n0, v);
        const double l3 = r1 ? (r2 ? d.x : c.x) : (r2 ? b.x : a.x);
        const int r3 = step_down(l3, v);
        index = (index << 3) + 4 * r1 + 2 * r2 + r3;
        depth += 3;
        if (leaf_out != nullptr && depth == lg_bound) {  // the eight nodes were leaves
            const double r_ = r1 ? (r2 ? d.y : c.y) : (r2 ? b.y : a.y);
            leaf = r3 ? r_ : l3;
            have_leaf = true;
        }
    }
    if (leaf_out != nullptr) *leaf_out = have_leaf ? leaf : __ldg(tree + index);
    return index - bound;                                       // :136
}

constexpr int DESC_NT = 256;

template <typename VT>
__global__ void __launch_bounds__(DESC_NT)
segtree_prefix_kernel(const double* __restrict__ tree, int64_t bound, int lg_bound,
                      VT* __restrict__ value, int64_t M, int scale_by_total,
                      int64_t* __restrict__ out, int32_t* __restrict__ err_flag) {
    __shared__ double s_top[kTopCached];
    const int tid = threadIdx.x;
    const int cached = stage_top<DESC_NT>(tree, bound, s_top, tid);
    const double total = s_top[1];
    const int64_t j = blockIdx.x * (int64_t)DESC_NT + tid;
    if (j >= M) return;
    VT v = value[j];
    if (scale_by_total) v = (VT)__dmul_rn((double)v, total);     // prio.py:61
    if (!((double)v >= 0.0 && (double)v < total)) atomicOr(err_flag, TS_FLAG_PREFIX_RANGE);  // :79
    out[j] = descend<VT>(tree, bound, lg_bound, s_top, cached, v);
    value[j] = v;
}

// ============================================================================================
// Fused PER sample: scalar = u * total, descent, importance weight, batch-max normalisation.
// prio.py:59-64 (sample_indices), :66-76 (get_weight), :101 (weight / max(weight)).
//
// The whole batch is handled by ONE thread-block cluster (8 CTAs x 512 threads = 4096 queries
// per sweep): the batch maximum that normalises the weights is then a block reduction plus an
// exchange through distributed shared memory behind the hardware cluster barrier -- no atomics,
// no scratch in global memory, no serial tail by a "last" CTA.
// ============================================================================================
// cluster shape: 8 x 512 (portable) or 16 x 256 (opt-in size) -- ts_tune_set("per_cluster")

// max over every thread of the cluster (all threads must call; result valid in all threads).
// The caller MUST execute cluster_wait() once afterwards, before returning from the kernel.
template <int NT>
__device__ __forceinline__ double cluster_max(double m, double* s_warp, double* s_cta) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, d));
    if ((tid & 31) == 0) s_warp[tid >> 5] = m;
    __syncthreads();
    if (tid < 32) {
        m = tid < NT / 32 ? s_warp[tid] : 0.0;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, d));
        if (tid == 0) *s_cta = m;
    }
    cluster_sync_all();  // every CTA's maximum is published
    // lane r reads CTA r's maximum (one remote read per lane, in parallel), then a warp reduction
    const uint32_t nrank = cluster_nctarank();
    const uint32_t lane = (uint32_t)tid & 31u;
    m = lane < nrank ? dsmem_ld_f64(s_cta, lane) : 0.0;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, d));
    // nobody may leave while peers still read its s_cta: arrive now, the caller waits
    // (cluster_wait) just before the kernel ends, under the normalising pass
    cluster_arrive();
    return m;
}

__device__ __forceinline__ double is_weight(double leaf, double min_prio, double beta) {
    return pow_fast(__ddiv_rn(leaf, min_prio), -beta);           // prio.py:76
}

// SAMPLE: idx from the descent; otherwise idx is given (get_weight on known indices)
// Sharded form: this GPU holds shard `rank` of G power-of-two-aligned subtrees of
// the reference's single global tree (manager.py:209-214).  `roots3` = G x {root, max_prio,
// min_prio} gathered from all ranks; the log2(G) levels above the shard roots are rebuilt here
// in the global tree's own pairwise order, every query is routed down them with the same strict
// compare / subtract, and only the queries that land in this shard descend locally -- the rest
// get idx = -1, weight = 0.  Normalisation needs the maximum over ALL ranks, so with G > 1 the
// kernel writes the local maximum to *wmax_out and leaves the weights raw.
//
// Two exchange modes.  NCCL mode: the caller all-gathers the records into `roots3` and
// MAX-all-reduces *wmax_out between this kernel and ts_per_normalize.  P2P mode (`slots` set):
// the exchange is INSIDE the kernels, over NVLink peer memory -- every rank owns a slot buffer
// that all peers can write (torch symmetric memory / cudaIpc); ts_per_publish stores
// {root, max_prio, min_prio, epoch} straight into every peer's buffer after an update, this
// kernel spins (bounded) until the G records of its epoch have landed, and the weight maximum
// is exchanged the same way before the normalising division -- no collective library call,
// no extra launch, capturable in a CUDA graph (epochs live on the device).
constexpr int kOwnedList = 768;  // owned draws a CTA keeps for its dense descent pass (12 KB)
struct ShardInfo {
    const double* roots3;  // NCCL mode: G x {root, max_prio, min_prio}; nullptr otherwise
    int G, rank;
    double* wmax_out;      // NCCL mode
    double* slots;                       // P2P mode: this rank's slot buffer, 16 * G doubles
    double* const* peer_slots;           // P2P mode: device array of the G ranks' slot buffers
    unsigned long long* state;           // P2P mode: {records published, samples drawn} by this rank
    int32_t* owned = nullptr;            // P2P mode, optional: see PublishInfo::owned
};
// SHARD: 0 = one GPU (no exchange code in the instantiation at all), 1 = NCCL mode, 2 = P2P mode
template <bool SAMPLE, int NT, int SHARD>
__global__ void __launch_bounds__(NT, 1)
per_weight_cluster_kernel(const double* __restrict__ tree, int64_t bound, int lg_bound,
                          const double* __restrict__ u, int64_t M,
                          const double* __restrict__ prio, double beta, int weight_norm,
                          int64_t* __restrict__ idx, double* __restrict__ weight,
                          int32_t* __restrict__ err_flag, ShardInfo sh, int rebuild_n) {
    __shared__ double s_top[SAMPLE ? kTopCached : 2];
    __shared__ double s_gtop[2 * kMaxShards];  // heap of the levels above the shard roots
    __shared__ double s_warp[NT / 32];
    __shared__ double s_cta;
    const int tid = threadIdx.x;
    __shared__ double s_gext[2 * kMaxShards];  // max_prio | min_prio of every rank
    __shared__ ulonglong2 s_list[SAMPLE && SHARD != 0 ? kOwnedList : 1];  // {position, routed v}
    __shared__ unsigned int s_nlist;
    constexpr bool p2p = SHARD == 2;
    constexpr bool sharded = SHARD != 0;
    pdl_launch_dependents();
    pdl_wait();  // the tree and the priorities come from the kernel before
    const int G = sharded ? sh.G : 1;
    int cached = 0;
    double total = 0.0;
    double min_prio = p2p ? INFINITY : __ldg(prio + 1);
    unsigned long long epoch = 0, draw = 0;
    if (sharded) {
        if (p2p) {
            epoch = __ldcg(sh.state);         // records this rank has published == everyone's
            draw = __ldcg(sh.state + 1) + 1;  // this sample's number
        }
        if (tid < G) {
            if (p2p) {
                const double* rec = sh.slots + slot_record(G, epoch, tid);
                double r0 = 0.0, r1 = 0.0, r2 = 0.0;
                if (!(ll_load_f64(rec, epoch, r0) && ll_load_f64(rec + 2, epoch, r1) &&
                      ll_load_f64(rec + 4, epoch, r2)))
                    atomicOr(err_flag, TS_FLAG_PEER_TIMEOUT);
                s_gtop[G + tid] = r0;
                s_gext[tid] = r1;
                s_gext[G + tid] = r2;
            } else {
                s_gtop[G + tid] = __ldcg(sh.roots3 + 3 * tid);
                s_gext[tid] = __ldcg(sh.roots3 + 3 * tid + 1);
                s_gext[G + tid] = __ldcg(sh.roots3 + 3 * tid + 2);
            }
        }
        __syncthreads();
        for (int w = G >> 1; w >= 1; w >>= 1) {
            if (tid < w) s_gtop[w + tid] = __dadd_rn(s_gtop[2 * (w + tid)], s_gtop[2 * (w + tid) + 1]);
            __syncthreads();
        }
        for (int r = 0; r < G; ++r) min_prio = fmin(min_prio, s_gext[G + r]);
    }
    const int64_t sweep = (int64_t)cluster_nctarank() * NT;
    const int64_t j0 = (int64_t)cluster_ctarank() * NT + tid;
    // the first query's draw is fetched while the top of the tree is being staged
    const double u_first = SAMPLE && j0 < M ? __ldg(u + j0) : 0.0;
    if (SAMPLE) {
        if (rebuild_n > 0) {
            // lazily updated tree (ts_per_update_lazy): the levels above the rebuild_n range roots
            // are stale in memory.  Warp 0 fetches the roots itself and rebuilds those levels in
            // shared memory (bit-identical to what the eager update writes; cluster rank 0
            // writes them back) WHILE the other warps stage the levels below: one barrier.
            const int n = rebuild_n;
            cached = (int)(2 * bound < kTopCached ? 2 * bound : kTopCached);
            if (tid < 32) {
                double r[4];
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    r[q] = tid + 32 * q < n ? __ldcg(tree + n + tid + 32 * q) : 0.0;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (tid + 32 * q < n) s_top[n + tid + 32 * q] = r[q];
                __syncwarp();
                warp_top_smem(s_top, n, tid,
                              cluster_ctarank() == 0 ? const_cast<double*>(tree) : nullptr);
            } else {
                for (int i = 2 * n + (tid - 32) * 2; i < cached; i += (NT - 32) * 2) {
                    const double2 t = __ldcg(reinterpret_cast<const double2*>(tree + i));
                    s_top[i] = t.x;
                    s_top[i + 1] = t.y;
                }
            }
            __syncthreads();
        } else {
            cached = stage_top<NT>(tree, bound, s_top, tid);
        }
        total = sharded ? s_gtop[1] : s_top[1];
    }
    double wmax = 0.0, w_first = 0.0;
    // Sharded sampling: a rank owns ~1/G of the draws, scattered over the batch.  Descending "if
    // mine" inside the sweep loop makes every warp walk a full dependent chain per sweep for the
    // one or two lanes that own something (G = 8, batch 4096 per rank: 8 chains per thread instead
    // of 1).  So the sweeps only ROUTE (shared-memory top of the global tree) and append the owned
    // draws to a CTA-wide list; the descents then run densely, one per thread.
    constexpr bool compact = SAMPLE && sharded;
    __shared__ unsigned int s_obase;
    // owned-position list for the update that follows (accumulator slot = parity of the draw number;
    // the lead CTA clears the other slot for the next draw: nobody touches it during this launch)
    unsigned int* const o_acc = p2p && sh.owned != nullptr
        ? reinterpret_cast<unsigned int*>(sh.owned) + (draw & 1ull) : nullptr;
    if (compact) {
        if (tid == 0) {
            s_nlist = 0u;
            if (o_acc != nullptr && cluster_ctarank() == 0)
                reinterpret_cast<unsigned int*>(sh.owned)[(draw & 1ull) ^ 1ull] = 0u;
        }
        __syncthreads();
        double u_next = u_first;
        for (int64_t j = j0; j < M; j += sweep) {
            const double uj = u_next;
            // the next sweep's draw is fetched under this one's routing (weak scaling: G sweeps)
            if (j + sweep < M) u_next = __ldg(u + j + sweep);
            double v = __dmul_rn(uj, total);  // prio.py:61
            if (!(v >= 0.0 && v < total)) atomicOr(err_flag, TS_FLAG_PREFIX_RANGE);
            int node = 1;
            while (node < G) {                                   // the global tree's top levels
                node <<= 1;
                node += step_down(s_gtop[node], v);
            }
            if (node - G != sh.rank) {
                idx[j] = -1;
                weight[j] = 0.0;
                continue;
            }
            const unsigned int pos = atomicAdd(&s_nlist, 1u);
            if (pos < (unsigned)kOwnedList) {
                s_list[pos] = make_ulonglong2((unsigned long long)j,
                                              (unsigned long long)__double_as_longlong(v));
            } else {  // far more than this CTA's share lands in this shard: finish it here
                double leaf_val = 0.0;
                if (o_acc != nullptr) sh.owned[4 + atomicAdd(o_acc, 1u)] = (int32_t)j;
                idx[j] = descend<double>(tree, bound, lg_bound, s_top, cached, v, &leaf_val);
                const double w = is_weight(leaf_val, min_prio, beta);
                weight[j] = w;
                wmax = fmax(wmax, w);
            }
        }
        __syncthreads();
        const unsigned int n = min(s_nlist, (unsigned)kOwnedList);
        if (o_acc != nullptr && tid == 0 && n > 0) s_obase = atomicAdd(o_acc, n);
        for (unsigned int i = tid; i < n; i += NT) {
            const ulonglong2 it = s_list[i];
            double v = __longlong_as_double((long long)it.y), leaf_val = 0.0;
            const int64_t j = (int64_t)it.x;
            idx[j] = descend<double>(tree, bound, lg_bound, s_top, cached, v, &leaf_val);
            const double w = is_weight(leaf_val, min_prio, beta);
            weight[j] = w;  // raw; normalised below by the thread whose sweep owns position j
            wmax = fmax(wmax, w);
        }
        __syncthreads();  // the raw weights of this CTA's positions are all in place
        if (o_acc != nullptr)
            for (unsigned int i = tid; i < n; i += NT) sh.owned[4 + s_obase + i] = (int32_t)s_list[i].x;
    }
    for (int64_t j = j0; !compact && j < M; j += sweep) {
        int64_t leaf;
        double leaf_val = 0.0;
        bool mine = true;
        if (SAMPLE) {
            double v = __dmul_rn(j == j0 ? u_first : u[j], total);  // prio.py:61
            if (!(v >= 0.0 && v < total)) atomicOr(err_flag, TS_FLAG_PREFIX_RANGE);
            leaf = descend<double>(tree, bound, lg_bound, s_top, cached, v, &leaf_val);
            idx[j] = leaf;
        } else {
            leaf = idx[j];
            mine = leaf >= 0;
            if (mine) leaf_val = __ldg(tree + bound + leaf);
        }
        const double w = mine ? is_weight(leaf_val, min_prio, beta) : 0.0;
        if (j == j0) w_first = w; else weight[j] = w;  // later sweeps park the raw weight
        wmax = fmax(wmax, w);
    }
    if (weight_norm) wmax = cluster_max<NT>(wmax, s_warp, &s_cta);
    bool scale = weight_norm && !sharded;
    if (sharded && !p2p && weight_norm && j0 == 0) *sh.wmax_out = wmax;
    if (p2p) {
        const bool lead = cluster_ctarank() == 0;
        if (weight_norm) {
            // maximum over ALL ranks: store ours into every peer's buffer, wait for theirs
            if (lead && tid < G) {
                ll_store_f64(sh.peer_slots[tid] + slot_wmax(G, draw, sh.rank), wmax, draw);
            }
            if (tid < G) {
                double m = 0.0;
                if (!ll_load_f64(sh.slots + slot_wmax(G, draw, tid), draw, m))
                    atomicOr(err_flag, TS_FLAG_PEER_TIMEOUT);
                s_gtop[tid] = m;
            }
            __syncthreads();
            wmax = 0.0;
            for (int r = 0; r < G; ++r) wmax = fmax(wmax, s_gtop[r]);
            scale = true;
        }
        if (!weight_norm) cluster_sync_all();  // every CTA has read `state` before it moves
        if (lead && tid == 0) {
            // (a cluster barrier lies behind us: every CTA's additions to the accumulator are done)
            if (o_acc != nullptr) sh.owned[2] = (int32_t)__ldcg(o_acc);
            sh.state[1] = draw;
            // the reference has ONE _max_prio / _min_prio (prio.py:35): adopt the global extrema
            double gmax = 0.0;
            for (int r = 0; r < G; ++r) gmax = fmax(gmax, s_gext[r]);
            const_cast<double*>(prio)[0] = gmax;
            const_cast<double*>(prio)[1] = min_prio;
        }
    }
    if (compact && s_nlist <= (unsigned)kOwnedList) {
        // only the owned positions carry a weight (the others were written as 0 while routing):
        // normalise those from the CTA's list instead of sweeping the whole batch again
        if (scale) {
            const unsigned int n = s_nlist;
            for (unsigned int i = tid; i < n; i += NT) {
                const int64_t j = (int64_t)s_list[i].x;
                weight[j] = __ddiv_rn(weight[j], wmax);          // prio.py:101
            }
        }
    } else {
        for (int64_t j = j0; j < M; j += sweep) {
            const double w = (!compact && j == j0) ? w_first : weight[j];
            weight[j] = scale ? __ddiv_rn(w, wmax) : w;          // prio.py:101
        }
    }
    if (weight_norm) cluster_wait();  // second half of cluster_max's closing barrier
}

// After a tree change: this rank's {root, max_prio, min_prio} goes straight into every peer's
// slot buffer over NVLink (G small stores), tagged with the new epoch.
__global__ void __launch_bounds__(kMaxShards)
per_publish_kernel(const double* __restrict__ tree, const double* __restrict__ prio,
                   unsigned long long* __restrict__ state, double* const* __restrict__ peer_slots,
                   int G, int rank) {
    publish_record(tree, prio, PublishInfo{state, peer_slots, G, rank}, threadIdx.x);
}

int32_t publish_launch(const double* tree, const double* prio, const PublishInfo& pub,
                       cudaStream_t st) {
    per_publish_kernel<<<1, kMaxShards, 0, st>>>(tree, prio, pub.state, pub.peer_slots, pub.G,
                                                 pub.rank);
    return check_launch();
}

// weight[j] /= *wmax (prio.py:101) once the maximum over all shards is known
__global__ void __launch_bounds__(256)
per_normalize_kernel(double* __restrict__ weight, int64_t M, const double* __restrict__ wmax) {
    const int64_t j = blockIdx.x * (int64_t)256 + threadIdx.x;
    if (j < M) weight[j] = __ddiv_rn(weight[j], __ldg(wmax));
}

}  // namespace
}  // namespace ts

using namespace ts;

// one cluster handles the whole batch
template <bool SAMPLE, int SHARD>
static int32_t launch_per_cluster_mode(const double* tree, int64_t bound, int lg, const double* u,
                                       int64_t M, const double* prio, double beta,
                                       int weight_norm, int64_t* idx, double* weight,
                                       int32_t* err_flag, ShardInfo sh, cudaStream_t st,
                                       int rebuild_n = 0) {
    // 16 CTAs x 256 threads spreads a large batch over twice the SMs (7.4 vs 8.7 us at 4096
    // queries); it is an opt-in cluster size, so it is used only where the device can place it.
    static int wide_ok = -1;  // per instantiation
    auto kern16 = per_weight_cluster_kernel<SAMPLE, 256, SHARD>;
    if (wide_ok < 0) {
        wide_ok = 0;
        if (cudaFuncSetAttribute(kern16, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) ==
            cudaSuccess) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(16);
            q.blockDim = dim3(256);
            cudaLaunchAttribute qa[1];
            qa[0].id = cudaLaunchAttributeClusterDimension;
            qa[0].val.clusterDim.x = 16;
            qa[0].val.clusterDim.y = 1;
            qa[0].val.clusterDim.z = 1;
            q.attrs = qa;
            q.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, kern16, &q) == cudaSuccess && n >= 1) wide_ok = 1;
        }
        (void)cudaGetLastError();
    }
    const bool wide = wide_ok == 1 && (g_tune_per_cluster == 16 || (g_tune_per_cluster == 0 && M > 1024));
    auto kern = wide ? kern16 : per_weight_cluster_kernel<SAMPLE, 512, SHARD>;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(wide ? 16 : 8);
    cfg.blockDim = dim3(wide ? 256 : 512);
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = wide ? 16 : 8;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_tune_pdl ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tree, bound, lg, u, M, prio, beta, weight_norm,
                                       idx, weight, err_flag, sh, rebuild_n);
    if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; (void)cudaGetLastError(); return TS_ERR_CUDA; }
    return check_launch();
}

template <bool SAMPLE>
static int32_t launch_per_cluster(const double* tree, int64_t bound, int lg, const double* u,
                                  int64_t M, const double* prio, double beta, int weight_norm,
                                  int64_t* idx, double* weight, int32_t* err_flag,
                                  ShardInfo sh, cudaStream_t st, int rebuild_n = 0) {
    if (sh.slots != nullptr)
        return launch_per_cluster_mode<SAMPLE, 2>(tree, bound, lg, u, M, prio, beta, weight_norm,
                                                  idx, weight, err_flag, sh, st);
    if (sh.roots3 != nullptr)
        return launch_per_cluster_mode<SAMPLE, 1>(tree, bound, lg, u, M, prio, beta, weight_norm,
                                                  idx, weight, err_flag, sh, st);
    return launch_per_cluster_mode<SAMPLE, 0>(tree, bound, lg, u, M, prio, beta, weight_norm, idx,
                                              weight, err_flag, sh, st, rebuild_n);
}

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
    if ((bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (!aligned(tree, 16)) return TS_ERR_ALIGN;
    const unsigned grid = (unsigned)((M + DESC_NT - 1) / DESC_NT);
    cudaStream_t st = (cudaStream_t)stream;
    if (value_dtype == TS_F64)
        segtree_prefix_kernel<double><<<grid, DESC_NT, 0, st>>>(
            tree, bound, ilog2(bound), (double*)value, M, scale_by_total, out, err_flag);
    else if (value_dtype == TS_F32 && !scale_by_total)
        segtree_prefix_kernel<float><<<grid, DESC_NT, 0, st>>>(tree, bound, ilog2(bound),
                                                              (float*)value, M, 0, out, err_flag);
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
    if (!tree || !u || !prio || !idx || !weight || !err_flag) return TS_ERR_BAD_ARG;
    if ((bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (!aligned(tree, 16)) return TS_ERR_ALIGN;
    (void)scratch16;  // kept in the ABI; the cluster kernel needs no global scratch
    return launch_per_cluster<true>(tree, bound, ilog2(bound), u, M, prio, beta, weight_norm, idx,
                                    weight, err_flag, ShardInfo{nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr},
                                    (cudaStream_t)stream);
}

extern "C" int32_t ts_per_weight(const double* tree, int64_t bound, const int64_t* idx, int64_t M,
                                 const double* prio, double beta, int32_t weight_norm,
                                 double* weight, void* scratch16, ts_stream_t stream) {
    if (M < 0 || bound <= 0) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!tree || !idx || !prio || !weight) return TS_ERR_BAD_ARG;
    (void)scratch16;
    return launch_per_cluster<false>(tree, bound, 0, nullptr, M, prio, beta, weight_norm,
                                     const_cast<int64_t*>(idx), weight, nullptr,
                                     ShardInfo{nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr}, (cudaStream_t)stream);
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

// ---- lazily finished tree: update stops at the range roots, the sampler rebuilds the top -------
extern "C" int32_t ts_segtree_lazy_top_nodes(int64_t bound) {
    if (bound <= 0 || (bound & (bound - 1)) != 0) return 0;
    return g_tune_segtree_ranged != 0 ? lazy_top_nodes(bound) : 0;
}

extern "C" int32_t ts_per_update_lazy(double* tree, int64_t size, int64_t bound,
                                      const int64_t* idx, const void* td, int32_t td_dtype,
                                      int64_t M, double alpha, double* prio, uint32_t* stamp,
                                      uint32_t* counter, int32_t* err_flag, ts_stream_t stream) {
    if (!td || !prio) return TS_ERR_BAD_ARG;
    if (td_dtype != TS_F32 && td_dtype != TS_F64) return TS_ERR_BAD_ARG;
    ValueSrc vs{td, 1, alpha, prio, td_dtype == TS_F64 ? VM_PER_F64 : VM_PER_F32};
    return update_impl(tree, size, bound, idx, vs, M, stamp, counter, err_flag,
                       (cudaStream_t)stream, PublishInfo{nullptr, nullptr, 1, 0}, true);
}

extern "C" int32_t ts_per_sample_lazy(double* tree, int64_t bound, const double* u, int64_t M,
                                      const double* prio, double beta, int32_t weight_norm,
                                      int64_t* idx, double* weight, int32_t* err_flag,
                                      ts_stream_t stream) {
    if (M < 0 || bound <= 0) return TS_ERR_BAD_ARG;
    if (!tree || !prio || (bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (!aligned(tree, 16)) return TS_ERR_ALIGN;
    const int n = ts_segtree_lazy_top_nodes(bound);
    if (M == 0) {  // nothing to draw: still leave a finished tree behind
        if (n > 0) segtree_finish_top_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(tree, n);
        return check_launch();
    }
    if (!u || !idx || !weight || !err_flag) return TS_ERR_BAD_ARG;
    return launch_per_cluster<true>(tree, bound, ilog2(bound), u, M, prio, beta, weight_norm, idx,
                                    weight, err_flag,
                                    ShardInfo{nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr},
                                    (cudaStream_t)stream, n);
}

extern "C" int32_t ts_segtree_finish_top(double* tree, int64_t bound, ts_stream_t stream) {
    if (!tree || bound <= 0 || (bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    const int n = ts_segtree_lazy_top_nodes(bound);
    if (n > 0) segtree_finish_top_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(tree, n);
    return check_launch();
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

extern "C" int32_t ts_debug_update_ctas(uint64_t* out512) {
    if (!out512) return TS_ERR_BAD_ARG;
    cudaError_t e = cudaMemcpyFromSymbol(out512, g_update_cta_ns, sizeof(unsigned long long) * 512);
    if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
    return TS_OK;
}

extern "C" int32_t ts_debug_update_phases(uint64_t* out8) {
    uint64_t* out6 = out8;
    if (!out6) return TS_ERR_BAD_ARG;
    unsigned long long h[8];
    cudaError_t e = cudaMemcpyFromSymbol(h, g_update_phase_ns, sizeof(h));
    if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
    for (int i = 0; i < 8; ++i) out6[i] = h[i];
    return TS_OK;
}

extern "C" int32_t ts_per_sample_sharded(const double* tree, int64_t bound, const double* u,
                                         int64_t M, const double* roots3, int32_t G,
                                         int32_t rank, const double* prio, double beta,
                                         int32_t weight_norm, int64_t* idx, double* weight,
                                         double* wmax_out, int32_t* err_flag,
                                         ts_stream_t stream) {
    if (M < 0 || bound <= 0 || G < 1 || G > kMaxShards || (G & (G - 1)) != 0 || rank < 0 ||
        rank >= G)
        return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!tree || !u || !roots3 || !prio || !idx || !weight || !wmax_out || !err_flag)
        return TS_ERR_BAD_ARG;
    if ((bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (!aligned(tree, 16)) return TS_ERR_ALIGN;
    return launch_per_cluster<true>(tree, bound, ilog2(bound), u, M, prio, beta, weight_norm, idx,
                                    weight, err_flag, ShardInfo{roots3, G, rank, wmax_out, nullptr, nullptr, nullptr},
                                    (cudaStream_t)stream);
}

extern "C" int32_t ts_per_normalize(double* weight, int64_t M, const double* wmax,
                                    ts_stream_t stream) {
    if (M < 0) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!weight || !wmax) return TS_ERR_BAD_ARG;
    per_normalize_kernel<<<(unsigned)((M + 255) / 256), 256, 0, (cudaStream_t)stream>>>(weight, M,
                                                                                       wmax);
    return check_launch();
}

extern "C" int32_t ts_per_sample_p2p_owned(const double* tree, int64_t bound, const double* u,
                                           int64_t M, double* slots, double* const* peer_slots,
                                           uint64_t* state, int32_t G, int32_t rank, double* prio,
                                           double beta, int32_t weight_norm, int64_t* idx,
                                           double* weight, int32_t* owned, int32_t* err_flag,
                                           ts_stream_t stream) {
    if (M < 0 || bound <= 0 || G < 1 || G > kMaxShards || (G & (G - 1)) != 0 || rank < 0 ||
        rank >= G)
        return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;  // NOTE: every rank passes the same M, so all skip together
    if (!tree || !u || !slots || !peer_slots || !state || !prio || !idx || !weight || !err_flag)
        return TS_ERR_BAD_ARG;
    if ((bound & (bound - 1)) != 0) return TS_ERR_BAD_ARG;
    if (!aligned(tree, 16) || !aligned(slots, 8)) return TS_ERR_ALIGN;
    return launch_per_cluster<true>(
        tree, bound, ilog2(bound), u, M, prio, beta, weight_norm, idx, weight, err_flag,
        ShardInfo{nullptr, G, rank, nullptr, slots, peer_slots, (unsigned long long*)state, owned},
        (cudaStream_t)stream);
}

extern "C" int32_t ts_per_sample_p2p(const double* tree, int64_t bound, const double* u, int64_t M,
                                     double* slots, double* const* peer_slots, uint64_t* state,
                                     int32_t G, int32_t rank, double* prio, double beta,
                                     int32_t weight_norm, int64_t* idx, double* weight,
                                     int32_t* err_flag, ts_stream_t stream) {
    return ts_per_sample_p2p_owned(tree, bound, u, M, slots, peer_slots, state, G, rank, prio, beta,
                                   weight_norm, idx, weight, nullptr, err_flag, stream);
}

// ---- measurement only: flag round trip between two GPUs over peer memory ------------------------
namespace ts { namespace {
// One thread per GPU.  Rank 0 sends i, waits for the echo; rank 1 waits for i, echoes it.  mode 0:
// the exchange's own primitives (st.release.sys flag, ld.acquire.sys poll with the 64 ns back-off of
// wait_epoch); mode 1: a full shard record as publish_record writes it (three relaxed stores, a
// system fence, the release flag); mode 2: mode 0 with a tight poll; mode 3: the record WITHOUT the
// fence (the release store alone orders this thread's earlier stores).
__global__ void pingpong_kernel(double* mine, double* const* peers, int peer, int iters,
                                int initiator, int mode, unsigned long long* out_ns) {
    double* remote = peers[peer];
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    bool ok = true;
    for (int i = 1; i <= iters && ok; ++i) {
        const unsigned long long tag = (unsigned long long)i;
        auto send = [&]() {
            if (mode == 1 || mode == 3) {
                st_relaxed_sys_f64(remote, (double)i);
                st_relaxed_sys_f64(remote + 1, 1.0);
                st_relaxed_sys_f64(remote + 2, 2.0);
                if (mode == 1) __threadfence_system();
            }
            st_release_sys_u64(remote + 3, tag);
        };
        auto recv = [&]() {
            if (mode == 2) {
                int spin = 0;
                while (ld_acquire_sys_u64(mine + 3) != tag)
                    if (++spin > (1 << 26)) return false;
                return true;
            }
            return wait_epoch(mine + 3, tag);
        };
        if (initiator) { send(); ok = recv(); } else { ok = recv(); if (ok) send(); }
    }
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    out_ns[0] = ok ? t1 - t0 : 0ull;
}
} }  // namespace ts::

extern "C" int32_t ts_debug_pingpong(double* mine, double* const* peers, int32_t peer,
                                     int32_t iters, int32_t initiator, int32_t mode,
                                     uint64_t* out_ns, ts_stream_t stream) {
    if (!mine || !peers || !out_ns || iters < 1 || mode < 0 || mode > 3) return TS_ERR_BAD_ARG;
    pingpong_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(mine, peers, peer, iters, initiator, mode,
                                                       (unsigned long long*)out_ns);
    return check_launch();
}

extern "C" int32_t ts_per_publish(const double* tree, const double* prio, uint64_t* state,
                                  double* const* peer_slots, int32_t G, int32_t rank,
                                  ts_stream_t stream) {
    if (!tree || !prio || !state || !peer_slots || G < 1 || G > kMaxShards || rank < 0 ||
        rank >= G)
        return TS_ERR_BAD_ARG;
    return publish_launch(tree, prio, PublishInfo{(unsigned long long*)state, peer_slots, G, rank},
                          (cudaStream_t)stream);
}

extern "C" int32_t ts_per_update_publish_owned(double* tree, int64_t size, int64_t bound,
                                               const int64_t* idx, const void* td,
                                               int32_t td_dtype, int64_t M, double alpha,
                                               double* prio, uint32_t* stamp, uint32_t* counter,
                                               int32_t* err_flag, uint64_t* state,
                                               double* const* peer_slots, int32_t G, int32_t rank,
                                               const int32_t* owned, ts_stream_t stream) {
    if (!td || !prio || !state || !peer_slots || G < 1 || G > kMaxShards || rank < 0 || rank >= G)
        return TS_ERR_BAD_ARG;
    if (td_dtype != TS_F32 && td_dtype != TS_F64) return TS_ERR_BAD_ARG;
    ValueSrc vs{td, 1, alpha, prio, td_dtype == TS_F64 ? VM_PER_F64 : VM_PER_F32};
    return update_impl(tree, size, bound, idx, vs, M, stamp, counter, err_flag,
                       (cudaStream_t)stream,
                       PublishInfo{(unsigned long long*)state, peer_slots, G, rank, owned});
}

extern "C" int32_t ts_per_update_publish(double* tree, int64_t size, int64_t bound,
                                         const int64_t* idx, const void* td, int32_t td_dtype,
                                         int64_t M, double alpha, double* prio, uint32_t* stamp,
                                         uint32_t* counter, int32_t* err_flag, uint64_t* state,
                                         double* const* peer_slots, int32_t G, int32_t rank,
                                         ts_stream_t stream) {
    return ts_per_update_publish_owned(tree, size, bound, idx, td, td_dtype, M, alpha, prio, stamp,
                                       counter, err_flag, state, peer_slots, G, rank, nullptr,
                                       stream);
}
