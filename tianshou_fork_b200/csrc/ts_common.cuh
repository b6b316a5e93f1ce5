// Shared device/host helpers for the sm_100a kernels behind include/ts_b200.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ts_b200.h"

namespace ts {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// Last cudaError_t seen by this host thread (ts_last_cuda_error()).
extern thread_local int32_t g_last_cuda_error;

inline int32_t check_launch() {
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        g_last_cuda_error = (int32_t)e;
        (void)cudaGetLastError();
        return TS_ERR_CUDA;
    }
    return TS_OK;
}

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// numpy-style modulo for a positive modulus.
__host__ __device__ __forceinline__ int64_t pymod(int64_t a, int64_t m) {
    int64_t r = a % m;
    return r < 0 ? r + m : r;
}

// Sub-buffer of a (already wrapped) global index: division when all sub-buffers are equal
// (vecbuf.py:35), else binary search over _extend_offset (manager.py:38).
__device__ __forceinline__ int64_t sub_of(int64_t idx, const int64_t* __restrict__ offset,
                                          int64_t B, int64_t sub_size) {
    // A single ring has nothing to look up.  Its `%` is by the CURRENT size (base.py:158-168),
    // which kernels read from the device bookkeeping (offset[1], lengths[0]); `sub_size` is a
    // by-value launch argument that a captured CUDA graph would freeze while the ring fills.
    if (B == 1) return 0;
    if (sub_size > 0) {
        if ((((uint64_t)idx | (uint64_t)sub_size) >> 32) == 0)  // the usual case: 32-bit divide
            return (int64_t)((uint32_t)idx / (uint32_t)sub_size);
        return idx / sub_size;
    }
    int64_t lo = 0, hi = B;  // offset[lo] <= idx < offset[hi]
    while (hi - lo > 1) {
        int64_t mid = (lo + hi) >> 1;
        if (__ldg(offset + mid) <= idx) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ double shfl_down_f64(double v, int d) {
    return __shfl_down_sync(0xffffffffu, v, d);
}
__device__ __forceinline__ double shfl_idx_f64(double v, int l) {
    return __shfl_sync(0xffffffffu, v, l);
}

// Streaming (evict-first) vector accesses: every transition is touched exactly once.
__device__ __forceinline__ float4 ld_stream(const float4* p) { return __ldcs(p); }
__device__ __forceinline__ double2 ld_stream(const double2* p) { return __ldcs(p); }
__device__ __forceinline__ uint2 ld_stream(const uint2* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(float4* p, float4 v) { __stcs(p, v); }
__device__ __forceinline__ void st_stream(double2* p, double2 v) { __stcs(p, v); }

// ---- programmatic dependent launch (griddepcontrol) ---------------------------------------------
// A kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start while its
// predecessor in the stream is still running: launch latency and the prologue that touches no
// global data overlap the predecessor's tail.  pdl_wait() returns once the predecessor grid has
// completed and its writes are visible; every global read of data an earlier kernel may have
// produced comes after it.  Both are no-ops under an ordinary launch.
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- sm_100a bulk-copy engine (TMA, 1-D form) + mbarrier -------------------------------------
// cp.async.bulk moves a contiguous run of bytes global<->shared without touching registers;
// loads complete on an mbarrier (transaction bytes), stores are tracked by bulk groups.
// Sizes are multiples of 16 B and both addresses are 16-B aligned.
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
// non-blocking: has the phase with this parity completed?
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void prefetch_l1(const void* p) {
    asm volatile("prefetch.global.L1 [%0];" :: "l"(p));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
// L2 eviction-priority policy for data that is touched exactly once (TS_HINT: bit 0 = loads,
// bit 1 = stores; build-time experiment knob)
#ifndef TS_HINT
#define TS_HINT 0
#endif
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                         uint32_t bar) {
#if TS_HINT & 1
    const uint64_t pol = l2_evict_first_policy();
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1], %2, [%3], %4;"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol)
        : "memory");
#else
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
        : "memory");
#endif
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
#if TS_HINT & 2
    const uint64_t pol = l2_evict_first_policy();
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 ::"l"(dst), "r"(src), "r"(bytes), "l"(pol)
                 : "memory");
#else
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
                 "r"(src), "r"(bytes)
                 : "memory");
#endif
}
__device__ __forceinline__ void bulk_commit() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// all earlier bulk stores have finished READING their shared-memory source
__device__ __forceinline__ void bulk_wait_read0() {
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait0() {
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (before a bulk store)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// named barriers (1..15; 0 is __syncthreads): `threads` = arrivals that complete it, a multiple of 32
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Tuning knobs (ts_tune_set): 0 = automatic choice.
extern int g_tune_gae_rows_variant;
extern int g_tune_gae_flat_variant;  // 0 auto, 1 decoupled look-back, 2 pipelined ranges
extern int g_tune_gae_flat_ctas;     // pipelined ranges: CTAs per SM (0 auto, 1..3)
extern int g_tune_gae_flat_stages;   // pipelined ranges: stages of the ring (0 auto, 2..6)
extern int g_tune_segtree_max_ctas;  // CTAs of the cooperative tree update (<= 148)
extern int g_tune_per_cluster;       // PER sampling cluster: 0 auto, 8 (x512 threads) or 16 (x256)
extern int g_tune_segtree_ranged;  // 1 (default): per-CTA leaf ranges with a shared-memory table; 0: tiered
extern int g_tune_gae_rows_packed;  // 1 (default): short contiguous rows go through the pipelined-ranges kernel
extern int g_tune_gae_balance;  // 1: persistent GAE grids sized so that all CTAs get equal work (measured slower: default 0)
extern int g_tune_pdl;  // 1 (default): PER and GAE rows kernels are launched with programmatic stream serialization
extern int g_tune_segtree_group;  // 128 (default) or 256 nodes per warp-rebuilt group

}  // namespace ts
