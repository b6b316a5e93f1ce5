// Row gather: out[j, :] = src[index[j], :]  (buffer[indices], data/buffer/base.py:360-389).
// Pure byte movement; one warp per row with 16-byte accesses when the row pitch allows,
// several rows per warp for short rows.
#include "ts_common.cuh"

namespace ts {
namespace {

template <typename V>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const V* __restrict__ src, int64_t n_rows, int64_t row_vecs,
                   const int64_t* __restrict__ index, int64_t M, V* __restrict__ out) {
    // lanes_per_row: smallest power of two >= row_vecs, capped at 32
    int lpr = 1;
    while (lpr < 32 && lpr < row_vecs) lpr <<= 1;
    const int rows_per_warp = 32 / lpr;
    const int lane = threadIdx.x & 31;
    const int sub = lane / lpr, l = lane % lpr;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t j = warp * rows_per_warp + sub; j < M; j += nwarps * rows_per_warp) {
        int64_t r = index[j];
        if (r < 0) r += n_rows;  // numpy-style negative index
        const V* s = src + r * row_vecs;
        V* d = out + j * row_vecs;
        for (int64_t c = l; c < row_vecs; c += lpr) d[c] = __ldg(s + c);
    }
}

// ============================================================================================
// Frame-stack gather (ReplayBuffer.get with stack_num > 1, tianshou/data/buffer/base.py:317-358,
// and the obs_next of ignore_obs_next buffers, :374-378): for sample j the frames
//     g[k-1] = index[j],  g[i-1] = prev(g[i])                 (episode-aware, manager.py:259-277)
// are stacked oldest first into out[j, 0..k-1, :].  With `out_next` the same launch also writes
// the stack that ends at n = next(index[j]): it is either the same stack (index[j] closes an
// episode, n == index[j]) or the stack shifted by one frame (prev(n) == index[j]), so k + 1 frames
// are read for 2k frames written.
//
// The prev chain moves back by at most one slot per step and stays inside one sub-buffer, so the
// k - 1 end flags it can meet are fetched in ONE parallel round (a lane each) and the chain is
// resolved in registers: two dependent round trips per sample (ring bookkeeping, flags).  The
// one case where the second stack is not a shift of the first -- index[j] is an unfilled slot
// beyond the ring's current length, so next() wraps into the valid part -- is resolved on its own.
// ============================================================================================
constexpr int kMaxStack = 16;

struct RingArgs {
    const int64_t* offset;      // [B + 1]
    const uint8_t* done;        // [total]
    const int64_t* last_index;  // [B]
    const int64_t* lengths;     // [B]
    int64_t B, sub_size;
};

// One warp: the k - 1 end flags a prev chain starting at ring position `rel` (slot `first`) can
// meet, plus the flag of `first` itself in bit 0 (for next()).
__device__ __forceinline__ unsigned int chain_flags(const RingArgs& r, int64_t start, int64_t len,
                                                    int64_t last, int64_t rel, int64_t first,
                                                    int k, int lane) {
    const int64_t pos = lane == 0 ? first : pymod(rel - lane, len) + start;
    int endflag = 0;
    if (lane < k) endflag = (__ldg(r.done + pos) != 0) | (pos == last);
    return __ballot_sync(0xffffffffu, endflag != 0);
}
// One lane: g[k-1] = first, g[i-1] = prev(g[i]) from the flags above (manager.py:274-276).
__device__ __forceinline__ void chain_resolve(int64_t start, int64_t len, int64_t rel,
                                              int64_t first, unsigned int ends, int k,
                                              int64_t* g) {
    int64_t c = rel;
    int t = 1;
    g[k - 1] = first;
    for (int i = k - 2; i >= 0; --i) {
        const int64_t sb = pymod(c - 1, len);              // :274
        if ((ends >> t) & 1u) {
            c = pymod(sb + 1, len);                        // :276 with end_flag = 1
        } else {
            c = sb;
            ++t;
        }
        g[i] = c + start;
    }
}

// One warp: frames of sample `raw` into s_g[0..k-1]; s_g[k] = next(idx).  Returns (in every lane)
// how the stack that ends at next(idx) relates to it: 0 = the same frames (idx closes an episode),
// 1 = shifted by one (s_g[1..k]), 2 = unrelated (idx is an unfilled slot beyond the ring's current
// length, so next() wraps into the valid part): that chain is resolved on its own into s_g2.
__device__ __forceinline__ int stack_chain(const RingArgs& r, int64_t raw, int k, bool want_next,
                                           int lane, int64_t* s_g, int64_t* s_g2) {
    const int64_t total = __ldg(r.offset + r.B);
    const int64_t idx = pymod(raw, total);
    const int64_t sub = sub_of(idx, r.offset, r.B, r.sub_size);
    const int64_t start = __ldg(r.offset + sub);
    int64_t len = __ldg(r.lengths + sub);
    len = len > 1 ? len : 1;                              // max(1, cur_len)  manager.py:271
    const int64_t last = __ldg(r.last_index + sub);
    const int64_t rel = idx - start;
    const unsigned int ends = chain_flags(r, start, len, last, rel, idx, k, lane);
    const int64_t n = pymod(rel + 1 - (int64_t)(ends & 1u), len) + start;  // next(): :295-296
    if (lane == 0) {
        chain_resolve(start, len, rel, idx, ends, k, s_g);
        s_g[k] = n;
    }
    int mode = n == idx ? 0 : 1;
    if (want_next && mode == 1 && rel >= len) {           // warp-uniform
        mode = 2;
        const unsigned int ends2 = chain_flags(r, start, len, last, n - start, n, k, lane);
        if (lane == 0) chain_resolve(start, len, n - start, n, ends2, k, s_g2);
    }
    __syncwarp();
    return mode;
}

// bulk-copy form: one warp per CTA, frames staged in shared memory
__global__ void __launch_bounds__(32)
gather_stack_tma_kernel(const unsigned char* __restrict__ src, int64_t row_bytes,
                        const int64_t* __restrict__ index, int64_t M, int k, RingArgs r,
                        unsigned char* __restrict__ out, unsigned char* __restrict__ out_next) {
    extern __shared__ __align__(128) unsigned char s_frames[];  // (k or k + 1) rows
    __shared__ int64_t s_g[kMaxStack + 1], s_g2[kMaxStack];
    __shared__ __align__(8) unsigned long long s_bar;
    const int lane = threadIdx.x;
    const uint32_t bar = smem_u32(&s_bar);
    const bool want_next = out_next != nullptr;
    if (lane == 0) {
        mbar_init(bar, 1u);
        fence_mbar_init();
    }
    __syncwarp();
    uint32_t parity = 0;
    for (int64_t j = blockIdx.x; j < M; j += gridDim.x) {
        const int mode = stack_chain(r, index[j], k, want_next, lane, s_g, s_g2);
        const int nf = want_next && mode == 1 ? k + 1 : k;
        if (lane == 0) {
            bulk_wait_read0();  // the previous sample's stores have left the staging buffer
            mbar_expect_tx(bar, (uint32_t)(nf * row_bytes));
            for (int f = 0; f < nf; ++f)
                bulk_g2s(smem_u32(s_frames + f * row_bytes), src + s_g[f] * row_bytes,
                         (uint32_t)row_bytes, bar);
        }
        mbar_wait(bar, parity);
        parity ^= 1u;
        if (lane == 0) {
            // frames 0..k-1 are contiguous in shared memory and in the output: one store each
            bulk_s2g(out + j * k * row_bytes, smem_u32(s_frames), (uint32_t)(k * row_bytes));
            if (want_next && mode != 2)
                bulk_s2g(out_next + j * k * row_bytes,
                         smem_u32(s_frames + (mode == 1 ? row_bytes : 0)),
                         (uint32_t)(k * row_bytes));
            bulk_commit();
        }
        if (want_next && mode == 2) {  // the rare unrelated stack: a second pass through the buffer
            if (lane == 0) {
                bulk_wait_read0();
                mbar_expect_tx(bar, (uint32_t)(k * row_bytes));
                for (int f = 0; f < k; ++f)
                    bulk_g2s(smem_u32(s_frames + f * row_bytes), src + s_g2[f] * row_bytes,
                             (uint32_t)row_bytes, bar);
            }
            mbar_wait(bar, parity);
            parity ^= 1u;
            if (lane == 0) {
                bulk_s2g(out_next + j * k * row_bytes, smem_u32(s_frames),
                         (uint32_t)(k * row_bytes));
                bulk_commit();
            }
        }
        __syncwarp();
    }
    if (lane == 0) bulk_wait0();
}

// register form for rows the bulk-copy engine cannot take (size or alignment)
template <typename V>
__global__ void __launch_bounds__(256)
gather_stack_kernel(const V* __restrict__ src, int64_t row_vecs, const int64_t* __restrict__ index,
                    int64_t M, int k, RingArgs r, V* __restrict__ out, V* __restrict__ out_next) {
    __shared__ int64_t s_g[kMaxStack + 1], s_g2[kMaxStack];
    __shared__ int s_mode;
    const int tid = threadIdx.x;
    for (int64_t j = blockIdx.x; j < M; j += gridDim.x) {
        if (tid < 32) {
            const int mode = stack_chain(r, index[j], k, out_next != nullptr, tid, s_g, s_g2);
            if (tid == 0) s_mode = mode;
        }
        __syncthreads();
        const int mode = s_mode;
        for (int64_t c = tid; c < (int64_t)k * row_vecs; c += 256) {
            const int f = (int)(c / row_vecs);
            const int64_t w = c - (int64_t)f * row_vecs;
            out[j * k * row_vecs + c] = __ldg(src + s_g[f] * row_vecs + w);
            if (out_next != nullptr) {
                const int64_t g = mode == 2 ? s_g2[f] : s_g[f + mode];
                out_next[j * k * row_vecs + c] = __ldg(src + g * row_vecs + w);
            }
        }
        __syncthreads();
    }
}

// ============================================================================================
// Multi-field gather: the small per-transition fields of buffer[indices] (act, rew, terminated,
// truncated, done, ...; data/buffer/base.py:379-389) in ONE launch instead of one indexing
// launch per key.  The field table travels in the kernel parameters.  One thread per
// (row, field); rows here are a few bytes (scalars, small action vectors), wide rows go through
// ts_gather_rows / ts_gather_stack.
// ============================================================================================
constexpr int kMaxFields = 8;
struct FieldArgs {
    const uint8_t* src[kMaxFields];
    uint8_t* dst[kMaxFields];
    int32_t row_bytes[kMaxFields];
    int32_t unit[kMaxFields];  // widest access the pointers and the pitch allow: 1, 2, 4 or 8
    int32_t n;
};

__global__ void __launch_bounds__(256)
gather_fields_kernel(FieldArgs a, int64_t n_rows, const int64_t* __restrict__ index, int64_t M) {
    const int64_t total = M * a.n;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
        const int f = (int)(t % a.n);
        const int64_t j = t / a.n;
        int64_t r = __ldg(index + j);
        if (r < 0) r += n_rows;  // numpy-style negative index
        const int rb = a.row_bytes[f];
        const uint8_t* s = a.src[f] + r * rb;
        uint8_t* d = a.dst[f] + j * rb;
        switch (a.unit[f]) {
            case 8:
                for (int c = 0; c < rb; c += 8)
                    *(uint2*)(d + c) = __ldg((const uint2*)(s + c));
                break;
            case 4:
                for (int c = 0; c < rb; c += 4)
                    *(uint32_t*)(d + c) = __ldg((const uint32_t*)(s + c));
                break;
            case 2:
                for (int c = 0; c < rb; c += 2)
                    *(uint16_t*)(d + c) = __ldg((const uint16_t*)(s + c));
                break;
            default:
                for (int c = 0; c < rb; ++c) d[c] = __ldg(s + c);
        }
    }
}

// The inverse of gather_fields_kernel: dst[f][index[j], :] = src[f][j, :] -- one vectorised env
// step written into the storage of every key in one launch (manager.py:160 does one fancy-index
// assignment per key).  Indices are distinct (one slot per sub-buffer per step).
__global__ void __launch_bounds__(256)
scatter_fields_kernel(FieldArgs a, int64_t n_rows, const int64_t* __restrict__ index, int64_t M) {
    // one WARP per (row, field): rows range from one flag byte to a 7 KB Atari frame
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t total = M * a.n;
    for (int64_t t = warp; t < total; t += nwarps) {
        const int f = (int)(t % a.n);
        const int64_t j = t / a.n;
        int64_t r = __ldg(index + j);
        if (r < 0) r += n_rows;
        if (r < 0 || r >= n_rows) continue;  // never write outside the storage
        const int rb = a.row_bytes[f];
        const uint8_t* s = a.src[f] + j * rb;
        uint8_t* d = a.dst[f] + r * rb;
        switch (a.unit[f]) {
            case 8:
                for (int c = lane * 8; c < rb; c += 256)
                    *(uint2*)(d + c) = __ldg((const uint2*)(s + c));
                break;
            case 4:
                for (int c = lane * 4; c < rb; c += 128)
                    *(uint32_t*)(d + c) = __ldg((const uint32_t*)(s + c));
                break;
            case 2:
                for (int c = lane * 2; c < rb; c += 64)
                    *(uint16_t*)(d + c) = __ldg((const uint16_t*)(s + c));
                break;
            default:
                for (int c = lane; c < rb; c += 32) d[c] = __ldg(s + c);
        }
    }
}

}  // namespace
}  // namespace ts

using namespace ts;

extern "C" int32_t ts_gather_stack(const void* src, int64_t row_bytes, const int64_t* index,
                                   int64_t M, int32_t stack_num, const int64_t* offset, int64_t B,
                                   int64_t sub_size, const uint8_t* done,
                                   const int64_t* last_index, const int64_t* lengths, void* out,
                                   void* out_next, ts_stream_t stream) {
    if (M < 0 || row_bytes < 0 || B <= 0 || stack_num < 1 || stack_num > kMaxStack)
        return TS_ERR_BAD_ARG;
    if (M == 0 || row_bytes == 0) return TS_OK;
    if (!src || !index || !offset || !done || !last_index || !lengths || !out) return TS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const RingArgs r{offset, done, last_index, lengths, B, sub_size};
    const int k = stack_num;
    const size_t smem = (size_t)(out_next ? k + 1 : k) * (size_t)row_bytes;
    const bool a16 = row_bytes % 16 == 0 && aligned(src, 16) && aligned(out, 16) &&
                     (out_next == nullptr || aligned(out_next, 16));
    if (a16 && smem <= 200 * 1024 && (size_t)k * (size_t)row_bytes < (1u << 20)) {
        static size_t attr_bytes = 0;
        if (smem > attr_bytes) {
            cudaError_t e = cudaFuncSetAttribute(gather_stack_tma_kernel,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 200 * 1024);
            if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
            attr_bytes = 200 * 1024;
        }
        // as many single-warp CTAs per SM as the staging buffers allow (at most 8)
        int per_sm = (int)((220 * 1024) / (smem + 1024));
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        int64_t grid = (int64_t)kNumSMs * per_sm;
        if (grid > M) grid = M;
        gather_stack_tma_kernel<<<(unsigned)grid, 32, smem, st>>>(
            (const unsigned char*)src, row_bytes, index, M, k, r, (unsigned char*)out,
            (unsigned char*)out_next);
        return check_launch();
    }
    int64_t grid = (int64_t)kNumSMs * 8;
    if (grid > M) grid = M;
#define TS_GS(V)                                                                               \
    gather_stack_kernel<V><<<(unsigned)grid, 256, 0, st>>>(                                    \
        (const V*)src, row_bytes / (int64_t)sizeof(V), index, M, k, r, (V*)out, (V*)out_next)
    if (a16) TS_GS(uint4);
    else if (row_bytes % 4 == 0 && aligned(src, 4) && aligned(out, 4) &&
             (out_next == nullptr || aligned(out_next, 4))) TS_GS(uint32_t);
    else TS_GS(uint8_t);
#undef TS_GS
    return check_launch();
}

extern "C" int32_t ts_gather_rows(const void* src, int64_t n_rows, int64_t row_bytes,
                                  const int64_t* index, int64_t M, void* out,
                                  ts_stream_t stream) {
    if (M < 0 || n_rows < 0 || row_bytes < 0) return TS_ERR_BAD_ARG;
    if (M == 0 || row_bytes == 0) return TS_OK;
    if (!src || !index || !out) return TS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t warps = M;
    int64_t blocks = (warps * 32 + 255) / 256;
    const int64_t cap = (int64_t)kNumSMs * 32;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (row_bytes % 16 == 0 && aligned(src, 16) && aligned(out, 16))
        gather_rows_kernel<uint4><<<(unsigned)blocks, 256, 0, st>>>(
            (const uint4*)src, n_rows, row_bytes / 16, index, M, (uint4*)out);
    else if (row_bytes % 8 == 0 && aligned(src, 8) && aligned(out, 8))
        gather_rows_kernel<uint2><<<(unsigned)blocks, 256, 0, st>>>(
            (const uint2*)src, n_rows, row_bytes / 8, index, M, (uint2*)out);
    else if (row_bytes % 4 == 0 && aligned(src, 4) && aligned(out, 4))
        gather_rows_kernel<uint32_t><<<(unsigned)blocks, 256, 0, st>>>(
            (const uint32_t*)src, n_rows, row_bytes / 4, index, M, (uint32_t*)out);
    else
        gather_rows_kernel<uint8_t><<<(unsigned)blocks, 256, 0, st>>>(
            (const uint8_t*)src, n_rows, row_bytes, index, M, (uint8_t*)out);
    return check_launch();
}

static int32_t launch_fields(bool scatter, int32_t n_fields, const void* const* src,
                             void* const* dst, const int64_t* row_bytes, int64_t n_rows,
                             const int64_t* index, int64_t M, ts_stream_t stream) {
    if (n_fields < 0 || n_fields > kMaxFields || M < 0 || n_rows < 0) return TS_ERR_BAD_ARG;
    if (n_fields == 0 || M == 0) return TS_OK;
    if (!src || !dst || !row_bytes || !index) return TS_ERR_BAD_ARG;
    FieldArgs a;
    a.n = 0;
    for (int f = 0; f < n_fields; ++f) {
        if (row_bytes[f] < 0 || row_bytes[f] > (1 << 20)) return TS_ERR_BAD_ARG;
        if (row_bytes[f] == 0) continue;
        if (!src[f] || !dst[f]) return TS_ERR_BAD_ARG;
        const int k = a.n++;
        a.src[k] = (const uint8_t*)src[f];
        a.dst[k] = (uint8_t*)dst[f];
        a.row_bytes[k] = (int32_t)row_bytes[f];
        int unit = 8;
        while (unit > 1 && (row_bytes[f] % unit != 0 || !aligned(src[f], unit) ||
                            !aligned(dst[f], unit)))
            unit >>= 1;
        a.unit[k] = unit;
    }
    if (a.n == 0) return TS_OK;
    const int64_t total = scatter ? M * a.n * 32 : M * a.n;  // scatter: a warp per item
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)kNumSMs * 8;
    if (blocks > cap) blocks = cap;
    if (scatter)
        scatter_fields_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, n_rows, index, M);
    else
        gather_fields_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, n_rows, index, M);
    return check_launch();
}

extern "C" int32_t ts_gather_fields(int32_t n_fields, const void* const* src, void* const* dst,
                                    const int64_t* row_bytes, int64_t n_rows,
                                    const int64_t* index, int64_t M, ts_stream_t stream) {
    return launch_fields(false, n_fields, src, dst, row_bytes, n_rows, index, M, stream);
}

extern "C" int32_t ts_scatter_fields(int32_t n_fields, const void* const* src, void* const* dst,
                                     const int64_t* row_bytes, int64_t n_rows,
                                     const int64_t* index, int64_t M, ts_stream_t stream) {
    return launch_fields(true, n_fields, src, dst, row_bytes, n_rows, index, M, stream);
}
