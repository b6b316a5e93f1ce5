// Episode-aware neighbour indices: prev / next / unfinished_index as integer kernels.
//
// Replaces numba _prev_index / _next_index (tianshou/data/buffer/manager.py:259-297), whose
// cost is O(buffer_num * M) (a boolean mask over all M indices per sub-buffer), with one
// thread per index: O(M), sub-buffer found by division (equal sub-buffers, vecbuf.py:35) or
// binary search over _extend_offset.  The single-ring ReplayBuffer.prev/next
// (tianshou/data/buffer/base.py:153-168) is the B == 1 case with lengths[0] = current size
// (for a non-empty ring both take `%` by the current size).
#include "ts_common.cuh"

namespace ts {
namespace {

template <bool NEXT>
__global__ void __launch_bounds__(256)
neighbour_kernel(const int64_t* __restrict__ index, int64_t M,
                 const int64_t* __restrict__ offset, int64_t B, int64_t sub_size,
                 const uint8_t* __restrict__ done, const int64_t* __restrict__ last_index,
                 const int64_t* __restrict__ lengths, int64_t* __restrict__ out) {
    const int64_t total = __ldg(offset + B);
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M;
         j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t idx = pymod(index[j], total);          // manager.py:267 / :288
        const int64_t k = sub_of(idx, offset, B, sub_size);
        const int64_t start = __ldg(offset + k);
        int64_t cur_len = __ldg(lengths + k);
        cur_len = cur_len > 1 ? cur_len : 1;                 // max(1, cur_len)  :271 / :292
        const int64_t last = __ldg(last_index + k);
        int64_t r;
        if (NEXT) {
            const int64_t end_flag = (done[idx] != 0) | (idx == last);           // :295
            r = pymod(idx - start + 1 - end_flag, cur_len) + start;              // :296
        } else {
            const int64_t sub = pymod(idx - start - 1, cur_len);                 // :274
            const int64_t end_flag = (done[sub + start] != 0) | (sub + start == last);  // :275
            r = pymod(sub + end_flag, cur_len) + start;                          // :276
        }
        out[j] = r;
    }
}

__global__ void unfinished_kernel(const uint8_t* __restrict__ done,
                                  const int64_t* __restrict__ last_index,
                                  const int64_t* __restrict__ lengths, int64_t B,
                                  int64_t* __restrict__ out) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < B;
         k += (int64_t)gridDim.x * blockDim.x) {
        // base.py:150-151: last = (_index - 1) % _size, which is last_index once size > 0
        const int64_t last = last_index[k];
        out[k] = (lengths[k] > 0 && !done[last]) ? last : -1;
    }
}

inline unsigned grid_for(int64_t n, int threads) {
    int64_t g = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * 16;
    return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

// Unfinished tails as cut flags at BATCH positions, for a batch that is the buffer's own
// chronological listing (sample_indices(0), manager.py:187-192): sub-buffer k occupies batch
// positions run_first[k] + j with buffer index (ring_start[k] + j) % lengths[k] + offset[k].
// Replaces `end_flag[np.isin(indices, buffer.unfinished_index())] = True`
// (policy/base.py:379-380) without touching the N indices: B look-ups instead of N.
__global__ void mark_unfinished_kernel(const uint8_t* __restrict__ done,
                                       const int64_t* __restrict__ last_index,
                                       const int64_t* __restrict__ lengths,
                                       const int64_t* __restrict__ offset,
                                       const int64_t* __restrict__ ring_start,
                                       const int64_t* __restrict__ run_first, int64_t B,
                                       int64_t N, uint8_t* __restrict__ flags) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= B) return;
    const int64_t len = lengths[k];
    if (len <= 0) return;
    const int64_t last = last_index[k];
    if (done[last] != 0) return;  // the newest episode of this sub-buffer is finished
    const int64_t rel = pymod(last - offset[k] - ring_start[k], len);
    const int64_t pos = run_first[k] + rel;
    if (pos >= 0 && pos < N) flags[pos] = 1;
}

}  // namespace
}  // namespace ts

using namespace ts;

static int32_t neighbour(bool next, const int64_t* index, int64_t M, const int64_t* offset,
                         int64_t B, int64_t sub_size, const uint8_t* done,
                         const int64_t* last_index, const int64_t* lengths, int64_t* out,
                         ts_stream_t stream) {
    if (M < 0 || B <= 0) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!index || !offset || !done || !last_index || !lengths || !out) return TS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (next)
        neighbour_kernel<true><<<grid_for(M, 256), 256, 0, st>>>(index, M, offset, B, sub_size,
                                                                 done, last_index, lengths, out);
    else
        neighbour_kernel<false><<<grid_for(M, 256), 256, 0, st>>>(index, M, offset, B, sub_size,
                                                                  done, last_index, lengths, out);
    return check_launch();
}

extern "C" int32_t ts_prev_index(const int64_t* index, int64_t M, const int64_t* offset,
                                 int64_t B, int64_t sub_size, const uint8_t* done,
                                 const int64_t* last_index, const int64_t* lengths, int64_t* out,
                                 ts_stream_t stream) {
    return neighbour(false, index, M, offset, B, sub_size, done, last_index, lengths, out, stream);
}

extern "C" int32_t ts_next_index(const int64_t* index, int64_t M, const int64_t* offset,
                                 int64_t B, int64_t sub_size, const uint8_t* done,
                                 const int64_t* last_index, const int64_t* lengths, int64_t* out,
                                 ts_stream_t stream) {
    return neighbour(true, index, M, offset, B, sub_size, done, last_index, lengths, out, stream);
}

extern "C" int32_t ts_unfinished_index(const uint8_t* done, const int64_t* last_index,
                                       const int64_t* lengths, int64_t B, int64_t* out,
                                       ts_stream_t stream) {
    if (B <= 0 || !done || !last_index || !lengths || !out) return TS_ERR_BAD_ARG;
    unfinished_kernel<<<grid_for(B, 128), 128, 0, (cudaStream_t)stream>>>(done, last_index,
                                                                          lengths, B, out);
    return check_launch();
}

extern "C" int32_t ts_mark_unfinished(const uint8_t* done, const int64_t* last_index,
                                      const int64_t* lengths, const int64_t* offset,
                                      const int64_t* ring_start, const int64_t* run_first,
                                      int64_t B, int64_t N, uint8_t* flags, ts_stream_t stream) {
    if (B <= 0 || N < 0 || !done || !last_index || !lengths || !offset || !ring_start ||
        !run_first || !flags)
        return TS_ERR_BAD_ARG;
    mark_unfinished_kernel<<<grid_for(B, 128), 128, 0, (cudaStream_t)stream>>>(
        done, last_index, lengths, offset, ring_start, run_first, B, N, flags);
    return check_launch();
}
