// Device-side add of one vectorised env step.
//
// Replaces, for rollouts that are produced on the GPU: the Python loop over sub-buffers in
// ReplayBufferManager.add (tianshou/data/buffer/manager.py:104-161) with ReplayBuffer._add_index
// per transition (tianshou/data/buffer/base.py:195-214), and the per-key fancy-index stores
// `self._meta[ptrs] = batch` (manager.py:160).
//
// One launch: a warp per added transition.  Lane 0 advances the ring of the transition's
// sub-buffer (pointer, size, last_index, lengths) and its running episode statistics exactly
// as _add_index does; then the warp copies the transition's row of every stored key (up to 16:
// obs, act, rew, terminated, truncated, done, obs_next, ...) to the slot the pointer names.
// Every sub-buffer appears at most once per call (one transition per env per step), so the
// thread that owns transition j is the only one that touches the state of its sub-buffer:
// no atomics, no second launch, no host round trip for `rew` / `done`.
#include "ts_common.cuh"

namespace ts {
namespace {

constexpr int kMaxKeys = 16;

struct AddKeys {
    const unsigned char* src[kMaxKeys];  // [M, row_bytes] rows of the step, device
    unsigned char* dst[kMaxKeys];        // [total, row_bytes] storage, device
    int64_t row_bytes[kMaxKeys];
    int n;
};

struct AddState {
    const int64_t* offset;  // [B + 1]
    int64_t* ring;          // [B][2] = (index, size) of every sub-buffer
    int64_t* last_index;    // [B] global index of the newest transition  (manager.py:144)
    int64_t* lengths;       // [B]                                        (manager.py:145)
    double* ep_rew;         // [B] running episode return                 (base.py:203)
    int64_t* ep_len;        // [B]
    int64_t* ep_idx;        // [B] local start index of the running episode
};

template <typename RewT>
__global__ void __launch_bounds__(256)
add_step_kernel(AddKeys keys, int64_t M, const int64_t* __restrict__ buffer_ids,
                const RewT* __restrict__ rew, const uint8_t* __restrict__ done, AddState st,
                int64_t B, int64_t* __restrict__ ptr_out, double* __restrict__ ep_rew_out,
                int64_t* __restrict__ ep_len_out, int64_t* __restrict__ ep_idx_out,
                int32_t* __restrict__ err_flag) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t j = warp; j < M; j += nwarps) {
        int64_t ptr = -1;
        if (lane == 0) {
            const int64_t k = buffer_ids != nullptr ? buffer_ids[j] : j;
            if (k < 0 || k >= B) {
                atomicOr(err_flag, TS_FLAG_INDEX_RANGE);
            } else {
                const int64_t start = st.offset[k];
                const int64_t cap = st.offset[k + 1] - start;
                const int64_t idx = st.ring[2 * k], size = st.ring[2 * k + 1];
                const int64_t nsize = size + 1 < cap ? size + 1 : cap;  // base.py:198
                const int64_t nidx = (idx + 1) % cap;                   // :199
                st.ring[2 * k] = nidx;
                st.ring[2 * k + 1] = nsize;
                ptr = idx + start;
                st.last_index[k] = ptr;
                st.lengths[k] = nsize;
                const double er = st.ep_rew[k] + (double)rew[j];        // :203
                const int64_t el = st.ep_len[k] + 1;                    // :204
                const int64_t ei = st.ep_idx[k];
                if (done[j] != 0) {                                     // :206-211
                    ep_rew_out[j] = er;
                    ep_len_out[j] = el;
                    st.ep_rew[k] = 0.0;
                    st.ep_len[k] = 0;
                    st.ep_idx[k] = nidx;
                } else {                                                // :213
                    ep_rew_out[j] = 0.0;
                    ep_len_out[j] = 0;
                    st.ep_rew[k] = er;
                    st.ep_len[k] = el;
                }
                ep_idx_out[j] = ei + start;                             // manager.py:143
                ptr_out[j] = ptr;
            }
        }
        ptr = __shfl_sync(0xffffffffu, ptr, 0);
        if (ptr < 0) continue;
        for (int q = 0; q < keys.n; ++q) {
            const int64_t rb = keys.row_bytes[q];
            const unsigned char* s = keys.src[q] + j * rb;
            unsigned char* d = keys.dst[q] + ptr * rb;
            if (((rb | (int64_t)(uintptr_t)s | (int64_t)(uintptr_t)d) & 15) == 0) {
                for (int64_t c = lane; c < rb / 16; c += 32)
                    reinterpret_cast<uint4*>(d)[c] = __ldg(reinterpret_cast<const uint4*>(s) + c);
            } else if (((rb | (int64_t)(uintptr_t)s | (int64_t)(uintptr_t)d) & 3) == 0) {
                for (int64_t c = lane; c < rb / 4; c += 32)
                    reinterpret_cast<uint32_t*>(d)[c] =
                        __ldg(reinterpret_cast<const uint32_t*>(s) + c);
            } else {
                for (int64_t c = lane; c < rb; c += 32) d[c] = __ldg(s + c);
            }
        }
    }
}

}  // namespace
}  // namespace ts

using namespace ts;

extern "C" int32_t ts_add_step(int32_t n_keys, const void* const* src, void* const* dst,
                               const int64_t* row_bytes, int64_t M, const int64_t* buffer_ids,
                               const void* rew, int32_t rew_dtype, const uint8_t* done,
                               const int64_t* offset, int64_t B, int64_t* ring,
                               int64_t* last_index, int64_t* lengths, double* ep_rew,
                               int64_t* ep_len, int64_t* ep_idx, int64_t* ptr_out,
                               double* ep_rew_out, int64_t* ep_len_out, int64_t* ep_idx_out,
                               int32_t* err_flag, ts_stream_t stream) {
    if (n_keys < 0 || n_keys > kMaxKeys || M < 0 || B <= 0) return TS_ERR_BAD_ARG;
    if (rew_dtype != TS_F32 && rew_dtype != TS_F64) return TS_ERR_BAD_ARG;
    if (M == 0) return TS_OK;
    if (!rew || !done || !offset || !ring || !last_index || !lengths || !ep_rew || !ep_len ||
        !ep_idx || !ptr_out || !ep_rew_out || !ep_len_out || !ep_idx_out || !err_flag)
        return TS_ERR_BAD_ARG;
    if (n_keys > 0 && (!src || !dst || !row_bytes)) return TS_ERR_BAD_ARG;
    AddKeys keys;
    keys.n = n_keys;
    for (int q = 0; q < n_keys; ++q) {
        if (!src[q] || !dst[q] || row_bytes[q] < 0) return TS_ERR_BAD_ARG;
        keys.src[q] = (const unsigned char*)src[q];
        keys.dst[q] = (unsigned char*)dst[q];
        keys.row_bytes[q] = row_bytes[q];
    }
    const AddState st{offset, ring, last_index, lengths, ep_rew, ep_len, ep_idx};
    int64_t blocks = (M * 32 + 255) / 256;  // a warp per transition
    const int64_t cap = (int64_t)kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    cudaStream_t s = (cudaStream_t)stream;
    if (rew_dtype == TS_F64)
        add_step_kernel<double><<<(unsigned)blocks, 256, 0, s>>>(
            keys, M, buffer_ids, (const double*)rew, done, st, B, ptr_out, ep_rew_out, ep_len_out,
            ep_idx_out, err_flag);
    else
        add_step_kernel<float><<<(unsigned)blocks, 256, 0, s>>>(
            keys, M, buffer_ids, (const float*)rew, done, st, B, ptr_out, ep_rew_out, ep_len_out,
            ep_idx_out, err_flag);
    return check_launch();
}
