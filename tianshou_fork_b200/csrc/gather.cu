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

}  // namespace
}  // namespace ts

using namespace ts;

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
