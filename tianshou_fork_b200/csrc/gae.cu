// GAE / TD(lambda) as a segmented reverse linear-recurrence scan (sm_100a).
//
// Replaces BasePolicy.compute_episodic_return + value_mask + numba _gae_return
// (tianshou/policy/base.py:317-384, 453-497).  The serial loop
//     gae = delta[i] + discount[i] * gae          (policy/base.py:494-496)
// is the affine map f_i(c) = delta_i + a_i * c with a_i = gamma*lambda*(1-end_i).  Maps
// compose associatively, (X1,P1) o (X2,P2) = (X1 + P1*X2, P1*P2), so a tile of the time
// axis is reduced with per-thread registers (8 consecutive steps), a warp-shuffle suffix
// scan and one shared-memory hop across warps; the value-mask, the end-flag logic and the
// `returns = adv + v_s` write are fused so each transition is read once and written once
// (13-14 B in, 8 B out for f32).  All arithmetic is f64 like the numba path.
#include "ts_common.cuh"

namespace ts {
namespace {

constexpr int VEC = 8;  // consecutive time steps per thread: 2 x LDG.128 per f32 array

struct Affine {  // c -> X + P * c
    double P, X;
};

// ---- tile loads ---------------------------------------------------------------------------
// 8 consecutive values starting at p (element i0 of its array), `nvalid` of them in range.
// `vec` says the address is 16-byte aligned and the caller allows vector access.
__device__ __forceinline__ void load8(const float* __restrict__ p, int nvalid, bool vec,
                                      double (&out)[VEC]) {
    if (vec && nvalid == VEC) {
        float4 a = ld_stream(reinterpret_cast<const float4*>(p));
        float4 b = ld_stream(reinterpret_cast<const float4*>(p) + 1);
        out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w;
        out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) out[k] = k < nvalid ? (double)__ldcs(p + k) : 0.0;
    }
}
__device__ __forceinline__ void load8(const double* __restrict__ p, int nvalid, bool vec,
                                      double (&out)[VEC]) {
    if (vec && nvalid == VEC) {
#pragma unroll
        for (int k = 0; k < VEC / 2; ++k) {
            double2 a = ld_stream(reinterpret_cast<const double2*>(p) + k);
            out[2 * k] = a.x; out[2 * k + 1] = a.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) out[k] = k < nvalid ? __ldcs(p + k) : 0.0;
    }
}
// 8 flag bytes -> bit mask (bit k set iff byte k != 0)
__device__ __forceinline__ uint32_t load8_flags(const uint8_t* __restrict__ p, int nvalid,
                                                bool vec) {
    uint32_t m = 0;
    if (vec && nvalid == VEC) {
        uint2 w = ld_stream(reinterpret_cast<const uint2*>(p));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            m |= ((w.x >> (8 * k)) & 0xffu) ? (1u << k) : 0u;
            m |= ((w.y >> (8 * k)) & 0xffu) ? (1u << (k + 4)) : 0u;
        }
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            if (k < nvalid && __ldcs(p + k)) m |= 1u << k;
    }
    return m;
}

__device__ __forceinline__ void store8(float* __restrict__ p, int nvalid, bool vec,
                                       const double (&v)[VEC]) {
    if (vec && nvalid == VEC) {
        st_stream(reinterpret_cast<float4*>(p),
                  make_float4((float)v[0], (float)v[1], (float)v[2], (float)v[3]));
        st_stream(reinterpret_cast<float4*>(p) + 1,
                  make_float4((float)v[4], (float)v[5], (float)v[6], (float)v[7]));
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            if (k < nvalid) p[k] = (float)v[k];
    }
}
__device__ __forceinline__ void store8(double* __restrict__ p, int nvalid, bool vec,
                                       const double (&v)[VEC]) {
    if (vec && nvalid == VEC) {
#pragma unroll
        for (int k = 0; k < VEC / 2; ++k)
            st_stream(reinterpret_cast<double2*>(p) + k, make_double2(v[2 * k], v[2 * k + 1]));
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            if (k < nvalid) p[k] = v[k];
    }
}

// ---- per-thread state for one 8-step strip -------------------------------------------------
struct Strip {
    double delta[VEC];  // rew + gamma * v_s_ * ~terminated - v_s      (policy/base.py:376, 491)
    double vs[VEC];     // v_s, kept for returns = adv + v_s            (:382)
    uint32_t cut;       // bit k: a_k == 0   (end flag, :492)
    uint32_t valid;     // bit k: element in range (out-of-range strips are the identity map)
};

__device__ __forceinline__ double coef(const Strip& s, int k, double gl) {
    return ((s.valid >> k) & 1u) ? (((s.cut >> k) & 1u) ? 0.0 : gl) : 1.0;
}

// Thread aggregate: composition of the strip's maps, first element outermost.
__device__ __forceinline__ Affine strip_aggregate(const Strip& s, double gl) {
    Affine f{1.0, 0.0};
#pragma unroll
    for (int k = VEC - 1; k >= 0; --k) {
        double a = coef(s, k, gl);
        f.X = fma(a, f.X, s.delta[k]);
        f.P *= a;
    }
    return f;
}

// Suffix composition over the lanes of a warp.  On return `incl` covers lanes [lane, 32) and
// `excl` covers lanes (lane, 32) (identity for lane 31).
__device__ __forceinline__ void warp_suffix_scan(Affine f, int lane, Affine& incl, Affine& excl) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double P2 = shfl_down_f64(f.P, d), X2 = shfl_down_f64(f.X, d);
        if (lane + d < 32) {
            f.X = fma(f.P, X2, f.X);
            f.P *= P2;
        }
    }
    incl = f;
    excl.P = shfl_down_f64(f.P, 1);
    excl.X = shfl_down_f64(f.X, 1);
    if (lane == 31) { excl.P = 1.0; excl.X = 0.0; }
}

// Block-level: given every thread's strip, returns the tile aggregate (all threads) and, via
// `excl_tile`, the map from the tile's carry-in to this thread's carry-in.
template <int NT>
__device__ __forceinline__ Affine block_suffix(const Affine& mine, int tid, Affine& excl_tile,
                                               double* sP, double* sX) {
    constexpr int NW = NT / 32;
    const int lane = tid & 31, warp = tid >> 5;
    Affine incl, excl;
    warp_suffix_scan(mine, lane, incl, excl);
    if (NW == 1) {
        excl_tile = excl;
        return Affine{shfl_idx_f64(incl.P, 0), shfl_idx_f64(incl.X, 0)};
    }
    if (lane == 0) { sP[warp] = incl.P; sX[warp] = incl.X; }
    __syncthreads();
    // maps of the later warps, composed: covers warps (warp, NW)
    Affine later{1.0, 0.0};
#pragma unroll
    for (int w = NW - 1; w >= 1; --w) {
        if (w > warp) {  // later <- f_w o later, built from the far end inwards
            later.X = fma(sP[w], later.X, sX[w]);
            later.P *= sP[w];
        }
    }
    // thread carry-in = excl( later( tile carry ) )
    excl_tile.X = fma(excl.P, later.X, excl.X);
    excl_tile.P = excl.P * later.P;
    Affine tile{1.0, 0.0};
#pragma unroll
    for (int w = NW - 1; w >= 0; --w) {
        tile.X = fma(sP[w], tile.X, sX[w]);
        tile.P *= sP[w];
    }
    return tile;
}

// Second pass: replay the strip serially from the thread's true carry-in.
__device__ __forceinline__ double strip_replay(const Strip& s, double gl, double carry,
                                               double (&adv)[VEC], double (&ret)[VEC]) {
    double A = carry;
#pragma unroll
    for (int k = VEC - 1; k >= 0; --k) {
        A = fma(coef(s, k, gl), A, s.delta[k]);
        adv[k] = A;
        ret[k] = A + s.vs[k];
    }
    return A;
}

// ============================================================================================
// Canonical [E, T] kernel: one CTA per env stream, chunks of NT*8 steps walked from the end.
// ============================================================================================
template <typename RewT, typename InT, typename OutT, int NT>
__global__ void __launch_bounds__(NT)
gae_rows_kernel(const RewT* __restrict__ rew, const InT* __restrict__ v_s,
                const InT* __restrict__ v_next, const uint8_t* __restrict__ term,
                const uint8_t* __restrict__ trunc, int64_t T, int64_t buf_stride,
                int64_t val_stride, double gamma, double gl, OutT* __restrict__ adv,
                OutT* __restrict__ ret, int vec_ok) {
    constexpr int TILE = NT * VEC;
    constexpr int NW = NT / 32;
    __shared__ double sP[NW > 1 ? NW : 1], sX[NW > 1 ? NW : 1];
    __shared__ double s_carry;

    const int tid = threadIdx.x;
    const int64_t row = blockIdx.x;
    const RewT* r_rew = rew + row * buf_stride;
    const uint8_t* r_term = term + row * buf_stride;
    const uint8_t* r_trunc = trunc + row * buf_stride;
    const InT* r_vs = v_s + row * val_stride;
    const InT* r_vn = v_next + row * val_stride;
    OutT* r_adv = adv + row * val_stride;
    OutT* r_ret = ret + row * val_stride;

    const int64_t nchunks = (T + TILE - 1) / TILE;
    double carry = 0.0;  // advantage beyond the end of the stream
    for (int64_t c = nchunks - 1; c >= 0; --c) {
        const int64_t t0 = c * TILE + (int64_t)tid * VEC;
        const int64_t left = T - t0;
        const int nvalid = left >= VEC ? VEC : (left > 0 ? (int)left : 0);
        const bool vec = vec_ok != 0;

        Strip s;
        double vn[VEC], rw[VEC];
        uint32_t mterm = 0, mtrunc = 0;
        if (nvalid > 0) {
            load8(r_rew + t0, nvalid, vec, rw);
            load8(r_vs + t0, nvalid, vec, s.vs);
            load8(r_vn + t0, nvalid, vec, vn);
            mterm = load8_flags(r_term + t0, nvalid, vec);
            mtrunc = load8_flags(r_trunc + t0, nvalid, vec);
        } else {
#pragma unroll
            for (int k = 0; k < VEC; ++k) { rw[k] = 0.0; s.vs[k] = 0.0; vn[k] = 0.0; }
        }
        s.valid = nvalid >= VEC ? 0xffu : ((1u << nvalid) - 1u);
        s.cut = mterm | mtrunc;
        // the stream's last step is done or the unfinished index (manager.py:68-74)
        if (t0 + VEC >= T && nvalid > 0) s.cut |= 1u << (nvalid - 1);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            double v1 = ((mterm >> k) & 1u) ? 0.0 : vn[k];
            s.delta[k] = fma(v1, gamma, rw[k]) - s.vs[k];
        }

        Affine mine = strip_aggregate(s, gl);
        Affine to_me;
        (void)block_suffix<NT>(mine, tid, to_me, sP, sX);
        const double my_carry = fma(to_me.P, carry, to_me.X);
        double a_out[VEC], r_out[VEC];
        const double head = strip_replay(s, gl, my_carry, a_out, r_out);
        if (nvalid > 0) {
            store8(r_adv + t0, nvalid, vec, a_out);
            store8(r_ret + t0, nvalid, vec, r_out);
        }
        if (c > 0) {  // uniform branch: hand the chunk's first advantage to the next chunk
            if (tid == 0) s_carry = head;
            __syncthreads();
            carry = s_carry;
            __syncthreads();
        }
    }
}

// ============================================================================================
// Canonical [E, T] kernel, bulk-copy pipelined (the C2 fast path).
// Persistent CTAs; each owns rows blockIdx.x, blockIdx.x + gridDim.x, ... and walks every row
// in time tiles of NT*8 steps from the end.  One elected thread keeps STAGES tiles of the five
// input arrays in flight with cp.async.bulk (completion on an mbarrier per stage), so HBM
// reads never wait for the scan arithmetic; outputs are staged in shared memory and leave as
// two bulk stores per tile (full-line writes).  Requires 16-byte aligned rows (T, strides
// multiples of 16 elements); anything else takes gae_rows_kernel above.
// ============================================================================================
// The arithmetic below is written for instruction count (the first version of this kernel was
// issue-bound at 610 warp instructions per 8-step strip, profiles/r01b): flag bytes are
// tested in place with one LOP3->predicate each, the end-of-episode cuts are applied with
// predicated DFMAs instead of selected coefficients, and the carry is added to all eight
// local advantages independently through gl^1..gl^8.
struct GlPow { double p[VEC + 1]; };  // p[j] = (gamma*lambda)^j, kernel parameter

// A thread's eight consecutive f32 values are two 16-byte chunks, 32 bytes apart from the next
// thread's: in program order the eight threads of a quarter warp hit every bank group twice (a
// 2-way conflict on every 128-bit access: 2.07 M conflicts per C2 launch, profiles/r01k_gae_ncu.md).
// TS_SMEM_SWZ=1 lets threads 4..7 of every eight take their SECOND chunk first (the quarter warp
// then covers all eight bank groups once) and swaps the halves back in registers: conflicts drop
// to 0.80 M, but the selects cost 8 registers (64 -> 72) and 8 % more instructions, and the
// kernel gets SLOWER -- A/B on one box, twice: 31.82 vs 30.80 us per C2 launch
// (profiles/r02_gae.md).  The conflicts were never the bound; the switch stays off.
#ifndef TS_SMEM_SWZ
#define TS_SMEM_SWZ 0
#endif
template <typename T>
__device__ __forceinline__ void smem_ld8(const T* p, T (&v)[VEC]) {
    if constexpr (sizeof(T) == 4) {
        const int sel = TS_SMEM_SWZ ? (threadIdx.x >> 2) & 1 : 0;
        const float4* q = reinterpret_cast<const float4*>(p);
        const float4 x = q[sel], y = q[sel ^ 1];
        const float4 a = sel ? y : x, b = sel ? x : y;
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < VEC / 2; ++k) {
            const double2 a = *reinterpret_cast<const double2*>(p + 2 * k);
            v[2 * k] = a.x; v[2 * k + 1] = a.y;
        }
    }
}
template <typename T>
__device__ __forceinline__ void smem_st8(T* p, const double (&v)[VEC]) {
    if constexpr (sizeof(T) == 4) {
        const int sel = TS_SMEM_SWZ ? (threadIdx.x >> 2) & 1 : 0;
        const float4 a = make_float4((float)v[0], (float)v[1], (float)v[2], (float)v[3]);
        const float4 b = make_float4((float)v[4], (float)v[5], (float)v[6], (float)v[7]);
        float4* q = reinterpret_cast<float4*>(p);
        q[sel] = sel ? b : a;
        q[sel ^ 1] = sel ? a : b;
    } else {
#pragma unroll
        for (int k = 0; k < VEC / 2; ++k)
            *reinterpret_cast<double2*>(p + 2 * k) = make_double2(v[2 * k], v[2 * k + 1]);
    }
}
// byte k (0..7) of the 8 flag bytes held in w is non-zero
__device__ __forceinline__ bool flag_byte(const uint2& w, int k) {
    return k < 4 ? (w.x & (0xffu << (8 * k))) != 0u : (w.y & (0xffu << (8 * (k - 4)))) != 0u;
}
// no flag byte at positions k..7
__device__ __forceinline__ bool no_flag_from(const uint2& w, int k) {
    return k < 4 ? (w.y == 0u && (w.x >> (8 * k)) == 0u) : (w.y >> (8 * (k - 4))) == 0u;
}

template <typename RewT, typename InT, typename OutT, int NT>
struct RowsTmaLayout {
    static constexpr int TILE = NT * VEC;
    static constexpr int OFF_VS = TILE * (int)sizeof(RewT);
    static constexpr int OFF_VN = OFF_VS + TILE * (int)sizeof(InT);
    static constexpr int OFF_TERM = OFF_VN + TILE * (int)sizeof(InT);
    static constexpr int OFF_TRUNC = OFF_TERM + TILE;
    static constexpr int IN_BYTES = OFF_TRUNC + TILE;          // one input stage
    static constexpr int OUT_BYTES = 2 * TILE * (int)sizeof(OutT);  // adv | ret
    static constexpr int IN_BYTES_PER_STEP = (int)sizeof(RewT) + 2 * (int)sizeof(InT) + 2;
    static size_t smem_bytes(int stages) {
        return (size_t)stages * IN_BYTES + OUT_BYTES + (size_t)stages * 8;
    }
};

template <typename RewT, typename InT, typename OutT, int NT>
// (768 / NT: ptxas allots the same 80 registers with or without the second argument, but
// schedules better with it -- 16384 x 2048 rows: 120 vs 128 us on one box, profiles/r02_gae.md)
__global__ void __launch_bounds__(NT, 768 / NT)
gae_rows_tma_kernel(const RewT* __restrict__ rew, const InT* __restrict__ v_s,
                    const InT* __restrict__ v_next, const uint8_t* __restrict__ term,
                    const uint8_t* __restrict__ trunc, int E, int64_t T, int64_t buf_stride,
                    int64_t val_stride, double gamma, const GlPow glp, OutT* __restrict__ adv,
                    OutT* __restrict__ ret, int stages) {
    using L = RowsTmaLayout<RewT, InT, OutT, NT>;
    constexpr int TILE = L::TILE;
    constexpr int NW = NT / 32;
    static_assert(NW > 1 && NW <= 8, "cross-warp step assumes 2..8 warps");
    extern __shared__ __align__(128) unsigned char smem[];
    // double-buffered by tile parity: only the storing warp waits at a tile's end (see below)
    __shared__ double sP[2][NW], sX[2][NW];
    __shared__ double s_carry[2];
    constexpr int STORER = NT - 32;  // lane 0 of the last warp issues the bulk stores

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = tid * VEC;
    unsigned char* out_base = smem + (size_t)stages * L::IN_BYTES;
    const uint32_t bar0 = smem_u32(out_base + L::OUT_BYTES);
    OutT* s_adv = reinterpret_cast<OutT*>(out_base);
    OutT* s_ret = s_adv + TILE;

    const int nchunks = (int)((T + TILE - 1) / TILE);
    const int n_first = (int)(T - (int64_t)(nchunks - 1) * TILE);  // steps in a row's last tile
    const int rows_mine = (int)blockIdx.x < E ? (E - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int ntiles = rows_mine * nchunks;  // E * nchunks < 2^31 checked by the host

    // producer cursor (elected thread): tile p_it = (row p_row, chunk p_c)
    int p_it = 0, p_row = blockIdx.x, p_c = nchunks - 1, p_stage = 0;
    auto issue = [&]() {
        const int n = p_c == nchunks - 1 ? n_first : TILE;
        const int64_t t0 = (int64_t)p_c * TILE;
        const uint32_t bar = bar0 + 8u * p_stage;
        const uint32_t dst = smem_u32(smem + (size_t)p_stage * L::IN_BYTES);
        mbar_expect_tx(bar, (uint32_t)n * L::IN_BYTES_PER_STEP);
        const int64_t b = (int64_t)p_row * buf_stride + t0, v = (int64_t)p_row * val_stride + t0;
        bulk_g2s(dst, rew + b, (uint32_t)n * sizeof(RewT), bar);
        bulk_g2s(dst + L::OFF_VS, v_s + v, (uint32_t)n * sizeof(InT), bar);
        bulk_g2s(dst + L::OFF_VN, v_next + v, (uint32_t)n * sizeof(InT), bar);
        bulk_g2s(dst + L::OFF_TERM, term + b, (uint32_t)n, bar);
        bulk_g2s(dst + L::OFF_TRUNC, trunc + b, (uint32_t)n, bar);
        ++p_it;
        if (--p_c < 0) { p_c = nchunks - 1; p_row += gridDim.x; }
        if (++p_stage == stages) p_stage = 0;
    };

    pdl_launch_dependents();
    if (tid == 0) {
        for (int s = 0; s < stages; ++s) mbar_init(bar0 + 8u * s, 1u);
        fence_mbar_init();
    }
    __syncthreads();
    pdl_wait();  // inputs may come from the kernel before (the critic's forward pass)
    if (tid == 0)
        while (p_it < stages && p_it < ntiles) issue();

    // consumer cursor (all threads)
    int row = blockIdx.x, c = nchunks - 1, stage = 0;
    uint32_t parity = 0;
    for (int it = 0; it < ntiles; ++it) {
        const int b = it & 1;
        const bool row_end = c == nchunks - 1;  // this tile holds the stream's last step
        const int n = row_end ? n_first : TILE;
        mbar_wait(bar0 + 8u * stage, parity);

        const unsigned char* in = smem + (size_t)stage * L::IN_BYTES;
        const bool live = e0 < n;  // n is a multiple of 16: strips are whole or empty
        RewT rw[VEC];
        InT vs[VEC], vn[VEC];
        smem_ld8(reinterpret_cast<const RewT*>(in) + e0, rw);
        smem_ld8(reinterpret_cast<const InT*>(in + L::OFF_VS) + e0, vs);
        smem_ld8(reinterpret_cast<const InT*>(in + L::OFF_VN) + e0, vn);
        const uint2 wt = *reinterpret_cast<const uint2*>(in + L::OFF_TERM + e0);
        uint2 wc = *reinterpret_cast<const uint2*>(in + L::OFF_TRUNC + e0);
        wc.x |= wt.x;
        wc.y |= wt.y;
        // the stream's last step is done or the unfinished index (manager.py:68-74)
        if (row_end && e0 + VEC == n) wc.y |= 0x01000000u;

        // d[k] = delta_k = rew + gamma * v_s_ * ~terminated - v_s   (policy/base.py:376, 491)
        double d[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const InT v1 = flag_byte(wt, k) ? InT(0) : vn[k];
            d[k] = fma((double)v1, gamma, (double)rw[k]) - (double)vs[k];
        }
        // local advantages (zero carry-in): d[k] += gl * d[k+1] unless step k ends an episode
#pragma unroll
        for (int k = VEC - 2; k >= 0; --k)
            if (!flag_byte(wc, k)) d[k] = fma(glp.p[1], d[k + 1], d[k]);
        // the strip as a map of its carry-in: c -> X + P * c
        double P = (wc.x | wc.y) == 0u ? glp.p[VEC] : 0.0;
        double X = d[0];
        if (!live) { P = 1.0; X = 0.0; }

        // ---- suffix scan over the warp: (P, X) <- this lane's strip .. lane 31's strip --------
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) {
            const double P2 = shfl_down_f64(P, dlt), X2 = shfl_down_f64(X, dlt);
            if (lane + dlt < 32) {
                X = fma(P, X2, X);
                P *= P2;
            }
        }
        // map of the lanes after this one
        double eP = shfl_down_f64(P, 1), eX = shfl_down_f64(X, 1);
        if (lane == 31) { eP = 1.0; eX = 0.0; }
        if (lane == 0) { sP[b][warp] = P; sX[b][warp] = X; }
        if (tid == STORER) bulk_wait_read0();  // previous tile's stores have left the out buffer
        __syncthreads();
        // every thread holds its strip in registers: the stage can be refilled
        if (tid == 0 && p_it < ntiles) issue();

        // carry into this thread's strip: the advantage at the previous tile's first step (nothing
        // beyond the end of a stream), through the later warps, then the later lanes
        double cin = row_end ? 0.0 : s_carry[b ^ 1];
#pragma unroll
        for (int w = NW - 1; w >= 1; --w)
            if (w > warp) cin = fma(sP[b][w], cin, sX[b][w]);
        cin = fma(eP, cin, eX);
        // A_k = d[k] + gl^(8-k) * cin while no episode ends in k..7
        double r_out[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            if (no_flag_from(wc, k)) d[k] = fma(glp.p[VEC - k], cin, d[k]);
            r_out[k] = d[k] + (double)vs[k];  // returns = advantage + v_s (:382)
        }
        if (live) {
            smem_st8(s_adv + e0, d);
            smem_st8(s_ret + e0, r_out);
        }
        fence_proxy_async_smem();
        if (tid == 0) s_carry[b] = d[0];  // advantage at the tile's first step
        // End of tile: only the storing warp waits for the staged outputs (named barrier 1); the
        // others arrive and go on to the next tile's loads and strip arithmetic.  The next tile's
        // __syncthreads orders s_carry / sP / sX (other parity) and the reuse of the out buffer.
        if (warp == NW - 1) {
            named_bar_sync(1, NT);
            if (lane == 0) {
                const int64_t v = (int64_t)row * val_stride + (int64_t)c * TILE;
                bulk_s2g(adv + v, smem_u32(s_adv), (uint32_t)n * sizeof(OutT));
                bulk_s2g(ret + v, smem_u32(s_ret), (uint32_t)n * sizeof(OutT));
                bulk_commit();
            }
        } else {
            named_bar_arrive(1, NT);
        }
        if (--c < 0) { c = nchunks - 1; row += gridDim.x; }
        if (++stage == stages) { stage = 0; parity ^= 1u; }
    }
    if (tid == STORER) bulk_wait0();
}

// ============================================================================================
// General flattened form, bulk-copy pipelined: one sequence of N transitions in the caller's
// order (batch == buffer[indices]), cut only by end flags, exactly _gae_return's contract.
//
// The 16-aligned part of the sequence is split into tiles of NT*8 steps; persistent CTA b owns
// a CONTIGUOUS range of tiles and walks it from its end with the same stage ring, scan and
// staged stores as gae_rows_tma_kernel, so the carry between tiles stays in registers.  What
// a CTA needs from outside is ONE number: the advantage of the first step after its range.
// It does not wait for the CTA that owns that step: it runs that CTA's bottom tile as a GHOST
// tile through its own pipeline first (scanned with zero carry-in, nothing stored).  An episode
// boundary anywhere in those 2048 steps, or a discount product below 1e-30 over them, makes the
// ghost tile's head advantage exact -- on real rollouts always.  Only with no episode end AND
// gamma*lambda ~ 1 does the CTA fall back to waiting for its neighbour's published head value;
// the chain starts at the CTA that owns the end of the sequence, which never waits.
// ============================================================================================
struct FlatGather {               // how value mask and unfinished tails are found (nullable)
    const int64_t* indices;       // batch position -> buffer index
    const uint8_t* buf_term;      // buffer.terminated (value_mask, policy/base.py:335)
    const int64_t* unfinished;    // per sub-buffer unfinished index or -1 (:380)
    const int64_t* offset;
    int64_t B, sub_size;
    int64_t seg;                  // > 0: the sequence is rows of `seg` >= 8 steps, cut at every row end
};

// the end of a row of `seg` steps, if it falls inside the 8-step strip that starts at row position r
__device__ __forceinline__ void seg_cut(uint32_t r, uint32_t seg, uint2& wc) {
    const uint32_t qs = seg - 1u - r;  // steps from the strip's first element to the row's last
    if (qs < (uint32_t)VEC) {
        const uint32_t bit = 1u << (8 * (qs & 3u));
        if (qs < 4u) wc.x |= bit; else wc.y |= bit;
    }
}

// mask / cut of one transition from its flags (and, with indices, from the buffer)
__device__ __forceinline__ void flat_flags(const FlatGather& g, int64_t idx, uint32_t t,
                                           uint32_t c, bool& masked, bool& cut) {
    masked = t != 0u;
    cut = (t | c) != 0u;
    if (g.indices != nullptr) {
        if (g.buf_term != nullptr) masked = __ldg(g.buf_term + idx) != 0;
        if (g.unfinished != nullptr &&
            __ldg(g.unfinished + sub_of(idx, g.offset, g.B, g.sub_size)) == idx)
            cut = true;
    }
}

// Discount product over a ghost tile below which the carry from beyond it is taken as 0: a
// relative error of 1e-30 of the carried advantage, fourteen orders below one f64 ulp.
constexpr double kCarryNegligible = 1e-30;

template <typename RewT, typename InT, typename OutT, int NT, bool HAS_IDX>
struct FlatTmaLayout {
    using R = RowsTmaLayout<RewT, InT, OutT, NT>;
    static constexpr int TILE = R::TILE;
    static constexpr int OFF_IDX = R::IN_BYTES;  // int64 indices behind the five arrays
    static constexpr int IN_BYTES = R::IN_BYTES + (HAS_IDX ? TILE * 8 : 0);
    static constexpr int OUT_BYTES = R::OUT_BYTES;
    static constexpr int IN_BYTES_PER_STEP = R::IN_BYTES_PER_STEP + (HAS_IDX ? 8 : 0);
    static size_t smem_bytes(int stages) {
        return (size_t)stages * IN_BYTES + OUT_BYTES + (size_t)stages * 8;
    }
};

template <typename RewT, typename InT, typename OutT, int NT, bool HAS_IDX>
// 3 x 256 threads per SM (<= 80 registers) without indices; with them a stage is 45 KB and only
// two CTAs fit an SM's shared memory anyway, so the index arithmetic gets 128 registers
__global__ void __launch_bounds__(NT, HAS_IDX ? 2 : 768 / NT)
gae_flat_tma_kernel(const RewT* __restrict__ rew, const InT* __restrict__ v_s,
                    const InT* __restrict__ v_next, const uint8_t* __restrict__ term,
                    const uint8_t* __restrict__ trunc, const FlatGather g, int64_t N,
                    double gamma, const GlPow glp, OutT* __restrict__ adv,
                    OutT* __restrict__ ret, int stages, double* __restrict__ heads,
                    unsigned int* __restrict__ head_flags, unsigned int* __restrict__ sync) {
    using L = FlatTmaLayout<RewT, InT, OutT, NT, HAS_IDX>;
    using R = RowsTmaLayout<RewT, InT, OutT, NT>;
    constexpr int TILE = L::TILE;
    constexpr int NW = NT / 32;
    static_assert(NW > 1 && NW <= 8, "cross-warp step assumes 2..8 warps");
    extern __shared__ __align__(128) unsigned char smem[];
    // double-buffered by tile parity: only the storing warp waits at a tile's end (gae_rows_tma_kernel)
    __shared__ double sP[2][NW], sX[2][NW];
    __shared__ double s_carry[2];
    constexpr int STORER = NT - 32;  // lane 0 of the last warp issues the bulk stores

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = tid * VEC;
    unsigned char* out_base = smem + (size_t)stages * L::IN_BYTES;
    const uint32_t bar0 = smem_u32(out_base + L::OUT_BYTES);
    OutT* s_adv = reinterpret_cast<OutT*>(out_base);
    OutT* s_ret = s_adv + TILE;

    const int64_t N16 = N & ~(int64_t)15;  // the bulk-copy part; [N16, N) is done by one thread
    const int64_t ntiles_all = (N16 + TILE - 1) / TILE;
    // balanced contiguous ranges; CTA 0 owns the END of the sequence (it never needs a neighbour)
    const int64_t base = ntiles_all / gridDim.x, rem = ntiles_all % gridDim.x;
    const int64_t b = blockIdx.x;
    const int64_t before = b * base + (b < rem ? b : rem);  // tiles owned by CTAs 0..b-1
    const int ntiles = (int)(base + (b < rem ? 1 : 0));
    const int64_t k_hi = ntiles_all - before;
    const int n_last = (int)(N16 - (ntiles_all - 1) * TILE);  // steps in the sequence's last tile

    // The carry INTO the range (the advantage of step k_hi * TILE) comes from a GHOST tile: every
    // CTA but the one that owns the end of the sequence runs its neighbour's bottom tile k_hi
    // through the same ring and the same scan first, with zero carry-in and without storing
    // anything.  If an episode ends anywhere in those 2048 steps (or the discount product over them
    // is negligible) the ghost tile's head advantage IS the carry; only otherwise (no episode end
    // and gamma * lambda ~ 1) does the CTA wait for its neighbour's published head value.  The
    // earlier design read ahead with plain loads BEFORE the first tile: two to six dependent DRAM
    // round trips queued behind the bulk loads of all CTAs, 4 us of every launch without indices
    // and 10 us with them (profiles/r02_gae.md); the ghost tile costs 1 / 14 more input reads.
    // (rows of `seg` steps packed back to back: a range that starts on a row boundary needs no
    // carry at all -- always the case when seg divides the tile)
    const bool ghost = blockIdx.x > 0 && ntiles > 0 &&
                       !(g.seg > 0 && (k_hi * (int64_t)TILE) % g.seg == 0);
    const int nit = ntiles + (ghost ? 1 : 0);
    // producer cursor (elected thread): tile p_k, descending
    int p_it = 0, p_stage = 0;
    int64_t p_k = k_hi - (ghost ? 0 : 1);
    auto issue = [&]() {
        const int n = p_k == ntiles_all - 1 ? n_last : TILE;
        const int64_t t0 = p_k * TILE;
        const uint32_t bar = bar0 + 8u * p_stage;
        const uint32_t dst = smem_u32(smem + (size_t)p_stage * L::IN_BYTES);
        mbar_expect_tx(bar, (uint32_t)n * L::IN_BYTES_PER_STEP);
        bulk_g2s(dst, rew + t0, (uint32_t)n * sizeof(RewT), bar);
        bulk_g2s(dst + R::OFF_VS, v_s + t0, (uint32_t)n * sizeof(InT), bar);
        bulk_g2s(dst + R::OFF_VN, v_next + t0, (uint32_t)n * sizeof(InT), bar);
        bulk_g2s(dst + R::OFF_TERM, term + t0, (uint32_t)n, bar);
        bulk_g2s(dst + R::OFF_TRUNC, trunc + t0, (uint32_t)n, bar);
        if (HAS_IDX) bulk_g2s(dst + L::OFF_IDX, g.indices + t0, (uint32_t)n * 8u, bar);
        ++p_it;
        --p_k;
        if (++p_stage == stages) p_stage = 0;
    };

    // ---- CTA 0: the ragged end [N16, N), at most 15 steps, one per lane, outputs included ------
    if (tid == 0) s_carry[1] = 0.0;  // slot of the tile "before" the first one
    const bool working = ntiles > 0 || blockIdx.x == 0;
    pdl_launch_dependents();
    // sync[1] = calls completed on this workspace, sync[2] = CTAs of this call that are done.
    // Head flags carry the call's tag, so nothing has to be cleared between calls (and CUDA-graph
    // replays, which cannot change an argument, still get a fresh tag each time).
    unsigned int tag = 0u;
    if (tid == 0) {
        for (int s = 0; s < stages; ++s) mbar_init(bar0 + 8u * s, 1u);
        fence_mbar_init();
    }
    pdl_wait();  // inputs and the workspace may come from the kernel before
    if (tid == 0) {
        while (p_it < stages && p_it < nit) issue();
        if (sync != nullptr) tag = (__ldcg(sync + 1) + 1u) | 0x80000000u;
    }
    __syncthreads();
    // a CTA's last act: the one that finishes last advances the call count
    auto leave = [&]() {
        if (tid == 0 && sync != nullptr) {
            __threadfence();
            if (atomicAdd(sync + 2, 1u) == gridDim.x - 1u) {
                sync[1] = tag & 0x7fffffffu;
                sync[2] = 0u;
            }
        }
    };
    if (working && blockIdx.x == 0) {
        if (warp == 0) {  // one step per lane: loads in parallel, the recurrence by shuffles
            const int nr = (int)(N - N16);
            const int64_t i = N16 + lane;
            double dl = 0.0, vsl = 0.0, mine = 0.0;
            bool cutl = true;
            if (lane < nr) {
                const int64_t idx = g.indices != nullptr ? g.indices[i] : i;
                bool masked;
                flat_flags(g, idx, term[i], trunc[i], masked, cutl);
                if (g.seg > 0 && (i + 1) % g.seg == 0) cutl = true;
                const double v1 = masked ? 0.0 : (double)v_next[i];
                vsl = (double)v_s[i];
                dl = fma(v1, gamma, (double)rew[i]) - vsl;
            }
            double A = 0.0;
            for (int j = nr - 1; j >= 0; --j) {
                const double dj = shfl_idx_f64(dl, j);
                const bool cj = __shfl_sync(0xffffffffu, (int)cutl, j) != 0;
                A = cj ? dj : fma(glp.p[1], A, dj);
                if (lane == j) mine = A;
            }
            if (lane < nr) {
                adv[i] = (OutT)mine;
                ret[i] = (OutT)(mine + vsl);
            }
            if (lane == 0) s_carry[1] = A;
        }
    }
    __syncthreads();
    if (ntiles == 0) {  // nothing but (for CTA 0) the ragged end: pass the value on
        if (blockIdx.x == 0 && tid == 0 && heads != nullptr) {
            heads[0] = s_carry[1];
            __threadfence();
            *(volatile unsigned int*)head_flags = tag;
        }
        leave();
        return;
    }

    int stage = 0;
    uint32_t parity = 0;
    int64_t k = k_hi - (ghost ? 0 : 1);
    // row position of the current tile's first step, kept incrementally (tiles go down by TILE)
    const uint32_t seg = (uint32_t)g.seg;
    uint32_t r_tile = seg ? (uint32_t)((k * TILE) % g.seg) : 0u;
    const uint32_t r_back = seg ? seg - (uint32_t)(TILE % g.seg) : 0u;
    for (int it = 0; it < nit; ++it, --k) {
        const int b = it & 1;
        const bool is_ghost = ghost && it == 0;  // the neighbour's tile: scanned, not stored
        const int n = k == ntiles_all - 1 ? n_last : TILE;
        mbar_wait(bar0 + 8u * stage, parity);

        const unsigned char* in = smem + (size_t)stage * L::IN_BYTES;
        const bool live = e0 < n;
        RewT rw[VEC];
        InT vs[VEC], vn[VEC];
        smem_ld8(reinterpret_cast<const RewT*>(in) + e0, rw);
        smem_ld8(reinterpret_cast<const InT*>(in + R::OFF_VS) + e0, vs);
        smem_ld8(reinterpret_cast<const InT*>(in + R::OFF_VN) + e0, vn);
        uint2 wt = *reinterpret_cast<const uint2*>(in + R::OFF_TERM + e0);
        uint2 wc = *reinterpret_cast<const uint2*>(in + R::OFF_TRUNC + e0);
        wc.x |= wt.x;
        wc.y |= wt.y;
        if (HAS_IDX && live) {
            // value mask and unfinished tails come from the buffer, through the indices
            int64_t id[VEC];
#pragma unroll
            for (int q = 0; q < VEC / 2; ++q) {
                const longlong2 a =
                    *(reinterpret_cast<const longlong2*>(in + L::OFF_IDX) + (e0 / 2 + q));
                id[2 * q] = a.x; id[2 * q + 1] = a.y;
            }
            // strips of consecutive indices are the rule (batch order is chronological inside
            // a sub-buffer, manager.py:187-192)
            bool consec = true;
#pragma unroll
            for (int q = 1; q < VEC; ++q) consec = consec && id[q] == id[0] + q;
            if (g.buf_term != nullptr) {
                const uintptr_t a0 = reinterpret_cast<uintptr_t>(g.buf_term + id[0]);
                if (consec && (a0 & 7) == 0) {
                    // the eight mask bytes are one aligned word: one dependent load, not eight
                    wt = __ldg(reinterpret_cast<const uint2*>(g.buf_term + id[0]));
                } else if (consec && (a0 & 3) == 0) {
                    wt.x = __ldg(reinterpret_cast<const uint32_t*>(g.buf_term + id[0]));
                    wt.y = __ldg(reinterpret_cast<const uint32_t*>(g.buf_term + id[0] + 4));
                } else {
                    uint32_t m_lo = 0, m_hi = 0;
#pragma unroll
                    for (int q = 0; q < VEC; ++q) {
                        const uint32_t bit =
                            __ldg(g.buf_term + id[q]) != 0 ? 1u << (8 * (q & 3)) : 0u;
                        if (q < 4) m_lo |= bit; else m_hi |= bit;
                    }
                    wt.x = m_lo; wt.y = m_hi;
                }
            }
            if (g.unfinished != nullptr) {
                // a strip of consecutive indices touches at most two sub-buffers of >= 8 steps:
                // two sub-buffer look-ups instead of eight
                const int64_t s0 = sub_of(id[0], g.offset, g.B, g.sub_size);
                // equal sub-buffers: the strip's last index is in s0 or, past its end, in s0 + 1
                // (one division per strip instead of two)
                const int64_t s7 = (consec && g.sub_size > 0)
                    ? s0 + (id[0] - s0 * g.sub_size + (VEC - 1) >= g.sub_size ? 1 : 0)
                    : sub_of(id[VEC - 1], g.offset, g.B, g.sub_size);
                if (consec && s7 - s0 <= 1) {
                    const int64_t u0 = __ldg(g.unfinished + s0) - id[0];
                    const int64_t u7 = s7 != s0 ? __ldg(g.unfinished + s7) - id[0] : -1;
                    if (u0 >= 0 && u0 < VEC) {
                        const uint32_t bit = 1u << (8 * ((int)u0 & 3));
                        if (u0 < 4) wc.x |= bit; else wc.y |= bit;
                    }
                    if (u7 >= 0 && u7 < VEC) {
                        const uint32_t bit = 1u << (8 * ((int)u7 & 3));
                        if (u7 < 4) wc.x |= bit; else wc.y |= bit;
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < VEC; ++q) {
                        if (__ldg(g.unfinished + sub_of(id[q], g.offset, g.B, g.sub_size)) == id[q]) {
                            const uint32_t bit = 1u << (8 * (q & 3));
                            if (q < 4) wc.x |= bit; else wc.y |= bit;
                        }
                    }
                }
            }
        }

        if (seg) {
            seg_cut((r_tile + (uint32_t)e0) % seg, seg, wc);
            r_tile += r_back;
            if (r_tile >= seg) r_tile -= seg;
        }

        double d[VEC];
#pragma unroll
        for (int q = 0; q < VEC; ++q) {
            const InT v1 = flag_byte(wt, q) ? InT(0) : vn[q];
            d[q] = fma((double)v1, gamma, (double)rw[q]) - (double)vs[q];
        }
#pragma unroll
        for (int q = VEC - 2; q >= 0; --q)
            if (!flag_byte(wc, q)) d[q] = fma(glp.p[1], d[q + 1], d[q]);
        double P = (wc.x | wc.y) == 0u ? glp.p[VEC] : 0.0;
        double X = d[0];
        if (!live) { P = 1.0; X = 0.0; }
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) {
            const double P2 = shfl_down_f64(P, dlt), X2 = shfl_down_f64(X, dlt);
            if (lane + dlt < 32) { X = fma(P, X2, X); P *= P2; }
        }
        double eP = shfl_down_f64(P, 1), eX = shfl_down_f64(X, 1);
        if (lane == 31) { eP = 1.0; eX = 0.0; }
        if (lane == 0) { sP[b][warp] = P; sX[b][warp] = X; }
        if (tid == STORER) bulk_wait_read0();
        __syncthreads();
        if (tid == 0 && p_it < nit) issue();

        double cin = s_carry[b ^ 1];  // the advantage at the previous tile's first step
#pragma unroll
        for (int w = NW - 1; w >= 1; --w)
            if (w > warp) cin = fma(sP[b][w], cin, sX[b][w]);
        cin = fma(eP, cin, eX);
        double r_out[VEC];
#pragma unroll
        for (int q = 0; q < VEC; ++q) {
            if (no_flag_from(wc, q)) d[q] = fma(glp.p[VEC - q], cin, d[q]);
            r_out[q] = d[q] + (double)vs[q];
        }
        if (live && !is_ghost) {
            smem_st8(s_adv + e0, d);
            smem_st8(s_ret + e0, r_out);
        }
        fence_proxy_async_smem();
        if (tid == 0) {
            double head = d[0];  // the advantage of the tile's first step
            if (is_ghost) {
                // exact iff the carry from beyond the ghost tile cannot reach its first step
                double Pt = 1.0;
#pragma unroll
                for (int w = 0; w < NW; ++w) Pt *= sP[b][w];
                if (fabs(Pt) >= kCarryNegligible && head_flags != nullptr) {
                    // no episode end in 2048 steps and no discount to speak of: the neighbour's
                    // published head value (the chain starts at CTA 0, which never waits; all
                    // CTAs of the launch are co-resident).  Never with row segments.
                    unsigned int f;
                    do {
                        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];"
                                     : "=r"(f) : "l"(head_flags + blockIdx.x - 1) : "memory");
                    } while (f != tag);
                    head = __ldcg(heads + blockIdx.x - 1);
                }
            }
            s_carry[b] = head;
        }
        // only the storing warp waits for the staged outputs; the next tile's __syncthreads orders
        // s_carry / sP / sX (other parity) and the reuse of the out buffer
        if (!is_ghost) {
            if (warp == NW - 1) {
                named_bar_sync(1, NT);
                if (lane == 0) {
                    const int64_t t0 = k * TILE;
                    bulk_s2g(adv + t0, smem_u32(s_adv), (uint32_t)n * sizeof(OutT));
                    bulk_s2g(ret + t0, smem_u32(s_ret), (uint32_t)n * sizeof(OutT));
                    bulk_commit();
                }
            } else {
                named_bar_arrive(1, NT);
            }
        }
        if (++stage == stages) { stage = 0; parity ^= 1u; }
    }
    if (tid == 0 && heads != nullptr) {
        // this range's first advantage (written by this thread), for a waiting neighbour
        heads[blockIdx.x] = s_carry[(nit - 1) & 1];
        __threadfence();
        *(volatile unsigned int*)(head_flags + blockIdx.x) = tag;
    }
    if (tid == STORER) bulk_wait0();
    leave();
}

// ============================================================================================
// General flattened kernel: single pass over N with decoupled look-back between tiles.
// Tiles are claimed through an atomic ticket so that a tile's successors in time (which it
// depends on) always started before it: ticket 0 is the LAST tile of the sequence.
// ============================================================================================
struct FlatWorkspace {
    size_t clear_bytes;    // look-back kernel: ticket + status are zeroed before every call
    unsigned int* ticket;  // [1]: look-back ticket, the word right before `status`
    unsigned int* sync;    // [3]: pipelined ranges: -, calls done, CTAs done.  NEVER cleared by the
                           // look-back path: the call count makes the head-flag tags unique for the
                           // lifetime of the workspace (stale flags of an earlier call never match)
    unsigned int* status;  // [tiles] 0 = nothing, 1 = aggregate published, 2 = value published
    double* aggP;          // [tiles]
    double* aggX;          // [tiles]
    double* value;         // [tiles] advantage at the tile's first element
};

constexpr int FLAT_NT = 256;
constexpr int FLAT_TILE = FLAT_NT * VEC;

__host__ __device__ inline int64_t flat_tiles(int64_t N) { return (N + FLAT_TILE - 1) / FLAT_TILE; }

template <typename RewT, typename InT, typename OutT>
__global__ void __launch_bounds__(FLAT_NT)
gae_flat_kernel(const RewT* __restrict__ rew, const InT* __restrict__ v_s,
                const InT* __restrict__ v_next, const uint8_t* __restrict__ term,
                const uint8_t* __restrict__ trunc, const int64_t* __restrict__ indices,
                const uint8_t* __restrict__ buf_term, const int64_t* __restrict__ unfinished,
                const int64_t* __restrict__ offset, int64_t B, int64_t sub_size, int64_t N,
                double gamma, double gl, OutT* __restrict__ adv, OutT* __restrict__ ret,
                FlatWorkspace ws, int vec_ok) {
    constexpr int NT = FLAT_NT;
    constexpr int NW = NT / 32;
    __shared__ double sP[NW], sX[NW];
    __shared__ double s_carry;
    __shared__ unsigned int s_ticket;

    const int tid = threadIdx.x;
    if (tid == 0) s_ticket = atomicAdd(ws.ticket, 1u);
    __syncthreads();
    const int64_t ticket = s_ticket;
    const int64_t ntiles = flat_tiles(N);
    const int64_t tile = ntiles - 1 - ticket;

    const int64_t i0 = tile * FLAT_TILE + (int64_t)tid * VEC;
    const int64_t left = N - i0;
    const int nvalid = left >= VEC ? VEC : (left > 0 ? (int)left : 0);
    const bool vec = vec_ok != 0;

    Strip s;
    double vn[VEC], rw[VEC];
    uint32_t mterm = 0, mtrunc = 0, mmask = 0, munf = 0;
    if (nvalid > 0) {
        load8(rew + i0, nvalid, vec, rw);
        load8(v_s + i0, nvalid, vec, s.vs);
        load8(v_next + i0, nvalid, vec, vn);
        mterm = load8_flags(term + i0, nvalid, vec);
        mtrunc = load8_flags(trunc + i0, nvalid, vec);
        mmask = mterm;
        if (indices != nullptr) {
            mmask = 0;
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                if (k < nvalid) {
                    const int64_t idx = __ldcs(indices + i0 + k);
                    // value_mask = ~buffer.terminated[indices]          (policy/base.py:335)
                    if (buf_term != nullptr ? __ldg(buf_term + idx) != 0 : ((mterm >> k) & 1u))
                        mmask |= 1u << k;
                    // end_flag[isin(indices, unfinished_index())] = True (:380)
                    if (unfinished != nullptr) {
                        const int64_t sub = sub_of(idx, offset, B, sub_size);
                        if (__ldg(unfinished + sub) == idx) munf |= 1u << k;
                    }
                }
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) { rw[k] = 0.0; s.vs[k] = 0.0; vn[k] = 0.0; }
    }
    s.valid = nvalid >= VEC ? 0xffu : ((1u << nvalid) - 1u);
    s.cut = mterm | mtrunc | munf;
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        double v1 = ((mmask >> k) & 1u) ? 0.0 : vn[k];
        s.delta[k] = fma(v1, gamma, rw[k]) - s.vs[k];
    }

    Affine mine = strip_aggregate(s, gl);
    Affine to_me;
    const Affine agg = block_suffix<NT>(mine, tid, to_me, sP, sX);

    // ---- decoupled look-back (warp 0) -------------------------------------------------------
    if (tid < 32) {
        const int lane = tid;
        double carry;
        if (ticket == 0) {
            carry = 0.0;
        } else {
            if (lane == 0) {
                ws.aggP[ticket] = agg.P;
                ws.aggX[ticket] = agg.X;
                __threadfence();
                *(volatile unsigned int*)(ws.status + ticket) = 1u;
            }
            Affine acc{1.0, 0.0};  // composition of the tiles already walked (nearest outermost)
            int64_t base = ticket - 1;
            bool finished = false;
            while (!finished) {
                const int64_t i = base - lane;
                Affine f{1.0, 0.0};
                unsigned int st;
                unsigned term_mask, ready_mask;
                int first_term;
                do {
                    st = i >= 0 ? *(volatile unsigned int*)(ws.status + i) : 2u;
                    bool is_term = false;
                    if (st != 0u) {
                        __threadfence();
                        if (i < 0) { f.P = 0.0; f.X = 0.0; is_term = true; }
                        else if (st == 2u) {
                            f.P = 0.0; f.X = *(volatile double*)(ws.value + i); is_term = true;
                        } else {
                            f.P = *(volatile double*)(ws.aggP + i);
                            f.X = *(volatile double*)(ws.aggX + i);
                            is_term = f.P == 0.0;
                        }
                    }
                    term_mask = __ballot_sync(0xffffffffu, is_term);
                    ready_mask = __ballot_sync(0xffffffffu, st != 0u);
                    first_term = term_mask ? __ffs(term_mask) - 1 : 32;
                    // need every lane up to (and including) the first terminating one
                } while ((ready_mask | ~((first_term >= 31) ? 0xffffffffu
                                                            : ((2u << first_term) - 1u))) !=
                         0xffffffffu);
                if (lane > first_term) { f.P = 1.0; f.X = 0.0; }  // not needed, maybe not ready
                // compose lanes 0..31, lane 0 outermost
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    double P2 = shfl_down_f64(f.P, d), X2 = shfl_down_f64(f.X, d);
                    if (lane + d < 32) { f.X = fma(f.P, X2, f.X); f.P *= P2; }
                }
                const double wP = shfl_idx_f64(f.P, 0), wX = shfl_idx_f64(f.X, 0);
                acc.X = fma(acc.P, wX, acc.X);
                acc.P *= wP;
                finished = first_term < 32;
                base -= 32;
            }
            carry = acc.X;  // acc.P == 0 here: the walk ended on a constant map
        }
        if (lane == 0) {
            const double v = fma(agg.P, carry, agg.X);
            ws.value[ticket] = v;
            __threadfence();
            *(volatile unsigned int*)(ws.status + ticket) = 2u;
            s_carry = carry;
        }
    }
    __syncthreads();
    const double my_carry = fma(to_me.P, s_carry, to_me.X);
    double a_out[VEC], r_out[VEC];
    (void)strip_replay(s, gl, my_carry, a_out, r_out);
    if (nvalid > 0) {
        store8(adv + i0, nvalid, vec, a_out);
        store8(ret + i0, nvalid, vec, r_out);
    }
}

// Variants of the canonical kernel (ts_tune_set("gae_rows_variant", id)); 0 = automatic.
struct RowsVariant { int nt, stages, ctas_per_sm; };  // nt == 0: register kernel
constexpr RowsVariant kRowsVariants[] = {
    {0, 0, 0},     // 0: automatic
    {0, 0, 0},     // 1: gae_rows_kernel (LDG -> registers, one CTA per row)
    {128, 3, 4},   // 2
    {128, 4, 3},   // 3
    {256, 3, 2},   // 4
    {256, 2, 3},   // 5
    {128, 2, 6},   // 6
    {256, 4, 1},   // 7
    {256, 6, 1},   // 8
    {128, 6, 2},   // 9
};
constexpr int kNumRowsVariants = (int)(sizeof(kRowsVariants) / sizeof(kRowsVariants[0]));
constexpr int kRowsAutoTma = 4;  // 256 threads x 3 stages x 2 CTAs/SM: best on distinct buffers (r01k_rows_variants.md)

// Persistent grid for `units` equal work items on at most `cap` co-resident CTAs.  Optional
// balancing ("gae_balance"): with w = ceil(units / cap) items per CTA, ceil(units / w) CTAs all
// get w or w - 1 items (4096 rows: 410 CTAs x 10 instead of 444 CTAs with 9 or 10).  Measured
// SLOWER at C2 (32.9 vs 31.7 us): the extra 34 CTAs' loads in flight are worth more than the
// even finish, so the default keeps the grid at the residency cap.
inline unsigned balanced_grid(int64_t units, int64_t cap) {
    if (units <= cap) return (unsigned)(units < 1 ? 1 : units);
    if (!g_tune_gae_balance) return (unsigned)cap;
    const int64_t w = (units + cap - 1) / cap;
    return (unsigned)((units + w - 1) / w);
}

// launch with programmatic stream serialization when enabled (see ts_common.cuh)
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem,
                       cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_tune_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

template <typename RewT, typename InT, typename OutT, int NT>
int32_t launch_rows_tma(const void* rew, const void* v_s, const void* v_next,
                        const uint8_t* term, const uint8_t* trunc, int64_t E, int64_t T,
                        int64_t bs, int64_t vs, double gamma, double lam, void* adv, void* ret,
                        int stages, int ctas_per_sm, cudaStream_t stream) {
    using L = RowsTmaLayout<RewT, InT, OutT, NT>;
    auto kern = gae_rows_tma_kernel<RewT, InT, OutT, NT>;
    while (stages > 2 && L::smem_bytes(stages) > 227 * 1024) --stages;  // wide dtypes
    const size_t smem = L::smem_bytes(stages);
    if (smem > 227 * 1024) return TS_ERR_BAD_ARG;
    static size_t attr_bytes = 0;  // per instantiation
    if (smem > attr_bytes) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
        attr_bytes = smem;
    }
    const unsigned grid = balanced_grid(E, (int64_t)kNumSMs * ctas_per_sm);
    GlPow glp;
    glp.p[0] = 1.0;
    for (int j = 1; j <= VEC; ++j) glp.p[j] = glp.p[j - 1] * (gamma * lam);
    cudaError_t e = launch_pdl(kern, grid, NT, smem, stream, (const RewT*)rew, (const InT*)v_s,
                               (const InT*)v_next, term, trunc, (int)E, T, bs, vs, gamma, glp,
                               (OutT*)adv, (OutT*)ret, stages);
    if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; (void)cudaGetLastError(); return TS_ERR_CUDA; }
    return check_launch();
}

template <typename RewT, typename InT, typename OutT, bool HAS_IDX>
int32_t launch_flat_tma(const void* rew, const void* v_s, const void* v_next, const uint8_t* term,
                        const uint8_t* trunc, const FlatGather& g, int64_t N, double gamma,
                        double lam, void* adv, void* ret, FlatWorkspace ws, cudaStream_t stream);

template <typename RewT, typename InT, typename OutT>
int32_t launch_rows(const void* rew, const void* v_s, const void* v_next, const uint8_t* term,
                    const uint8_t* trunc, int64_t E, int64_t T, int64_t bs, int64_t vs,
                    double gamma, double lam, void* adv, void* ret, cudaStream_t stream) {
    const int vec_ok = aligned(rew, 16) && aligned(v_s, 16) && aligned(v_next, 16) &&
                       aligned(term, 8) && aligned(trunc, 8) && aligned(adv, 16) &&
                       aligned(ret, 16) && bs % VEC == 0 && vs % VEC == 0;
    // bulk-copy path: every row segment starts and ends on a 16-byte boundary
    const bool tma_ok = vec_ok && aligned(term, 16) && aligned(trunc, 16) && T % 16 == 0 &&
                        bs % 16 == 0 && vs % 16 == 0 &&
                        E * ((T + 1023) / 1024) < 0x7fffffffLL;  // 32-bit tile counters
    int variant = g_tune_gae_rows_variant;
    if (variant < 0 || variant >= kNumRowsVariants) return TS_ERR_BAD_ARG;
    // Short rows stored back to back (T < one 2048-step tile): the row-per-tile kernel would run
    // its tiles mostly empty (T = 512: a quarter full).  The whole [E, T] block is instead walked
    // as ONE sequence of E * T steps by the pipelined-ranges kernel, with a cut at every row end;
    // a row end always lies inside a CTA's 2048-step ghost tile, so the carry into its range is
    // always exact and it needs no workspace (no neighbour flags).
    if (variant == 0 && g_tune_gae_rows_packed && bs == T && vs == T && T >= VEC && T < 2048 &&
        E * T >= (1 << 20) && vec_ok && aligned(term, 16) && aligned(trunc, 16) &&
        E * T < (1LL << 40)) {
        const FlatGather g{nullptr, nullptr, nullptr, nullptr, 0, 0, T};
        FlatWorkspace ws{};
        return launch_flat_tma<RewT, InT, OutT, false>(rew, v_s, v_next, term, trunc, g, E * T,
                                                       gamma, lam, adv, ret, ws, stream);
    }
    // automatic: the pipelined kernel pays off once there are enough rows to keep the
    // persistent CTAs busy with several tiles each; short batches stay on the register kernel
    if (variant == 0) variant = (tma_ok && E * T >= (1 << 20)) ? kRowsAutoTma : 1;
    const RowsVariant rv = kRowsVariants[variant];
    if (rv.nt != 0 && tma_ok) {
        if (rv.nt == 128)
            return launch_rows_tma<RewT, InT, OutT, 128>(rew, v_s, v_next, term, trunc, E, T, bs,
                                                         vs, gamma, lam, adv, ret, rv.stages,
                                                         rv.ctas_per_sm, stream);
        return launch_rows_tma<RewT, InT, OutT, 256>(rew, v_s, v_next, term, trunc, E, T, bs, vs,
                                                     gamma, lam, adv, ret, rv.stages,
                                                     rv.ctas_per_sm, stream);
    }
    const double gl = gamma * lam;
#define TS_ROWS(NT)                                                                            \
    gae_rows_kernel<RewT, InT, OutT, NT><<<(unsigned)E, NT, 0, stream>>>(                      \
        (const RewT*)rew, (const InT*)v_s, (const InT*)v_next, term, trunc, T, bs, vs, gamma, \
        gl, (OutT*)adv, (OutT*)ret, vec_ok)
    if (T > 1024) TS_ROWS(256);
    else if (T > 512) TS_ROWS(128);
    else if (T > 256) TS_ROWS(64);
    else TS_ROWS(32);
#undef TS_ROWS
    return check_launch();
}

template <typename RewT, typename InT, typename OutT, bool HAS_IDX>
int32_t launch_flat_tma(const void* rew, const void* v_s, const void* v_next, const uint8_t* term,
                        const uint8_t* trunc, const FlatGather& g, int64_t N, double gamma,
                        double lam, void* adv, void* ret, FlatWorkspace ws, cudaStream_t stream) {
    constexpr int NT = 256;
    using L = FlatTmaLayout<RewT, InT, OutT, NT, HAS_IDX>;
    auto kern = gae_flat_tma_kernel<RewT, InT, OutT, NT, HAS_IDX>;
    int want = g_tune_gae_flat_ctas > 0 ? g_tune_gae_flat_ctas : 3;
    int stages = g_tune_gae_flat_stages > 0 ? g_tune_gae_flat_stages : 2;
    if (stages < 2) stages = 2;
    while (stages > 2 && L::smem_bytes(stages) > 227 * 1024) --stages;
    const size_t smem = L::smem_bytes(stages);
    if (smem > 227 * 1024) return TS_ERR_BAD_ARG;
    // what actually fits (the CTAs of one launch may wait on each other: all must be resident)
    static int fits[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // per instantiation, by stage count
    static int attr_stages = 0;
    if (attr_stages != stages) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e == cudaSuccess && fits[stages] == 0)
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fits[stages], kern, NT, smem);
        if (e != cudaSuccess || fits[stages] < 1) {
            fits[stages] = 0;
            g_last_cuda_error = (int32_t)e;
            return TS_ERR_CUDA;
        }
        attr_stages = stages;
    }
    const int ctas_per_sm = fits[stages] < want ? fits[stages] : want;
    const int64_t tiles = ((N & ~(int64_t)15) + L::TILE - 1) / L::TILE;
    const unsigned grid = balanced_grid(tiles, (int64_t)kNumSMs * ctas_per_sm);
    GlPow glp;
    glp.p[0] = 1.0;
    for (int j = 1; j <= VEC; ++j) glp.p[j] = glp.p[j - 1] * (gamma * lam);
    cudaError_t e = launch_pdl(kern, grid, NT, smem, stream, (const RewT*)rew, (const InT*)v_s,
                               (const InT*)v_next, term, trunc, g, N, gamma, glp, (OutT*)adv,
                               (OutT*)ret, stages, ws.aggP, ws.status, ws.sync);
    if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; (void)cudaGetLastError(); return TS_ERR_CUDA; }
    return check_launch();
}

template <typename RewT, typename InT, typename OutT>
int32_t launch_flat(const void* rew, const void* v_s, const void* v_next, const uint8_t* term,
                    const uint8_t* trunc, const int64_t* indices, const uint8_t* buf_term,
                    const int64_t* unfinished, const int64_t* offset, int64_t B,
                    int64_t sub_size, int64_t N, double gamma, double lam, void* adv, void* ret,
                    FlatWorkspace ws, cudaStream_t stream) {
    const int vec_ok = aligned(rew, 16) && aligned(v_s, 16) && aligned(v_next, 16) &&
                       aligned(term, 8) && aligned(trunc, 8) && aligned(adv, 16) &&
                       aligned(ret, 16);
    const bool tma_ok = vec_ok && aligned(term, 16) && aligned(trunc, 16) &&
                        (indices == nullptr || aligned(indices, 16)) && N < (1LL << 40);
    int variant = g_tune_gae_flat_variant;  // 0 auto, 1 decoupled look-back, 2 pipelined ranges
    // short sequences have too few tiles per CTA to amortise the ghost tile: look-back kernel
    if (variant == 0) variant = N >= (1 << 20) ? 2 : 1;
    if (variant == 2 && tma_ok) {
        const FlatGather g{indices, buf_term, unfinished, offset, B, sub_size, 0};
        if (indices != nullptr)
            return launch_flat_tma<RewT, InT, OutT, true>(rew, v_s, v_next, term, trunc, g, N, gamma,
                                                          lam, adv, ret, ws, stream);
        return launch_flat_tma<RewT, InT, OutT, false>(rew, v_s, v_next, term, trunc, g, N, gamma,
                                                       lam, adv, ret, ws, stream);
    }
    {
        cudaError_t e = cudaMemsetAsync(ws.ticket, 0, ws.clear_bytes, stream);
        if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
    }
    gae_flat_kernel<RewT, InT, OutT><<<(unsigned)flat_tiles(N), FLAT_NT, 0, stream>>>(
        (const RewT*)rew, (const InT*)v_s, (const InT*)v_next, term, trunc, indices, buf_term,
        unfinished, offset, B, sub_size, N, gamma, gamma * lam, (OutT*)adv, (OutT*)ret, ws,
        vec_ok);
    return check_launch();
}

}  // namespace
}  // namespace ts

using namespace ts;

// dtype dispatch: (rew, value, out) in {f32, f64}^3
#define TS_DISPATCH3(R, V, O, CALL)                                                           \
    do {                                                                                      \
        const int key = ((R) == TS_F64 ? 4 : 0) | ((V) == TS_F64 ? 2 : 0) | ((O) == TS_F64 ? 1 : 0); \
        switch (key) {                                                                        \
            case 0: return CALL(float, float, float);                                         \
            case 1: return CALL(float, float, double);                                        \
            case 2: return CALL(float, double, float);                                        \
            case 3: return CALL(float, double, double);                                       \
            case 4: return CALL(double, float, float);                                        \
            case 5: return CALL(double, float, double);                                       \
            case 6: return CALL(double, double, float);                                       \
            default: return CALL(double, double, double);                                     \
        }                                                                                     \
    } while (0)

static bool dtype_ok(int32_t d) { return d == TS_F32 || d == TS_F64; }

extern "C" int32_t ts_gae_rows(const void* rew, const void* v_s, const void* v_s_next,
                               const uint8_t* terminated, const uint8_t* truncated, int64_t E,
                               int64_t T, int64_t buf_stride, int64_t val_stride, double gamma,
                               double gae_lambda, int32_t rew_dtype, int32_t val_dtype,
                               int32_t out_dtype, void* adv, void* ret, ts_stream_t stream) {
    if (E < 0 || T < 0 || buf_stride < T || val_stride < T) return TS_ERR_BAD_ARG;
    if (!dtype_ok(rew_dtype) || !dtype_ok(val_dtype) || !dtype_ok(out_dtype))
        return TS_ERR_BAD_ARG;
    if (E == 0 || T == 0) return TS_OK;
    if (!rew || !v_s || !v_s_next || !terminated || !truncated || !adv || !ret)
        return TS_ERR_BAD_ARG;
    if (E > 0x7fffffffLL) return TS_ERR_BAD_ARG;
    const size_t rsz = rew_dtype == TS_F64 ? 8 : 4, isz = val_dtype == TS_F64 ? 8 : 4,
                 osz = out_dtype == TS_F64 ? 8 : 4;
    if (!aligned(rew, rsz) || !aligned(v_s, isz) || !aligned(v_s_next, isz) ||
        !aligned(adv, osz) || !aligned(ret, osz))
        return TS_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
#define TS_GO(R, I, O)                                                                        \
    launch_rows<R, I, O>(rew, v_s, v_s_next, terminated, truncated, E, T, buf_stride,        \
                         val_stride, gamma, gae_lambda, adv, ret, st)
    TS_DISPATCH3(rew_dtype, val_dtype, out_dtype, TS_GO);
#undef TS_GO
}

extern "C" size_t ts_gae_flat_workspace_bytes(int64_t N) {
    const size_t tiles = (size_t)flat_tiles(N < 1 ? 1 : N);
    // ticket (16 B) + status (u32, padded to 16) + 3 double arrays
    return 16 + ((tiles * 4 + 15) / 16) * 16 + 3 * tiles * 8;
}

extern "C" int32_t ts_gae_flat(const void* rew, const void* v_s, const void* v_s_next,
                               const uint8_t* terminated, const uint8_t* truncated,
                               const int64_t* indices, const uint8_t* buf_terminated,
                               const int64_t* unfinished, const int64_t* offset, int64_t B,
                               int64_t sub_size, int64_t N, double gamma, double gae_lambda,
                               int32_t rew_dtype, int32_t val_dtype, int32_t out_dtype,
                               void* adv, void* ret, void* workspace, size_t workspace_bytes,
                               ts_stream_t stream) {
    if (N < 0) return TS_ERR_BAD_ARG;
    if (!dtype_ok(rew_dtype) || !dtype_ok(val_dtype) || !dtype_ok(out_dtype))
        return TS_ERR_BAD_ARG;
    if (N == 0) return TS_OK;
    if (!rew || !v_s || !v_s_next || !terminated || !truncated || !adv || !ret || !workspace)
        return TS_ERR_BAD_ARG;
    if (unfinished != nullptr && (indices == nullptr || B <= 0 || (sub_size <= 0 && !offset)))
        return TS_ERR_BAD_ARG;
    if (workspace_bytes < ts_gae_flat_workspace_bytes(N)) return TS_ERR_WORKSPACE;
    if (!aligned(workspace, 16)) return TS_ERR_ALIGN;
    const size_t rsz = rew_dtype == TS_F64 ? 8 : 4, isz = val_dtype == TS_F64 ? 8 : 4,
                 osz = out_dtype == TS_F64 ? 8 : 4;
    if (!aligned(rew, rsz) || !aligned(v_s, isz) || !aligned(v_s_next, isz) ||
        !aligned(adv, osz) || !aligned(ret, osz))
        return TS_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t tiles = (size_t)flat_tiles(N);
    const size_t status_bytes = ((tiles * 4 + 15) / 16) * 16;
    char* w = (char*)workspace;
    FlatWorkspace ws;
    ws.sync = (unsigned int*)w;           // words 0..2
    ws.ticket = (unsigned int*)(w + 12);  // word 3, contiguous with status: one memset clears both
    ws.status = (unsigned int*)(w + 16);
    ws.aggP = (double*)(w + 16 + status_bytes);
    ws.aggX = ws.aggP + tiles;
    ws.value = ws.aggX + tiles;
    ws.clear_bytes = 4 + status_bytes;
#define TS_GO(R, I, O)                                                                        \
    launch_flat<R, I, O>(rew, v_s, v_s_next, terminated, truncated, indices, buf_terminated, \
                         unfinished, offset, B, sub_size, N, gamma, gae_lambda, adv, ret, ws, \
                         st)
    TS_DISPATCH3(rew_dtype, val_dtype, out_dtype, TS_GO);
#undef TS_GO
}


// ---------------------------------------------------------------------------------------------
// ts_gae_flat_static: ts_gae_flat for callers that cannot provide (or allocate) a workspace.
// No counterpart in the reference.
//
// The scratch of the flat kernels (call counter, inter-CTA flags, tile aggregates) lives in a
// file-scope device array of fixed size, held by every GPU that loads this library until the
// process exits.  Zero at module load, which is the "zero-initialised once" the kernels need;
// afterwards it keeps the last call's contents like any caller-owned workspace (the look-back
// kernel clears its own ticket and status words per call, the pipelined kernel tags its flags
// with a call count), so back-to-back calls on ONE stream are safe.  Every call shares the
// array: calls that overlap on the GPU (two streams, two host threads) race on it.
// ---------------------------------------------------------------------------------------------
#include <mutex>

namespace ts {
namespace {
constexpr int64_t kStaticGaeMaxN = TS_GAE_FLAT_STATIC_MAX_N;
constexpr size_t kStaticGaeTiles = (size_t)((kStaticGaeMaxN + FLAT_TILE - 1) / FLAT_TILE);
constexpr size_t kStaticGaeBytes =
    16 + ((kStaticGaeTiles * 4 + 15) / 16) * 16 + 3 * kStaticGaeTiles * 8;  // 229 392 B
}  // namespace
}  // namespace ts

static __device__ __align__(256) unsigned char g_gae_flat_scratch[ts::kStaticGaeBytes];

// In host code `g_gae_flat_scratch` is a host-side stand-in, not a device address: the address
// the kernels get comes from cudaGetSymbolAddress, once per device, with that device current.
static int32_t gae_static_scratch(void** out) {
    constexpr int kMaxDevices = 64;
    static void* ptrs[kMaxDevices] = {};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        g_last_cuda_error = (int32_t)(e != cudaSuccess ? e : cudaErrorInvalidDevice);
        return TS_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_gae_flat_scratch);
        if (e != cudaSuccess || p == nullptr) {
            g_last_cuda_error = (int32_t)e;
            (void)cudaGetLastError();
            return TS_ERR_CUDA;
        }
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return TS_OK;
}

extern "C" int64_t ts_gae_flat_static_max_n(void) { return ts::kStaticGaeMaxN; }

extern "C" int32_t ts_gae_flat_static(const void* rew, const void* v_s, const void* v_s_next,
                                      const uint8_t* terminated, const uint8_t* truncated,
                                      const int64_t* indices, const uint8_t* buf_terminated,
                                      const int64_t* unfinished, const int64_t* offset, int64_t B,
                                      int64_t sub_size, int64_t N, double gamma,
                                      double gae_lambda, int32_t rew_dtype, int32_t val_dtype,
                                      int32_t out_dtype, void* adv, void* ret,
                                      ts_stream_t stream) {
    if (N > ts::kStaticGaeMaxN) return TS_ERR_TOO_LARGE;
    if (N <= 0) return N == 0 ? TS_OK : TS_ERR_BAD_ARG;
    static_assert(ts::kStaticGaeBytes <= sizeof(g_gae_flat_scratch), "scratch covers the maximum");
    void* scratch = nullptr;
    const int32_t rc = gae_static_scratch(&scratch);
    if (rc != TS_OK) return rc;
    return ts_gae_flat(rew, v_s, v_s_next, terminated, truncated, indices, buf_terminated,
                       unfinished, offset, B, sub_size, N, gamma, gae_lambda, rew_dtype,
                       val_dtype, out_dtype, adv, ret, scratch, ts::kStaticGaeBytes, stream);
}
