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
// General flattened kernel: single pass over N with decoupled look-back between tiles.
// Tiles are claimed through an atomic ticket so that a tile's successors in time (which it
// depends on) always started before it: ticket 0 is the LAST tile of the sequence.
// ============================================================================================
struct FlatWorkspace {
    unsigned int* ticket;  // [4] (padding)
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

template <typename RewT, typename InT, typename OutT>
int32_t launch_rows(const void* rew, const void* v_s, const void* v_next, const uint8_t* term,
                    const uint8_t* trunc, int64_t E, int64_t T, int64_t bs, int64_t vs,
                    double gamma, double lam, void* adv, void* ret, cudaStream_t stream) {
    const int vec_ok = aligned(rew, 16) && aligned(v_s, 16) && aligned(v_next, 16) &&
                       aligned(term, 8) && aligned(trunc, 8) && aligned(adv, 16) &&
                       aligned(ret, 16) && bs % VEC == 0 && vs % VEC == 0;
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

template <typename RewT, typename InT, typename OutT>
int32_t launch_flat(const void* rew, const void* v_s, const void* v_next, const uint8_t* term,
                    const uint8_t* trunc, const int64_t* indices, const uint8_t* buf_term,
                    const int64_t* unfinished, const int64_t* offset, int64_t B,
                    int64_t sub_size, int64_t N, double gamma, double lam, void* adv, void* ret,
                    FlatWorkspace ws, cudaStream_t stream) {
    const int vec_ok = aligned(rew, 16) && aligned(v_s, 16) && aligned(v_next, 16) &&
                       aligned(term, 8) && aligned(trunc, 8) && aligned(adv, 16) &&
                       aligned(ret, 16);
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
    ws.ticket = (unsigned int*)w;
    ws.status = (unsigned int*)(w + 16);
    ws.aggP = (double*)(w + 16 + status_bytes);
    ws.aggX = ws.aggP + tiles;
    ws.value = ws.aggX + tiles;
    cudaError_t e = cudaMemsetAsync(w, 0, 16 + status_bytes, st);
    if (e != cudaSuccess) { g_last_cuda_error = (int32_t)e; return TS_ERR_CUDA; }
#define TS_GO(R, I, O)                                                                        \
    launch_flat<R, I, O>(rew, v_s, v_s_next, terminated, truncated, indices, buf_terminated, \
                         unfinished, offset, B, sub_size, N, gamma, gae_lambda, adv, ret, ws, \
                         st)
    TS_DISPATCH3(rew_dtype, val_dtype, out_dtype, TS_GO);
#undef TS_GO
}
