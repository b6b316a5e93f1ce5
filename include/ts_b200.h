/*
 * ts_b200.h -- C-ABI of the B200 (sm_100a) return / replay hot path.
 *
 * The reference (aai-institute/tianshou-fork, Tianshou 0.5.1) has no FFI layer on this
 * path: its boundary is plain Python (tianshou/data/__init__.py:4-41 and the static methods
 * of tianshou/policy/base.py:BasePolicy).  The compiled code it calls is five numba @njit
 * functions; each entry point below replaces one of them (or one NumPy expression around
 * them) and cites it.  The Python host package `tianshou_fork_b200` binds these symbols with
 * ctypes and mirrors the reference classes; INTEGRATION.md shows the stub a reference
 * maintainer would add.
 *
 * Conventions
 *   - Plain pointers and sizes only; no torch / C++ types.  All data pointers are DEVICE
 *     pointers owned by the caller and valid for the stream-ordered duration of the call.
 *   - Every entry point enqueues work on `stream` (a cudaStream_t passed as void*) and
 *     returns immediately: it never allocates, never synchronises and never reads device
 *     memory on the host, so all of them are legal inside CUDA-graph capture.
 *   - Return value: TS_OK or a negative TS_ERR_* code.  ts_last_cuda_error() gives the
 *     cudaError_t behind the most recent TS_ERR_CUDA on the calling thread.
 *   - Range contracts that the reference enforces with `assert` on host arrays
 *     (segtree.py:53, :79) are evaluated on the device into a caller-provided int32
 *     `err_flag` word (bit mask TS_FLAG_*) which the host shim checks lazily.
 *   - Stateless and re-entrant; ordering is the stream's.
 *   - Booleans are uint8 (numpy bool_ / torch.bool layout).  Indices are int64.
 */
#ifndef TS_B200_H_
#define TS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TS_B200_VERSION 100 /* major*1000 + minor*100 + patch */

#define TS_OK 0
#define TS_ERR_BAD_ARG (-1)   /* null pointer, negative size, unsupported dtype code */
#define TS_ERR_WORKSPACE (-2) /* workspace_bytes smaller than ts_*_workspace_bytes() */
#define TS_ERR_CUDA (-3)      /* launch failed; see ts_last_cuda_error() */
#define TS_ERR_ALIGN (-4)     /* pointer not aligned to its element size */
#define TS_ERR_TOO_LARGE (-5) /* size beyond the fixed capacity of a *_static entry point */

/* dtype codes for the float-typed arguments */
#define TS_F32 0
#define TS_F64 1

/* bits of the device-side error flag word */
#define TS_FLAG_INDEX_RANGE 1  /* SegmentTree.__setitem__: index outside [0, size)   */
#define TS_FLAG_PREFIX_RANGE 2 /* get_prefix_sum_idx: value outside [0, tree[1])      */
#define TS_FLAG_PEER_TIMEOUT 4 /* sharded PER: a peer's record never arrived (bounded spin) */

typedef void* ts_stream_t; /* cudaStream_t */

int32_t ts_version(void);
const char* ts_error_string(int32_t code);
int32_t ts_last_cuda_error(void);
/* Kernel-selection knobs for benchmarking and for the tests that cover every kernel
 * (process-wide, not part of the reference-facing surface).  Keys:
 *   "gae_rows_variant"  0 = automatic, else an entry of csrc/gae.cu:kRowsVariants
 *   "gae_flat_variant"  0 = automatic, 1 = decoupled look-back, 2 = pipelined ranges
 *   "gae_flat_ctas"     pipelined ranges: CTAs per SM (0 = automatic = 3, 1..3)
 *   "gae_flat_stages"   pipelined ranges: stages of the bulk-copy ring (0 = automatic = 2, 2..6)
 *   "segtree_ranged"    1 (default) = ranged update kernel for trees <= 2^20 leaves, 0 = tiered
 *   "segtree_group"     tiered kernel: 128 (default) or 256 nodes per warp-rebuilt group
 *   "segtree_max_ctas"  tiered kernel: CTAs of the cooperative launch (<= 148)
 *   "per_cluster"       PER sampling cluster: 0 = automatic, 8 (x512 threads) or 16 (x256)
 *   "gae_rows_packed"   1 (default) = contiguous rows shorter than a 2048-step tile are walked as
 *                       one sequence with row cuts (pipelined ranges), 0 = one row per tile
 *   "gae_balance"       1 = persistent GAE grids sized for equal work per CTA (default 0)
 *   "pdl"               1 (default) = PER and GAE rows kernels use programmatic dependent
 *                       launch, 0 = off */
int32_t ts_tune_set(const char* key, int64_t value);
/* Debug: globaltimer ns at the phase boundaries of the most recent small-batch tree update
 * (synchronises the device; host buffer of 8 uint64: start, batch scanned, groups rebuilt,
 * last-CTA ticket, -, end, table cleared, batch loads landed). */
int32_t ts_debug_update_phases(uint64_t* out8);
/* Debug: per-CTA globaltimer ns {start, batch scanned, groups rebuilt, ticket taken} of the most
 * recent ranged tree update, 128 CTAs x 4 (synchronises the device). */
int32_t ts_debug_update_ctas(uint64_t* out512);

/* ------------------------------------------------------------------------------------------
 * GAE / TD(lambda) returns.
 * Replaces: BasePolicy.compute_episodic_return (tianshou/policy/base.py:337-384) with
 *           value_mask (:317-335) and the numba kernel _gae_return (:453-497) fused:
 *   v_s_  <- v_s_ * ~terminated                       (:375-376)
 *   end   <- terminated | truncated | unfinished-tail (:379-380)
 *   A_t   <- delta_t + gamma*lambda*(1-end_t) * A_{t+1},  delta_t = rew + gamma*v_s_ - v_s
 *   ret   <- A + v_s                                  (:382)
 * Arithmetic is f64 (as the reference's numba path); inputs/outputs are f32 or f64.
 * ---------------------------------------------------------------------------------------- */

/*
 * Canonical rollout layout: E independent env streams (VectorReplayBuffer sub-buffers,
 * vecbuf.py:33-37) of T steps each, row-major.  rew/terminated/truncated rows are
 * `buf_stride` elements apart (the sub-buffer capacity), v_s/v_s_next/adv/ret rows are
 * `val_stride` apart.  The last step of every row is an episode boundary: it is either done
 * or the sub-buffer's unfinished index (manager.py:68-74), so rows never exchange data.
 * rew_dtype applies to rew, val_dtype to v_s and v_s_next, out_dtype to adv and ret.
 */
int32_t ts_gae_rows(const void* rew, const void* v_s, const void* v_s_next,
                    const uint8_t* terminated, const uint8_t* truncated, int64_t E, int64_t T,
                    int64_t buf_stride, int64_t val_stride, double gamma, double gae_lambda,
                    int32_t rew_dtype, int32_t val_dtype, int32_t out_dtype, void* adv,
                    void* ret, ts_stream_t stream);

/*
 * General form: one flattened batch of N transitions in the caller's order, exactly the
 * reference's contract (batch == buffer[indices]); the recurrence runs across the whole
 * array and is cut only by end flags, as _gae_return does.
 *   rew, terminated, truncated, v_s, v_s_next : batch order, length N.
 *   indices (nullable)        : batch position -> buffer index.  When given,
 *                               buf_terminated[indices[i]] is the value mask (:335) and
 *                               indices[i] == unfinished[sub(indices[i])] forces end (:380).
 *   unfinished[B]             : per sub-buffer global index of the unfinished tail or -1
 *                               (output of ts_unfinished_index).
 *   offset[B+1], sub_size     : sub-buffer layout (_extend_offset, manager.py:38);
 *                               sub_size > 0 means equal sub-buffers (k = idx / sub_size).
 * workspace: ts_gae_flat_workspace_bytes(N) bytes, 16-byte aligned, zero-initialised ONCE by
 *   the caller and opaque afterwards (one stream at a time): the pipelined kernel keeps a call
 *   count in it that tags its inter-CTA flags, so no call has to clear anything (calls of any
 *   size may alternate on one workspace: the short-sequence kernel clears only its own words).
 * Approximation (documented, bounded): a CTA of the pipelined kernel obtains the carry into its
 *   range from a "ghost" pass over the next 2048 steps; when no episode ends there it treats a
 *   discount product (gamma*lambda)^2048 < 1e-30 as zero (csrc/gae.cu:kCarryNegligible) -- a
 *   relative error below 1e-30 of the carried advantage, fourteen orders under one f64 ulp; every
 *   other operation is the reference's f64 recurrence.
 */
size_t ts_gae_flat_workspace_bytes(int64_t N);
int32_t ts_gae_flat(const void* rew, const void* v_s, const void* v_s_next,
                    const uint8_t* terminated, const uint8_t* truncated, const int64_t* indices,
                    const uint8_t* buf_terminated, const int64_t* unfinished,
                    const int64_t* offset, int64_t B, int64_t sub_size, int64_t N, double gamma,
                    double gae_lambda, int32_t rew_dtype, int32_t val_dtype, int32_t out_dtype,
                    void* adv, void* ret, void* workspace, size_t workspace_bytes,
                    ts_stream_t stream);

/*
 * ts_gae_flat without a caller workspace and without any allocation.
 * NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and results as ts_gae_flat minus `workspace`; the scratch
 * is a file-scope `static __device__` array of 229 392 bytes inside the library (resident on
 * every GPU that loads it until the process exits), which bounds the call at
 * N <= TS_GAE_FLAT_STATIC_MAX_N transitions: larger N returns TS_ERR_TOO_LARGE before any
 * launch.  The call performs no cudaMalloc / cudaFree / allocating library call.
 * Every call shares that array and it keeps the last call's contents (the kernels clear or
 * re-tag what must start fresh), so calls issued back to back on ONE stream are safe, while
 * calls that overlap on the GPU -- other streams, other host threads -- race on it: use
 * ts_gae_flat with one workspace per stream for those.
 */
#define TS_GAE_FLAT_STATIC_MAX_N (1 << 24)
int64_t ts_gae_flat_static_max_n(void);
int32_t ts_gae_flat_static(const void* rew, const void* v_s, const void* v_s_next,
                           const uint8_t* terminated, const uint8_t* truncated,
                           const int64_t* indices, const uint8_t* buf_terminated,
                           const int64_t* unfinished, const int64_t* offset, int64_t B,
                           int64_t sub_size, int64_t N, double gamma, double gae_lambda,
                           int32_t rew_dtype, int32_t val_dtype, int32_t out_dtype, void* adv,
                           void* ret, ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Episode-aware index arithmetic.
 * Replaces: numba _prev_index / _next_index (tianshou/data/buffer/manager.py:259-297),
 *           ReplayBuffer.prev / next (tianshou/data/buffer/base.py:153-168, the B == 1 case
 *           with lengths[0] == current size) and unfinished_index (base.py:148-151,
 *           manager.py:68-74).  Integer work: results are bit-exact.
 *   index[M] any int64 (reduced mod offset[B] first, so -1 maps to the last slot);
 *   offset[B+1]; done[cap]; last_index[B]; lengths[B]; sub_size as above.
 * ---------------------------------------------------------------------------------------- */
int32_t ts_prev_index(const int64_t* index, int64_t M, const int64_t* offset, int64_t B,
                      int64_t sub_size, const uint8_t* done, const int64_t* last_index,
                      const int64_t* lengths, int64_t* out, ts_stream_t stream);
int32_t ts_next_index(const int64_t* index, int64_t M, const int64_t* offset, int64_t B,
                      int64_t sub_size, const uint8_t* done, const int64_t* last_index,
                      const int64_t* lengths, int64_t* out, ts_stream_t stream);
/* out[k] = last_index[k] if lengths[k] > 0 and !done[last_index[k]] else -1 */
int32_t ts_unfinished_index(const uint8_t* done, const int64_t* last_index,
                            const int64_t* lengths, int64_t B, int64_t* out,
                            ts_stream_t stream);

/* Unfinished tails as cut flags at BATCH positions, for a batch that is the buffer's own
 * chronological listing (sample_indices(0), tianshou/data/buffer/manager.py:187-192): sub-buffer
 * k occupies batch positions run_first[k] + j, j < lengths[k], with buffer index
 * (ring_start[k] + j) % lengths[k] + offset[k].  Sets flags[pos] = 1 (uint8[N], otherwise
 * untouched) where the reference sets end_flag[np.isin(indices, unfinished_index())]
 * (tianshou/policy/base.py:379-380): B look-ups instead of a pass over the N indices, so that
 * ts_gae_flat can run without them. */
int32_t ts_mark_unfinished(const uint8_t* done, const int64_t* last_index,
                           const int64_t* lengths, const int64_t* offset,
                           const int64_t* ring_start, const int64_t* run_first, int64_t B,
                           int64_t N, uint8_t* flags, ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * n-step TD target.
 * Replaces: BasePolicy.compute_nstep_return (tianshou/policy/base.py:386-440) and numba
 *           _nstep_return (:500-524).  Two launches around the caller's target network:
 *   chain  : walks buffer.next n_step-1 times (:422-425, fused, no [n,bsz] matrix), with
 *            end = done | unfinished (:433-434); emits terminal[bsz] (:428), the partial
 *            discounted reward sum[bsz] (f64) and gamma^k * value_mask(terminal) coef[bsz].
 *   combine: out[b,d] = target_q[b,d] * coef[b] + sum[b]            (:432, :523)
 * ---------------------------------------------------------------------------------------- */
int32_t ts_nstep_chain(const int64_t* indice, int64_t bsz, const void* rew, int32_t rew_dtype,
                       const uint8_t* done, const uint8_t* terminated, const int64_t* offset,
                       int64_t B, int64_t sub_size, const int64_t* last_index,
                       const int64_t* lengths, double gamma, int32_t n_step, int64_t* terminal,
                       double* ret_sum, double* coef, ts_stream_t stream);
int32_t ts_nstep_combine(const void* target_q, int32_t q_dtype, int64_t bsz, int64_t D,
                         const double* ret_sum, const double* coef, void* out,
                         ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Segment tree (f64 binary heap of 2*bound nodes, leaf i at bound+i).
 * Replaces: numba _setitem / _reduce / _get_prefix_sum_idx
 *           (tianshou/data/utils/segtree.py:98-137) behind SegmentTree.__setitem__ (:37-54),
 *           reduce (:56-64) and get_prefix_sum_idx (:66-85).  Every node is the IEEE f64 sum
 *           left + right (no FMA), so node values are bit-identical to the reference's.
 * ---------------------------------------------------------------------------------------- */

/* Scratch owned by the tree object, both zero-initialised ONCE by the caller and opaque
 * afterwards (one tree, one stream at a time): `stamp` is uint32[size], 16-byte aligned
 * (last-writer-wins arbitration of duplicated indices, segtree.py:40-49; entries carry a
 * per-call epoch so they never need resetting), `counter` is uint32[4] (grid-barrier
 * arrivals, epoch).  Batches of up to 16383 indices run as ONE launch: on trees of 4097 ..
 * 2^20 leaves an ordinary launch whose CTAs never wait on each other (only counter[0] is
 * used: "last CTA out finishes the top"); on larger trees a cooperative launch (<= 148 CTAs
 * that wait on each other: do not run two of them concurrently on one device from
 * different streams). */
/* tree[index[j]] = value[j] for j < M (M < 2^31), then every ancestor recomputed bottom-up.
 * value is f32 or f64 (cast on store, segtree.py:91-92); value_stride 0 broadcasts value[0].
 * index outside [0,size) sets TS_FLAG_INDEX_RANGE in *err_flag and that item is skipped. */
int32_t ts_segtree_update(double* tree, int64_t size, int64_t bound, const int64_t* index,
                          const void* value, int32_t value_dtype, int64_t value_stride,
                          int64_t M, uint32_t* stamp, uint32_t* counter, int32_t* err_flag,
                          ts_stream_t stream);
/* *out = sum(leaf[start:end]) with the reference's climb order (:107-119); start==0 and
 * end==size is tree[1] up to that order -- the caller handles the :58-59 fast path. */
int32_t ts_segtree_reduce(const double* tree, int64_t bound, int64_t start, int64_t end,
                          double* out, ts_stream_t stream);
/* Batched prefix-sum descent (:122-137): strict `<`, subtract-on-right, `value` mutated in
 * place like the reference (:134).  value_dtype f64 or f32.  If scale_by_total != 0 the
 * query is value[j] * tree[1] (prio.py:61) computed on device in f64 (f64 values only). */
int32_t ts_segtree_prefix_idx(const double* tree, int64_t bound, void* value,
                              int32_t value_dtype, int64_t M, int32_t scale_by_total,
                              int64_t* out, int32_t* err_flag, ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Prioritized replay, fused steps.
 * Replaces: PrioritizedReplayBuffer.sample_indices / get_weight / __getitem__ weight
 *           normalisation / update_weight (tianshou/data/buffer/prio.py:59-102).
 * `prio` is double[2] on the device: {max_prio, min_prio} (prio.py:35, :88-89).
 * ---------------------------------------------------------------------------------------- */
/* idx = prefix_idx(u * tree[1]); weight = (leaf/min_prio)^(-beta); if weight_norm,
 * weight /= max(weight) (done by the last CTA to finish, same launch).  u is NOT mutated.
 * The batch is handled by one 8-CTA thread-block cluster; the maximum is exchanged through
 * distributed shared memory.  scratch16 is unused (kept for ABI stability; may be NULL). */
int32_t ts_per_sample(const double* tree, int64_t bound, const double* u, int64_t M,
                      const double* prio, double beta, int32_t weight_norm, int64_t* idx,
                      double* weight, void* scratch16, int32_t* err_flag, ts_stream_t stream);
/* Sharded form (one process per GPU): this GPU holds shard `rank` of G
 * power-of-two-aligned subtrees of the reference's single global tree (manager.py:209-214).
 * roots3 = double[G][3] = {shard root, max_prio, min_prio} of every rank (all-gathered by the
 * caller after each update).  The log2(G) levels above the shard roots are rebuilt in the global
 * tree's pairwise order, queries are routed down them, and those that land in this shard
 * descend locally: idx[j] = LOCAL leaf, or -1 (weight 0) for queries owned by another rank.
 * min_prio is the minimum over all ranks.  With weight_norm the kernel writes the maximum raw
 * weight of this rank to *wmax_out; after a MAX all-reduce the caller applies ts_per_normalize. */
int32_t ts_per_sample_sharded(const double* tree, int64_t bound, const double* u, int64_t M,
                              const double* roots3, int32_t G, int32_t rank, const double* prio,
                              double beta, int32_t weight_norm, int64_t* idx, double* weight,
                              double* wmax_out, int32_t* err_flag, ts_stream_t stream);
/* The same step with the exchange INSIDE the kernels, over NVLink peer memory (no collective
 * library call, CUDA-graph capturable).  Every rank owns `slots`, a zero-initialised buffer of
 * 16 * G doubles that all peers can write (symmetric memory / cudaIpc); `peer_slots` is a DEVICE
 * array of the G ranks' buffer addresses as mapped in this process; `state` is uint64[2] on the
 * device, zero-initialised: {records published, samples drawn}.
 *   ts_per_publish    after every tree change (fill, add, update_weight): stores this rank's
 *                     {root, max_prio, min_prio} into every peer's buffer as six self-validating
 *                     64-bit words (32 bits of payload + the low 32 bits of the epoch each: no
 *                     flag word and no ordering between the stores, one NVLink hop).
 *   ts_per_sample_p2p waits (bounded; TS_FLAG_PEER_TIMEOUT) for the G records of the current
 *                     epoch, routes and descends as ts_per_sample_sharded, exchanges the weight
 *                     maximum the same way and writes NORMALISED weights; adopts the global
 *                     priority extrema into `prio` (the reference has one pair, prio.py:35).
 * All ranks issue the same sequence of calls (SPMD). */
int32_t ts_per_publish(const double* tree, const double* prio, uint64_t* state,
                       double* const* peer_slots, int32_t G, int32_t rank, ts_stream_t stream);
/* ts_per_update followed by ts_per_publish as ONE launch: the CTA that writes the new root also
 * stores this rank's record into every peer's buffer (compute + exchange fused). */
int32_t ts_per_update_publish(double* tree, int64_t size, int64_t bound, const int64_t* idx,
                              const void* td, int32_t td_dtype, int64_t M, double alpha,
                              double* prio, uint32_t* stamp, uint32_t* counter,
                              int32_t* err_flag, uint64_t* state, double* const* peer_slots,
                              int32_t G, int32_t rank, ts_stream_t stream);
int32_t ts_per_sample_p2p(const double* tree, int64_t bound, const double* u, int64_t M,
                          double* slots, double* const* peer_slots, uint64_t* state, int32_t G,
                          int32_t rank, double* prio, double beta, int32_t weight_norm,
                          int64_t* idx, double* weight, int32_t* err_flag, ts_stream_t stream);
/* The same pair with a hand-over for batches larger than one update pass (4096 entries): with G
 * shards a rank owns ~M / G of the M draws, but ts_per_update_publish has to scan all M entries
 * of `idx` (in every CTA) to find them.  `owned` is int32[4 + M] on the device, zero-initialised
 * once: the sampling kernel appends the batch positions it owns and leaves their count in
 * owned[2]; the update of the SAME, unmodified `idx` then visits only those.  owned == NULL gives
 * the plain entry points.  (No reference counterpart: prio.py:78-89 is one NumPy assignment.) */
int32_t ts_per_sample_p2p_owned(const double* tree, int64_t bound, const double* u, int64_t M,
                                double* slots, double* const* peer_slots, uint64_t* state,
                                int32_t G, int32_t rank, double* prio, double beta,
                                int32_t weight_norm, int64_t* idx, double* weight, int32_t* owned,
                                int32_t* err_flag, ts_stream_t stream);
int32_t ts_per_update_publish_owned(double* tree, int64_t size, int64_t bound, const int64_t* idx,
                                    const void* td, int32_t td_dtype, int64_t M, double alpha,
                                    double* prio, uint32_t* stamp, uint32_t* counter,
                                    int32_t* err_flag, uint64_t* state,
                                    double* const* peer_slots, int32_t G, int32_t rank,
                                    const int32_t* owned, ts_stream_t stream);
/* Measurement only (scripts/nvlink_pingpong.py): `iters` flag round trips between this GPU and rank
 * `peer` through 4-double slots in peer-writable memory, with the primitives of the exchange above
 * (mode 0: release flag / acquire poll; 1: a full shard record + system fence; 2: tight poll).
 * out_ns[0] = nanoseconds for all round trips (0: timed out).  No reference counterpart. */
int32_t ts_debug_pingpong(double* mine, double* const* peers, int32_t peer, int32_t iters,
                          int32_t initiator, int32_t mode, uint64_t* out_ns, ts_stream_t stream);
/* weight[j] /= *wmax  (prio.py:101) */
int32_t ts_per_normalize(double* weight, int64_t M, const double* wmax, ts_stream_t stream);
/* get_weight on given indices (prio.py:66-76), optional batch-max normalisation (:101);
 * idx[j] == -1 gives weight 0. */
int32_t ts_per_weight(const double* tree, int64_t bound, const int64_t* idx, int64_t M,
                      const double* prio, double beta, int32_t weight_norm, double* weight,
                      void* scratch16, ts_stream_t stream);
/* p = |td| + eps computed in td's dtype (f32 eps, prio.py:38,86); leaf = p^alpha rounded to
 * td's dtype (:87); tree update as ts_segtree_update; prio <- {max(max,p), min(min,p)}.
 * Entries with idx[j] == -1 are skipped silently (queries of another shard). */
int32_t ts_per_update(double* tree, int64_t size, int64_t bound, const int64_t* idx,
                      const void* td, int32_t td_dtype, int64_t M, double alpha, double* prio,
                      uint32_t* stamp, uint32_t* counter, int32_t* err_flag,
                      ts_stream_t stream);
/*
 * The replay step with a LAZILY finished tree (same reference lines as ts_per_update /
 * ts_per_sample: prio.py:78-89, :59-64, :66-76, :101 over segtree.py:98-104, :122-137).
 * ts_per_update_lazy writes the new leaves and every ancestor up to the roots of its per-CTA
 * leaf ranges and stops: the levels above them -- heap indices below
 * ts_segtree_lazy_top_nodes(bound), at most 128 nodes -- stay stale.  ts_per_sample_lazy stages
 * the top of the tree in shared memory anyway; it rebuilds those levels there from the range
 * roots in the tree's pairwise order (node = left + right, the bits the reference holds), samples
 * from them and writes them back, so after it the tree in memory is complete again.  This takes
 * the update's last-CTA ticket (fence + atomic), its serial top rebuild and their latency off the
 * step.  Any OTHER reader of the inner nodes between the two calls (ts_segtree_reduce,
 * ts_segtree_prefix_idx, ts_per_sample, a host copy) must call ts_segtree_finish_top first;
 * leaves are always current.  Trees outside the ranged-update regime (<= 4096 or > 2^20 leaves,
 * batches > 262144) are updated completely by ts_per_update_lazy; ts_segtree_lazy_top_nodes
 * returns 0 for them and the other two calls then behave like their eager counterparts.
 */
int32_t ts_segtree_lazy_top_nodes(int64_t bound);
int32_t ts_per_update_lazy(double* tree, int64_t size, int64_t bound, const int64_t* idx,
                           const void* td, int32_t td_dtype, int64_t M, double alpha,
                           double* prio, uint32_t* stamp, uint32_t* counter, int32_t* err_flag,
                           ts_stream_t stream);
int32_t ts_per_sample_lazy(double* tree, int64_t bound, const double* u, int64_t M,
                           const double* prio, double beta, int32_t weight_norm, int64_t* idx,
                           double* weight, int32_t* err_flag, ts_stream_t stream);
int32_t ts_segtree_finish_top(double* tree, int64_t bound, ts_stream_t stream);

/* init_weight (prio.py:42-43): tree[idx] = max_prio^alpha, max_prio read from the device
 * scalar; leaf_dtype TS_F32 rounds the power to f32 (the reference's result type once
 * _max_prio has become an np.float32, prio.py:88). */
int32_t ts_per_init_weight(double* tree, int64_t size, int64_t bound, const int64_t* idx,
                           int64_t M, double alpha, const double* prio, int32_t leaf_dtype,
                           uint32_t* stamp, uint32_t* counter, int32_t* err_flag,
                           ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * PPO clipped-surrogate loss, fused with advantage normalisation.
 * Replaces: the loss arithmetic of PPOPolicy.learn (tianshou/policy/modelfree/ppo.py:109-139)
 *           over one minibatch given by `mb_index` (Batch.split, data/batch.py:896-912):
 *   adv' = (adv - mean) / (std_unbiased + eps)         if norm_adv   (:109-113)
 *   ratio = exp(logp - logp_old); surr1/surr2; dual clip (:114-128); value (clip) loss (:130-139)
 * logp[m], value[m] are minibatch-ordered network outputs; logp_old/adv/returns/v_s are
 * full-batch arrays gathered through mb_index (nullable => identity).
 * Forward writes out[4] = {clip_loss, vf_loss, adv_mean, adv_std} (f32 on device) and the
 * gradients dclip/dlogp[m], dvf/dvalue[m] (f32) used by the autograd backward.
 * workspace: ts_ppo_loss_workspace_bytes(m).
 * ---------------------------------------------------------------------------------------- */
size_t ts_ppo_loss_workspace_bytes(int64_t m);
int32_t ts_ppo_loss(const float* logp, const float* value, const float* logp_old,
                    const float* adv, const float* returns, const float* v_s,
                    const int64_t* mb_index, int64_t m, float eps_clip, float dual_clip,
                    int32_t value_clip, int32_t norm_adv, float eps, float* out,
                    float* dlogp, float* dvalue, void* workspace, size_t workspace_bytes,
                    ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Row gather (buffer[indices], tianshou/data/buffer/base.py:360-389): out[j,:] =
 * src[index[j],:] for rows of row_bytes bytes.  Pure byte copy.
 * ---------------------------------------------------------------------------------------- */
int32_t ts_gather_rows(const void* src, int64_t n_rows, int64_t row_bytes, const int64_t* index,
                       int64_t M, void* out, ts_stream_t stream);

/* Multi-field gather: the small per-transition fields of buffer[indices] -- act, rew,
 * terminated, truncated, done, flat info / policy entries (tianshou/data/buffer/base.py:379-389,
 * one NumPy fancy-index copy per key there) -- in one launch: dst[f][j,:] = src[f][index[j],:]
 * for n_fields <= 8 fields of row_bytes[f] bytes each.  src / dst / row_bytes are HOST arrays of
 * device pointers / sizes; every storage has n_rows rows; negative indices wrap like NumPy's. */
int32_t ts_gather_fields(int32_t n_fields, const void* const* src, void* const* dst,
                         const int64_t* row_bytes, int64_t n_rows, const int64_t* index,
                         int64_t M, ts_stream_t stream);

/* Multi-field scatter, the inverse of ts_gather_fields: dst[f][index[j],:] = src[f][j,:].
 * Replaces `self._meta[ptrs] = batch` of ReplayBufferManager.add
 * (tianshou/data/buffer/manager.py:160; one fancy-index assignment per key there): the rows of
 * one vectorised env step, uploaded as one packed buffer, go into the storage of every key in
 * one launch.  Indices must be distinct; an index outside [-n_rows, n_rows) is skipped. */
int32_t ts_scatter_fields(int32_t n_fields, const void* const* src, void* const* dst,
                          const int64_t* row_bytes, int64_t n_rows, const int64_t* index,
                          int64_t M, ts_stream_t stream);

/* Frame-stack gather (ReplayBuffer.get with stack_num > 1, base.py:317-358; Atari-style
 * save_only_last_obs buffers, atari_dqn.py:135-140): out[j, i, :] = src[g_i, :] with
 * g_{k-1} = index[j] % total and g_{i-1} = prev(g_i) (episode-aware, manager.py:259-277; one
 * ring is the B == 1 case), oldest frame first, k = stack_num <= 16.  When out_next is not NULL
 * the stack ending at next(index[j]) (base.py:378, manager.py:280-297) is written there by the
 * same launch: k + 1 frames read for 2k written.  Pure byte copy of rows of row_bytes bytes. */
int32_t ts_gather_stack(const void* src, int64_t row_bytes, const int64_t* index, int64_t M,
                        int32_t stack_num, const int64_t* offset, int64_t B, int64_t sub_size,
                        const uint8_t* done, const int64_t* last_index, const int64_t* lengths,
                        void* out, void* out_next, ts_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Device-side add of one vectorised env step (rollouts produced on the GPU).
 * Replaces: the loop over sub-buffers in ReplayBufferManager.add
 *           (tianshou/data/buffer/manager.py:104-161), ReplayBuffer._add_index per transition
 *           (tianshou/data/buffer/base.py:195-214) and `self._meta[ptrs] = batch` (manager.py:160).
 * Transition j goes to sub-buffer buffer_ids[j] (NULL: j), each sub-buffer at most once per call.
 * src/dst/row_bytes are HOST arrays of n_keys <= 16 entries: rows of the step [M, row_bytes]
 * and the storage they are stored into [total, row_bytes] (device pointers).  ring = int64
 * [B][2] (index, size); ep_rew/ep_len/ep_idx = running episode statistics; all device-resident
 * and advanced in place.  Outputs (device, [M]): the global slot, and the episode return /
 * length (0 unless done[j]) / start index, as _add_index returns them.  A sub-buffer id outside
 * [0, B) sets TS_FLAG_INDEX_RANGE and skips the transition.
 * ---------------------------------------------------------------------------------------- */
int32_t ts_add_step(int32_t n_keys, const void* const* src, void* const* dst,
                    const int64_t* row_bytes, int64_t M, const int64_t* buffer_ids,
                    const void* rew, int32_t rew_dtype, const uint8_t* done,
                    const int64_t* offset, int64_t B, int64_t* ring, int64_t* last_index,
                    int64_t* lengths, double* ep_rew, int64_t* ep_len, int64_t* ep_idx,
                    int64_t* ptr_out, double* ep_rew_out, int64_t* ep_len_out,
                    int64_t* ep_idx_out, int32_t* err_flag, ts_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TS_B200_H_ */
