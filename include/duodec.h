/*
 * duodec.h - C ABI of the B200-native speculative-decoding verification path.
 *
 * Drop-in boundary for the hot path of Nanji-Huaji/DuoDecoding.  The reference
 * has no FFI of its own: the path is reached through Python module functions
 * (src/utils.py, src/decoding_ops.py, src/communication.py, src/model_gpu.py).
 * Each entry point below names the reference function(s) it replaces; the
 * Python host layer in duodecoding_b200/ keeps those functions' names and
 * signatures and calls this library through ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless
 *     the name ends in _host; nothing here allocates or synchronises;
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued on it;
 *   - strides are in ELEMENTS; the vocabulary dimension is contiguous;
 *   - dtype codes: DUODEC_F32 = 0, DUODEC_BF16 = 1;
 *   - return value: 0 on success, a negative DUODEC_E* code for a rejected
 *     argument, or a positive cudaError_t from the launch;
 *   - re-entrant: a call depends on its arguments only.  Process-global state is limited to one-time
 *     kernel attribute setup and the two tuning / test hooks duodec_set_team_size and
 *     duodec_set_stats_buffer below (not meant for concurrent use).
 *   - a verify workspace holds the records and per-sequence cells of ONE launch in flight: do not
 *     share it between launches that may overlap (different streams or threads); launches on one
 *     stream may reuse it.
 */
#ifndef DUODEC_H_
#define DUODEC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DUODEC_F32 0
#define DUODEC_BF16 1

#define DUODEC_EINVAL (-1)   /* bad shape / null pointer / unsupported dtype */
#define DUODEC_ETOOBIG (-2)  /* row does not fit the largest team (V too large) */
#define DUODEC_EWORKSPACE (-3)
#define DUODEC_ESTATICMAX (-4) /* duodec_verify_fused_static: the call needs more scratch than the built-in array holds */

/* Library / build identification: "duodec-b200 <version> sm_100a". */
const char* duodec_version(void);

/* Team (thread-block cluster) size the library would pick for a row of V
 * entries when `rows_resident` rows must sit in shared memory together
 * (1: norm_logits / top-k ops, 2: fused verify).  0 if V is too large. */
int duodec_team_size(int V, int rows_resident);

/* Override the team-size heuristic (0 = automatic).  Tuning / test hook. */
void duodec_set_team_size(int team);

/* Tuning hook: device int[32] that the scan kernels increment (16: rows finished by the fast select,
 * 17: rows handed to the general search, 19: general searches run).  NULL (default) disables it. */
void duodec_set_stats_buffer(int* device_counters);

/* ---- logits -> probabilities --------------------------------------------
 * Replaces utils.norm_logits + utils.top_k_top_p_filter
 * (/root/reference/src/utils.py:650-709).
 *   temperature == 0 : one-hot of the first argmax
 *   else             : softmax(top_k_top_p_filter(logits / temperature))
 * logits [rows, V] (in_dtype), out [rows, V] (out_dtype). */
int duodec_norm_logits(const void* logits, int in_dtype, int64_t in_row_stride, void* out, int out_dtype,
                       int64_t out_row_stride, int rows, int V, float temperature, int top_k, float top_p,
                       void* stream);
/* The same over [batch, inner, V] views with two row strides each: what KVCacheModel's history
 * buffers are ([B, max_len, V] rows [start, start+S), src/model_gpu.py:186-203) - one launch. */
int duodec_norm_logits_batched(const void* logits, int in_dtype, int64_t in_batch_stride, int64_t in_row_stride,
                               void* out, int out_dtype, int64_t out_batch_stride, int64_t out_row_stride, int batch,
                               int inner, int V, float temperature, int top_k, float top_p, void* stream);

/* ---- fused verification on logits -----------------------------------------
 * One call = norm_logits on the gamma draft rows and gamma+1 target rows
 * (src/model_gpu.py:253), the accept test and leading-run scan
 * (src/decoding_ops.py:127-147, applied to EVERY sequence of the batch), the
 * residual resample at the first rejected position (src/decoding_ops.py:190-206)
 * or the bonus sample (src/decoding_ops.py:209-220).
 *
 *   draft   [B, >=gamma,   Vd] logits, strides (d_sb, d_ss, 1)
 *   target  [B, >=gamma+1, Vt] logits, strides (t_sb, t_ss, 1)
 *   tokens  [B, gamma] int64 drafted ids (row stride tok_sb)
 *   r       [B, gamma] float32 uniforms of the accept test (row stride r_sb)
 *   noise   [B, >=max(Vd,Vt)] float32 Exp(1) draws (row stride noise_sb): the
 *           sampled token of sequence b is argmax_i prob_i / noise[b,i], which
 *           is what torch.multinomial computes internally.
 * outputs
 *   accept_mask [B,gamma] uint8, sel_target_p / sel_draft_p [B,gamma] float32,
 *   accepted_len [B] int64, token [B] int64, all_accepted [B] uint8.
 * flags: DUODEC_VERIFY_ROW0_RULE - the reference's batch rule (src/decoding_ops.py:144-147): the
 *   accepted count of sequence 0 positions EVERY sequence (accepted_len[b] = accepted_len[0], every
 *   token drawn at that position); without it every sequence is verified on its own count.  Only
 *   the fast paths implement it (equal vocabularies, no top-k, 16-byte aligned rows; T > 0 or greedy T == 0):
 *   DUODEC_EINVAL otherwise when B > 1.
 *   DUODEC_VERIFY_PATH(p) - test hook: 0 automatic (row-pair scan kernel when the call qualifies,
 *   else the general team kernel), 1 general team kernel, 5 row-pair scan kernel.
 *   DUODEC_VERIFY_PHASES(m) - measurement hook: 1 scan kernel only, 2 sample kernel only, 0 / 3 both.
 * workspace: duodec_verify_workspace_bytes(B,gamma) bytes, 16-byte aligned, zero-filled ONCE by
 *   the caller before first use (the kernels restore what they need to zero).  One launch in flight
 *   per workspace. */
#define DUODEC_VERIFY_ROW0_RULE 1
#define DUODEC_VERIFY_PATH(p) (((p) & 15) << 4)
#define DUODEC_VERIFY_PHASES(m) (((m) & 3) << 8)
size_t duodec_verify_workspace_bytes(int B, int gamma);

int duodec_verify_fused(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                        int64_t t_ss, int B, int gamma, int Vd, int Vt, const int64_t* tokens, int64_t tok_sb,
                        const float* r, int64_t r_sb, const float* noise, int64_t noise_sb, float temperature,
                        int top_k, float top_p, float draft_temperature, int draft_top_k, float draft_top_p,
                        int flags, uint8_t* accept_mask, float* sel_target_p, float* sel_draft_p,
                        int64_t* accepted_len, int64_t* token, uint8_t* all_accepted, void* workspace,
                        size_t workspace_bytes, void* stream);

/* No counterpart in the reference: added for callers that
 * cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  duodec_verify_fused with the same arguments and results, minus the workspace: the
 * scratch is a file-scope `static __device__ __align__(256)` array of DUODEC_VERIFY_STATIC_BYTES
 * (12 MiB) inside the library, which every GPU that loads the library holds until exit.  The entry
 * point allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); it resolves the
 * array's device address once per device (cudaGetSymbolAddress) and launches the same kernels.
 *   - limit: duodec_verify_workspace_bytes(B, gamma) <= DUODEC_VERIFY_STATIC_BYTES (about 30 000 rows
 *     B * (2 * gamma + 1); C2 needs 10.5 MB, C5 10.7 MB), else DUODEC_ESTATICMAX and nothing is launched;
 *     duodec_verify_static_max_bytes() returns the limit.
 *   - EVERY call on a device shares the one array and it keeps the last call's contents: the call
 *     zero-fills the part it uses (cudaMemsetAsync on `stream`, ahead of the kernels) whenever
 *     (B, gamma) differ from the previous call on that device; with equal shapes the kernels restore
 *     what must be zero themselves and the exponent-offset hints of the previous call are reused.
 *   - calls that overlap on the GPU (different streams, or graphs replayed concurrently) RACE on the
 *     array: serialise them, or use duodec_verify_fused with one workspace per stream. */
#define DUODEC_VERIFY_STATIC_BYTES (12u << 20)
size_t duodec_verify_static_max_bytes(void);
int duodec_verify_fused_static(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss,
                               int64_t t_sb, int64_t t_ss, int B, int gamma, int Vd, int Vt, const int64_t* tokens,
                               int64_t tok_sb, const float* r, int64_t r_sb, const float* noise, int64_t noise_sb,
                               float temperature, int top_k, float top_p, float draft_temperature, int draft_top_k,
                               float draft_top_p, int flags, uint8_t* accept_mask, float* sel_target_p,
                               float* sel_draft_p, int64_t* accepted_len, int64_t* token, uint8_t* all_accepted,
                               void* stream);

/* ---- probability-space operators -----------------------------------------
 * Accept test on stored probability histories; replaces the tensor part of
 * decoding_ops.compute_acceptance_result (src/decoding_ops.py:127-147) for
 * every sequence.  draft_probs/target_probs [B, gamma, V] with strides
 * (sb, ss, 1); sel_* have the probability dtype. */
int duodec_accept_probs(const void* draft_probs, int64_t d_sb, int64_t d_ss, const void* target_probs,
                        int64_t t_sb, int64_t t_ss, int dtype, const int64_t* tokens, int64_t tok_sb,
                        const float* r, int64_t r_sb, int B, int gamma, int V, uint8_t* accept_mask,
                        void* sel_draft_p, void* sel_target_p, int64_t* accepted_len, void* stream);

/* utils.sample with injected noise (src/utils.py:727-748): out[row] int64. */
int duodec_sample_rows(const void* probs, int dtype, int64_t row_stride, const float* noise, int64_t noise_stride,
                       int rows, int V, int64_t* out, void* stream);

/* decoding_ops.sample_reject_token = sample(max_fn(p - q))
 * (src/decoding_ops.py:190-206, src/utils.py:787-810). */
int duodec_residual_sample(const void* target_probs, int64_t t_stride, const void* draft_probs, int64_t d_stride,
                           int dtype, const float* noise, int64_t noise_stride, int rows, int V, int64_t* out,
                           void* stream);

/* utils.max_fn (src/utils.py:787-810): out float32 [rows, V]. */
int duodec_max_fn(const void* x, int dtype, int64_t row_stride, int rows, int V, float* out, int64_t out_stride,
                  void* stream);

/* ---- transfer_top_k compression --------------------------------------------
 * mode 0: utils.rebuild_topk_probs / rebuild_topk_uniform_probs (src/utils.py:751-784)
 * mode 1: CommunicationSimulator._apply_top_k_compression (src/communication.py:288-308)
 * mode 2: CommunicationSimulator.compress_rebuild_probs   (src/communication.py:348-395)
 * mode 3: CommunicationSimulator.rebuild_full_probs       (src/communication.py:310-346; k ignored)
 * probs/out [rows, V] of `dtype`; requires 0 < k < V for modes 0-2. */
int duodec_topk_rebuild(const void* probs, int dtype, int64_t in_stride, void* out, int64_t out_stride, int rows,
                        int V, int k, int mode, void* stream);

/* ---- KV rollback --------------------------------------------------------
 * KVCacheModel.rollback (src/model_gpu.py:362-383) for a cache that keeps
 * per-sequence length counters: seq_len[b] = min(seq_len[b], end_pos) in
 * place; no cache bytes move.  end_pos_vec (int64 [B]) may be NULL, then the
 * scalar applies to every sequence. */
int duodec_kv_truncate(int32_t* seq_len, const int64_t* end_pos_vec, int64_t end_pos, int B, void* stream);

/* ---- vocabulary-sharded verification (V = 151936 over `world` GPUs) -------------------------
 * Every rank holds the slice [v_begin, v_begin + Vs) of all draft / target rows (strides of the
 * LOCAL slices).  The step is stats -> all-gather -> accept -> sample -> all-gather -> finish; the
 * two all-gathers are the caller's (NCCL over NVLink in duodecoding_b200/sharded.py).  Replaces
 * the same reference code as duodec_verify_fused, without top-k; top-p through the duodec_shard_nucleus_* calls below.
 *   stats    [B*(2*gamma+1), 4] float32 per rank: slice max, sum exp(x - max), token logit, flag
 *   plan     [B] 48-byte records (device scratch between accept and sample)
 *   partial  [B, nslice] 24-byte records per rank (the slice is cut in nslice parts for occupancy):
 *            residual mass, best race key, zero-mass fallback key; gathered as [world, B, nslice] */
#define DUODEC_SHARD_PLAN_BYTES 48
#define DUODEC_SHARD_PARTIAL_BYTES 24
/* Optional fused exchange (NULL: the caller gathers, e.g. with NCCL).  `peers` is a DEVICE array of `world`
 * base pointers of peer-mapped (symmetric) buffers, own buffer at [rank]; every buffer starts with a 1 KB
 * header (zero-filled once: flags, block counters, step number) followed by the gather areas at the given
 * byte offsets: gstats [world][B*(2*gamma+1)] x 16 B and gpartial [world][B*nslice] x 24 B.  With it
 * duodec_shard_stats / _sample store their results into every peer's areas and publish the step;
 * duodec_shard_accept / _finish wait for all ranks and read their own areas (the stats / gathered / partial
 * pointers may then be NULL).  All ranks must issue the four calls of a step in the same order. */
typedef struct {
  void* const* peers;
  int rank, world;
  int64_t off_stats, off_partial;
  int mode; /* 0: every block stores its result into every peer; 1: the last block copies the rank's slab;
             * 2 ("pull"): results stay in the rank's own buffer, the CALLER separates the producing and the consuming
             * call by a barrier over all ranks (e.g. torch symmetric memory's hdl.barrier() on the same stream) and the
             * consumers read every peer's slab in place - no flags, the header is unused */
} duodec_shard_xchg;
#define DUODEC_SHARD_XCHG_HEADER 1024
int duodec_shard_stats(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                       int64_t t_ss, int B, int gamma, int Vs, int v_begin, const int64_t* tokens, int64_t tok_sb,
                       float temperature, float draft_temperature, float* stats, const duodec_shard_xchg* xchg,
                       void* stream);
int duodec_shard_accept(const float* gathered_stats, int world, int B, int gamma, const float* r, int64_t r_sb,
                        uint8_t* accept_mask, float* sel_target_p, float* sel_draft_p, int64_t* accepted_len,
                        uint8_t* all_accepted, void* plan, const duodec_shard_xchg* xchg, const int64_t* tokens,
                        int64_t tok_sb, const void* cuts, void* stream);
int duodec_shard_sample(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                        int64_t t_ss, int B, int gamma, int Vs, int v_begin, float temperature,
                        float draft_temperature, const void* plan, const float* noise, int64_t noise_sb,
                        int nslice, void* partial, const duodec_shard_xchg* xchg, int filtered, void* stream);
int duodec_shard_finish(const void* gathered_partials, int world, int B, int nslice, int64_t* token,
                        const duodec_shard_xchg* xchg, void* stream);

/* Top-p under vocabulary sharding (reference: top_k_top_p_filter, src/utils.py:650-683, as applied by
 * norm_logits, :686-709): a distributed radix select of the nucleus boundary over the order-preserving key of the
 * scaled logit, 4 bits per level, between exchange 1 and duodec_shard_accept.  Masses are 40-bit fixed point
 * relative to the row maximum and summed as integers: the answer does not depend on the world size.
 *   levels  = duodec_shard_nucleus_levels(dtype, T, draft T): 4 (bf16 rows at temperature 1) or 8
 *   for level in 0 .. levels-1:
 *     duodec_shard_nucleus_level   -> hist [rows, DUODEC_SHARD_HIST_BYTES] of this rank
 *     (all-gather hist -> [world, rows, DUODEC_SHARD_HIST_BYTES])
 *     duodec_shard_nucleus_reduce  -> sel [rows, DUODEC_SHARD_SEL_BYTES] updated; at the last level quota [rows]
 *   duodec_shard_nucleus_tie       -> cand [rows] int32: global index of this rank's last kept tie
 *   (all-gather cand -> [world, rows])
 *   duodec_shard_nucleus_close     -> cuts [rows, DUODEC_SHARD_CUT_BYTES]: (x*, i*, kept mass, state)
 * then duodec_shard_accept(..., tokens, tok_sb, cuts, ...) and duodec_shard_sample(..., filtered = 1, ...).
 * top_p / draft_top_p <= 0 leave that model's rows unfiltered.  Not available with the fused exchange. */
#define DUODEC_SHARD_HIST_BYTES 192
#define DUODEC_SHARD_SEL_BYTES 32
#define DUODEC_SHARD_CUT_BYTES 16
int duodec_shard_nucleus_levels(int dtype, float temperature, float draft_temperature);
int duodec_shard_nucleus_level(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss,
                               int64_t t_sb, int64_t t_ss, int B, int gamma, int Vs, int v_begin, float temperature,
                               float draft_temperature, float top_p, float draft_top_p, const float* gathered_stats,
                               int world, void* sel, int level, void* hist, void* stream);
int duodec_shard_nucleus_reduce(const void* gathered_hist, int world, int rank, int B, int gamma, float top_p,
                                float draft_top_p, void* sel, int level, int levels, int32_t* quota, void* stream);
int duodec_shard_nucleus_tie(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                             int64_t t_ss, int B, int gamma, int Vs, int v_begin, float temperature,
                             float draft_temperature, const void* sel, const int32_t* quota, int levels,
                             int32_t* cand, void* stream);
int duodec_shard_nucleus_close(const void* sel, const float* gathered_stats, const int32_t* gathered_cand, int world,
                               int B, int gamma, int levels, void* cuts, void* stream);

/* ---- transfer_top_k under vocabulary sharding (merge of per-shard top-k candidates) ------------------
 * Replaces utils.rebuild_topk_probs (src/utils.py:751-777) / communication._apply_top_k_compression
 * (src/communication.py:288-308) when every rank holds a vocabulary slice of the probability rows:
 *   1. duodec_topk_rebuild(mode 1) on the local slice keeps the slice's own top-min(k, Vs) entries;
 *   2. duodec_compact_nonzero lists them as (value, GLOBAL index) in index order, kcap slots per row, free
 *      slots (-1, INT_MAX), count[row] = number of non-zero entries found;
 *   3. the caller gathers the lists of all ranks (rank order = global index order), runs
 *      duodec_topk_rebuild(mode 1) over the gathered values to pick the global top-k (lowest index on ties)
 *      and derives the boundary (tau, istar), the kept mass and the number of holes;
 *   4. duodec_shard_topk_apply writes the rebuilt slice: kept entries (v > tau, or v == tau and global index
 *      <= istar, v != 0) keep their value, every other entry takes fill[row]; all divided by total[row]. */
int duodec_compact_nonzero(const void* x, int dtype, int64_t stride, int rows, int Vs, int v_begin, int kcap,
                           float* cand_val, int32_t* cand_idx, int32_t* count, void* stream);
int duodec_shard_topk_apply(const void* x, int dtype, int64_t in_stride, void* out, int64_t out_stride, int rows,
                            int Vs, int v_begin, const float* tau, const int32_t* istar, const float* fill,
                            const float* total, void* stream);

/* Rank bookkeeping of accepted tokens: _compute_token_vocab_rank / _compute_transfer_topk_rank /
 * _record_accepted_token_ranks, src/baselines.py:121-168 (per token: a count with .item(), a sorted torch.topk
 * and a .nonzero()).  Row i of `probs` (rows x V, row_stride elements apart) and tokens[i * tok_stride]:
 *   out[2*i]   = number of entries greater than the token's probability (vocab rank - 1); -1 for an invalid id
 *   out[2*i+1] = number of entries EQUAL to it at a lower index (sorted top-k rank = out[2*i] + out[2*i+1] + 1,
 *                equal values in index order - torch.topk leaves that order unspecified) */
int duodec_token_ranks(const void* probs, int dtype, int64_t row_stride, int rows, int V, const int64_t* tokens,
                       int64_t tok_stride, int32_t* out, void* stream);

/* ---- CUHLM helpers (CommunicationSimulator) --------------------------------------------
 * duodec_uncertainty_tokens: calculate_uncertainty (src/communication.py:541-575) - for each of M
 *   temperatures temps[m] > 0 one draw from softmax(logits / temps[m]) of ONE row logits[V], as
 *   argmax_i p_i / noise[m, i] (noise: Exp(1), what torch.multinomial draws); tokens[M] int64.  The
 *   uncertainty is the fraction of tokens that differ from the draft token (one reduction on M values).
 * duodec_compressed_vocab_size: _calculate_compressed_vocab_size (src/communication.py:681-755) - with the
 *   probabilities sorted descending (float32) and their inclusive float64 prefix sums, the smallest
 *   k in [1, V) with sum_{i>=k} |s_i - (1 - sum_{i<k} s_i)/(V-k)| / denominator <= theta; *k_out is
 *   0x7f7f7f7f when no k qualifies (the caller applies the reference's conservative default). */
int duodec_uncertainty_tokens(const void* logits, int dtype, int V, const float* temps, const float* noise,
                              int64_t noise_stride, int M, int64_t* tokens, void* stream);
int duodec_compressed_vocab_size(const float* sorted_probs, const double* prefix, int V, double denominator,
                                 double theta, int32_t* k_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DUODEC_H_ */
