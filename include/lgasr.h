/*
 * lgasr.h - C ABI of the B200-native (sm_100a) fused CTC / attention / LM beam-search scorer.
 *
 * The reference (an ESPnet fork, pure Python on this path) has no FFI here: the path sits behind
 * the duck-typed Python scorer API (espnet/nets/scorer_interface.py:9-186).  This header is the
 * C-ABI a binding of that API sits on; each entry point names the reference code it replaces.
 * The Python binding (ctypes) is llm_guided_asr_b200/_lib.py; INTEGRATION.md shows the stub a
 * reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, a negative LGASR_E* code otherwise; no exceptions
 *     cross the ABI; lgasr_last_error() returns a message for the calling thread's last failure.
 *   - pointers named d_* are CUDA device pointers, h_* are host pointers. The caller owns all
 *     I/O buffers; the handle owns only its own state / workspace.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - a handle is single-stream and not re-entrant (like the reference scorer object, which keeps
 *     a mutable per-utterance `self.impl`, espnet/nets/scorers/ctc.py:24,98); different handles
 *     are independent.
 *   - fp32 throughout (espnet/nets/ctc_prefix_score.py:38 follows x.dtype; fp32 is the default
 *     of espnet2/bin/asr_inference.py), log-zero is -1e10 (ctc_prefix_score.py:32), blank id 0
 *     (espnet/nets/scorers/ctc.py:98).
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef LGASR_H_
#define LGASR_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LGASR_VERSION 100

#define LGASR_OK 0
#define LGASR_EINVAL -1   /* bad argument / unsupported configuration */
#define LGASR_ECUDA -2    /* a CUDA runtime call or kernel failed      */
#define LGASR_ESTATE -3   /* call order violated (e.g. step before begin) */
#define LGASR_ENOMEM -4

#define LGASR_MAX_TERMS 8      /* full-scorer terms fused per step (decoder, length_bonus, lm, ...) */
#define LGASR_MAX_PREBEAM 128  /* pre-beam size S = int(1.5*beam) <= 128 */

typedef struct lgasr_handle lgasr_handle;

/* Search configuration = the knobs of BeamSearch.__init__ (espnet/nets/beam_search.py:37-121)
 * plus the batch extents. */
typedef struct {
  int32_t n_utts;        /* B: utterances decoded in lock-step by this handle                     */
  int32_t max_frames;    /* T: encoder frames of the longest utterance                            */
  int32_t vocab;         /* V                                                                     */
  int32_t beam;          /* beam_size                                                             */
  int32_t pre_beam;      /* S = int(pre_beam_ratio*beam)  (beam_search.py:105)                    */
  int32_t n_terms;       /* K: full-scorer terms, in the dict order of self.full_scorers          */
  int32_t pre_beam_key;  /* -1: "full" (weighted sum); k>=0: rank by term k; -2: None (no pre-beam)*/
  int32_t use_ctc;       /* 1 if a CTC partial scorer with non-zero weight is present             */
  float ctc_weight;      /* weights["ctc"]                                                        */
  int32_t blank;         /* 0 in espnet2                                                          */
  int32_t sos;
  int32_t eos;
  int32_t max_len_cap;   /* upper bound of maxlen over the batch (decode positions)               */
  int32_t primer_len;    /* len(hyp_primer) or 1 ([sos])  (batch_beam_search.py:146-147)          */
  int32_t end_detect;    /* 1 if maxlenratio == 0.0 (beam_search.py:443)                          */
} lgasr_config;

/* Library / device ------------------------------------------------------------------------- */
int lgasr_version(void);
const char* lgasr_last_error(void);
/* number of kernels launched by this library since load (bench.py "gpu_launches") */
int64_t lgasr_launch_count(void);

/* Handle life cycle ------------------------------------------------------------------------ */
int lgasr_create(const lgasr_config* cfg, lgasr_handle** out);
int lgasr_destroy(lgasr_handle* h);

/* Replaces CTCPrefixScoreTH.__init__ (ctc_prefix_score.py:21-65) for B utterances:
 * d_logp [B][T][V] fp32 log-posteriors (output of CTC.log_softmax, espnet2/asr/ctc.py:196-204),
 * h_xlens[B] valid frames.  Frames >= xlens are padded as :46-49 (blank 0, others log-zero).
 * The posteriors are re-laid out as xT[B][V][Tp] (frames contiguous) inside the handle. */
int lgasr_set_posteriors(lgasr_handle* h, const float* d_logp, const int32_t* h_xlens, void* stream);
/* same, from a host buffer (pinned or pageable); the H2D copy is issued on `stream` */
int lgasr_set_posteriors_host(lgasr_handle* h, const float* h_logp, const int32_t* h_xlens, void* stream);

/* Replaces BatchBeamSearch.init_hyp + CTCPrefixScorer.batch_init_state
 * (batch_beam_search.py:130-159, scorers/ctc.py:87-99) and the length bounds of
 * BeamSearch.forward (beam_search.py:417-427): h_maxlen[B], h_minlen[B] decode positions,
 * h_primer[primer_len] initial tokens (int64). */
int lgasr_begin(lgasr_handle* h, const int32_t* h_maxlen, const int32_t* h_minlen,
                const int64_t* h_primer, void* stream);

/* Continuous batching (SURVEY.md 7.3-6, 8e "continuous refill"): utterance slot `utt` of a running handle, whose previous
 * utterance has finished (done flag of lgasr_poll; read its n-best with lgasr_gather_ended first), takes a NEW utterance:
 * d_logp [T][V] fp32 log-posteriors of that utterance alone (T = max_frames of the handle; frames >= xlen are padded as
 * ctc_prefix_score.py:46-49), its length bounds, the handle's primer.  The other slots are untouched: every slot is at
 * its own decode position (the kernels derive it from the slot's prefix length).  Replaces starting a new
 * CTCPrefixScoreTH / init_hyp for the next utterance (scorers/ctc.py:87-99, batch_beam_search.py:130-159) while the rest
 * of the batch keeps decoding.  The reference decodes strictly one utterance at a time (asr_inference.py:759-764). */
int lgasr_refill(lgasr_handle* h, int utt, const float* d_logp, int xlen, int maxlen, int minlen, void* stream);

/* One decode position for every unfinished utterance = BatchBeamSearch.search + post_process +
 * end_detect (batch_beam_search.py:259-429, e2e_asr_common.py:17-48) with the CTC prefix
 * recursion of CTCPrefixScoreTH.__call__ / select_state (ctc_prefix_score.py:67-187,
 * scorers/ctc.py:40-63) fused in.
 *   h_term_ptr[K]   device pointers to the (B*beam, V) fp32 log-prob matrix of each full scorer,
 *                   row = utt*beam + running-hypothesis slot, row stride ld elements;
 *                   NULL means a constant term (LengthBonus, scorers/length_bonus.py:38-60)
 *   h_term_const[K] the constant of a NULL term
 *   h_term_weight[K] weights[k]
 * The fp32 accumulation order is the reference's: ((0 + w0*s0) + w1*s1 ...) + w_ctc*ctc + score. */
int lgasr_step(lgasr_handle* h, const float* const* h_term_ptr, const float* h_term_const,
               const float* h_term_weight, int64_t ld, void* stream);

/* Device-resident views of the running hypotheses after the last step (valid until the next
 * step): what the full scorers need to score the next position and to re-order their states.
 *   which: 0 yseq int64 [B][beam][Lcap] (Lcap returned in *ld), 1 n_running int32 [B],
 *          2 parent int32 [B][beam] (slot of the parent among the previous running hyps),
 *          3 token int32 [B][beam], 4 done int32 [B], 5 score fp32 [B][beam], 6 cur_len int32 [B] */
int lgasr_view(lgasr_handle* h, int which, void** d_ptr, int64_t* ld);

/* Small synchronising read of n_running[B], done[B] and (optionally, may be NULL) parent[B*beam],
 * token[B*beam]: one pinned D2H copy + one stream sync per step is all the host loop needs (the
 * reference syncs O(beam) times per step). Returns the number of unfinished utterances in *n_active. */
int lgasr_poll(lgasr_handle* h, int32_t* h_n_running, int32_t* h_done, int32_t* h_parent, int32_t* h_token,
               int32_t* n_active, void* stream);

/* N-best read-out = the sort of BeamSearch.forward (beam_search.py:452-459) input: all ended
 * hypotheses of utterance `utt`, in the order they ended.  Call with h_tokens == NULL to get
 * the counts: *n_ended and *max_len.  Otherwise fills h_tokens[n_ended][max_len] (int64, padded
 * with -1), h_lens[n_ended], h_score[n_ended], h_scores[n_ended][K+use_ctc] (cumulative
 * un-weighted per-scorer log-probs, beam_search.py:263-293; CTC last). Synchronises `stream`. */
int lgasr_gather_ended(lgasr_handle* h, int utt, int32_t* n_ended, int32_t* max_len, int64_t* h_tokens,
                       int32_t* h_lens, float* h_score, float* h_scores, void* stream);

/* The best ended hypothesis of EVERY utterance in one call = head of the sorted n-best of BeamSearch.forward
 * (beam_search.py:452-459; score, or score / (len - 1) with normalize_length; the first of equal scores, as
 * Python's stable sort keeps it).  h_tokens [B][max_len] int64 padded with -1 (max_len >= primer_len +
 * max_len_cap + 1), h_lens[B] (0: no hypothesis ended), h_score[B] (un-normalised).  Synchronises `stream`.
 * This is the record an utterance-sharded driver gathers at the end. */
int lgasr_gather_best(lgasr_handle* h, int normalize_length, int max_len, int64_t* h_tokens, int32_t* h_lens,
                      float* h_score, void* stream);

/* Per-kernel device timing (bench.py "roofline"): when on, every step records CUDA events around
 * its two kernels on the launching stream. lgasr_get_profile synchronises the events and returns
 * the accumulated durations (ms) since profiling was switched on, and the number of steps. */
int lgasr_set_profiling(lgasr_handle* h, int on);
int lgasr_get_profile(lgasr_handle* h, float* ms_score_rows, float* ms_select, int32_t* n_steps);

/* Launch geometry of the fused step kernel chosen for this handle (bench.py reports it): cluster = 1 when the
 * `beam` CTAs of an utterance are launched as one thread-block cluster (records exchanged through distributed
 * shared memory), 0 for the ticket variant (records through L2); dynamic shared memory per CTA; floats per
 * staged row chunk; stages. */
int lgasr_launch_info(lgasr_handle* h, int32_t* cluster, int32_t* smem_bytes, int32_t* chunk_floats, int32_t* stages);

/* Debug time line of the fused step kernel: when the environment variable LGASR_TRACE is set at
 * lgasr_create, every CTA of k_step stamps clock64() at its phase boundaries; this copies the
 * [B][beam][2 groups][16] stamps of the most recent step to h_out (n = capacity in elements). */
int lgasr_debug_trace(lgasr_handle* h, int64_t* h_out, int64_t n);

/* Whole-batch replay with pre-computed full-scorer matrices, device resident:
 * d_terms[step][k] -> (B*beam, V) matrices (flattened h_term_ptr[n_steps*K]); runs lgasr_begin's
 * state forward for up to n_steps positions without host round trips (bench.py "value" leg). */
int lgasr_decode_replay(lgasr_handle* h, int n_steps, const float* const* h_term_ptr,
                        const float* h_term_const, const float* h_term_weight, int64_t ld, void* stream);

/* Same workload through HOST buffers (bench.py "e2e" leg): posteriors and every step's matrices
 * are copied host->device inside the call, results stay in the handle (gather_ended copies
 * device->host).  h_term_ptr[n_steps*K] are HOST pointers (pinned for full speed). */
int lgasr_decode_replay_host(lgasr_handle* h, const float* h_logp, const int32_t* h_xlens,
                             const int32_t* h_maxlen, const int32_t* h_minlen, const int64_t* h_primer,
                             int n_steps, const float* const* h_term_ptr, const float* h_term_const,
                             const float* h_term_weight, int64_t ld, void* stream);

/* Stand-alone scorer ABI = CTCPrefixScorer.batch_score_partial + select_state for ONE utterance
 * (scorers/ctc.py:40-63,101-126 -> ctc_prefix_score.py:67-187), so the drop-in scorer class can be
 * driven by an unmodified BatchBeamSearch.
 *   d_xT [V][Tp] transposed posteriors of the utterance (lgasr_transpose_posteriors), T frames
 *   d_r_prev [n][2][Tp], d_s_prev[n], d_last[n] (int32 last tokens), olen = len(y)-1,
 *   d_ids [n][S] int64 candidate tokens (S == 0: full vocabulary)
 *   out: d_scores [n][V] = log_psi - s_prev (log-zero - s_prev outside the candidates),
 *        d_log_psi [n][V] */
int lgasr_transpose_posteriors(const float* d_logp, int T, int V, int xlen, int blank, float* d_xT, int Tp,
                               void* stream);
int lgasr_ctc_initial_state(const float* d_xT, int T, int Tp, int blank, float* d_r0, void* stream);
int lgasr_ctc_prefix_score(const float* d_xT, int T, int Tp, int V, int blank, int eos, int n,
                           const float* d_r_prev, const float* d_s_prev, const int32_t* d_last, int olen,
                           const int64_t* d_ids, int S, float* d_scores, float* d_log_psi, void* stream);
/* survivor states: for m selections (hyp d_sel_hyp[i], token d_sel_tok[i]) compute
 * d_r_new [m][2][Tp] = r[:, :, hyp, idmap[hyp, tok]] of the reference */
int lgasr_ctc_select_state(const float* d_xT, int T, int Tp, int blank, int m, const float* d_r_prev,
                           const int32_t* d_last, const int32_t* d_sel_hyp, const int32_t* d_sel_tok, int olen,
                           float* d_r_new, void* stream);

/* The same two entry points with the attention window of CTCPrefixScoreTH (margin > 0 with att_w,
 * ctc_prefix_score.py:56-61,143-148; the ESPnet1 decoders pass attention weights, rnn/decoders.py:665,824-826): the
 * recursion and the prefix sum run over frames [win_start, win_end) only; (0, 0) = no window.  The caller derives the
 * window from the attention weights exactly as :143-148 does (CTCPrefixScorer.batch_score_partial(..., att_w=...)). */
int lgasr_ctc_prefix_score_win(const float* d_xT, int T, int Tp, int V, int blank, int eos, int n, const float* d_r_prev,
                               const float* d_s_prev, const int32_t* d_last, int olen, const int64_t* d_ids, int S,
                               float* d_scores, float* d_log_psi, int win_start, int win_end, void* stream);
int lgasr_ctc_select_state_win(const float* d_xT, int T, int Tp, int blank, int m, const float* d_r_prev,
                               const int32_t* d_last, const int32_t* d_sel_hyp, const int32_t* d_sel_tok, int olen,
                               float* d_r_new, int win_start, int win_end, void* stream);


/* Streaming extension = CTCPrefixScoreTH.extend_state (ctc_prefix_score.py:246-269) for n survivor states:
 * d_r_old [n][2][Tp_old] valid for T_old frames -> d_r_new [n][2][Tp_new] valid for T_new >= T_old frames of the
 * (extended) posteriors d_xT [V][Tp_new]: old frames copied, new frames r^n = log-zero and
 * r^b[t] = r^b[t-1] + x[t, blank], added one frame at a time like the reference.  (extend_prob itself is
 * lgasr_transpose_posteriors of the longer utterance plus a copy of the frames already held.) */
int lgasr_ctc_extend_state(const float* d_xT, int T_old, int T_new, int Tp_old, int Tp_new, int blank, int n,
                           const float* d_r_old, float* d_r_new, void* stream);

/* ---- either side of the step ---------------------------------------------- */

/* Re-order a caller-owned per-hypothesis cache of a full scorer (decoder / LM / LLM key-value or layer
 * output cache) by the parents chosen at the LAST lgasr_step, reading them where the step left them on
 * the device: no lgasr_poll / host round trip is needed before the next scorer call.  Replaces
 * BatchBeamSearch.search's select_state per survivor (espnet/nets/batch_beam_search.py:347-355) and the
 * per-layer torch.stack / unstack of BaseTransformerDecoder.batch_score
 * (espnet2/asr/decoder/transformer_decoder.py:268-272,285-286).
 *   d_src, d_dst: [n_outer][B*beam][row stride] bytes (outer = layer; strides may differ: the source may
 *   be the tightly packed tensor a decoder just produced, the destination a padded persistent buffer);
 *   dst[o][b*beam + j][0:row_bytes) = src[o][b*beam + parent[b][j]][0:row_bytes) for j < n_running[b];
 *   other destination rows are left untouched.  row_bytes % 4 == 0; 16-byte accesses are used when every
 *   pointer / stride / size allows.  d_dst != d_src (ping-pong buffers).  No-op before the first step. */
int lgasr_reorder_rows(lgasr_handle* h, void* d_dst, const void* d_src, int64_t n_outer, int64_t dst_outer_bytes,
                       int64_t src_outer_bytes, int64_t dst_row_bytes, int64_t src_row_bytes, int64_t row_bytes,
                       void* stream);

#define LGASR_F32 0
#define LGASR_BF16 1
/* d_out[r][0:vocab) = log_softmax(d_logits[r][0:vocab)) in fp32 for `rows` rows; d_logits fp32 (LGASR_F32)
 * or bf16 (LGASR_BF16) with leading dimension ld_in elements, d_out fp32 with leading dimension ld_out:
 * the (n, V) hand-off of the full scorers (torch.log_softmax at transformer_decoder.py:212,
 * espnet2/asr/llm/llama.py:260,352) written straight into the padded row layout lgasr_step reads.
 * vocab <= 131072 (rows above 16 K entries are split over a thread-block cluster of up to 8 CTAs). */
int lgasr_log_softmax_rows(const void* d_logits, int dtype, int64_t ld_in, float* d_out, int64_t ld_out, int rows,
                           int vocab, void* stream);

/* Normaliser only: d_lse[r] = logsumexp(d_logits[r][0:vocab)) in fp32 (same kernel, no (n, V) output).  Together with
 * lgasr_step_ex this is the LLM logits hand-off: the 32 k - 128 k wide row of
 * LLMGuidedTransformerDecoder.batch_score (espnet2/asr/decoder/transformer_decoder.py:1115-1132; bf16 logits,
 * espnet2/asr/llm/llama.py:27,47-52) is read ONCE by the step, at 2 bytes per entry, instead of
 * bf16 read + fp32 write (log_softmax) + fp32 read (step) = 10 bytes per entry. */
int lgasr_row_logsumexp(const void* d_logits, int dtype, int64_t ld_in, float* d_lse, int rows, int vocab, void* stream);

/* lgasr_row_logsumexp for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same
 * inputs, output and tolerance as lgasr_row_logsumexp, but a row is split over plain CTAs (no thread-block cluster)
 * whose partial (max, sum) pairs meet in a FIXED scratch array inside the library (a file-scope __device__ array of
 * 278 528 bytes that every GPU loading the library holds until exit); the call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine) - it queues one cudaMemsetAsync of the scratch tickets and one kernel on
 * `stream`.  Limits: rows <= LGASR_NOALLOC_MAX_ROWS, vocab <= LGASR_NOALLOC_MAX_VOCAB, else LGASR_EINVAL with a message.
 * Every call on a device shares the one scratch array: calls that OVERLAP on the GPU (different streams, or graph
 * branches that may run concurrently) race on it - serialise them.  The first call on each device resolves the
 * array's device address (cudaGetSymbolAddress); make that call once outside a capture. */
#define LGASR_NOALLOC_MAX_ROWS 4096
#define LGASR_NOALLOC_MAX_VOCAB (1 << 20)
int lgasr_row_logsumexp_noalloc(const void* d_logits, int dtype, int64_t ld_in, float* d_lse, int rows, int vocab,
                                void* stream);

/* lgasr_step with typed term matrices: h_term_dtype[k] is LGASR_F32 or LGASR_BF16 (storage of matrix k);
 * h_term_norm[k] (may be NULL, and the array itself may be NULL) is a device array of B*beam fp32 normalisers:
 * the score of entry v of row r is float(matrix[r][v]) - norm[r], i.e. log_softmax(logits) when norm = logsumexp.
 * The fp32 fusion order is that of lgasr_step.  Rows must be 16-byte aligned (ld % 4 == 0 for fp32, % 8 for bf16 is
 * not required: 8-byte loads) and vocab % 4 == 0.  The pre-beam key term, if any, must be plain fp32. */
int lgasr_step_ex(lgasr_handle* h, const void* const* h_term_ptr, const int32_t* h_term_dtype,
                  const float* const* h_term_norm, const float* h_term_const, const float* h_term_weight, int64_t ld,
                  void* stream);

/* Timestamp-rule full scorer of the OWSM recipe = ScoreFilter.batch_score
 * (espnet2/bin/s2t_inference.py:47-146): d_out[r][0:vocab) = 0.0 or -inf from the prefix
 * d_yseq[r][0:cur_len) (int64, leading dimension ld_y; e.g. the handle's yseq view).  Where the reference
 * raises IndexError (cur_len < 3 without <notimestamps> in the prefix) y[-3] is taken as "not sos". */
int lgasr_score_filter(const int64_t* d_yseq, int64_t ld_y, int cur_len, int rows, int vocab, int notimestamps,
                       int first_time, int last_time, int sos, int eos, float* d_out, int64_t ld_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LGASR_H_ */
