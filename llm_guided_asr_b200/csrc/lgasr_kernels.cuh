// Kernels of the fused step.  Two launches per decode position:
//
//   k_score_rows   grid (beam, B): one CTA per running hypothesis
//                    warp 0      : CTC state of the hypothesis = forward recursion of its parent's
//                                  state extended by its last token (select_state-on-read), then phi
//                    warps 1..7  : stream the K full-scorer rows once, fuse sum_k w_k*s_k, keep the
//                                  pre-beam top-S in registers (no (n,V) temporaries, no topk launch)
//                    all warps   : log psi of the S candidates (+eos) by warp-shuffle log-sum-exp,
//                                  + w_ctc*(log psi - s_prev) + running score -> candidate records
//   k_select       grid (B): global top-beam over the <= beam*(S+1) records of an utterance,
//                    hypothesis bookkeeping, eos / ended pool, end detection, next yseq column.
#pragma once
#include "lgasr_device.cuh"

namespace lgasr {

constexpr int kMaxTerms = 8;
constexpr int kRowThreads = 256;
constexpr int kRowWarps = kRowThreads / 32;
constexpr int kSelThreads = 256;

struct Terms {
  const float* ptr[kMaxTerms];   // (B*beam, ld) matrix of term k (fp32, or bf16 when bf16[k]); NULL: constant term
  const float* norm[kMaxTerms];  // optional per-row normaliser: log-probability = entry - norm[row] (logits + logsumexp)
  float cst[kMaxTerms];
  float w[kMaxTerms];
  int slot[kMaxTerms];  // index of term k among the terms that have a matrix (staging order)
  unsigned char bf16[kMaxTerms];
  int nptr;             // number of terms with a matrix
  int K;
  long long ld;
};

struct Dev {
  int B, T, Tp, V, W, S, R, K, key, use_ctc, mode, blank, sos, eos, Lcap, Ecap, maxlen_cap, end_detect, NS;
  int P;  // primer length: decode position of an utterance = cur_len - P
  float w_ctc;
  double d_end;
  const float* xT;  // [B][V][Tp]
  const int* xlen;  // [B]
  float* r[2];      // [B][W][2][Tp]
  float* s_prev[2];
  float* score[2];
  float* scores[2];      // [B][W][NS]
  long long* yseq[2];    // [B][W][Lcap]
  int* run_slot[2];      // [B][W] slot of a running hyp in the best list of the step that made it
  int* cur_len[2];       // [B]  ping-pong like every per-step state: a step reads [pp], writes [pp^1]
  int* n_run[2];         // [B]
  int* done[2];          // [B]
  int4* rmeta[2];        // [B][W] (parent slot or -1, last token, parent's last token or -1, 0)
  int* arrive;           // [B] CTAs of the utterance that finished scoring in the current launch
  int* step_flag;        // [B] decode positions of the utterance whose selection is complete (release / acquire)
  int* step_parent;      // [B][W]
  int* step_tok;         // [B][W]
  const int* maxlen;     // [B]
  const int* minlen;     // [B]
  float* cand_total;     // [B][W][R]
  int* cand_tok;         // [B][W][R]
  float* cand_psi;       // [B][W][R]
  int* bp_par;           // [B][maxlen_cap][W]
  int* bp_tok;           // [B][maxlen_cap][W]
  int* n_ended;          // [B]
  float* end_score;      // [B][Ecap]
  int4* end_meta;        // [B][Ecap]  (step, slot, len, extra_eos)
  float* end_scores;     // [B][Ecap][NS]
  float* best_by_len;    // [B][Lcap + 2]
  float* best_all;       // [B]
  long long* trace;      // debug: [B][W][2][16] clock64 stamps of k_step phases (NULL = off)
  // packed copies of the state k_step2 reads at the top of a position (one 16-byte load each instead of a dozen
  // scattered 4-byte loads): written next to the arrays above by k_begin / utt_select / utt_carry_done
  int4* urec[2];         // [B]      (done, n_run, cur_len, 0)
  int4* hrec[2];         // [B][W][2] (parent slot, last token, parent's last token, run_slot), (s_prev, score, 0, 0) as bits
  int4* ustat;           // [B]      (xlen, maxlen, minlen, 0), fixed for the utterance
  int* status;           // device status word: bit 0 = an eligible ended hypothesis did not fit the ended pool
};

// modes
constexpr int kModePreBeam = 0;  // CTC with pre-beam: S candidates + eos per hypothesis
constexpr int kModeNoCtc = 1;    // full scorers only: top-beam per hypothesis
constexpr int kModeFullCtc = 2;  // CTC over the whole vocabulary (ctc_weight == 1.0 / tiny V)

// one full-scorer entry: fp32 or bf16 storage, optional per-row normaliser (log-probability = logit - logsumexp)
__device__ __forceinline__ float term_at(const Terms& t, int k, int row, long long rowoff, int v) {
  float s;
  if (t.bf16[k]) {
    const unsigned short u = __ldg(reinterpret_cast<const unsigned short*>(t.ptr[k]) + rowoff + v);
    s = __uint_as_float((unsigned)u << 16);
  } else {
    s = __ldg(t.ptr[k] + rowoff + v);
  }
  if (t.norm[k]) s = __fsub_rn(s, __ldg(t.norm[k] + row));
  return s;
}

// sum_k w_k * s_k[v] in the reference's fp32 order: ((0 + w0*s0) + w1*s1) + ...  (no FMA contraction)
__device__ __forceinline__ float fuse_at(const Terms& t, long long rowoff, int v) {
  float acc = 0.f;
#pragma unroll 1
  for (int k = 0; k < t.K; ++k) {
    float s = t.ptr[k] ? __ldg(t.ptr[k] + rowoff + v) : t.cst[k];
    acc = __fadd_rn(acc, __fmul_rn(t.w[k], s));
  }
  return acc;
}

__device__ __forceinline__ float4 fuse_at4(const Terms& t, long long rowoff, int v) {
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < kMaxTerms; ++k) {
    if (k < t.K) {
      float4 s;
      if (t.ptr[k])
        s = __ldcs(reinterpret_cast<const float4*>(t.ptr[k] + rowoff + v));
      else
        s = make_float4(t.cst[k], t.cst[k], t.cst[k], t.cst[k]);
      acc.x = __fadd_rn(acc.x, __fmul_rn(t.w[k], s.x));
      acc.y = __fadd_rn(acc.y, __fmul_rn(t.w[k], s.y));
      acc.z = __fadd_rn(acc.z, __fmul_rn(t.w[k], s.z));
      acc.w = __fadd_rn(acc.w, __fmul_rn(t.w[k], s.w));
    }
  }
  return acc;
}

// ------------------------------------------------------------------------------------------
// CTC state of hypothesis (b, h): recursion of the parent's state by the hypothesis' last token
// (what CTCPrefixScoreTH.__call__ computes for every candidate, ctc_prefix_score.py:122-160, and
// CTCPrefixScorer.select_state picks, scorers/ctc.py:56-62), done here only for the survivors.
// One warp.  Shared buffers (Tp floats each): bA, bB, bC, bD.
// On exit: bA[t] = r_sum[t] (log(r^n + r^b)), bD[t] = r^b[t] of the hypothesis for t in [0, T).
__device__ __forceinline__ void warp_ctc_state(const float* __restrict__ xT_b,  // [V][Tp] of the utterance
                                               const float* __restrict__ r_par,  // [2][Tp] parent state (or own)
                                               float* __restrict__ r_out,        // [2][Tp] own state (written)
                                               int has_parent, int tok, int par_last, int olen_par, int T, int Tp,
                                               int blank, float* bA, float* bB, float* bC, float* bD, int lane,
                                               int win_s = 0, int win_e = 0) {
  // attention window (ctc_prefix_score.py:143-148): the recursion runs over frames [win_s, win_e) only (0, 0: the whole
  // utterance from max(output_length, 1)); frames outside keep log-zero
  const int Tall = T;
  if (win_e > 0 && win_e < T) T = win_e;
  if (!has_parent) {
    for (int t = lane; t < T; t += 32) {
      float rn = r_par[t], rb = r_par[Tp + t];
      r_out[t] = rn;
      r_out[Tp + t] = rb;
      bA[t] = lse2(rn, rb);
      bD[t] = rb;
    }
    __syncwarp();
    return;
  }
  const int start = win_s > 0 ? win_s : (olen_par > 1 ? olen_par : 1);
  const float* xc = xT_b + (long long)tok * Tp;
  const float* x0 = xT_b + (long long)blank * Tp;
  const bool same = (tok == par_last);
  // phi of the parent, x rows
  for (int t = lane; t < T; t += 32) {
    float rn = r_par[t], rb = r_par[Tp + t];
    bD[t] = same ? rb : lse2(rn, rb);  // phi_prev[t]
    bA[t] = __ldg(xc + t);
    bC[t] = __ldg(x0 + t);
  }
  __syncwarp();
  const float rn0 = (olen_par == 0 && start == 1) ? bA[0] : LGASR_LOGZERO;  // r[start-1, 0]  (:128-129)
  for (int t = start + lane; t < T; t += 32) bB[t] = bD[t - 1] + bA[t];
  __syncwarp();
  warp_affine_scan(bA, bB, start, T, rn0, lane);  // bB = r^n
  __syncwarp();
  // r^b: a = x0, b[t] = r^n[t-1] + x0[t]
  for (int t = start + lane; t < T; t += 32) {
    float prev = (t - 1 >= start) ? bB[t - 1] : rn0;
    bD[t] = prev + bC[t];
  }
  __syncwarp();
  warp_affine_scan(bC, bD, start, T, LGASR_LOGZERO, lane);  // bD = r^b
  __syncwarp();
  for (int t = lane; t < Tall; t += 32) {
    float rn, rb;
    if (t >= T) {
      rn = LGASR_LOGZERO;
      rb = LGASR_LOGZERO;
    } else if (t >= start) {
      rn = bB[t];
      rb = bD[t];
    } else if (t == start - 1) {
      rn = rn0;
      rb = LGASR_LOGZERO;
    } else {
      rn = LGASR_LOGZERO;
      rb = LGASR_LOGZERO;
    }
    r_out[t] = rn;
    r_out[Tp + t] = rb;
    bA[t] = lse2(rn, rb);
    bD[t] = rb;
  }
  __syncwarp();
}

template <int NL>
__global__ void __launch_bounds__(kRowThreads) k_score_rows(Dev d, Terms tm, int pp, int rp) {
  const int b = blockIdx.y, h = blockIdx.x;
  if (d.done[pp][b] || h >= d.n_run[pp][b]) return;
  extern __shared__ float smem[];
  const int Tp = d.Tp;
  float* bA = smem;  // r_sum
  float* bB = bA + Tp;
  float* bC = bB + Tp;
  float* bD = bC + Tp;  // r^b
  float* m_val = bD + Tp;                              // [kRowWarps][capS]
  const int capS = 32 * NL;
  int* m_idx = reinterpret_cast<int*>(m_val + kRowWarps * capS);
  float* sel_val = reinterpret_cast<float*>(m_idx + kRowWarps * capS);  // [capS]
  int* sel_tok = reinterpret_cast<int*>(sel_val + capS);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int T = d.xlen[b];
  const int row = b * d.W + h;
  const long long rowoff = (long long)row * tm.ld;
  const int cur_len = d.cur_len[pp][b];
  const int olen = cur_len - 1;
  const long long* yrow = d.yseq[pp] + (long long)row * d.Lcap;
  const int last = (int)yrow[cur_len - 1];
  const float* xT_b = d.xT + (long long)b * d.V * Tp;
  const int mode = d.mode;
  const int n_sel = (mode == kModePreBeam) ? d.S : d.W;

  WarpTop<NL> top;
  top.init(n_sel);

  const int first_stream_warp = d.use_ctc ? 1 : 0;
  if (d.use_ctc && warp == 0) {
    const int par = d.rmeta[pp][row].x;
    const int has_parent = par >= 0;
    const float* r_par = d.r[rp] + ((long long)(b * d.W + (has_parent ? par : h)) * 2) * Tp;
    float* r_out = d.r[rp ^ 1] + ((long long)row * 2) * Tp;
    const int par_last = (has_parent && cur_len >= 2) ? (int)yrow[cur_len - 2] : -1;
    warp_ctc_state(xT_b, r_par, r_out, has_parent, last, par_last, olen - 1, T, Tp, d.blank, bA, bB, bC, bD, lane);
  } else if (mode != kModeFullCtc) {
    // ---- stream the fused row, keep the top n_sel --------------------------------------------
    const int V = d.V;
    const int nw = kRowWarps - first_stream_warp;
    const int w = warp - first_stream_warp;
    bool vec_ok = (tm.ld % 4 == 0) && (d.key < 0 || true);
#pragma unroll
    for (int k = 0; k < kMaxTerms; ++k)
      if (k < tm.K && tm.ptr[k]) vec_ok = vec_ok && ((reinterpret_cast<uintptr_t>(tm.ptr[k]) & 15) == 0);
    if (vec_ok) {
      const int nvec = V >> 2;
      for (int q0 = w * 32; q0 < nvec; q0 += nw * 32) {
        const int q = q0 + lane;
        const bool ok = q < nvec;
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) {
          if (d.key < 0)
            f = fuse_at4(tm, rowoff, q * 4);
          else
            f = __ldcs(reinterpret_cast<const float4*>(tm.ptr[d.key] + rowoff + q * 4));
        }
        top.offer(f.x, q * 4 + 0, ok, lane);
        top.offer(f.y, q * 4 + 1, ok, lane);
        top.offer(f.z, q * 4 + 2, ok, lane);
        top.offer(f.w, q * 4 + 3, ok, lane);
      }
      // tail
      for (int v0 = (nvec << 2) + w * 32; v0 < V; v0 += nw * 32) {
        const int v = v0 + lane;
        const bool ok = v < V;
        float f = 0.f;
        if (ok) f = (d.key < 0) ? fuse_at(tm, rowoff, v) : __ldg(tm.ptr[d.key] + rowoff + v);
        top.offer(f, v, ok, lane);
      }
    } else {
      for (int v0 = w * 32; v0 < V; v0 += nw * 32) {
        const int v = v0 + lane;
        const bool ok = v < V;
        float f = 0.f;
        if (ok) f = (d.key < 0) ? fuse_at(tm, rowoff, v) : __ldg(tm.ptr[d.key] + rowoff + v);
        top.offer(f, v, ok, lane);
      }
    }
  }
  __syncthreads();

  const float s_prev = d.use_ctc ? d.s_prev[pp][row] : 0.f;
  const float hyp_score = d.score[pp][row];
  const int start = olen > 1 ? olen : 1;
  const float psi_eos = d.use_ctc ? bA[T - 1] : 0.f;  // r_sum at the end frame (:180-181)

  if (mode == kModeFullCtc) {
    // ---- CTC over the whole vocabulary: total[v] for every v, keep the top beam ---------------
    for (int v = warp; v < d.V; v += kRowWarps) {
      const float* phi = (v == last) ? bD : bA;
      const float* xr = xT_b + (long long)v * Tp;
      float rn0 = (olen == 0) ? __ldg(xr) : LGASR_LOGZERO;
      float psi = warp_prefix_lse(phi, xr, start, T, rn0, lane);
      if (v == d.eos) psi = psi_eos;
      if (d.eos != d.blank && v == d.blank) psi = LGASR_LOGZERO;
      float wfull = fuse_at(tm, rowoff, v);
      float total = __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi, s_prev))), hyp_score);
      if (top.passes(total, v)) top.insert(total, v, lane);
    }
  }

  // ---- merge the per-warp lists --------------------------------------------------------------
  {
    const bool has_list = (mode == kModeFullCtc) || (warp >= first_stream_warp);
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      m_val[warp * capS + j * 32 + lane] = has_list ? top.v[j] : -INFINITY;
      m_idx[warp * capS + j * 32 + lane] = has_list ? top.i[j] : 0x7fffffff;
    }
  }
  __syncthreads();
  if (warp == 0) {
    WarpTop<NL> fin;
    fin.init(n_sel);
    for (int e0 = 0; e0 < kRowWarps * capS; e0 += 32) {
      float v = m_val[e0 + lane];
      int i = m_idx[e0 + lane];
      fin.offer(v, i, i != 0x7fffffff, lane);
    }
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      sel_val[j * 32 + lane] = fin.v[j];
      sel_tok[j * 32 + lane] = fin.i[j];
    }
  }
  __syncthreads();

  // ---- candidate records -----------------------------------------------------------------------
  const int R = d.R;
  float* c_total = d.cand_total + (long long)row * R;
  int* c_tok = d.cand_tok + (long long)row * R;
  float* c_psi = d.cand_psi + (long long)row * R;
  const int n_rec = (mode == kModePreBeam) ? d.S + 1 : d.W;
  for (int j = warp; j < n_rec; j += kRowWarps) {
    int c;
    bool valid = true;
    if (j < n_sel) {
      c = sel_tok[j];
      if (c == 0x7fffffff) valid = false;  // fewer than n_sel tokens in the vocabulary
    } else {
      // eos is always scored by the CTC scorer (:180-181); skip if it already is a candidate
      c = d.eos;
      bool dup = false;
      for (int e = lane; e < n_sel; e += 32) dup |= (sel_tok[e] == c);
      if (__any_sync(FULLMASK, dup)) valid = false;
    }
    float total = -INFINITY, psi = 0.f;
    if (valid) {
      const float wfull = fuse_at(tm, rowoff, c);
      if (d.use_ctc) {
        const float* phi = (c == last) ? bD : bA;
        const float* xr = xT_b + (long long)c * Tp;
        float rn0 = (olen == 0) ? __ldg(xr) : LGASR_LOGZERO;
        psi = warp_prefix_lse(phi, xr, start, T, rn0, lane);
        if (c == d.eos) psi = psi_eos;
        if (d.eos != d.blank && c == d.blank) psi = LGASR_LOGZERO;
        total = __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi, s_prev))), hyp_score);
      } else {
        total = __fadd_rn(wfull, hyp_score);
      }
    }
    if (lane == 0) {
      c_total[j] = total;
      c_tok[j] = valid ? c : -1;
      c_psi[j] = psi;
    }
  }
}

// ------------------------------------------------------------------------------------------
// total score back from a packed record key
__device__ __forceinline__ float key_total(unsigned long long k) {
  const unsigned u = (unsigned)(k >> 32);
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// Selection of one utterance (all threads of a CTA): global top-beam over the candidate records of its
// running hypotheses (batch_beam, batch_beam_search.py:101-128), hypothesis bookkeeping (:325-363),
// post_process (:365-429) and end_detect (e2e_asr_common.py:17-48).
// Reads state [pp], writes state [pp^1].  smem: >= utt_select_smem_floats(W, R, NS) floats, 16-byte aligned.
// Every global operand whose address is known up front (records, the parents' fields, the ended-pool
// scalars) is requested in ONE batch of loads before the first barrier, so the whole selection costs a
// single global round trip plus the scattered reads of the chosen tokens' full-scorer scores.
constexpr int kSelListCap = 256;  // survivors of the row-maximum prefilter of utt_select
__host__ __device__ inline long long utt_select_smem_floats(int W, int R, int NS) {
  // records (key u64 + psi) | sel, dest | parents' fields | row maxima (u64) | survivor list (key u64 + index) | scalars
  return (long long)W * R * 3 + 2LL * W + (long long)W * (3 + NS) + 2LL * W + 3LL * kSelListCap + 8;
}

// PRE: the records and the parents' fields are already in `smem` (written there through distributed shared
// memory by the CTAs of the cluster, k_step<.., true>); otherwise they are fetched from global memory.
template <bool PRE>
__device__ __forceinline__ void utt_select(const Dev& d, const Terms& tm, int b, int pp, int step, float* smem, int n_run,
                                           int cur_len, int maxlen, int minlen, long long* tr = nullptr) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int R = d.R, W = d.W, NS = d.NS;
  const int n_rec_row = (d.mode == kModePreBeam) ? d.S + 1 : d.W;
  const int nrec = n_run * n_rec_row;
  // packed records: key = (ordered total << 32) | ~(h*V + tok): larger key = higher total, then lower flat index
  unsigned long long* r_key = reinterpret_cast<unsigned long long*>(smem);  // [W*R]   0 = invalid
  float* r_psi = reinterpret_cast<float*>(r_key + W * R);                   // [W*R]
  int* sel = reinterpret_cast<int*>(r_psi + W * R);                         // [W] record index per rank
  int* dest = sel + W;             // [W] >=0 running idx, -1-e ended idx e, INT_MIN dropped
  int* p_slot = dest + W;          // [W] parents: slot in the best list of the step that made them
  int* p_last = p_slot + W;        // [W] parents: last token
  float* p_sprev = reinterpret_cast<float*>(p_last + W);  // [W]
  float* p_scores = p_sprev + W;   // [W][NS]
  // 8-byte aligned tail: W*R*3 + 2W + 3W + W*NS floats precede it; pad to an even count
  float* tail = p_scores + W * NS;
  if ((tail - smem) & 1) ++tail;
  unsigned* rowmax = reinterpret_cast<unsigned*>(tail);  // [W] best ordered total of each row (upper key halves); 2W floats reserved
  unsigned long long* c_key = reinterpret_cast<unsigned long long*>(tail) + W;  // [kSelListCap]
  int* c_e = reinterpret_cast<int*>(c_key + kSelListCap);                    // [kSelListCap]
  int* sh_nsel = c_e + kSelListCap;
  int* sh_cnt = sh_nsel + 1;
  const bool last = (step == maxlen - 1);
  const int new_len = cur_len + 1 + (last ? 1 : 0);
  float* bbl = d.best_by_len + (long long)b * (d.Lcap + 2);
  int n_end0 = 0;
  float best_all0 = 0.f, bbl_new0 = 0.f, bbl_m[3] = {-INFINITY, -INFINITY, -INFINITY};
  // row-maximum prefilter: with W running rows, the W row maxima are W distinct records, so nothing below
  // their minimum can be among the W best; typically ~2W of the W*(S+1) records survive
  const bool prefilter = (n_run >= W) && (nrec <= 2 * nthr);
  if (tid < W) rowmax[tid] = 0u;
  if (tid == 0) *sh_cnt = 0;
  if (prefilter) __syncthreads();
  if (PRE) {
    if (tid == 0) {
      n_end0 = d.n_ended[b];
      best_all0 = d.best_all[b];
      bbl_new0 = bbl[new_len];
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        const int len = step - m;
        if (len >= 0 && len <= d.Lcap + 1) bbl_m[m] = bbl[len];
      }
    }
    if (prefilter)
      for (int e = tid; e < nrec; e += nthr) {
        const unsigned long long ke = r_key[e];
        if (ke) atomicMax(rowmax + e / n_rec_row, (unsigned)(ke >> 32));
      }
  } else {
  // first pass of both loops: every global load is issued before anything is consumed (one round trip)
  const int npar = n_run * (3 + NS);
  int tok0 = -1, h0 = 0;
  float tot0 = 0.f, psi0 = 0.f;
  if (tid < nrec) {
    h0 = tid / n_rec_row;
    const long long o = (long long)(b * W + h0) * R + (tid - h0 * n_rec_row);
    tok0 = __ldcg(d.cand_tok + o);  // written by other CTAs of this launch: read through L2
    tot0 = __ldcg(d.cand_total + o);
    psi0 = __ldcg(d.cand_psi + o);
  }
  unsigned pv0 = 0u;
  int ph0 = 0, pf0 = 0;
  if (tid < npar) {
    ph0 = tid / (3 + NS);
    pf0 = tid - ph0 * (3 + NS);
    const int prow = b * W + ph0;
    if (pf0 == 0) pv0 = (unsigned)d.run_slot[pp][prow];
    else if (pf0 == 1) pv0 = (unsigned)d.rmeta[pp][prow].y;
    else if (pf0 == 2) pv0 = __float_as_uint(d.s_prev[pp][prow]);
    else pv0 = __float_as_uint(d.scores[pp][(long long)prow * NS + (pf0 - 3)]);
  }
  // scalars of the serial part (only this utterance's selection ever writes them)
  if (tid == 0) {
    n_end0 = d.n_ended[b];
    best_all0 = d.best_all[b];
    bbl_new0 = bbl[new_len];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      const int len = step - m;
      if (len >= 0 && len <= d.Lcap + 1) bbl_m[m] = bbl[len];
    }
  }
  auto put_record = [&](int e, int h, int tok, float tot, float psi) {
    r_psi[e] = psi;
    unsigned u = __float_as_uint(tot + 0.0f);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    const unsigned long long ke = tok >= 0 ? (((unsigned long long)u << 32) | (unsigned)(~(h * d.V + tok))) : 0ull;
    r_key[e] = ke;
    if (prefilter && ke) atomicMax(rowmax + h, (unsigned)(ke >> 32));  // native 32-bit shared atomic
  };
  auto put_parent = [&](int h, int f, unsigned v) {
    if (f == 0) p_slot[h] = (int)v;
    else if (f == 1) p_last[h] = (int)v;
    else if (f == 2) p_sprev[h] = __uint_as_float(v);
    else p_scores[h * NS + (f - 3)] = __uint_as_float(v);
  };
  if (tid < nrec) put_record(tid, h0, tok0, tot0, psi0);
  if (tid < npar) put_parent(ph0, pf0, pv0);
  for (int e = tid + nthr; e < nrec; e += nthr) {  // wide beams only
    const int h = e / n_rec_row, j = e - h * n_rec_row;
    const long long o = (long long)(b * W + h) * R + j;
    put_record(e, h, __ldcg(d.cand_tok + o), __ldcg(d.cand_total + o), __ldcg(d.cand_psi + o));
  }
  for (int i = tid + nthr; i < npar; i += nthr) {
    const int h = i / (3 + NS), f = i - h * (3 + NS);
    const int prow = b * W + h;
    unsigned v;
    if (f == 0) v = (unsigned)d.run_slot[pp][prow];
    else if (f == 1) v = (unsigned)d.rmeta[pp][prow].y;
    else if (f == 2) v = __float_as_uint(d.s_prev[pp][prow]);
    else v = __float_as_uint(d.scores[pp][(long long)prow * NS + (f - 3)]);
    put_parent(h, f, v);
  }
  }
  if (tid < W) sel[tid] = -1;
  __syncthreads();
  if (tr && tid == 0) tr[9 + 3] = clock64();
  bool ranked = false;
  if (prefilter) {
    unsigned tau = 0xffffffffu;  // minimum over the rows of the best ordered total
    for (int h = 0; h < W; ++h) tau = min(tau, rowmax[h]);
    for (int e = tid; e < nrec; e += nthr) {
      const unsigned long long ke = r_key[e];
      if (ke != 0ull && (unsigned)(ke >> 32) >= tau) {
        const int slot = atomicAdd(sh_cnt, 1);
        if (slot < kSelListCap) {
          c_key[slot] = ke;
          c_e[slot] = e;
        }
      }
    }
    __syncthreads();
    const int n = *sh_cnt;
    if (n <= kSelListCap) {
      for (int i = tid; i < n; i += nthr) {
        const unsigned long long ke = c_key[i];
        int rank = 0;
#pragma unroll 4
        for (int o = 0; o < n; ++o) rank += (c_key[o] > ke) ? 1 : 0;
        if (rank < W) sel[rank] = c_e[i];
      }
      ranked = true;
      __syncthreads();
    }
  }
  if (ranked) {
  } else if (nrec <= 2 * nthr) {
    // rank by counting (two keys per shared-memory load)
    const ulonglong2* k2 = reinterpret_cast<const ulonglong2*>(r_key);
    const int n2 = nrec >> 1;
    for (int e = tid; e < nrec; e += nthr) {
      const unsigned long long ke = r_key[e];
      if (ke == 0ull) continue;
      int rank = 0;
#pragma unroll 4
      for (int o = 0; o < n2; ++o) {
        const ulonglong2 kk = k2[o];
        rank += (kk.x > ke) ? 1 : 0;
        rank += (kk.y > ke) ? 1 : 0;
      }
      if (nrec & 1) rank += (r_key[nrec - 1] > ke) ? 1 : 0;
      if (rank < W) sel[rank] = e;
    }
    __syncthreads();
  } else {
    // wide beams: W rounds of block arg-best (each round takes the best record below the previous pick)
    __shared__ unsigned long long sh_redk[32];
    __shared__ int sh_rede[32];
    const int nw = nthr >> 5;
    unsigned long long pk = ~0ull;
    for (int r = 0; r < W; ++r) {
      unsigned long long bk = 0ull;
      int be = -1;
      for (int e = tid; e < nrec; e += nthr) {
        const unsigned long long ke = r_key[e];
        if (ke != 0ull && ke < pk && ke > bk) {
          bk = ke;
          be = e;
        }
      }
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) {
        const unsigned long long ok = __shfl_xor_sync(FULLMASK, bk, o);
        const int oe = __shfl_xor_sync(FULLMASK, be, o);
        if (ok > bk) {
          bk = ok;
          be = oe;
        }
      }
      if ((tid & 31) == 0) {
        sh_redk[tid >> 5] = bk;
        sh_rede[tid >> 5] = be;
      }
      __syncthreads();
      bk = sh_redk[0];
      be = sh_rede[0];
      for (int w = 1; w < nw; ++w) {
        if (sh_redk[w] > bk) {
          bk = sh_redk[w];
          be = sh_rede[w];
        }
      }
      if (tid == 0) sel[r] = be;
      __syncthreads();
      if (be < 0) break;
      pk = bk;
    }
  }

  if (tr && tid == 0) tr[10] = clock64();
  // bookkeeping by warp 0, 32 ranks at a time: where each kept record goes (running slot / ended pool /
  // dropped), in rank order exactly like the reference's loop over the best list (:325-362, :400-429)
  if (tid < 32) {
    const int lane = tid;
    const unsigned lt = (1u << lane) - 1u;
    int nsel = 0, nrun_new = 0, n_end = __shfl_sync(FULLMASK, n_end0, 0);
    float best_new = -INFINITY;  // best ended score of this step: all of them have length new_len
    for (int base = 0; base < W; base += 32) {
      const int j = base + lane;
      const int e = (j < W) ? sel[j] : -1;
      const bool valid = e >= 0;
      int tok = -1;
      if (valid) tok = (int)(~(unsigned)r_key[e]) - (e / n_rec_row) * d.V;
      const bool is_eos = valid && (last || tok == d.eos);
      const bool elig = is_eos && step >= minlen;
      const bool run = valid && !is_eos;
      const unsigned m_valid = __ballot_sync(FULLMASK, valid);
      const unsigned m_el = __ballot_sync(FULLMASK, elig);
      const unsigned m_run = __ballot_sync(FULLMASK, run);
      float sc = -INFINITY;
      if (run) dest[j] = nrun_new + __popc(m_run & lt);
      if (is_eos) {
        const int my_end = n_end + __popc(m_el & lt);
        if (elig && my_end < d.Ecap) {
          dest[j] = -1 - my_end;
          sc = key_total(r_key[e]);
          d.end_score[(long long)b * d.Ecap + my_end] = sc;
          d.end_meta[(long long)b * d.Ecap + my_end] = make_int4(step, j, new_len, last ? 1 : 0);
        } else {
          dest[j] = INT_MIN;
          if (elig) atomicOr(d.status, 1);  // pool overflow: reported by lgasr_poll / the gather calls, never silent
        }
      }
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) sc = fmaxf(sc, __shfl_xor_sync(FULLMASK, sc, o));
      best_new = fmaxf(best_new, sc);
      n_end = min(n_end + __popc(m_el), d.Ecap);
      nrun_new += __popc(m_run);
      nsel += __popc(m_valid);
    }
    if (lane == 0) {
      float best_all = best_all0;
      if (best_new > -INFINITY) {
        if (best_new > best_all) best_all = best_new;
        if (best_new > bbl_new0) bbl[new_len] = best_new;
        d.n_ended[b] = n_end;
        d.best_all[b] = best_all;
      }
      // end detection (e2e_asr_common.py:17-48): lengths step, step-1, step-2 (none of them is new_len)
      bool fin = (nrun_new == 0) || last;
      if (!fin && d.end_detect && n_end > 0) {
        int count = 0;
#pragma unroll
        for (int m = 0; m < 3; ++m) {
          const float bs = bbl_m[m];
          if (bs > -INFINITY && ((double)bs - (double)best_all) < d.d_end) ++count;
        }
        if (count == 3) fin = true;
      }
      d.n_run[pp ^ 1][b] = nrun_new;
      d.cur_len[pp ^ 1][b] = cur_len + 1;
      d.done[pp ^ 1][b] = fin ? 1 : 0;
      d.urec[pp ^ 1][b] = make_int4(fin ? 1 : 0, nrun_new, cur_len + 1, 0);
      *sh_nsel = nsel;
    }
  }
  if (tr && tid == 0) tr[11] = clock64();
  __syncthreads();
  const int nsel = *sh_nsel;
  // per-hypothesis fields
  if (tid < nsel) {
    const int j = tid;
    const int e = sel[j];
    const int h = e / n_rec_row;
    const int prow = b * W + h;
    const int tok = (int)(~(unsigned)r_key[e]) - h * d.V;
    const float total = key_total(r_key[e]);
    const float psi = r_psi[e];
    // trellis
    const long long bpo = ((long long)b * d.maxlen_cap + step) * W + j;
    d.bp_par[bpo] = p_slot[h];
    d.bp_tok[bpo] = tok;
    const int dj = dest[j];
    float* sc_out = nullptr;
    if (dj >= 0) {
      const int nrow = b * W + dj;
      d.score[pp ^ 1][nrow] = total;
      d.s_prev[pp ^ 1][nrow] = psi;
      d.run_slot[pp ^ 1][nrow] = j;
      d.rmeta[pp ^ 1][nrow] = make_int4(h, tok, p_last[h], 0);
      d.hrec[pp ^ 1][2 * nrow] = make_int4(h, tok, p_last[h], j);
      d.hrec[pp ^ 1][2 * nrow + 1] = make_int4(__float_as_int(psi), __float_as_int(total), 0, 0);
      d.step_parent[nrow] = h;
      d.step_tok[nrow] = tok;
      sc_out = d.scores[pp ^ 1] + (long long)nrow * NS;
    } else if (dj != INT_MIN) {
      sc_out = d.end_scores + ((long long)b * d.Ecap + (-1 - dj)) * NS;
    }
    if (sc_out) {
      const float* sc_in = p_scores + h * NS;
      const long long rowoff = (long long)prow * tm.ld;
      for (int k = 0; k < tm.K; ++k) {
        const float s = tm.ptr[k] ? term_at(tm, k, prow, rowoff, tok) : tm.cst[k];
        sc_out[k] = __fadd_rn(sc_in[k], s);
      }
      if (d.use_ctc) sc_out[tm.K] = __fadd_rn(sc_in[tm.K], __fsub_rn(psi, p_sprev[h]));
    }
  }
  // next yseq rows of the running hypotheses
  const int warp = tid >> 5, lane = tid & 31;
  for (int j = warp; j < nsel; j += nthr >> 5) {
    const int dj = dest[j];
    if (dj < 0) continue;
    const int e = sel[j];
    const int h = e / n_rec_row;
    const long long* src = d.yseq[pp] + (long long)(b * W + h) * d.Lcap;
    long long* dst = d.yseq[pp ^ 1] + (long long)(b * W + dj) * d.Lcap;
    for (int t = lane; t < cur_len; t += 32) dst[t] = src[t];
    if (lane == 0) dst[cur_len] = (long long)((int)(~(unsigned)r_key[e]) - h * d.V);
  }
}

// a finished utterance carries its flags over to the next state buffer
__device__ __forceinline__ void utt_carry_done(const Dev& d, int b, int pp) {
  d.n_run[pp ^ 1][b] = 0;
  d.cur_len[pp ^ 1][b] = d.cur_len[pp][b];
  d.done[pp ^ 1][b] = 1;
  d.urec[pp ^ 1][b] = make_int4(1, 0, d.cur_len[pp][b], 0);
}

// stand-alone selection launch (used after k_score_rows, i.e. for the full-vocabulary CTC mode)
__global__ void __launch_bounds__(kSelThreads) k_select(Dev d, Terms tm, int pp, int step) {
  const int b = blockIdx.x;
  extern __shared__ float smem[];
  if (d.done[pp][b] || d.n_run[pp][b] <= 0) {
    if (threadIdx.x == 0) utt_carry_done(d, b, pp);
    return;
  }
  utt_select<false>(d, tm, b, pp, d.cur_len[pp][b] - d.P, smem, d.n_run[pp][b], d.cur_len[pp][b], d.maxlen[b], d.minlen[b]);
}

// ------------------------------------------------------------------------------------------
// x[B][T][V] -> xT[B][V][Tp] with the padding rule of CTCPrefixScoreTH.__init__ (:46-49)
// 64 x 64 tiles, block (32, 8): 16 independent coalesced loads, then 16 coalesced stores per thread.
__global__ void __launch_bounds__(256) k_transpose_pad(const float* __restrict__ x, float* __restrict__ xT,
                                                       const int* __restrict__ xlen, int T, int Tp, int V, int blank,
                                                       int xlen_scalar) {
  __shared__ float tile[64][65];
  const int b = blockIdx.z;
  const int v0 = blockIdx.x * 64, t0 = blockIdx.y * 64;
  const int len = min(xlen ? xlen[b] : xlen_scalar, T);
  const float* xb = x + (long long)b * T * V;
  float* xo = xT + (long long)b * V * Tp;
  const int tx = threadIdx.x, ty = threadIdx.y;
  float val[16];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int t = t0 + ty + 8 * k;
#pragma unroll
    for (int hv = 0; hv < 2; ++hv) {
      const int v = v0 + tx + 32 * hv;
      float r = (v == blank) ? 0.f : LGASR_LOGZERO;  // padding rule of ctc_prefix_score.py:46-49
      if (v < V && t < len) r = __ldcs(xb + (long long)t * V + v);
      val[2 * k + hv] = r;
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    tile[ty + 8 * k][tx] = val[2 * k];
    tile[ty + 8 * k][tx + 32] = val[2 * k + 1];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int v = v0 + ty + 8 * k;
    if (v < V) {
      const int t = t0 + tx;
      if (t < Tp) xo[(long long)v * Tp + t] = tile[tx][ty + 8 * k];
      if (t + 32 < Tp) xo[(long long)v * Tp + t + 32] = tile[tx + 32][ty + 8 * k];
    }
  }
}

__global__ void k_fill(float* p, long long n, float v) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    p[i] = v;
}

// initial CTC state of hypothesis 0 (ctc_prefix_score.py:82-93): r^n = log-zero, r^b = cumsum_t x[t, blank]
__global__ void k_ctc_init_state(const float* __restrict__ xT, float* __restrict__ r0, int T, int Tp, int V, int W,
                                 int blank, int per_utt_stride_rows) {
  const int b = blockIdx.x;
  const float* x0 = xT + ((long long)b * V + blank) * Tp;
  float* r = r0 + (long long)b * per_utt_stride_rows * 2 * Tp;
  extern __shared__ float cs[];  // [Tp]
  for (int t = threadIdx.x; t < T; t += blockDim.x) cs[t] = x0[t];  // all loads in flight at once
  __syncthreads();
  if (threadIdx.x == 0) {  // the sum itself stays sequential: same fp32 roundings as torch.cumsum on the CPU
    float acc = 0.f;
    for (int t = 0; t < T; ++t) {
      acc = __fadd_rn(acc, cs[t]);
      cs[t] = acc;
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    r[t] = LGASR_LOGZERO;
    r[Tp + t] = cs[t];
  }
}

// lengths of one utterance slot (lgasr_refill): what lgasr_set_posteriors / lgasr_begin copy for a whole batch
__global__ void k_set_utt(Dev d, int utt, int xlen, int maxlen, int minlen) {
  const_cast<int*>(d.xlen)[utt] = xlen;
  const_cast<int*>(d.maxlen)[utt] = maxlen;
  const_cast<int*>(d.minlen)[utt] = minlen;
}

__global__ void k_begin(Dev d, const long long* __restrict__ primer, int primer_len, int b0) {
  const int b = b0 + blockIdx.x;
  const int tid = threadIdx.x;
  const int W = d.W;
  if (tid == 0) {
    for (int p = 0; p < 2; ++p) {
      d.cur_len[p][b] = primer_len;
      d.n_run[p][b] = 1;
      d.done[p][b] = (d.maxlen[b] <= 0) ? 1 : 0;
    }
    d.arrive[b] = 0;
    d.step_flag[b] = 0;
    d.n_ended[b] = 0;
    d.best_all[b] = -INFINITY;
    const int dn = (d.maxlen[b] <= 0) ? 1 : 0;
    d.urec[0][b] = make_int4(dn, 1, primer_len, 0);
    d.urec[1][b] = make_int4(dn, 1, primer_len, 0);
    d.ustat[b] = make_int4(d.xlen[b], d.maxlen[b], d.minlen[b], 0);
  }
  for (int i = tid; i < d.Lcap + 2; i += blockDim.x) d.best_by_len[(long long)b * (d.Lcap + 2) + i] = -INFINITY;
  for (int i = tid; i < W; i += blockDim.x) {
    for (int p = 0; p < 2; ++p) {
      d.score[p][b * W + i] = 0.f;
      d.s_prev[p][b * W + i] = 0.f;
      d.run_slot[p][b * W + i] = -1;
      for (int k = 0; k < d.NS; ++k) d.scores[p][(long long)(b * W + i) * d.NS + k] = 0.f;
    }
    d.step_parent[b * W + i] = -1;
    d.step_tok[b * W + i] = -1;
    const int last = (int)primer[primer_len - 1];
    d.rmeta[0][b * W + i] = make_int4(-1, last, -1, 0);
    d.rmeta[1][b * W + i] = make_int4(-1, last, -1, 0);
    for (int p = 0; p < 2; ++p) {
      d.hrec[p][2 * (b * W + i)] = make_int4(-1, last, -1, -1);
      d.hrec[p][2 * (b * W + i) + 1] = make_int4(0, 0, 0, 0);  // s_prev = score = 0.0f
    }
  }
  for (int i = tid; i < primer_len; i += blockDim.x) {
    d.yseq[0][(long long)(b * W) * d.Lcap + i] = primer[i];
    d.yseq[1][(long long)(b * W) * d.Lcap + i] = primer[i];
  }
}

}  // namespace lgasr
