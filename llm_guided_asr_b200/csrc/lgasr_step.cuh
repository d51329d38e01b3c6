// k_step: ONE launch per decode position (sm_100a).
//
//   grid (beam, B), 256 threads: one CTA per running hypothesis, warp-specialised.
//
//   warps 0-1  "ctc"   : CTC state of the hypothesis = forward recursion of its parent's state by its
//                         last token (what CTCPrefixScoreTH.__call__ computes for EVERY candidate,
//                         ctc_prefix_score.py:122-160, and select_state then picks, scorers/ctc.py:56-62),
//                         done only for the survivor, as a 64-thread log-semiring affine scan over the
//                         frames.  Inputs (parent r, the token's and the blank's posterior rows) arrive by
//                         1-D TMA bulk copies (cp.async.bulk + mbarrier).
//   warps 2-7  "sel"   : stream the K full-scorer rows (TMA bulk copies, up to 4 stages in flight), fuse
//                         sum_k w_k*s_k in the reference's fp32 order and keep the exact pre-beam top-S
//                         (threshold from lane maxima, ballot compaction, rank-by-counting on packed
//                         64-bit keys; no (n,V) temporaries, no topk launch; batch_beam_search.py:279-308).
//   all warps          : the S candidates' posterior rows xT[b][c][:] are gathered by TMA into shared
//                         memory; log psi = logsumexp_t(phi[t-1] + x[t,c]) (+) r^n_{start-1} per warp
//                         (ctc_prefix_score.py:163-185), then w_ctc*(psi - s_prev) + fused + running score
//                         (batch_beam_search.py:313-318) -> candidate records.
//   last CTA of an utterance to finish (atomic ticket): utt_select = global top-beam + bookkeeping.
#pragma once
#include "lgasr_kernels.cuh"

namespace lgasr {

constexpr int kStepThreads = 256;
constexpr int kCtcWarps = 2, kCtcThreads = 32 * kCtcWarps;    // warps 0-1
constexpr int kSelWarps = 6, kSelGroup = 32 * kSelWarps;       // warps 2-7
constexpr int kMaxStages = 4;
constexpr int kPendCap = 256;
constexpr int kTopCap = 128;  // == LGASR_MAX_PREBEAM
constexpr int kStepWarps = kStepThreads / 32;

struct StepLaunch {
  int CH;              // floats per chunk per term (EPT * kSelGroup)
  int NST;             // stages in flight
  int tma_rows;        // 1: full-scorer rows are 16-byte aligned -> bulk copies; 0: plain loads
  int arena_floats;    // union buffer: row stages | candidate rows | selection scratch
  int rows_per_group;  // candidate rows that fit the arena
};

// ---- mbarrier / bulk-copy PTX -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  const uint32_t a = smem_u32(bar);
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(a), "r"(parity)
        : "memory");
  } while (!ok);
}
// global -> shared 1-D TMA copy; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void group_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// monotone map float -> uint32 (for redux.sync max)
__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// (value, index) packed into one sortable key: larger key = better = larger value, then smaller index
// (the total order of every selection, identical to lgasr::better).  +0.0f canonicalises -0.
__device__ __forceinline__ unsigned long long pack_key(float v, int idx) {
  return ((unsigned long long)f2ord(v + 0.0f) << 32) | (unsigned)(~idx);
}
__device__ __forceinline__ float key_val(unsigned long long k) { return ord2f((unsigned)(k >> 32)); }
__device__ __forceinline__ int key_idx(unsigned long long k) { return (int)(~(unsigned)k); }


// ---- group affine scan ------------------------------------------------------------------------------
// y[t] = lse2(y[t-1] + a[t], b(t)) for t in [start, T), y[start-1] = y0 (the recursion of
// ctc_prefix_score.py:155-160) by the kCtcThreads threads of the ctc group.  a, y: shared memory;
// b(t): functor evaluated once per frame by the thread that owns the frame.
// g: thread index in the group; ws: 2*kCtcWarps floats of shared scratch.  Ends with a group barrier.
template <typename BF>
__device__ __forceinline__ void group_affine_scan(const float* a, float* y, BF bfun, int start, int T, float y0, int g,
                                                  float* ws, int bar_id) {
  const int n = T - start;
  if (n <= 0) return;  // group-uniform
  const int lane = g & 31, w = g >> 5;
  int L = (n + kCtcThreads - 1) / kCtcThreads;
  L |= 1;  // odd segment length: the lanes of a warp hit distinct banks
  int s0 = start + g * L;
  int s1 = s0 + L;
  if (s0 > T) s0 = T;
  if (s1 > T) s1 = T;
  // pass 1: local recursion from the scan identity
  float cur = LGASR_NEG, A = 0.f;
  for (int t = s0; t < s1; ++t) {
    const float at = a[t];
    cur = lse2(cur + at, bfun(t));
    y[t] = cur;
    A += at;
  }
  // inclusive scan of the segment summaries inside the warp: (A1,B1) then (A2,B2) = (A1+A2, lse2(B1+A2, B2))
  float SA = A, SB = cur;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const float pa = __shfl_up_sync(FULLMASK, SA, d);
    const float pb = __shfl_up_sync(FULLMASK, SB, d);
    if (lane >= d) {
      SB = lse2(pb + SA, SB);
      SA = pa + SA;
    }
  }
  if (lane == 31) {
    ws[w] = SA;
    ws[kCtcWarps + w] = SB;
  }
  float ea = __shfl_up_sync(FULLMASK, SA, 1);
  float eb = __shfl_up_sync(FULLMASK, SB, 1);
  if (lane == 0) {
    ea = 0.f;
    eb = LGASR_NEG;
  }
  group_bar(bar_id, kCtcThreads);
  // summaries of the preceding warps, then this thread's exclusive prefix inside the warp
  float PA = 0.f, PB = LGASR_NEG;
  for (int i = 0; i < w; ++i) {
    const float wa = ws[i], wb = ws[kCtcWarps + i];
    PB = lse2(PB + wa, wb);
    PA += wa;
  }
  const float XA = PA + ea;
  const float XB = lse2(PB + ea, eb);
  const float yin = (g == 0) ? y0 : lse2(y0 + XA, XB);
  // pass 2: fix-up (independent iterations)
  float run = 0.f;
  for (int t = s0; t < s1; ++t) {
    run += a[t];
    y[t] = lse2(yin + run, y[t]);
  }
  group_bar(bar_id, kCtcThreads);
}

// CTC state of hypothesis (b, h) by the ctc group.
//
// State format of the fused path (handle-internal): row 0 = r^n, row 1 = r_sum = lse2(r^n, r^b).  r^b is
// not stored: r^b[t] = r_sum[t-1] + x_blank[t] for t >= start (exactly how :155-160 produces it), log-zero
// before; for the initial hypothesis r^n is log-zero and r^b == r_sum (:82-93).  That makes the second
// recursion produce r_sum directly (s_t = lse2(s_{t-1} + x_blank[t], r^n_t)) and saves two log-add-exp per frame.
//
// smem rows (Tp floats each): bR0/bR1 = parent r^n / r_sum (TMA), bA = x[:, tok] (TMA), bC = x[:, blank]
// (TMA), bB, bD scratch.  r_out (global) receives the state.  On exit, for u in [0, Tp):
//   bR0[u] = r_sum[u-1]   and   bR1[u] = r^b[u-1]   for pstart <= u < T, LGASR_NEG elsewhere
// (pstart = max(olen, 1) of THIS hypothesis): the phi arrays of the psi phase, shifted by one frame so that
// phi[t-1] + x[t] is an aligned element-wise sum.  *psi_eos = r_sum[T-1].
__device__ __forceinline__ void group_ctc_state(float* bR0, float* bR1, const float* bA, float* bB, const float* bC,
                                                float* bD, float* __restrict__ r_out, int has_parent, int par_initial,
                                                int same, int olen_par, int T, int Tp, int g, float* ws, float* psi_eos,
                                                int bar_id) {
  const int start = olen_par > 1 ? olen_par : 1;           // first frame of the recursion that makes this state
  const int pstart = olen_par + 1 > 1 ? olen_par + 1 : 1;  // max(olen, 1) of this hypothesis
  float rn0 = LGASR_LOGZERO;
  if (has_parent) {
    const int start_par = olen_par - 1 > 1 ? olen_par - 1 : 1;  // the parent's own recursion start
    rn0 = (olen_par == 0) ? bA[0] : LGASR_LOGZERO;               // r[start-1, 0]  (:128-129)
    // r^n: a = x_tok, b[t] = phi_parent[t-1] + x_tok[t]; phi = r_sum, or r^b for a repeated token (:131-140)
    group_affine_scan(
        bA, bB,
        [&](int t) {
          const int u = t - 1;
          float phi = bR1[u];
          if (same && !par_initial) phi = (u >= start_par) ? bR1[u - 1] + bC[u] : LGASR_LOGZERO;
          return phi + bA[t];
        },
        start, T, rn0, g, ws, bar_id);
    // r_sum: a = x_blank, b[t] = r^n[t]   (s_t = lse2(r^b_t, r^n_t), r^b_t = s_{t-1} + x_blank[t])
    group_affine_scan(
        bC, bD, [&](int t) { return bB[t]; }, start, T, rn0, g, ws, bar_id);
  }
  for (int u = g; u <= Tp; u += kCtcThreads) {
    const int t = u - 1;  // state of frame t
    float rn = LGASR_LOGZERO, rs = LGASR_LOGZERO, rb = LGASR_LOGZERO;
    if (t >= 0 && t < T) {
      if (!has_parent) {
        rn = bR0[t];
        rs = bR1[t];
        rb = rs;
      } else if (t >= start) {
        rn = bB[t];
        rs = bD[t];
        rb = ((t - 1 >= start) ? bD[t - 1] : rn0) + bC[t];
      } else if (t == start - 1) {
        rn = rn0;
        rs = rn0;
      }
      r_out[t] = rn;
      r_out[Tp + t] = rs;
    }
    if (t == T - 1) *psi_eos = rs;
    if (u < Tp) {
      const bool in = (u >= pstart && u < T);
      // !has_parent: bR0/bR1[t] are still read by the thread that owns u = t + 1: stage in bB/bD
      float* o0 = has_parent ? bR0 : bB;
      float* o1 = has_parent ? bR1 : bD;
      o0[u] = in ? rs : LGASR_NEG;
      o1[u] = in ? rb : LGASR_NEG;
    }
  }
  if (!has_parent) {
    group_bar(bar_id, kCtcThreads);
    for (int u = g; u < Tp; u += kCtcThreads) {
      bR0[u] = bB[u];
      bR1[u] = bD[u];
    }
  }
}

// ---- shared-memory header of k_step -------------------------------------------------------------------
struct __align__(16) StepSmem {
  unsigned long long bar_full[kMaxStages];  // full-scorer row stages
  unsigned long long bar_ctc;               // parent state + token row + blank row
  unsigned long long bar_cand[kStepWarps];  // candidate rows, one barrier per consuming warp
  unsigned long long keys[2][kTopCap + kPendCap];  // [cur]: sorted top list (n_top) followed by the pending pushes
  int pend_cnt, overflow, n_top, cur_list;
  unsigned long long tau;  // push threshold (key of the n_sel-th best so far; 0 = none)
  float w_eos, psi_eos;
  int last_flag, pad0;
  float ws[2 * kCtcWarps];
  unsigned wkth[kSelWarps];
};

// debug time line: stamp i of group grp (0 ctc / all, 1 sel) of this CTA
#define LGASR_TR(grp, i)                                                                       \
  do {                                                                                          \
    if (d.trace && tid == ((grp) ? kCtcThreads : 0))                                            \
      d.trace[((long long)(blockIdx.y * d.W + blockIdx.x) * 2 + (grp)) * 16 + (i)] = clock64(); \
  } while (0)

constexpr int kBarCtc = 1, kBarSel = 2;  // named barriers of the two groups (0 = __syncthreads)

// issue the TMA copies of chunk c of this hypothesis' full-scorer rows into its stage
__device__ __forceinline__ void issue_row_chunk(const Terms& tm, long long rowoff, int c, int CH, int NST, int V, float* arena,
                                                StepSmem* S) {
  const int stage = c % NST;
  const int n = min(CH, V - c * CH);
  const uint32_t bytes = (uint32_t)n * 4u;
  unsigned long long* bar = &S->bar_full[stage];
  mbar_expect_tx(reinterpret_cast<uint64_t*>(bar), bytes * (uint32_t)tm.nptr);
#pragma unroll 1
  for (int k = 0; k < tm.K; ++k)
    if (tm.ptr[k])
      bulk_g2s(arena + ((long long)stage * tm.nptr + tm.slot[k]) * CH, tm.ptr[k] + rowoff + (long long)c * CH, bytes,
               reinterpret_cast<uint64_t*>(bar));
}

// gather the posterior rows of candidates [g0, g1) of the sorted top list into the arena: lane w < 8 serves
// the barrier of consuming warp w (candidates g0+w, g0+w+8, ...).  Called by one full warp.
__device__ __forceinline__ void issue_cand_rows(const Dev& d, const unsigned long long* keys, int g0, int g1,
                                                const float* xT_b, float* arena, StepSmem* S, int lane) {
  const int Tp = d.Tp;
  const uint32_t rb = (uint32_t)Tp * 4u;
  fence_proxy_async();
  if (lane < kStepWarps) {
    uint32_t bytes = 0;
    for (int i = g0 + lane; i < g1; i += kStepWarps) {
      const int cc = key_idx(keys[i]);
      if (cc != d.eos && !(d.eos != d.blank && cc == d.blank)) bytes += rb;
    }
    uint64_t* bar = reinterpret_cast<uint64_t*>(&S->bar_cand[lane]);
    mbar_expect_tx(bar, bytes);
    for (int i = g0 + lane; i < g1; i += kStepWarps) {
      const int cc = key_idx(keys[i]);
      if (cc != d.eos && !(d.eos != d.blank && cc == d.blank))
        bulk_g2s(arena + (long long)(i - g0) * Tp, xT_b + (long long)cc * Tp, rb, bar);
    }
  }
}

template <int EPT>
__global__ void __launch_bounds__(kStepThreads, 5) k_step(Dev d, Terms tm, StepLaunch sl, int pp, int rp, int step) {
  const int b = blockIdx.y, h = blockIdx.x;
  const int tid = threadIdx.x;
  {
    const int done = d.done[pp][b], nr = d.n_run[pp][b];
    if (done || h >= nr) {
      if (h == 0 && tid == 0 && (done || nr <= 0)) utt_carry_done(d, b, pp);
      return;
    }
  }
  extern __shared__ __align__(16) unsigned char smem_raw[];
  StepSmem* S = reinterpret_cast<StepSmem*>(smem_raw);
  const int Tp = d.Tp, W = d.W, V = d.V;
  float* bR0 = reinterpret_cast<float*>(smem_raw + sizeof(StepSmem));
  float* bR1 = bR0 + Tp;
  float* bA = bR1 + Tp;
  float* bB = bA + Tp;
  float* bC = bB + Tp;
  float* bD = bC + Tp;
  float* arena = bD + Tp;
  constexpr int CH = EPT * kSelGroup;
  const int NST = sl.NST;

  const int lane = tid & 31, warp = tid >> 5;
  const int row = b * W + h;
  const long long rowoff = (long long)row * tm.ld;
  const int n_run = d.n_run[pp][b];
  const int T = d.xlen[b];
  const int cur_len = d.cur_len[pp][b];
  const int olen = cur_len - 1;
  const int4 rm = d.rmeta[pp][row];  // parent slot, last token, parent's last token
  const int last = rm.y;
  const float s_prev = d.use_ctc ? d.s_prev[pp][row] : 0.f;
  const float hyp_score = d.score[pp][row];
  const float* xT_b = d.xT + (long long)b * V * Tp;
  const int mode = d.mode;
  const int n_sel = (mode == kModePreBeam) ? d.S : d.W;

  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < kMaxStages; ++i) mbar_init(reinterpret_cast<uint64_t*>(&S->bar_full[i]), 1);
    mbar_init(reinterpret_cast<uint64_t*>(&S->bar_ctc), 1);
#pragma unroll
    for (int i = 0; i < kStepWarps; ++i) mbar_init(reinterpret_cast<uint64_t*>(&S->bar_cand[i]), 1);
    S->pend_cnt = 0;
    S->overflow = 0;
    S->n_top = 0;
    S->cur_list = 0;
    S->tau = 0ull;
    S->w_eos = 0.f;
    S->psi_eos = 0.f;
    mbar_fence_init();
  }
  __syncthreads();
  LGASR_TR(0, 0);
  LGASR_TR(1, 0);

  if (tid < kCtcThreads) {
    // ======================= ctc group: CTC state of this hypothesis ==================================
    if (d.use_ctc) {
      const int g = tid;
      const int has_parent = rm.x >= 0;
      if (g == 0) {
        const float* r_par = d.r[rp] + ((long long)(b * W + (has_parent ? rm.x : h)) * 2) * Tp;
        const uint32_t rb = (uint32_t)Tp * 4u;
        mbar_expect_tx(reinterpret_cast<uint64_t*>(&S->bar_ctc), has_parent ? 4u * rb : 2u * rb);
        bulk_g2s(bR0, r_par, 2u * rb, reinterpret_cast<uint64_t*>(&S->bar_ctc));
        if (has_parent) {
          bulk_g2s(bA, xT_b + (long long)last * Tp, rb, reinterpret_cast<uint64_t*>(&S->bar_ctc));
          bulk_g2s(bC, xT_b + (long long)d.blank * Tp, rb, reinterpret_cast<uint64_t*>(&S->bar_ctc));
        }
      }
      mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_ctc), 0);
      LGASR_TR(0, 1);
      float* r_out = d.r[rp ^ 1] + ((long long)row * 2) * Tp;
      group_ctc_state(bR0, bR1, bA, bB, bC, bD, r_out, has_parent, step == 1, last == rm.z, olen - 1, T, Tp, g, S->ws,
                      &S->psi_eos, kBarCtc);
      LGASR_TR(0, 2);
    }
  } else if (mode != kModeFullCtc) {
    // ======================= sel group: fused row stream + exact top-n_sel ============================
    const int j = tid - kCtcThreads;
    const int wj = j >> 5;
    const int nchunks = (V + CH - 1) / CH;
    const bool tma = sl.tma_rows != 0;
    if (tma && j == 0) {
      const int pre = nchunks < NST ? nchunks : NST;
      for (int c = 0; c < pre; ++c) issue_row_chunk(tm, rowoff, c, CH, NST, V, arena, S);
    }
    float tau_v = -INFINITY;
    int tau_i = 0x7fffffff;
    for (int c = 0; c < nchunks; ++c) {
      const int stage = c % NST;
      const int e0 = c * CH;
      float val[EPT];
      // ---- values of this thread: elements e0 + 4*(j + kSelGroup*i) + u -----------------------------
      const int n_ch = min(CH, V - e0);
      const float* sb;
      if (tma) {
        mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_full[stage]), (uint32_t)((c / NST) & 1));
        sb = arena + (long long)stage * tm.nptr * CH;
        if (c == 0) LGASR_TR(1, 1);
      } else {
        // rows that are not 16-byte aligned: the group stages the chunk with plain loads
        sb = arena;
#pragma unroll 1
        for (int k = 0; k < tm.K; ++k) {
          const float* src = tm.ptr[k];
          if (!src) continue;
          src += rowoff + e0;
          float* dst = arena + tm.slot[k] * CH;
          for (int i = j; i < n_ch; i += kSelGroup) dst[i] = __ldg(src + i);
        }
        group_bar(kBarSel, kSelGroup);
      }
      {
        float4 acc[EPT / 4];
        if (d.key >= 0) {
          const float* base = sb + tm.slot[d.key] * CH;
#pragma unroll
          for (int i = 0; i < EPT / 4; ++i) acc[i] = *reinterpret_cast<const float4*>(base + 4 * (j + kSelGroup * i));
        } else {
#pragma unroll
          for (int i = 0; i < EPT / 4; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
          for (int k = 0; k < tm.K; ++k) {
            const float wk = tm.w[k];
            if (tm.ptr[k]) {
              const float* base = sb + tm.slot[k] * CH;
#pragma unroll
              for (int i = 0; i < EPT / 4; ++i) {
                const float4 sv = *reinterpret_cast<const float4*>(base + 4 * (j + kSelGroup * i));
                acc[i].x = __fadd_rn(acc[i].x, __fmul_rn(wk, sv.x));
                acc[i].y = __fadd_rn(acc[i].y, __fmul_rn(wk, sv.y));
                acc[i].z = __fadd_rn(acc[i].z, __fmul_rn(wk, sv.z));
                acc[i].w = __fadd_rn(acc[i].w, __fmul_rn(wk, sv.w));
              }
            } else {
              const float cv = __fmul_rn(wk, tm.cst[k]);
#pragma unroll
              for (int i = 0; i < EPT / 4; ++i) {
                acc[i].x = __fadd_rn(acc[i].x, cv);
                acc[i].y = __fadd_rn(acc[i].y, cv);
                acc[i].z = __fadd_rn(acc[i].z, cv);
                acc[i].w = __fadd_rn(acc[i].w, cv);
              }
            }
          }
        }
#pragma unroll
        for (int i = 0; i < EPT / 4; ++i) {
          val[4 * i + 0] = acc[i].x;
          val[4 * i + 1] = acc[i].y;
          val[4 * i + 2] = acc[i].z;
          val[4 * i + 3] = acc[i].w;
        }
      }
      // the fused score of eos is needed even when eos is not a pre-beam candidate (:180-181)
      if (d.key < 0 && d.eos >= e0 && d.eos < e0 + CH) {
        const int q = (d.eos - e0) >> 2;
        if (q % kSelGroup == j) {
          const int ii = 4 * (q / kSelGroup) + (d.eos & 3);
#pragma unroll
          for (int i = 0; i < EPT; ++i)
            if (i == ii) S->w_eos = val[i];
        }
      }
      // elements past the end of the vocabulary never qualify: NaN fails every comparison below
      if (n_ch < CH) {
#pragma unroll
        for (int i = 0; i < EPT; ++i)
          if (4 * (j + kSelGroup * (i >> 2)) + (i & 3) >= n_ch) val[i] = __int_as_float(0x7fc00000);
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < EPT; ++i) mx = fmaxf(mx, val[i]);
      // ---- threshold of the first chunk: at least n_sel elements are >= the k-th largest lane maximum
      //      of every warp (k = ceil(n_sel / kSelWarps)) ------------------------------------------------
      if (c == 0) {
        unsigned cur = (mx > -INFINITY) ? f2ord(mx) : 0u;
        unsigned kth = 0u;
        const int kk = (n_sel + kSelWarps - 1) / kSelWarps;
        for (int r = 0; r < kk; ++r) {
          kth = __reduce_max_sync(FULLMASK, cur);
          const unsigned holders = __ballot_sync(FULLMASK, cur == kth);
          if (lane == __ffs(holders) - 1) cur = 0u;
        }
        if (lane == 0) S->wkth[wj] = kth;
        group_bar(kBarSel, kSelGroup);
        unsigned mn = S->wkth[0];
#pragma unroll
        for (int w2 = 1; w2 < kSelWarps; ++w2) mn = min(mn, S->wkth[w2]);
        tau_v = (mn == 0u) ? -INFINITY : ord2f(mn);
        tau_i = 0x7fffffff;
      }
      // ---- push / merge until no element of the chunk is left behind ---------------------------------
      unsigned pushed = 0u;
      bool first_round = true;
      while (true) {
        // cheap filter first (one compare per thread), the exact (value, index) test only in warps that pass
        if (__any_sync(FULLMASK, mx >= tau_v)) {
          unsigned pmask = 0u;
#pragma unroll
          for (int i = 0; i < EPT; ++i) {
            const int e = e0 + 4 * (j + kSelGroup * (i >> 2)) + (i & 3);
            if (better(val[i], e, tau_v, tau_i)) pmask |= 1u << i;
          }
          pmask &= ~pushed;
          const int ct = __popc(pmask);
          int inc = ct;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(FULLMASK, inc, o);
            if (lane >= o) inc += v;
          }
          const int total = __shfl_sync(FULLMASK, inc, 31);
          if (total > 0) {
            int base = 0;
            if (lane == 31) base = atomicAdd(&S->pend_cnt, total);
            base = __shfl_sync(FULLMASK, base, 31);
            int slot = base + inc - ct;
            unsigned long long* kp = S->keys[S->cur_list] + S->n_top;
#pragma unroll
            for (int i = 0; i < EPT; ++i) {
              if ((pmask >> i) & 1u) {
                if (slot < kPendCap) {
                  kp[slot] = pack_key(val[i], e0 + 4 * (j + kSelGroup * (i >> 2)) + (i & 3));
                  pushed |= 1u << i;
                } else {
                  S->overflow = 1;
                }
                ++slot;
              }
            }
          }
        }
        group_bar(kBarSel, kSelGroup);  // also: every thread has read its stage values
        if (first_round && tma && j == 0 && c + NST < nchunks) {
          fence_proxy_async();
          issue_row_chunk(tm, rowoff, c + NST, CH, NST, V, arena, S);
        }
        first_round = false;
        const int cnt = S->pend_cnt;
        const int ovf = S->overflow;
        if (cnt == 0) {
          group_bar(kBarSel, kSelGroup);  // nobody may push into the next chunk's list before everyone has read cnt
          break;
        }
        const int cl = S->cur_list;
        const int n_l = S->n_top;
        const int n = n_l + (cnt < kPendCap ? cnt : kPendCap);
        const unsigned long long* ks = S->keys[cl];
        unsigned long long* kn = S->keys[cl ^ 1];
        for (int e = j; e < n; e += kSelGroup) {
          const unsigned long long ke = ks[e];
          int rank = 0;
#pragma unroll 4
          for (int o = 0; o < n; ++o) rank += (ks[o] > ke) ? 1 : 0;
          if (rank < n_sel) kn[rank] = ke;
        }
        group_bar(kBarSel, kSelGroup);
        if (j == 0) {
          const int nt = n < n_sel ? n : n_sel;
          S->n_top = nt;
          S->cur_list = cl ^ 1;
          S->pend_cnt = 0;
          S->overflow = 0;
          S->tau = (nt == n_sel) ? kn[n_sel - 1] : 0ull;
        }
        group_bar(kBarSel, kSelGroup);
        {
          const unsigned long long tk2 = S->tau;
          if (tk2) {
            tau_v = key_val(tk2);
            tau_i = key_idx(tk2);
          } else {
            tau_v = -INFINITY;
            tau_i = 0x7fffffff;
          }
        }
        if (!ovf) break;
      }
    }
    LGASR_TR(1, 2);
    // ---- gather the candidates' posterior rows (first group) by TMA ------------------------------------
    if (mode == kModePreBeam && wj == 0) {
      const int nt = S->n_top;
      issue_cand_rows(d, S->keys[S->cur_list], 0, nt < sl.rows_per_group ? nt : sl.rows_per_group, xT_b, arena, S, lane);
    }
  }
  __syncthreads();
  LGASR_TR(0, 3);

  // ======================= all warps: candidate records ===============================================
  const int n_top = S->n_top;
  const unsigned long long* tk = S->keys[S->cur_list];
  const int R = d.R;
  float* c_total = d.cand_total + (long long)row * R;
  int* c_tok = d.cand_tok + (long long)row * R;
  float* c_psi = d.cand_psi + (long long)row * R;

  if (mode == kModeNoCtc) {
    if (lane == 0) {
      for (int jr = warp; jr < d.W; jr += kStepWarps) {
        float total = -INFINITY;
        int c = -1;
        if (jr < n_top) {
          c = key_idx(tk[jr]);
          const float wfull = (d.key < 0) ? key_val(tk[jr]) : fuse_at(tm, rowoff, c);
          total = __fadd_rn(wfull, hyp_score);
        }
        c_total[jr] = total;
        c_tok[jr] = c;
        c_psi[jr] = 0.f;
      }
    }
  } else {
    const float psi_eos = S->psi_eos;  // r_sum at the end frame (ctc_prefix_score.py:180-181)
    const int RPG = sl.rows_per_group;
    const int nq = Tp >> 2;
    constexpr float kL2E = 1.4426950408889634f;
    int grp = 0;
    for (int g0 = 0; g0 < n_top; g0 += RPG, ++grp) {
      const int g1 = (g0 + RPG < n_top) ? g0 + RPG : n_top;
      if (g0 > 0) {
        __syncthreads();  // the arena rows of the previous group have been consumed
        if (warp == 0) issue_cand_rows(d, tk, g0, g1, xT_b, arena, S, lane);
      }
      bool waited = false;
      for (int jr = g0 + warp; jr < g1; jr += kStepWarps) {
        const int c = key_idx(tk[jr]);
        const float wfull = (d.key < 0) ? key_val(tk[jr]) : fuse_at(tm, rowoff, c);
        float psi;
        if (c == d.eos) {
          psi = psi_eos;
        } else if (d.eos != d.blank && c == d.blank) {
          psi = LGASR_LOGZERO;  // :183-185
        } else {
          if (!waited) {
            mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand[warp]), (uint32_t)(grp & 1));
            waited = true;
            if (jr == 0) LGASR_TR(0, 4);
          }
          // log psi = logsumexp({phi[t-1] + x[t] : start <= t < T} U {r^n[start-1]})  (:163-171), the way
          // torch.logsumexp does it (max, sum of exp, log).  phi is pre-shifted and LGASR_NEG outside
          // [start, T), so the sum runs over whole float4 rows.
          const float4* x4 = reinterpret_cast<const float4*>(arena + (long long)(jr - g0) * Tp);
          const float4* p4 = reinterpret_cast<const float4*>((c == last) ? bR1 : bR0);
          const float rn0 = (olen == 0) ? arena[(long long)(jr - g0) * Tp] : LGASR_LOGZERO;
          float m = rn0;
          for (int q = lane; q < nq; q += 32) {
            const float4 xv = x4[q], pv = p4[q];
            m = fmaxf(fmaxf(fmaxf(m, pv.x + xv.x), fmaxf(pv.y + xv.y, pv.z + xv.z)), pv.w + xv.w);
          }
#pragma unroll
          for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULLMASK, m, o));
          // (e - m) first: the subtraction is exact near the maximum, where the terms matter
          float sum = (lane == 0) ? exp2f((rn0 - m) * kL2E) : 0.f;
          for (int q = lane; q < nq; q += 32) {
            const float4 xv = x4[q], pv = p4[q];
            sum += exp2f((pv.x + xv.x - m) * kL2E) + exp2f((pv.y + xv.y - m) * kL2E) +
                   exp2f((pv.z + xv.z - m) * kL2E) + exp2f((pv.w + xv.w - m) * kL2E);
          }
#pragma unroll
          for (int o = 16; o >= 1; o >>= 1) sum += __shfl_xor_sync(FULLMASK, sum, o);
          psi = m + logf(sum);
        }
        if (lane == 0) {
          c_total[jr] = __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi, s_prev))), hyp_score);
          c_tok[jr] = c;
          c_psi[jr] = psi;
        }
      }
      // a warp that had nothing to wait for in this group still has to consume the barrier phase
      if (!waited && (g0 + RPG < n_top)) mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand[warp]), (uint32_t)(grp & 1));
    }
    if (lane == 0) {
      // unused pre-beam slots (fewer than S tokens in the vocabulary)
      for (int jr = n_top + warp; jr < n_sel; jr += kStepWarps) {
        c_total[jr] = -INFINITY;
        c_tok[jr] = -1;
        c_psi[jr] = 0.f;
      }
    }
    if (warp == kStepWarps - 1) {
      // eos is always scored by the CTC scorer (:180-181); skip if it already is a candidate
      bool dup = false;
      for (int e = lane; e < n_top; e += 32) dup |= (key_idx(tk[e]) == d.eos);
      dup = __any_sync(FULLMASK, dup);
      if (lane == 0) {
        float total = -INFINITY;
        if (!dup) {
          const float wfull = (d.key < 0) ? S->w_eos : fuse_at(tm, rowoff, d.eos);
          total = __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi_eos, s_prev))), hyp_score);
        }
        c_total[n_sel] = total;
        c_tok[n_sel] = dup ? -1 : d.eos;
        c_psi[n_sel] = psi_eos;
      }
    }
  }

  // ======================= ticket: the last CTA of the utterance runs the selection ==================
  LGASR_TR(0, 5);
  if (lane == 0) __threadfence();  // only lane 0 of each warp wrote records
  __syncthreads();
  if (tid == 0) {
    const int old = atomicAdd(d.arrive + b, 1);
    S->last_flag = (old == n_run - 1) ? 1 : 0;
  }
  __syncthreads();
  LGASR_TR(0, 6);
  if (!S->last_flag) return;
  if (tid == 0) d.arrive[b] = 0;
  __threadfence();
  utt_select(d, tm, b, pp, step, arena);
  LGASR_TR(0, 7);
}

}  // namespace lgasr
