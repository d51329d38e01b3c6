// k_step: ONE launch per decode position (sm_100a).
//
//   grid (beam, B), 256 threads: one CTA per running hypothesis, warp-specialised.
//
//   warps 0-3  "ctc"   : CTC state of the hypothesis = forward recursion of its parent's state by its
//                         last token (what CTCPrefixScoreTH.__call__ computes for EVERY candidate,
//                         ctc_prefix_score.py:122-160, and select_state then picks, scorers/ctc.py:56-62),
//                         done only for the survivor, as a 128-thread log-semiring affine scan over the
//                         frames.  Inputs (parent r, the token's and the blank's posterior rows) arrive by
//                         1-D TMA bulk copies (cp.async.bulk + mbarrier).
//   warps 4-7  "sel"   : stream the K full-scorer rows (TMA bulk copies, up to 4 stages in flight), fuse
//                         sum_k w_k*s_k in the reference's fp32 order and keep the exact pre-beam top-S
//                         (threshold + rank-by-counting; no (n,V) temporaries, no topk launch;
//                         batch_beam_search.py:279-308).
//   all warps          : the S candidates' posterior rows xT[b][c][:] are gathered by TMA into shared
//                         memory; log psi = logsumexp_t(phi[t-1] + x[t,c]) (+) r^n_{start-1} per warp
//                         (ctc_prefix_score.py:163-185), then w_ctc*(psi - s_prev) + fused + running score
//                         (batch_beam_search.py:313-318) -> candidate records.
//   last CTA of an utterance to finish (atomic ticket): utt_select = global top-beam + bookkeeping.
#pragma once
#include "lgasr_kernels.cuh"

namespace lgasr {

constexpr int kStepThreads = 256;
constexpr int kGroupThreads = 128;  // threads of the ctc group and of the sel group
constexpr int kGroupWarps = 4;
constexpr int kMaxStages = 4;
constexpr int kPendCap = 256;
constexpr int kTopCap = 128;  // == LGASR_MAX_PREBEAM

struct StepLaunch {
  int CH;              // floats per chunk per term (1024 or 2048)
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
__device__ __forceinline__ void group_bar(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(kGroupThreads) : "memory"); }

// monotone map float -> uint32 (for redux.sync max)
__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// ---- 128-thread affine scan -------------------------------------------------------------------------
// y[t] = lse2(y[t-1] + a[t], b[t]) for t in [start, T), y[start-1] = y0 (the recursion of
// ctc_prefix_score.py:155-160).  a, y: shared memory; y holds b[t] on entry, y[t] on exit.
// g: thread index in the group; ws: 2*kGroupWarps floats of shared scratch.  Ends with a group barrier.
__device__ __forceinline__ void group_affine_scan(const float* a, float* y, int start, int T, float y0, int g, float* ws,
                                                  int bar_id) {
  const int n = T - start;
  if (n <= 0) return;  // group-uniform
  const int lane = g & 31, w = g >> 5;
  int L = (n + kGroupThreads - 1) / kGroupThreads;
  L |= 1;  // odd segment length: the lanes of a warp hit distinct banks
  int s0 = start + g * L;
  int s1 = s0 + L;
  if (s0 > T) s0 = T;
  if (s1 > T) s1 = T;
  // pass 1: local recursion from the scan identity
  float cur = LGASR_NEG, A = 0.f;
  for (int t = s0; t < s1; ++t) {
    const float at = a[t];
    cur = lse2(cur + at, y[t]);
    y[t] = cur;
    A += at;
  }
  // inclusive scan of the segment summaries (A, B) inside the warp: (A1,B1) then (A2,B2) = (A1+A2, lse2(B1+A2, B2))
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
    ws[kGroupWarps + w] = SB;
  }
  float ea = __shfl_up_sync(FULLMASK, SA, 1);
  float eb = __shfl_up_sync(FULLMASK, SB, 1);
  if (lane == 0) {
    ea = 0.f;
    eb = LGASR_NEG;
  }
  group_bar(bar_id);
  // summaries of the preceding warps, then this thread's exclusive prefix inside the warp
  float PA = 0.f, PB = LGASR_NEG;
  for (int i = 0; i < w; ++i) {
    const float wa = ws[i], wb = ws[kGroupWarps + i];
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
  group_bar(bar_id);
}

// CTC state of hypothesis (b, h) by the 128 threads of the ctc group.
// smem rows (Tp floats each): bR0/bR1 = parent r^n / r^b (TMA), bA = x[:, tok] (TMA), bC = x[:, blank] (TMA),
// bB, bD scratch.  On exit: bA[t] = r_sum[t] = lse2(r^n, r^b), bD[t] = r^b[t] for t in [0, T); r_out written.
__device__ __forceinline__ void group_ctc_state(const float* bR0, const float* bR1, float* bA, float* bB, float* bC,
                                                float* bD, float* __restrict__ r_out, int has_parent, int same,
                                                int olen_par, int T, int Tp, int g, float* ws, int bar_id) {
  if (!has_parent) {
    for (int t = g; t < T; t += kGroupThreads) {
      const float rn = bR0[t], rb = bR1[t];
      r_out[t] = rn;
      r_out[Tp + t] = rb;
      bA[t] = lse2(rn, rb);
      bD[t] = rb;
    }
    return;
  }
  const int start = olen_par > 1 ? olen_par : 1;
  for (int t = g; t < T; t += kGroupThreads) {
    const float rn = bR0[t], rb = bR1[t];
    bD[t] = same ? rb : lse2(rn, rb);  // phi of the parent (ctc_prefix_score.py:131-140)
  }
  group_bar(bar_id);
  const float rn0 = (olen_par == 0) ? bA[0] : LGASR_LOGZERO;  // r[start-1, 0]  (:128-129)
  for (int t = start + g; t < T; t += kGroupThreads) bB[t] = bD[t - 1] + bA[t];
  group_bar(bar_id);
  group_affine_scan(bA, bB, start, T, rn0, g, ws, bar_id);  // bB = r^n
  for (int t = start + g; t < T; t += kGroupThreads) {
    const float prev = (t - 1 >= start) ? bB[t - 1] : rn0;
    bD[t] = prev + bC[t];
  }
  group_bar(bar_id);
  group_affine_scan(bC, bD, start, T, LGASR_LOGZERO, g, ws, bar_id);  // bD = r^b
  for (int t = g; t < T; t += kGroupThreads) {
    float rn, rb;
    if (t >= start) {
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
}

// ---- shared-memory header of k_step -------------------------------------------------------------------
struct __align__(16) StepSmem {
  unsigned long long bar_full[kMaxStages];  // full-scorer row stages
  unsigned long long bar_ctc;               // parent state + token row + blank row
  unsigned long long bar_cand;              // candidate rows
  int pend_cnt, overflow, n_top, cur_list;
  float tau_v;
  int tau_i;
  float w_eos;
  int last_flag;
  float ws[2 * kGroupWarps];
  unsigned wkth[kGroupWarps];
  float top_val[2][kTopCap];
  int top_idx[2][kTopCap];
  float pend_val[kPendCap];
  int pend_idx[kPendCap];
};

// debug time line: stamp i of group grp (0 ctc / all, 1 sel) of this CTA
#define LGASR_TR(grp, i)                                                                      \
  do {                                                                                         \
    if (d.trace && (tid & (kGroupThreads - 1)) == 0 && (tid >> 7) == (grp))                    \
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

template <int EPT>
__global__ void __launch_bounds__(kStepThreads) k_step(Dev d, Terms tm, StepLaunch sl, int pp, int rp, int step) {
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
  constexpr int CH = EPT * kGroupThreads;
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
    mbar_init(reinterpret_cast<uint64_t*>(&S->bar_cand), 1);
    S->pend_cnt = 0;
    S->overflow = 0;
    S->n_top = 0;
    S->cur_list = 0;
    S->tau_v = -INFINITY;
    S->tau_i = 0x7fffffff;
    S->w_eos = 0.f;
    mbar_fence_init();
  }
  __syncthreads();
  LGASR_TR(0, 0);
  LGASR_TR(1, 0);

  if (tid < kGroupThreads) {
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
      group_ctc_state(bR0, bR1, bA, bB, bC, bD, r_out, has_parent, last == rm.z, olen - 1, T, Tp, g, S->ws, kBarCtc);
      LGASR_TR(0, 2);
    }
  } else if (mode != kModeFullCtc) {
    // ======================= sel group: fused row stream + exact top-n_sel ============================
    const int j = tid - kGroupThreads;
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
      // ---- values of this thread: elements e0 + 4*(j + 128*i) + u ---------------------------------
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
          for (int i = j; i < n_ch; i += kGroupThreads) dst[i] = __ldg(src + i);
        }
        group_bar(kBarSel);
      }
      {
        float4 acc[EPT / 4];
        if (d.key >= 0) {
          const float* base = sb + tm.slot[d.key] * CH;
#pragma unroll
          for (int i = 0; i < EPT / 4; ++i) acc[i] = *reinterpret_cast<const float4*>(base + 4 * (j + kGroupThreads * i));
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
                const float4 sv = *reinterpret_cast<const float4*>(base + 4 * (j + kGroupThreads * i));
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
      if (d.key < 0) {
#pragma unroll
        for (int i = 0; i < EPT; ++i)
          if (e0 + 4 * (j + kGroupThreads * (i >> 2)) + (i & 3) == d.eos) S->w_eos = val[i];
      }
      // ---- threshold of the first chunk: at least n_sel elements are >= the k-th largest lane maximum
      //      of every warp (k = ceil(n_sel / 4)) --------------------------------------------------------
      if (c == 0) {
        float mx = -INFINITY;
        bool any = false;
#pragma unroll
        for (int i = 0; i < EPT; ++i)
          if (e0 + 4 * (j + kGroupThreads * (i >> 2)) + (i & 3) < V) {
            mx = fmaxf(mx, val[i]);
            any = true;
          }
        unsigned cur = any ? f2ord(mx) : 0u;
        unsigned kth = 0u;
        const int kk = (n_sel + kGroupWarps - 1) / kGroupWarps;
        for (int r = 0; r < kk; ++r) {
          kth = __reduce_max_sync(FULLMASK, cur);
          const unsigned holders = __ballot_sync(FULLMASK, cur == kth);
          if (lane == __ffs(holders) - 1) cur = 0u;
        }
        if (lane == 0) S->wkth[wj] = kth;
        group_bar(kBarSel);
        unsigned mn = S->wkth[0];
#pragma unroll
        for (int w2 = 1; w2 < kGroupWarps; ++w2) mn = min(mn, S->wkth[w2]);
        tau_v = (mn == 0u) ? -INFINITY : ord2f(mn);
        tau_i = 0x7fffffff;
      }
      // ---- push / merge until no element of the chunk is left behind ---------------------------------
      unsigned pushed = 0u;
      bool first_round = true;
      while (true) {
#pragma unroll
        for (int i = 0; i < EPT; ++i) {
          const int e = e0 + 4 * (j + kGroupThreads * (i >> 2)) + (i & 3);
          if (e < V && !((pushed >> i) & 1u) && better(val[i], e, tau_v, tau_i)) {
            const int slot = atomicAdd(&S->pend_cnt, 1);
            if (slot < kPendCap) {
              S->pend_val[slot] = val[i];
              S->pend_idx[slot] = e;
              pushed |= 1u << i;
            } else {
              S->overflow = 1;
            }
          }
        }
        group_bar(kBarSel);  // also: every thread has read its stage values
        if (first_round && tma && j == 0 && c + NST < nchunks) {
          fence_proxy_async();
          issue_row_chunk(tm, rowoff, c + NST, CH, NST, V, arena, S);
        }
        first_round = false;
        const int cnt = S->pend_cnt;
        const int ovf = S->overflow;
        if (cnt == 0) {
          group_bar(kBarSel);  // nobody may push into the next chunk's list before everyone has read cnt
          break;
        }
        const int cl = S->cur_list;
        const int n_l = S->n_top;
        const int n_p = cnt < kPendCap ? cnt : kPendCap;
        const int n = n_l + n_p;
        const float* lv = S->top_val[cl];
        const int* li = S->top_idx[cl];
        float* nv = S->top_val[cl ^ 1];
        int* ni = S->top_idx[cl ^ 1];
        for (int e = j; e < n; e += kGroupThreads) {
          const float ve = e < n_l ? lv[e] : S->pend_val[e - n_l];
          const int ie = e < n_l ? li[e] : S->pend_idx[e - n_l];
          int rank = 0;
          for (int o = 0; o < n_l; ++o) rank += better(lv[o], li[o], ve, ie) ? 1 : 0;
          for (int o = 0; o < n_p; ++o) rank += better(S->pend_val[o], S->pend_idx[o], ve, ie) ? 1 : 0;
          if (rank < n_sel) {
            nv[rank] = ve;
            ni[rank] = ie;
          }
        }
        group_bar(kBarSel);
        if (j == 0) {
          const int nt = n < n_sel ? n : n_sel;
          S->n_top = nt;
          S->cur_list = cl ^ 1;
          S->pend_cnt = 0;
          S->overflow = 0;
          if (nt == n_sel) {
            S->tau_v = nv[n_sel - 1];
            S->tau_i = ni[n_sel - 1];
          } else {
            S->tau_v = -INFINITY;
            S->tau_i = 0x7fffffff;
          }
        }
        group_bar(kBarSel);
        tau_v = S->tau_v;
        tau_i = S->tau_i;
        if (!ovf) break;
      }
    }
    LGASR_TR(1, 2);
    // ---- gather the candidates' posterior rows (first group) by TMA ------------------------------------
    if (mode == kModePreBeam && j == 0) {
      const int cl = S->cur_list;
      const int nt = S->n_top;
      const int g1 = nt < sl.rows_per_group ? nt : sl.rows_per_group;
      const uint32_t rb = (uint32_t)Tp * 4u;
      uint32_t bytes = 0;
      for (int i = 0; i < g1; ++i) {
        const int cc = S->top_idx[cl][i];
        if (cc != d.eos && !(d.eos != d.blank && cc == d.blank)) bytes += rb;
      }
      fence_proxy_async();
      mbar_expect_tx(reinterpret_cast<uint64_t*>(&S->bar_cand), bytes);
      for (int i = 0; i < g1; ++i) {
        const int cc = S->top_idx[cl][i];
        if (cc != d.eos && !(d.eos != d.blank && cc == d.blank))
          bulk_g2s(arena + (long long)i * Tp, xT_b + (long long)cc * Tp, rb, reinterpret_cast<uint64_t*>(&S->bar_cand));
      }
    }
  }
  __syncthreads();
  LGASR_TR(0, 3);

  // ======================= all warps: candidate records ===============================================
  const int cl = S->cur_list;
  const int n_top = S->n_top;
  const float* tv = S->top_val[cl];
  const int* ti = S->top_idx[cl];
  const int R = d.R;
  float* c_total = d.cand_total + (long long)row * R;
  int* c_tok = d.cand_tok + (long long)row * R;
  float* c_psi = d.cand_psi + (long long)row * R;
  constexpr int kWarps = kStepThreads / 32;

  if (mode == kModeNoCtc) {
    for (int jr = tid; jr < d.W; jr += kStepThreads) {
      float total = -INFINITY;
      int c = -1;
      if (jr < n_top) {
        c = ti[jr];
        const float wfull = (d.key < 0) ? tv[jr] : fuse_at(tm, rowoff, c);
        total = __fadd_rn(wfull, hyp_score);
      }
      c_total[jr] = total;
      c_tok[jr] = c;
      c_psi[jr] = 0.f;
    }
  } else {
    const int start = olen > 1 ? olen : 1;
    const float psi_eos = bA[T - 1];  // r_sum at the end frame (ctc_prefix_score.py:180-181)
    const int RPG = sl.rows_per_group;
    const uint32_t rb = (uint32_t)Tp * 4u;
    int grp = 0;
    for (int g0 = 0; g0 < n_top; g0 += RPG, ++grp) {
      const int g1 = (g0 + RPG < n_top) ? g0 + RPG : n_top;
      if (g0 > 0) {
        __syncthreads();  // the arena rows of the previous group have been consumed
        if (tid == 0) {
          uint32_t bytes = 0;
          for (int i = g0; i < g1; ++i) {
            const int cc = ti[i];
            if (cc != d.eos && !(d.eos != d.blank && cc == d.blank)) bytes += rb;
          }
          fence_proxy_async();
          mbar_expect_tx(reinterpret_cast<uint64_t*>(&S->bar_cand), bytes);
          for (int i = g0; i < g1; ++i) {
            const int cc = ti[i];
            if (cc != d.eos && !(d.eos != d.blank && cc == d.blank))
              bulk_g2s(arena + (long long)(i - g0) * Tp, xT_b + (long long)cc * Tp, rb,
                       reinterpret_cast<uint64_t*>(&S->bar_cand));
          }
        }
      }
      mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand), (uint32_t)(grp & 1));
      LGASR_TR(0, 4);
      for (int jr = g0 + warp; jr < g1; jr += kWarps) {
        const int c = ti[jr];
        const float wfull = (d.key < 0) ? tv[jr] : fuse_at(tm, rowoff, c);
        float psi;
        if (c == d.eos) {
          psi = psi_eos;
        } else if (d.eos != d.blank && c == d.blank) {
          psi = LGASR_LOGZERO;  // :183-185
        } else {
          // log psi = logsumexp({phi[t-1] + x[t] : start <= t < T} U {r^n[start-1]})  (:163-171), the way
          // torch.logsumexp does it: max, sum of exp, log
          const float* x = arena + (long long)(jr - g0) * Tp;
          const float* phi = (c == last) ? bD : bA;
          const float rn0 = (olen == 0) ? x[0] : LGASR_LOGZERO;
          float m = rn0;
          for (int t = start + lane; t < T; t += 32) m = fmaxf(m, phi[t - 1] + x[t]);
#pragma unroll
          for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULLMASK, m, o));
          float sum = (lane == 0) ? expf(rn0 - m) : 0.f;
          for (int t = start + lane; t < T; t += 32) sum += expf(phi[t - 1] + x[t] - m);
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
    }
    // unused pre-beam slots (fewer than S tokens in the vocabulary) and the eos slot
    for (int jr = n_top + tid; jr < n_sel; jr += kStepThreads) {
      c_total[jr] = -INFINITY;
      c_tok[jr] = -1;
      c_psi[jr] = 0.f;
    }
    if (warp == kWarps - 1) {
      // eos is always scored by the CTC scorer (:180-181); skip if it already is a candidate
      bool dup = false;
      for (int e = lane; e < n_top; e += 32) dup |= (ti[e] == d.eos);
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
  __threadfence();
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
