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
//   record exchange: through rank 0's shared memory when the `beam` CTAs of an utterance run as one
//                         thread-block cluster (CL), else through L2 with an atomic ticket; the receiving CTA
//                         runs utt_select = global top-beam + bookkeeping.
//
//   Instances k_step<TR, CL, SEQ>: TR adds clock64/globaltimer stamps (LGASR_TRACE=1); SEQ is the instance of
//   programmatic dependent launches inside a replay: the sel work runs on all eight warps BEFORE
//   griddepcontrol.wait (it depends only on inputs of the step), the ctc work on four warps after it.
#pragma once
#include "lgasr_kernels.cuh"

namespace lgasr {

constexpr int kStepThreads = 256;
constexpr int kCtcWarps = 2, kCtcThreads = 32 * kCtcWarps;    // warps 0-1
constexpr int kSelWarps = 6, kSelGroup = 32 * kSelWarps;       // warps 2-7
constexpr int kMaxStages = 2;
constexpr int kPendCap = 256;
constexpr int kTopCap = 128;  // == LGASR_MAX_PREBEAM
constexpr int kStepWarps = kStepThreads / 32;

struct StepLaunch {
  int CH;              // floats per chunk per term (multiple of 4)
  int NST;             // stages in flight (1 or 2)
  int nchunks;         // ceil(V / CH)
  int kk_full;         // ceil(n_sel / kSelWarps): lane-maximum rounds of the chunk threshold
  int flags;           // replay with LGASR_FLAGS=1: publish step_flag[b] after the selection; SEQ launches acquire it
                       // instead of griddepcontrol.wait (utterances advance at their own pace)
  int pdl;             // launched with programmatic stream serialization: rows are requested before griddepcontrol.wait
  int tma_rows;        // 1: full-scorer rows are 16-byte aligned -> bulk copies; 0: plain loads
  int arena_floats;    // union buffer: row stages | candidate rows | selection scratch
  int rows_per_group;  // candidate rows that fit the arena
  int row_path;        // k_step2: 1 = one fp32 matrix (or the key term), 2 = two fp32 matrices, 0 = general term loop
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
// y[t] = lse2(y[t-1] + a[t], b[t]) for t in [start, T), y[start-1] = y0 (the recursion of
// ctc_prefix_score.py:155-160) by the kCtcThreads threads of the ctc group.  a, b, y: shared memory (b may
// alias y).  g: thread index in the group; ws: 2*kCtcWarps floats of shared scratch.  Ends with a group
// barrier.  Deliberately ONE out-of-line copy with rolled loops: the step kernel must stay small enough for
// the instruction cache (straight-line code runs at fetch speed, not at issue speed).
template <int NCW>
__device__ __noinline__ void group_affine_scan(const float* a, const float* b, float* y, int start, int T, float y0, int g,
                                               float* ws, int bar_id) {
  constexpr int NCT = 32 * NCW;
  const int n = T - start;
  if (n <= 0) return;  // group-uniform
  const int lane = g & 31, w = g >> 5;
  int L = (n + NCT - 1) / NCT;
  L |= 1;  // odd segment length: the lanes of a warp hit distinct banks
  int s0 = start + g * L;
  int s1 = s0 + L;
  if (s0 > T) s0 = T;
  if (s1 > T) s1 = T;
  // pass 1: local recursion from the scan identity
  float cur = LGASR_NEG, A = 0.f;
#pragma unroll 1
  for (int t = s0; t < s1; ++t) {
    const float at = a[t];
    cur = lse2(cur + at, b[t]);
    y[t] = cur;
    A += at;
  }
  // inclusive scan of the segment summaries inside the warp: (A1,B1) then (A2,B2) = (A1+A2, lse2(B1+A2, B2))
  float SA = A, SB = cur;
#pragma unroll 1
  for (int dd = 1; dd < 32; dd <<= 1) {
    const float pa = __shfl_up_sync(FULLMASK, SA, dd);
    const float pb = __shfl_up_sync(FULLMASK, SB, dd);
    if (lane >= dd) {
      SB = lse2(pb + SA, SB);
      SA = pa + SA;
    }
  }
  if (lane == 31) {
    ws[w] = SA;
    ws[NCW + w] = SB;
  }
  float ea = __shfl_up_sync(FULLMASK, SA, 1);
  float eb = __shfl_up_sync(FULLMASK, SB, 1);
  if (lane == 0) {
    ea = 0.f;
    eb = LGASR_NEG;
  }
  group_bar(bar_id, NCT);
  // summaries of the preceding warps followed by this thread's exclusive prefix inside the warp
  float XA = 0.f, XB = LGASR_NEG;
#pragma unroll 1
  for (int i = 0; i <= w; ++i) {
    const float wa = (i < w) ? ws[i] : ea, wb = (i < w) ? ws[NCW + i] : eb;
    XB = lse2(XB + wa, wb);
    XA += wa;
  }
  const float yin = (g == 0) ? y0 : lse2(y0 + XA, XB);
  // pass 2: fix-up (independent iterations)
  float run = 0.f;
#pragma unroll 1
  for (int t = s0; t < s1; ++t) {
    run += a[t];
    y[t] = lse2(yin + run, y[t]);
  }
  group_bar(bar_id, NCT);
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
template <int NCW>
__device__ __forceinline__ void group_ctc_scans(const float* bR1, const float* bA, float* bB, const float* bC, float* bD,
                                                int has_parent, int par_initial, int same, int olen_par, int T, int g,
                                                float* ws, int bar_id, long long* tr) {
  const int start = olen_par > 1 ? olen_par : 1;           // first frame of the recursion that makes this state
  float rn0 = LGASR_LOGZERO;
  if (has_parent) {
    const int start_par = olen_par - 1 > 1 ? olen_par - 1 : 1;  // the parent's own recursion start
    rn0 = (olen_par == 0) ? bA[0] : LGASR_LOGZERO;               // r[start-1, 0]  (:128-129)
    // r^n: a = x_tok, b[t] = phi_parent[t-1] + x_tok[t]; phi = r_sum, or r^b for a repeated token (:131-140)
#pragma unroll 1
    for (int t = start + g; t < T; t += 32 * NCW) {
      const int u = t - 1;
      float phi = bR1[u];
      if (same && !par_initial) phi = (u >= start_par) ? bR1[u - 1] + bC[u] : LGASR_LOGZERO;
      bB[t] = phi + bA[t];
    }
    group_bar(bar_id, 32 * NCW);
    group_affine_scan<NCW>(bA, bB, bB, start, T, rn0, g, ws, bar_id);
    if (tr && g == 0) tr[8] = clock64();
    // r_sum: a = x_blank, b[t] = r^n[t]   (s_t = lse2(r^b_t, r^n_t), r^b_t = s_{t-1} + x_blank[t])
    group_affine_scan<NCW>(bC, bB, bD, start, T, rn0, g, ws, bar_id);
    if (tr && g == 0) tr[9] = clock64();
  }
}

// second half, by ALL threads of the CTA after the groups have joined: write the state to global memory and
// lay out the shifted phi arrays of the psi phase (see above).  Ends with a CTA barrier.
__device__ __forceinline__ void cta_ctc_epilogue(float* bR0, float* bR1, const float* bA, float* bB, const float* bC,
                                                 float* bD, float* __restrict__ r_out, int has_parent, int olen_par, int T,
                                                 int Tp, int tid, float* psi_eos) {
  const int start = olen_par > 1 ? olen_par : 1;
  const int pstart = olen_par + 1 > 1 ? olen_par + 1 : 1;  // max(olen, 1) of this hypothesis
  const float rn0 = (has_parent && olen_par == 0) ? bA[0] : LGASR_LOGZERO;
#pragma unroll 1
  for (int u = tid; u <= Tp; u += kStepThreads) {
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
  __syncthreads();
  if (!has_parent) {
    for (int u = tid; u < Tp; u += kStepThreads) {
      bR0[u] = bB[u];
      bR1[u] = bD[u];
    }
    __syncthreads();
  }
}

// ---- shared-memory header of k_step -------------------------------------------------------------------
struct __align__(16) StepSmem {
  unsigned long long keys[2][kTopCap + kPendCap];  // [cur]: sorted top list (n_top) followed by the pending pushes
  unsigned long long bar_full[kMaxStages];  // full-scorer row stages
  unsigned long long bar_ctc;               // parent state + token row + blank row
  unsigned long long bar_cand[kStepWarps];  // candidate rows, one barrier per consuming warp
  int pend_cnt[2], overflow[2];             // double-buffered by merge iteration (see the sel group)
  int n_top, cur_list;
  float w_eos, psi_eos;
  int last_flag, pad0;
  float ws[2 * kStepWarps];
  unsigned wkth[kStepWarps];
};

// debug time line: stamp i of group grp (0 ctc / all, 1 sel) of this CTA
#define LGASR_TR(grp, i)                                                                       \
  do {                                                                                          \
    if (TR && tid == ((grp) ? SEL0 : 0))                                                 \
      d.trace[((long long)(blockIdx.y * d.W + blockIdx.x) * 2 + (grp)) * 16 + (i)] = clock64(); \
  } while (0)

__device__ __forceinline__ long long gtimer() {
  long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define LGASR_GT(i)                                                                                     \
  do {                                                                                                  \
    if (TR && tid == 0) d.trace[((long long)(blockIdx.y * d.W + blockIdx.x) * 2 + (step & 1)) * 16 + (i)] = gtimer(); \
  } while (0)

// ---- thread-block cluster helpers (one cluster = the `beam` CTAs of an utterance) -------------------------
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// shared::cluster address of `p` (an address in this CTA's shared memory) in the CTA of rank `rank`
__device__ __forceinline__ uint32_t map_rank(const void* p, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster(uint32_t addr, unsigned long long v) {
  asm volatile("st.shared::cluster.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void st_cluster(uint32_t addr, unsigned v) {
  asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// per-utterance hand-off between consecutive positions (SEQ launches): the selection of position s publishes
// step_flag[b] = s + 1 after its last write; the CTAs of position s + 1 acquire it instead of waiting for the
// whole previous grid (griddepcontrol.wait), so utterances advance at their own pace
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

constexpr int kBarCtc = 1, kBarSel = 2;  // named barriers of the two groups (0 = __syncthreads)

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// issue the TMA copies of chunk c of this hypothesis' full-scorer rows into its stage
__device__ __forceinline__ void issue_row_chunk(const Terms& tm, long long rowoff, int c, int CH, int NST, int V, float* arena,
                                                StepSmem* S) {
  const int stage = c & (NST - 1);
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
  if (g1 - g0 <= 2 * kStepWarps) {
    // one candidate per lane: lanes w and w + 8 feed the barrier of consuming warp w
    const int i = g0 + lane;
    int cc = -1;
    if (lane < 2 * kStepWarps && i < g1) cc = key_idx(keys[i]);
    const bool need = cc >= 0 && cc != d.eos && !(d.eos != d.blank && cc == d.blank);
    const uint32_t mine = need ? rb : 0u;
    const uint32_t both = mine + __shfl_xor_sync(FULLMASK, mine, kStepWarps);
    uint64_t* bar = reinterpret_cast<uint64_t*>(&S->bar_cand[lane & (kStepWarps - 1)]);
    if (lane < kStepWarps) mbar_expect_tx(bar, both);
    __syncwarp();
    if (need) bulk_g2s(arena + (long long)(i - g0) * Tp, xT_b + (long long)cc * Tp, rb, bar);
    return;
  }
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

// fused value of the 4 consecutive vocabulary entries at float offset 4*q of the staged chunk `sb`
// ([nptr][CH] floats): sum_k w_k * s_k in the reference's fp32 order (batch_beam_search.py:279-300), or the
// raw scores of the pre-beam key term (beam_search.py:302-308 with pre_beam_score_key != "full").
__device__ __forceinline__ float4 fused4(const Terms& tm, const float* sb, int CH, int key, int q) {
  if (key >= 0) return *reinterpret_cast<const float4*>(sb + tm.slot[key] * CH + 4 * q);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
  for (int k = 0; k < tm.K; ++k) {
    const float wk = tm.w[k];
    if (tm.ptr[k]) {
      const float4 sv = *reinterpret_cast<const float4*>(sb + tm.slot[k] * CH + 4 * q);
      acc.x = __fadd_rn(acc.x, __fmul_rn(wk, sv.x));
      acc.y = __fadd_rn(acc.y, __fmul_rn(wk, sv.y));
      acc.z = __fadd_rn(acc.z, __fmul_rn(wk, sv.z));
      acc.w = __fadd_rn(acc.w, __fmul_rn(wk, sv.w));
    } else {
      const float cv = __fmul_rn(wk, tm.cst[k]);
      acc.x = __fadd_rn(acc.x, cv);
      acc.y = __fadd_rn(acc.y, cv);
      acc.z = __fadd_rn(acc.z, cv);
      acc.w = __fadd_rn(acc.w, cv);
    }
  }
  return acc;
}

// CL: launched as clusters of `beam` CTAs (cluster rank == hypothesis slot): the candidate records and the
// parents' fields travel to the shared memory of the rank-0 CTA (distributed shared memory) and a cluster
// barrier replaces the fence + atomic ticket and the exchange through L2; rank 0 runs the selection.
// SEQ (launches with programmatic stream serialization, i.e. the positions after the first of a replay): the
// row stream + top-S selection depends only on inputs of the step, so ALL eight warps run it BEFORE
// griddepcontrol.wait, overlapped with the previous position's selection tail; after the wait the CTC scans get
// four warps instead of two.  !SEQ: the two groups run side by side after the wait (2 + 6 warps).
template <bool TR, bool CL, bool SEQ>
__global__ void __launch_bounds__(kStepThreads, 5) k_step(Dev d, Terms tm, StepLaunch sl, int pp, int rp, int step) {
  constexpr int NCW = SEQ ? 4 : kCtcWarps, NCT = 32 * NCW;      // ctc group: warps [0, NCW)
  constexpr int NSW = SEQ ? kStepWarps : kSelWarps, NSG = 32 * NSW;  // sel group: threads [SEL0, SEL0 + NSG)
  constexpr int SEL0 = SEQ ? 0 : kCtcThreads;
  const int b = blockIdx.y, h = blockIdx.x;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int W = d.W, V = d.V, Tp = d.Tp;
  const int row = b * W + h;
  const long long rowoff = (long long)row * tm.ld;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  StepSmem* S = reinterpret_cast<StepSmem*>(smem_raw);
  float* bR0 = reinterpret_cast<float*>(smem_raw + sizeof(StepSmem));
  float* bR1 = bR0 + Tp;
  float* bA = bR1 + Tp;
  float* bB = bA + Tp;
  float* bC = bB + Tp;
  float* bD = bC + Tp;
  float* arena = bD + Tp;
  const int CH = sl.CH, NST = sl.NST;
  const int nchunks = sl.nchunks;
  const bool tma = sl.tma_rows != 0;
  const int mode = d.mode;
  const float* xT_b = d.xT + (long long)b * V * Tp;
  const int n_sel = (mode == kModePreBeam) ? d.S : d.W;
  // ---- programmatic dependent launch: everything up to griddepcontrol.wait touches only kernel parameters,
  // this CTA's shared memory and inputs of the step (the full-scorer matrices and the posteriors, never written
  // by a k_step), so under PDL it overlaps the previous position's selection tail: the next grid's CTAs take the
  // slots of CTAs that have left.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  LGASR_GT(13);  // kernel entry (before griddepcontrol.wait)
  if (CL) cluster_arrive();  // phase 1: "this CTA runs"; waited for before the first remote store
  if (tid == SEL0) {
#pragma unroll
    for (int i = 0; i < kMaxStages; ++i) mbar_init(reinterpret_cast<uint64_t*>(&S->bar_full[i]), 1);
#pragma unroll
    for (int i = 0; i < kStepWarps; ++i) mbar_init(reinterpret_cast<uint64_t*>(&S->bar_cand[i]), 1);
    S->pend_cnt[0] = S->pend_cnt[1] = 0;
    S->overflow[0] = S->overflow[1] = 0;
    S->n_top = 0;
    S->cur_list = 0;
    S->w_eos = 0.f;
    S->psi_eos = 0.f;
    mbar_fence_init();
    if (sl.pdl && tma && mode != kModeFullCtc) {
      const int pre = nchunks < NST ? nchunks : NST;
      for (int c = 0; c < pre; ++c) issue_row_chunk(tm, rowoff, c, CH, NST, V, arena, S);
    }
  }

  // ======================= sel group: fused row stream + exact top-n_sel ================================
  auto sel_phase = [&]() {
    // ======================= sel group: fused row stream + exact top-n_sel ============================
    // Two sweeps over each staged chunk: lane maxima -> threshold (only while the list is not full), then
    // push everything above the threshold into a pending list (warp-aggregated shared atomics) and merge
    // by rank counting on packed (value, ~index) keys.  A pending-list overflow (degenerate rows, e.g. all
    // scores equal) is handled by extra rounds over the same chunk with the tightened threshold.
    const int j = tid - SEL0;
    const int wj = j >> 5;
    const int key = d.key;
    const int srows = tm.nptr > 0 ? tm.nptr : 1;
    int cur = 0, n_top = 0, it = 0;
    unsigned long long tau = 0ull;  // key of the n_sel-th best so far (0 = list not full)
    for (int c = 0; c < nchunks; ++c) {
      const int stage = tma ? (c & (NST - 1)) : 0;  // NST is 1 or 2
      const int e0 = c * CH;
      const int n_ch = min(CH, V - e0);
      const int nq = (n_ch + 3) >> 2;
      float* sb = arena + (long long)stage * srows * CH;
      if (tma) {
        mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_full[stage]), (uint32_t)((c >> (NST - 1)) & 1));
      } else {
        // rows that are not 16-byte aligned: the group stages the chunk with plain loads
#pragma unroll 1
        for (int k = 0; k < tm.K; ++k) {
          const float* src = tm.ptr[k];
          if (!src) continue;
          src += rowoff + e0;
          float* dst = sb + tm.slot[k] * CH;
          for (int i = j; i < n_ch; i += NSG) dst[i] = __ldg(src + i);
        }
        group_bar(kBarSel, NSG);
      }
      if (c == 0) LGASR_TR(1, 1);
      // ---- sweep 1: fuse the K terms in place (term-major, so the fp32 order is the reference's
      //      ((0 + w0*s0) + w1*s1) + ..., batch_beam_search.py:279-300) and take the lane maxima.
      //      Thread j owns float4s j, j+128, ...: it only ever overwrites what it alone reads. ------------
      float4* fz = reinterpret_cast<float4*>(key >= 0 ? sb + tm.slot[key] * CH : sb);
      float mx = -INFINITY;
      const bool one_term = (key < 0) && (tm.K == 1) && (tm.nptr == 1) && ((n_ch & 3) == 0);
      if (one_term) {  // the common shape (decoder only): w0 * s0, no tail
        const float w0 = tm.w[0];
#pragma unroll 2
        for (int q = j; q < nq; q += NSG) {
          float4 v = fz[q];
          v.x = __fmul_rn(w0, v.x);  // 0 + w0*s0 == w0*s0 (up to the sign of zero)
          v.y = __fmul_rn(w0, v.y);  // 0 + w0*s0 == w0*s0 (up to the sign of zero)
          v.z = __fmul_rn(w0, v.z);  // 0 + w0*s0 == w0*s0 (up to the sign of zero)
          v.w = __fmul_rn(w0, v.w);  // 0 + w0*s0 == w0*s0 (up to the sign of zero)
          fz[q] = v;
          mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
        }
      } else if ((key < 0) && (tm.K == 2) && (tm.nptr == 2) && ((n_ch & 3) == 0)) {
        // two matrices (decoder + LM/LLM): (0 + w0*s0) + w1*s1, the first sum being exact
        const float w0 = tm.w[0], w1 = tm.w[1];
        const float4* f1 = reinterpret_cast<const float4*>(sb + tm.slot[1] * CH);
        const float4* f0 = reinterpret_cast<const float4*>(sb + tm.slot[0] * CH);
#pragma unroll 2
        for (int q = j; q < nq; q += NSG) {
          const float4 a = f0[q], b2 = f1[q];
          float4 v;
          v.x = __fadd_rn(__fmul_rn(w0, a.x), __fmul_rn(w1, b2.x));
          v.y = __fadd_rn(__fmul_rn(w0, a.y), __fmul_rn(w1, b2.y));
          v.z = __fadd_rn(__fmul_rn(w0, a.z), __fmul_rn(w1, b2.z));
          v.w = __fadd_rn(__fmul_rn(w0, a.w), __fmul_rn(w1, b2.w));
          fz[q] = v;
          mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
        }
      } else
#pragma unroll 1
      for (int q = j; q < nq; q += NSG) {
        float4 acc;
        if (key >= 0) {
          acc = fz[q];
        } else {
          acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
          for (int k = 0; k < tm.K; ++k) {
            const float wk = tm.w[k];
            if (tm.ptr[k]) {
              const float4 sv = reinterpret_cast<const float4*>(sb + tm.slot[k] * CH)[q];
              acc.x = __fadd_rn(acc.x, __fmul_rn(wk, sv.x));
              acc.y = __fadd_rn(acc.y, __fmul_rn(wk, sv.y));
              acc.z = __fadd_rn(acc.z, __fmul_rn(wk, sv.z));
              acc.w = __fadd_rn(acc.w, __fmul_rn(wk, sv.w));
            } else {
              const float cv = __fmul_rn(wk, tm.cst[k]);
              acc.x = __fadd_rn(acc.x, cv);
              acc.y = __fadd_rn(acc.y, cv);
              acc.z = __fadd_rn(acc.z, cv);
              acc.w = __fadd_rn(acc.w, cv);
            }
          }
        }
        if (4 * q + 3 >= n_ch) {  // tail of a vocabulary that is not a multiple of 4
          if (4 * q + 1 >= n_ch) acc.y = -INFINITY;
          if (4 * q + 2 >= n_ch) acc.z = -INFINITY;
          acc.w = -INFINITY;
        }
        if (key < 0) fz[q] = acc;
        mx = fmaxf(fmaxf(mx, fmaxf(acc.x, acc.y)), fmaxf(acc.z, acc.w));
      }
      // the fused score of eos is needed even when eos is not a pre-beam candidate (:180-181)
      if (key < 0 && d.eos >= e0 && d.eos < e0 + n_ch && (((d.eos - e0) >> 2) % NSG) == j)
        S->w_eos = reinterpret_cast<const float*>(fz)[d.eos - e0];
      unsigned long long thr = tau;
      if (n_top < n_sel) {
        // ---- threshold of a chunk met with a non-full list: at least n_sel elements of the chunk are >=
        //      the kk-th largest lane maximum of every active warp (kk = ceil(n_sel / active warps)) ------
        const int nw_act = min(NSW, (nq + 31) >> 5);
        const int kk = (!SEQ && nw_act == NSW) ? sl.kk_full : (n_sel + nw_act - 1) / nw_act;
        unsigned co = (mx > -INFINITY) ? f2ord(mx + 0.0f) : 0u;
        unsigned kth = 0u;
        for (int r = 0; r < kk; ++r) {
          kth = __reduce_max_sync(FULLMASK, co);
          const unsigned holders = __ballot_sync(FULLMASK, co == kth);
          if (lane == __ffs(holders) - 1) co = 0u;
        }
        if (lane == 0) S->wkth[wj] = (wj < nw_act) ? kth : 0xffffffffu;  // idle warps do not vote
        group_bar(kBarSel, NSG);
        unsigned mn = S->wkth[0];
#pragma unroll
        for (int w2 = 1; w2 < NSW; ++w2) mn = min(mn, S->wkth[w2]);
        const unsigned long long thr_c = (mn == 0u) ? 0ull : (((unsigned long long)mn << 32) | 0x80000000ull);
        thr = thr_c > thr ? thr_c : thr;
        if (c == 0) LGASR_TR(1, 3);
      }
      // ---- sweep 2: push everything above the threshold, merge; repeat while the pending list overflows ----
      for (int round = 0;; ++round) {
        const int p = it & 1;
        unsigned long long* kp = S->keys[cur] + n_top;
        const float thr_v = thr ? key_val(thr) : -INFINITY;
        // rare path: a lane with a hit pushes on its own (shared atomics, no warp aggregation)
        auto push = [&](float val, int el) {
          if (val >= thr_v && el < n_ch) {
            const unsigned long long ke = pack_key(val, e0 + el);
            bool pu = ke > thr;
            if (round > 0 && pu) {  // an overflow round: skip what the list already holds
#pragma unroll 1
              for (int t = 0; t < n_top; ++t) pu = pu && (S->keys[cur][t] != ke);
            }
            if (pu) {
              const int slot = atomicAdd(&S->pend_cnt[p], 1);
              if (slot < kPendCap)
                kp[slot] = ke;
              else
                S->overflow[p] = 1;
            }
          }
        };
        // a thread whose own maximum of sweep 1 is below the threshold has nothing to push: on the chunks after
        // the list has filled (long rows) almost every thread skips the second pass over shared memory
        if (mx >= thr_v)
#pragma unroll 1
        for (int q = j; q < nq; q += NSG) {
          const float4 v = fz[q];
          if (fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)) >= thr_v) {
            push(v.x, 4 * q);
            push(v.y, 4 * q + 1);
            push(v.z, 4 * q + 2);
            push(v.w, 4 * q + 3);
          }
        }
        group_bar(kBarSel, NSG);  // A: pushes done, every thread has read its stage values
        const int cnt = S->pend_cnt[p];
        const int ovf = S->overflow[p];
        if (!ovf && tma && j == 0 && c + NST < nchunks) {
          fence_proxy_async();
          issue_row_chunk(tm, rowoff, c + NST, CH, NST, V, arena, S);
        }
        if (cnt > 0) {
          const int n = n_top + (cnt < kPendCap ? cnt : kPendCap);
          const unsigned long long* ks = S->keys[cur];
          unsigned long long* kn = S->keys[cur ^ 1];
          const ulonglong2* ks2 = reinterpret_cast<const ulonglong2*>(ks);
          const int n2 = n >> 1;
          for (int e = j; e < n; e += NSG) {
            const unsigned long long ke = ks[e];
            int rank = 0;
#pragma unroll 4
            for (int o = 0; o < n2; ++o) {
              const ulonglong2 k2 = ks2[o];
              rank += (k2.x > ke) ? 1 : 0;
              rank += (k2.y > ke) ? 1 : 0;
            }
            if (n & 1) rank += (ks[n - 1] > ke) ? 1 : 0;
            if (rank < n_sel) kn[rank] = ke;
          }
          group_bar(kBarSel, NSG);  // B: the merged list is complete
          if (j == 0) {
            S->pend_cnt[p] = 0;
            S->overflow[p] = 0;
          }
          n_top = n < n_sel ? n : n_sel;
          cur ^= 1;
          ++it;
          if (n_top == n_sel) {
            tau = S->keys[cur][n_sel - 1];
            thr = tau > thr ? tau : thr;
          }
        }
        if (!ovf) break;
      }
      if (c == 0) LGASR_TR(1, 4);
    }
    if (j == 0) {
      S->n_top = n_top;
      S->cur_list = cur;
    }
    LGASR_TR(1, 2);
    // ---- gather the candidates' posterior rows (first group) by TMA ------------------------------------
    if (mode == kModePreBeam && wj == 0)
      issue_cand_rows(d, S->keys[cur], 0, n_top < sl.rows_per_group ? n_top : sl.rows_per_group, xT_b, arena, S, lane);
  };
  if (SEQ) {
    __syncthreads();  // the header and the mbarriers are initialised
    LGASR_TR(1, 0);
    if (mode != kModeFullCtc) sel_phase();
  }
  if (SEQ && sl.flags) {
    // every producer of this utterance's state is a CTA of an earlier grid, resident or gone: no deadlock
    if (tid == 0) {
      int spins = 0;
      while (ld_acquire_gpu(d.step_flag + b) < step)
        if (++spins > (1 << 22)) __trap();  // a broken chain must fail, not hang the device
      asm volatile("fence.proxy.async.global;" ::: "memory");  // the parent state is fetched by TMA
    }
    __syncthreads();
  } else {
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  // per-hypothesis metadata: one batch of independent loads (rows h >= n_run hold stale but valid memory)
  const int done = d.done[pp][b], n_run = d.n_run[pp][b];
  const int T = d.xlen[b];
  const int cur_len = d.cur_len[pp][b];
  const int4 rm = d.rmeta[pp][row];  // parent slot, last token, parent's last token
  const float s_prev = d.use_ctc ? d.s_prev[pp][row] : 0.f;
  const float hyp_score = d.score[pp][row];
  const int maxlen_b = d.maxlen[b], minlen_b = d.minlen[b];  // for the selection (no dependent load there)
  // cluster variant: this hypothesis' fields the selection needs of its parents, sent to rank 0 later
  const int my_slot = CL ? d.run_slot[pp][row] : 0;
  const float my_sc = (CL && tid < d.NS) ? d.scores[pp][(long long)row * d.NS + tid] : 0.f;
  if (done || h >= n_run) {
    if (h == 0 && tid == 0 && (done || n_run <= 0)) {
      utt_carry_done(d, b, pp);
      if (sl.flags) st_release_gpu(d.step_flag + b, step + 1);
    }
    // copies already in flight target this CTA's shared memory: let them land before leaving
    if (SEQ) {
      if (mode == kModePreBeam) mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand[warp]), 0);
    } else if (tid == SEL0 && sl.pdl && tma && mode != kModeFullCtc) {
      const int pre = nchunks < NST ? nchunks : NST;
      for (int c = 0; c < pre; ++c) mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_full[c]), 0);
    }
    if (CL) {  // an idle CTA still takes part in both phases of the cluster barrier
      __syncwarp();
      cluster_wait();
      cluster_arrive();
    }
    return;
  }

  const int olen = cur_len - 1;
  const int last = rm.y;
  const int has_parent = rm.x >= 0;
  // decode position of THIS utterance (the utterances of a handle need not be at the same position: slot refill)
  const int stepu = cur_len - d.P;

  // the two issuing threads initialise their own mbarriers and start their TMA copies before the CTA barrier
  if (tid == 0) {
    mbar_init(reinterpret_cast<uint64_t*>(&S->bar_ctc), 1);
    mbar_fence_init();
    if (d.use_ctc) {
      const float* r_par = d.r[rp] + ((long long)(b * W + (has_parent ? rm.x : h)) * 2) * Tp;
      const uint32_t rb = (uint32_t)Tp * 4u;
      mbar_expect_tx(reinterpret_cast<uint64_t*>(&S->bar_ctc), has_parent ? 4u * rb : 2u * rb);
      bulk_g2s(bR0, r_par, 2u * rb, reinterpret_cast<uint64_t*>(&S->bar_ctc));
      if (has_parent) {
        bulk_g2s(bA, xT_b + (long long)last * Tp, rb, reinterpret_cast<uint64_t*>(&S->bar_ctc));
        bulk_g2s(bC, xT_b + (long long)d.blank * Tp, rb, reinterpret_cast<uint64_t*>(&S->bar_ctc));
      }
    }
  } else if (!SEQ && tid == SEL0 && !sl.pdl && tma && mode != kModeFullCtc) {
    const int pre = nchunks < NST ? nchunks : NST;
    for (int c = 0; c < pre; ++c) issue_row_chunk(tm, rowoff, c, CH, NST, V, arena, S);
  }
  __syncthreads();
  LGASR_TR(0, 0);
  if (!SEQ) LGASR_TR(1, 0);
  LGASR_GT(14);

  if (tid < NCT) {
    // ======================= ctc group: CTC state of this hypothesis ==================================
    if (d.use_ctc) {
      mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_ctc), 0);
      LGASR_TR(0, 1);
      group_ctc_scans<NCW>(bR1, bA, bB, bC, bD, has_parent, stepu == 1, last == rm.z, olen - 1, T, tid, S->ws, kBarCtc,
                           TR ? d.trace + ((long long)(blockIdx.y * d.W + blockIdx.x) * 2) * 16 : nullptr);
      LGASR_TR(0, 2);
    }
  } else if (!SEQ && mode != kModeFullCtc) {
    sel_phase();
  }
  __syncthreads();
  if (d.use_ctc)
    cta_ctc_epilogue(bR0, bR1, bA, bB, bC, bD, d.r[rp ^ 1] + ((long long)row * 2) * Tp, has_parent, olen - 1, T, Tp, tid,
                     &S->psi_eos);
  LGASR_TR(0, 3);

  // ======================= all warps: candidate records ===============================================
  const int n_top = S->n_top;
  const unsigned long long* tk = S->keys[S->cur_list];
  const int R = d.R;
  float* c_total = d.cand_total + (long long)row * R;
  int* c_tok = d.cand_tok + (long long)row * R;
  float* c_psi = d.cand_psi + (long long)row * R;
  // selection buffer of the cluster's rank-0 CTA (layout of utt_select): keys [W*R] u64 | psi [W*R] | sel, dest
  // [2W] | p_slot, p_last, p_sprev [3W] | p_scores [W*NS]
  float* selbuf = arena + sl.arena_floats;
  const uint32_t sel0 = CL ? map_rank(selbuf, 0) : 0u;
  if (CL) {
    cluster_wait();  // phase 1: every CTA of the cluster is running, its shared memory may be written
    const int WR = W * R;
    if (tid == 0) {
      st_cluster(sel0 + 4u * (3 * WR + 2 * W + h), (unsigned)my_slot);
      st_cluster(sel0 + 4u * (3 * WR + 3 * W + h), (unsigned)last);
      st_cluster(sel0 + 4u * (3 * WR + 4 * W + h), __float_as_uint(s_prev));
    }
    if (tid < d.NS) st_cluster(sel0 + 4u * (3 * WR + 5 * W + h * d.NS + tid), __float_as_uint(my_sc));
  }
  auto put_rec = [&](int jr, float total, int tok, float psi) {
    if (CL) {
      const int e = h * R + jr;
      unsigned u = __float_as_uint(total + 0.0f);
      u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
      const unsigned long long ke = tok >= 0 ? (((unsigned long long)u << 32) | (unsigned)(~(h * V + tok))) : 0ull;
      st_cluster(sel0 + 8u * e, ke);
      st_cluster(sel0 + 4u * (2 * W * R + e), __float_as_uint(psi));
    } else {
      c_total[jr] = total;
      c_tok[jr] = tok;
      c_psi[jr] = psi;
    }
  };

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
        put_rec(jr, total, c, 0.f);
      }
    }
  } else {
    const float psi_eos = S->psi_eos;  // r_sum at the end frame (ctc_prefix_score.py:180-181)
    const int RPG = sl.rows_per_group;
    const int nq = Tp >> 2;
    constexpr float kL2E = 1.4426950408889634f;
    // one candidate per half warp: candidate i of a group belongs to warp i % 8 (the warp whose mbarrier its
    // row was issued on), half (i / 8) % 2, round i / 16
    const int half = lane >> 4, hl = lane & 15;
    int grp = 0;
    for (int g0 = 0; g0 < n_top; g0 += RPG, ++grp) {
      const int g1 = (g0 + RPG < n_top) ? g0 + RPG : n_top;
      if (g0 > 0) {
        __syncthreads();  // the arena rows of the previous group have been consumed
        if (warp == 0) issue_cand_rows(d, tk, g0, g1, xT_b, arena, S, lane);
      }
      mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand[warp]), (uint32_t)(grp & 1));
      if (grp == 0) LGASR_TR(0, 4);
      for (int i0 = warp; i0 < g1 - g0; i0 += 2 * kStepWarps) {  // warp-uniform
        const int i = i0 + half * kStepWarps;
        const bool have = i < g1 - g0;
        const int jr = g0 + (have ? i : i0);
        const int c = key_idx(tk[jr]);
        const bool special = (c == d.eos) || (d.eos != d.blank && c == d.blank);
        // log psi = logsumexp({phi[t-1] + x[t] : start <= t < T} U {r^n[start-1]})  (:163-171), the way
        // torch.logsumexp does it (max, sum of exp, log).  phi is pre-shifted and LGASR_NEG outside
        // [start, T), so the sum runs over whole float4 rows.  (A special candidate runs the same code on
        // whatever its arena slot holds and discards the result.)
        const float4* x4 = reinterpret_cast<const float4*>(arena + (long long)(jr - g0) * Tp);
        const float4* p4 = reinterpret_cast<const float4*>((c == last) ? bR1 : bR0);
        const float rn0 = (olen == 0) ? arena[(long long)(jr - g0) * Tp] : LGASR_LOGZERO;
        float m = rn0;
#pragma unroll 1
        for (int q = hl; q < nq; q += 16) {
          const float4 xa = x4[q], pa = p4[q];
          m = fmaxf(fmaxf(fmaxf(m, pa.x + xa.x), fmaxf(pa.y + xa.y, pa.z + xa.z)), pa.w + xa.w);
        }
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULLMASK, m, o));
        // (e - m) first: the subtraction is exact near the maximum, where the terms matter
        float sum = (hl == 0) ? ex2_approx((rn0 - m) * kL2E) : 0.f;
#pragma unroll 1
        for (int q = hl; q < nq; q += 16) {
          const float4 xa = x4[q], pa = p4[q];
          sum += ex2_approx((pa.x + xa.x - m) * kL2E) + ex2_approx((pa.y + xa.y - m) * kL2E) +
                 ex2_approx((pa.z + xa.z - m) * kL2E) + ex2_approx((pa.w + xa.w - m) * kL2E);
        }
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) sum += __shfl_xor_sync(FULLMASK, sum, o);
        float psi = m + logf(sum);
        if (c == d.eos)
          psi = psi_eos;
        else if (special)
          psi = LGASR_LOGZERO;  // :183-185
        if (have && hl == 0) {
          const float wfull = (d.key < 0) ? key_val(tk[jr]) : fuse_at(tm, rowoff, c);
          put_rec(jr, __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi, s_prev))), hyp_score), c, psi);
        }
      }
    }
    if (lane == 0) {
      // unused pre-beam slots (fewer than S tokens in the vocabulary)
      for (int jr = n_top + warp; jr < n_sel; jr += kStepWarps) {
        put_rec(jr, -INFINITY, -1, 0.f);
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
        put_rec(n_sel, total, dup ? -1 : d.eos, psi_eos);
      }
    }
  }

  // ======================= ticket: the last CTA of the utterance runs the selection ==================
  // CTA barrier, then ONE fence + atomic by thread 0 (the cooperative-groups grid-sync pattern): the
  // barrier orders every warp's record stores before thread 0's cumulative gpu-scope fence.
  LGASR_TR(0, 5);
  if (CL) {
    // phase 2: every thread's remote stores are released by its own arrive; only rank 0 waits
    __syncwarp();
    cluster_arrive();
    LGASR_TR(0, 6);
    LGASR_GT(15);
    if (h != 0) return;
    cluster_wait();
    utt_select<true>(d, tm, b, pp, stepu, selbuf, n_run, cur_len, maxlen_b, minlen_b,
                     TR ? d.trace + ((long long)(blockIdx.y * d.W + blockIdx.x) * 2) * 16 : nullptr);
  } else {
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      const int old = atomicAdd(d.arrive + b, 1);
      const int is_last = (old == n_run - 1) ? 1 : 0;
      if (is_last) {
        __threadfence();
        d.arrive[b] = 0;
      }
      S->last_flag = is_last;
    }
    __syncthreads();
    LGASR_TR(0, 6);
    LGASR_GT(15);
    if (!S->last_flag) return;
    utt_select<false>(d, tm, b, pp, stepu, arena, n_run, cur_len, maxlen_b, minlen_b,
                      TR ? d.trace + ((long long)(blockIdx.y * d.W + blockIdx.x) * 2) * 16 : nullptr);
  }
  LGASR_TR(0, 7);
  LGASR_GT(15);
  if (sl.flags) {
    __syncthreads();  // every thread's writes of the selection, then one (cumulative) release by thread 0
    if (tid == 0) st_release_gpu(d.step_flag + b, step + 1);
  }
}

}  // namespace lgasr
