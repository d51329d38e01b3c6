// k_step2: the decode position in ONE launch, second generation (sm_100a).
//
//   grid (beam, B), 256 threads: one CTA per running hypothesis (b, h).  The kernel is issue-bound at the batch sizes
//   of interest (all CTAs of an SM walk through the same phase at the same time), so every phase is written for the
//   fewest issued instructions; all eight warps walk through the phases together:
//
//   1. row stream + exact top-S   the K full-scorer rows of the hypothesis are read ONCE, straight from global memory
//      (input-only: runs BEFORE      into registers (16-byte loads, 5 per thread in flight; long rows: groups of loads
//      griddepcontrol.wait)          behind one L2 bulk prefetch per row issued through the TMA engine), fused as
//                                    sum_k w_k*s_k in the reference's fp32 order (batch_beam_search.py:279-300) and
//                                    reduced to the exact pre-beam top-S (:302-308): lane maxima -> a value threshold
//                                    at least S elements pass -> the few elements above it are pushed from registers
//                                    -> rank counting on packed (value, ~index) keys.  No staging of the row in
//                                    shared memory, no second sweep, no topk launch.  fp32 or bf16 rows; an optional
//                                    per-row normaliser (logits + logsumexp) is folded in.
//   2. candidate gather           the S candidates' posterior rows xT[b][c][:] are requested by 1-D TMA bulk copies the
//                                    moment the list is final; they land while phase 3 runs.
//   3. CTC state                  parent r_sum row, token row and blank row arrive by TMA; two log-semiring affine
//      (ctc_prefix_score.py          scans over the frames (r^n, then r_sum; three frames per thread) give the
//       :122-160, scorers/           survivor's state.  The fix-up pass of the second scan writes the state AND the
//       ctc.py:56-62)                frame-shifted phi arrays of phase 4: there is no epilogue pass.
//   4. log psi + records          one candidate per half warp, max / sum-of-exp sweeps over float4 rows of
//      (:163-185,                    phi[t-1] + x[t] (torch.logsumexp's order of operations);
//       batch_beam_search.py         total = (fused + w_ctc*(psi - s_prev)) + score.
//       :313-318)
//   5. exchange + selection       thread-block cluster (distributed shared memory) or L2 ticket, then utt_select
//                                    (global top-beam, post_process, end_detect).
//
//   The decode position of an utterance is derived from its own prefix length (cur_len - primer_len), so the
//   utterances of a handle need not be at the same position (slot refill / continuous batching).
#pragma once
#include "lgasr_step.cuh"

namespace lgasr {

constexpr int kPend2 = 512;  // pending pushes of the row stream (u64 keys, aliased with the candidate-row arena)

struct __align__(16) Step2Smem {
  unsigned long long keys[2][kTopCap];  // sorted top list, ping-pong by merge
  unsigned long long bar_ctc;           // parent state + token row + blank row
  unsigned long long bar_cand;          // candidate rows of the current group
  // state of the previous position (four 16-byte records, loaded by threads 0..3)
  int4 m_u;                             // (done, n_run, cur_len, 0)
  int4 m_s;                             // (xlen, maxlen, minlen, 0)
  int4 m_h0;                            // (parent slot, last token, parent's last token, run_slot)
  int4 m_h1;                            // (s_prev, score, 0, 0) as bits
  float m_sc[kMaxTerms + 4];            // per-scorer cumulative scores (cluster exchange only)
  int pend_cnt[2], overflow[2];         // double-buffered by merge iteration
  float w_eos, psi_eos;
  int last_flag, pad0;
  float ws[2][2 * kStepWarps];          // (A, B) segment summaries per warp, one set per scan
  unsigned wkth[kStepWarps];            // vote of the row stream: smallest lane maximum extracted by each warp
  int wgot[kStepWarps];                 // ... and how many it extracted
};

// sum_k w_k * s_k[v] in the reference's fp32 order, any storage type
__device__ __noinline__ float fuse_at2(const Terms& t, int row, long long rowoff, int v) {
  float acc = 0.f;
#pragma unroll 1
  for (int k = 0; k < t.K; ++k) {
    const float s = t.ptr[k] ? term_at(t, k, row, rowoff, v) : t.cst[k];
    acc = __fadd_rn(acc, __fmul_rn(t.w[k], s));
  }
  return acc;
}

// quad q (vocabulary entries 4q .. 4q+3) of term k of this hypothesis' row (16-byte aligned rows, V % 4 == 0):
// fp32 or bf16 storage, optional per-row normaliser
__device__ __forceinline__ float4 load_quad_any(const Terms& t, int k, long long rowoff, float nrm, int q) {
  float4 s;
  if (t.bf16[k]) {
    const uint2 u = __ldcs(reinterpret_cast<const uint2*>(reinterpret_cast<const unsigned short*>(t.ptr[k]) + rowoff) + q);
    s.x = __uint_as_float(u.x << 16);
    s.y = __uint_as_float(u.x & 0xffff0000u);
    s.z = __uint_as_float(u.y << 16);
    s.w = __uint_as_float(u.y & 0xffff0000u);
  } else {
    s = __ldcs(reinterpret_cast<const float4*>(t.ptr[k] + rowoff) + q);
  }
  if (t.norm[k]) {
    s.x = __fsub_rn(s.x, nrm);
    s.y = __fsub_rn(s.y, nrm);
    s.z = __fsub_rn(s.z, nrm);
    s.w = __fsub_rn(s.w, nrm);
  }
  return s;
}

__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

__device__ __forceinline__ float max4(const float4& v) { return fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)); }

// load + fuse the (up to) G quads of this thread in one group: quads q0 + i*256, i < nv valid.  FULL: nv == G.
// Invalid quads come back as -inf.
template <int NTH, int G, int NT, bool FULL>
__device__ __forceinline__ void load_fuse(const Terms& tm, int row, long long rowoff, int key, int q0, int nv, float4 (&v)[G]) {
  if (NT == 1) {
    const int k0 = key >= 0 ? key : 0;
    const float4* p = reinterpret_cast<const float4*>(tm.ptr[k0] + rowoff) + q0;
    const float w0 = key >= 0 ? 1.0f : tm.w[0];  // the key term is ranked by its raw scores (beam_search.py:302-308)
    // an invalid quad must come out of the multiplication as -inf
    const float fill = w0 > 0.f ? -INFINITY : INFINITY;
#pragma unroll
    for (int i = 0; i < G; ++i)
      v[i] = (FULL || i < nv) ? __ldcs(p + i * NTH) : make_float4(fill, fill, fill, fill);
#pragma unroll
    for (int i = 0; i < G; ++i) {  // 0 + w0*s0 == w0*s0 (up to the sign of zero, which pack_key canonicalises)
      v[i].x = __fmul_rn(w0, v[i].x);
      v[i].y = __fmul_rn(w0, v[i].y);
      v[i].z = __fmul_rn(w0, v[i].z);
      v[i].w = __fmul_rn(w0, v[i].w);
    }
  } else if (NT == 2) {
    const float w0 = tm.w[0], w1 = tm.w[1];
    const float4* pa = reinterpret_cast<const float4*>(tm.ptr[0] + rowoff) + q0;
    const float4* pb = reinterpret_cast<const float4*>(tm.ptr[1] + rowoff) + q0;
    float4 u[G];
#pragma unroll
    for (int i = 0; i < G; ++i) {
      const bool ok = FULL || i < nv;
      v[i] = ok ? __ldcs(pa + i * NTH) : make_float4(0.f, 0.f, 0.f, 0.f);
      u[i] = ok ? __ldcs(pb + i * NTH) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < G; ++i) {  // (0 + w0*s0) + w1*s1, the first sum being exact
      v[i].x = __fadd_rn(__fmul_rn(w0, v[i].x), __fmul_rn(w1, u[i].x));
      v[i].y = __fadd_rn(__fmul_rn(w0, v[i].y), __fmul_rn(w1, u[i].y));
      v[i].z = __fadd_rn(__fmul_rn(w0, v[i].z), __fmul_rn(w1, u[i].z));
      v[i].w = __fadd_rn(__fmul_rn(w0, v[i].w), __fmul_rn(w1, u[i].w));
      if (!FULL && i >= nv) v[i] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
  } else {
#pragma unroll
    for (int i = 0; i < G; ++i) v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int k = 0; k < tm.K; ++k) {
      const float wk = tm.w[k];
      if (tm.ptr[k]) {
        const float nk = tm.norm[k] ? __ldg(tm.norm[k] + row) : 0.f;
#pragma unroll
        for (int i = 0; i < G; ++i) {
          if (FULL || i < nv) {
            const float4 s = load_quad_any(tm, k, rowoff, nk, q0 + i * NTH);
            v[i].x = __fadd_rn(v[i].x, __fmul_rn(wk, s.x));
            v[i].y = __fadd_rn(v[i].y, __fmul_rn(wk, s.y));
            v[i].z = __fadd_rn(v[i].z, __fmul_rn(wk, s.z));
            v[i].w = __fadd_rn(v[i].w, __fmul_rn(wk, s.w));
          }
        }
      } else {
        const float cv = __fmul_rn(wk, tm.cst[k]);
#pragma unroll
        for (int i = 0; i < G; ++i) {
          v[i].x = __fadd_rn(v[i].x, cv);
          v[i].y = __fadd_rn(v[i].y, cv);
          v[i].z = __fadd_rn(v[i].z, cv);
          v[i].w = __fadd_rn(v[i].w, cv);
        }
      }
    }
    if (!FULL) {
#pragma unroll
      for (int i = 0; i < G; ++i)
        if (i >= nv) v[i] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
  }
}

// ---- phase 1: fused row stream + exact top-n_sel ---------------------------------------------------------------
// G quads (4 consecutive vocabulary entries) per thread and group; thread tid owns quads (g*G + i)*256 + tid.
// NT: 1 = one fp32 matrix without normaliser (or the pre-beam key term), 2 = two of them, 0 = the general term loop
// (constants, bf16 storage, normalisers, any K).  Rows are 16-byte aligned and V % 4 == 0 (lgasr.cu routes anything
// else to k_step).
// Result: S->keys[cur][0 .. n_top) sorted best first (returned through n_top / cur); S->w_eos = fused score of eos.
// Every thread of the CTA calls this; ends with a CTA barrier after which the list is complete.
//
// Sweep kinds: kFirst (a group met with a non-full list takes its threshold from a vote over lane maxima, later
// groups use the n_sel-th key so far), kRedo (the pending list overflowed, e.g. a row of equal scores: sweep again
// with the tightened threshold; the merge drops what the list already holds), kFill (rows with fewer than n_sel
// finite entries, e.g. under ScoreFilter masks: -inf entries complete the list, lowest index first, one window of
// kPend2 entries per sweep).
template <int NTH, int G, int NT>
__device__ __forceinline__ void row_topk(const Dev& d, const Terms& tm, const StepLaunch& sl, int row, long long rowoff,
                                         int n_sel, Step2Smem* S, unsigned long long* pend, int tid, int& n_top_out,
                                         int& cur_out, long long* tr = nullptr) {
  constexpr int kFirst = 0, kRedo = 1, kFill = 2;
  constexpr int PG = G * NTH;  // quads per group
  constexpr int NW = NTH / 32;
  constexpr int LOGN = NTH == 256 ? 8 : NTH == 128 ? 7 : 6;
  static_assert(NTH == 256 || NTH == 128 || NTH == 64, "CTA sizes of k_step2");
  const int lane = tid & 31, warp = tid >> 5;
  const int key = d.key;
  const int nq = d.V >> 2;
  const int ngroups = (nq + PG - 1) / PG;
  int cur = 0, n_top = 0, it = 0;
  unsigned long long tau = 0ull;  // key of the n_sel-th best so far (0: list not full)
  if (ngroups > 1 && tid < tm.K && tm.ptr[tid]) {
    // long rows: one L2 bulk prefetch per 32 KB of each term's row (TMA engine), the loads below then hit L2
    const int es = tm.bf16[tid] ? 2 : 4;
    const char* base = reinterpret_cast<const char*>(tm.ptr[tid]) + rowoff * es;
    const long long bytes = ((long long)d.V * es) & ~15LL;
    if ((reinterpret_cast<uintptr_t>(base) & 15) == 0)
      for (long long o = 0; o < bytes; o += 32768)
        bulk_prefetch_l2(base + o, (uint32_t)(bytes - o < 32768 ? bytes - o : 32768));
  }
  // where eos lives: (group, thread, quad of the thread)
  const int eos_q = d.eos >> 2;
  const int eos_g = eos_q / PG, eos_t = eos_q & (NTH - 1), eos_i = (eos_q - eos_g * PG) >> LOGN;
  int kind = kFirst, fill_lo = 0;
  bool redo = false;
  for (;;) {  // sweeps over the row (normally one)
#pragma unroll 1
    for (int g = 0; g < ngroups; ++g) {
      float4 v[G];
      const int q0 = g * PG + tid;
      const bool lastg = g + 1 == ngroups;
      // quads of this thread in the group: q0 + i*NTH < nq (all G of them except at the end of the row)
      int nv = G;
      if (lastg) {
        nv = (nq - q0 + NTH - 1) >> LOGN;
        nv = nv < 0 ? 0 : nv;
      }
      if (nv >= G)
        load_fuse<NTH, G, NT, true>(tm, row, rowoff, key, q0, G, v);
      else
        load_fuse<NTH, G, NT, false>(tm, row, rowoff, key, q0, nv, v);
      float mx = max4(v[0]);
#pragma unroll
      for (int i = 1; i < G; ++i) mx = fmaxf(mx, max4(v[i]));
      if (tr && tid == 0 && g == 0) tr[8] = clock64() + (long long)(mx == 12345.f);  // rows in registers
      if (key < 0 && g == eos_g && tid == eos_t) {
        // the thread that holds eos publishes its fused score (needed even when eos is no candidate, :180-181)
        float4 ve = v[0];
#pragma unroll
        for (int i = 1; i < G; ++i)
          if (i == eos_i) ve = v[i];
        S->w_eos = (d.eos & 2) ? ((d.eos & 1) ? ve.w : ve.z) : ((d.eos & 1) ? ve.y : ve.x);
      }
      const int p = it & 1;
      const bool vote = (kind == kFirst) && (n_top < n_sel);
      unsigned long long thr = tau;
      if (vote) {
        // ---- threshold of a group met with a non-full list: the voting warps (the fully populated ones) each
        //      extract up to kk of their largest finite lane maxima; if together they hold >= n_sel of them, at least
        //      n_sel elements of the group are >= the smallest one extracted; otherwise every finite entry is pushed --
        int nvote = NW, kk = (n_sel + NW - 1) / NW;
        if (lastg) {
          const int left = nq - g * PG;
          const int act = left < NTH ? left : NTH;
          nvote = act >> 5;
          nvote = nvote > 0 ? nvote : 1;
          if (nvote < NW) kk = (n_sel + nvote - 1) / nvote;
        }
        unsigned co = (mx > -INFINITY) ? f2ord(mx + 0.0f) : 0u;
        unsigned kth = 0xffffffffu;
        int got = 0;
#pragma unroll 1
        for (int r = 0; r < kk; ++r) {
          const unsigned m2 = __reduce_max_sync(FULLMASK, co);
          if (m2 == 0u) break;  // no finite lane maximum left
          kth = m2;
          ++got;
          const unsigned holders = __ballot_sync(FULLMASK, co == m2);
          if (lane == __ffs(holders) - 1) co = 0u;
        }
        if (lane == 0) {
          S->wkth[warp] = (warp < nvote) ? kth : 0xffffffffu;
          S->wgot[warp] = (warp < nvote) ? got : 0;
        }
        __syncthreads();
        unsigned mn = S->wkth[0];
        int total = S->wgot[0];
#pragma unroll
        for (int w2 = 1; w2 < NW; ++w2) {
          mn = min(mn, S->wkth[w2]);
          total += S->wgot[w2];
        }
        const unsigned long long thr_c = (total >= n_sel) ? (((unsigned long long)mn << 32) | 0x80000000ull) : 0ull;
        thr = thr_c > thr ? thr_c : thr;
        if (tr && tid == 0 && g == 0) tr[9] = clock64();  // threshold known
      }
      if (kind != kFill) {
        // ---- push everything above the threshold (registers -> pending list); -inf entries never travel here.
        //      A quad rarely holds more than one such entry: take its maximum, push, blank it, look again. -----------
        const float lo_v = thr ? key_val(thr) : -3.4028234e38f;
        if (mx >= lo_v) {
#pragma unroll
          for (int i = 0; i < G; ++i) {
            float m = max4(v[i]);
            while (m >= lo_v) {
              const int j = (v[i].x == m) ? 0 : (v[i].y == m) ? 1 : (v[i].z == m) ? 2 : 3;
              const unsigned long long ke = pack_key(m, 4 * (q0 + i * NTH) + j);
              if (ke > thr) {
                const int slot = atomicAdd(&S->pend_cnt[p], 1);
                if (slot < kPend2)
                  pend[slot] = ke;
                else
                  S->overflow[p] = 1;
              }
              if (j == 0) v[i].x = -INFINITY;
              if (j == 1) v[i].y = -INFINITY;
              if (j == 2) v[i].z = -INFINITY;
              if (j == 3) v[i].w = -INFINITY;
              m = max4(v[i]);
            }
          }
        }
      } else {
        // -inf entries of the window [fill_lo, fill_lo + kPend2): at most kPend2 pushes
#pragma unroll
        for (int i = 0; i < G; ++i) {
          const int e = 4 * (q0 + i * NTH);
          if (e + 3 >= fill_lo && e < fill_lo + kPend2 && e < d.V) {
            const float vv[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (vv[j] == -INFINITY && e + j >= fill_lo && e + j < fill_lo + kPend2)
                pend[atomicAdd(&S->pend_cnt[p], 1)] = pack_key(vv[j], e + j);
          }
        }
      }
      // ---- merge the pending pushes into the sorted list: after a voted group (the list fills, later groups use
      //      its n_sel-th key) and at the end of the sweep.  Rank counting, two CTA barriers. --------------------------
      if (vote || lastg) {
        if (tr && tid == 0 && lastg) tr[13] = clock64();  // own pushes done
        __syncthreads();  // pushes done
        if (tr && tid == 0 && lastg) tr[14] = clock64();
        const int cnt_raw = S->pend_cnt[p];
        if (S->overflow[p]) redo = true;
        const int cnt = cnt_raw < kPend2 ? cnt_raw : kPend2;
        if (cnt > 0) {
          const int n = n_top + cnt;
          const unsigned long long* ks = S->keys[cur];
          unsigned long long* kn = S->keys[cur ^ 1];
          if (kind == kRedo) {
            // a repeated sweep pushes what the list already holds once more: blank those entries (key 0 ranks last)
            for (int e = tid; e < cnt; e += NTH) {
              const unsigned long long ke = pend[e];
              bool dup = false;
#pragma unroll 1
              for (int o = 0; o < n_top; ++o) dup = dup || (ks[o] == ke);
              if (dup) pend[e] = 0ull;
            }
            __syncthreads();
          }
          for (int e = tid; e < n; e += NTH) {
            const unsigned long long ke = e < n_top ? ks[e] : pend[e - n_top];
            int rank = 0;
#pragma unroll 1
            for (int o = 0; o < n_top; ++o) rank += (ks[o] > ke) ? 1 : 0;
#pragma unroll 4
            for (int o = 0; o < cnt; ++o) rank += (pend[o] > ke) ? 1 : 0;
            if (rank < n_sel && ke != 0ull) kn[rank] = ke;
          }
          __syncthreads();  // merged list complete
          if (tid == 0) {
            S->pend_cnt[p] = 0;
            S->overflow[p] = 0;
          }
          n_top = n < n_sel ? n : n_sel;
          cur ^= 1;
          if (n_top == n_sel) tau = S->keys[cur][n_sel - 1];
        } else {
          __syncthreads();
        }
        ++it;
      }
    }
    if (redo) {
      redo = false;
      kind = kRedo;
      continue;
    }
    if (n_top < n_sel && n_top < d.V) {  // fewer than n_sel finite entries: -inf entries fill up, lowest index first
      fill_lo = (kind == kFill) ? fill_lo + kPend2 : 0;
      kind = kFill;
      if (fill_lo < d.V) continue;
    }
    break;
  }
  n_top_out = n_top;
  cur_out = cur;
}

// gather the posterior rows of candidates [g0, g1) of the sorted list into the arena (one warp, one mbarrier)
__device__ __forceinline__ void issue_cand_rows2(const Dev& d, const unsigned long long* keys, int g0, int g1, const float* xT_b,
                                                 float* arena, uint64_t* bar, int lane) {
  const int Tp = d.Tp;
  const uint32_t rb = (uint32_t)Tp * 4u;
  fence_proxy_async();
  uint32_t n_need = 0;
  for (int i0 = g0; i0 < g1; i0 += 32) {
    const int i = i0 + lane;
    int cc = -1;
    if (i < g1) cc = key_idx(keys[i]);
    const bool need = cc >= 0 && cc != d.eos && !(d.eos != d.blank && cc == d.blank);
    n_need += __popc(__ballot_sync(FULLMASK, need));
  }
  if (lane == 0) mbar_expect_tx(bar, n_need * rb);
  __syncwarp();
  for (int i0 = g0; i0 < g1; i0 += 32) {
    const int i = i0 + lane;
    if (i < g1) {
      const int cc = key_idx(keys[i]);
      if (cc != d.eos && !(d.eos != d.blank && cc == d.blank))
        bulk_g2s(arena + (i - g0) * Tp, xT_b + (long long)cc * Tp, rb, bar);
    }
  }
}

// ---- phase 3 helper: prefix of one affine scan  y[t] = lse2(y[t-1] + a[t], b[t]) ------------------------------------
// `cur`/`A`: this thread's pass-1 result (local recursion from the identity; sum of a over its segment).  Only the
// first `nwa` warps own frames.  Returns y[s0 - 1], the state entering the thread's segment.  One CTA barrier inside.
__device__ __forceinline__ float scan_prefix(float cur, float A, float y0, int tid, int nwa, float* ws) {
  const int lane = tid & 31, w = tid >> 5;
  float SA = A, SB = cur, ea = 0.f, eb = LGASR_NEG;
  if (w < nwa) {
#pragma unroll
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
      ws[kStepWarps + w] = SB;
    }
    ea = __shfl_up_sync(FULLMASK, SA, 1);
    eb = __shfl_up_sync(FULLMASK, SB, 1);
    if (lane == 0) {
      ea = 0.f;
      eb = LGASR_NEG;
    }
  }
  __syncthreads();
  if (w >= nwa) return y0;
  float XA = 0.f, XB = LGASR_NEG;
#pragma unroll 1
  for (int i = 0; i < w; ++i) {  // summaries of the preceding warps
    const float wa = ws[i];
    XB = lse2(XB + wa, ws[kStepWarps + i]);
    XA += wa;
  }
  XB = lse2(XB + ea, eb);  // then this thread's exclusive prefix inside its warp
  XA += ea;
  return (tid == 0) ? y0 : lse2(y0 + XA, XB);
}

// NT: row path of phase 1 (1 = one fp32 matrix or the pre-beam key term, 2 = two fp32 matrices, 0 = general loop)
// TR: debug instance with clock64 / globaltimer stamps (LGASR_TRACE=1)
// NTH: threads per CTA (256, 128 or 64).  The glue around the streaming loops is paid per warp, so fewer, fatter warps
// issue fewer instructions per hypothesis (the choice is made in lgasr.cu from the grid size).
// MB: resident CTAs per SM the register allocation aims at (5 with 256 threads, 6 or 8 with 128; lgasr.cu picks 8 - with
// the candidate rows in two groups - for grids of several waves over streamed rows)
template <int NTH, bool CL, int NT, bool TR, int MB>
__global__ void __launch_bounds__(NTH, MB)
    k_step2(const __grid_constant__ Dev d, const __grid_constant__ Terms tm, const __grid_constant__ StepLaunch sl, int pp, int rp) {
  constexpr int NW = NTH / 32;
  constexpr int G1 = (NTH == 256 || MB >= 8) ? 5 : 10, G2 = NTH == 256 ? 2 : 5, G0 = NTH == 256 ? 2 : 4;
  constexpr int GG = NT == 1 ? G1 : (NT == 2 ? G2 : G0);
  const int b = blockIdx.y, h = blockIdx.x;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int W = d.W, Tp = d.Tp;
  const int TpS = Tp + 4;  // row stride of the scan buffers (room for the one-frame shift of the phi arrays)
  const int row = b * W + h;
  const long long rowoff = (long long)row * tm.ld;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Step2Smem* S = reinterpret_cast<Step2Smem*>(smem_raw);
  float* bR0 = reinterpret_cast<float*>(smem_raw + sizeof(Step2Smem));  // parent r^n (initial hypothesis only) -> phi P0
  float* bR1 = bR0 + TpS;                                               // parent r_sum                     -> phi P1
  float* bA = bR1 + TpS;                                                // x[:, last token]
  float* bC = bA + TpS;                                                 // x[:, blank]
  float* bN = bC + TpS;                                                 // r^n of this hypothesis
  float* bS = bN + TpS;                                                 // r_sum of this hypothesis
  float* arena = bS + TpS;                                              // pending keys | candidate rows | selection scratch
  const int mode = d.mode;
  const bool pdl = sl.pdl != 0;
  const float* xT_b = d.xT + (long long)b * d.V * Tp;
  const int n_sel = (mode == kModePreBeam) ? d.S : d.W;
#define LGASR_T2(i, val)                                          \
  do {                                                            \
    if (TR && tid == 0) d.trace[(long long)row * 32 + (i)] = val; \
  } while (0)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  LGASR_T2(16 + 13, gtimer());
  if (CL) cluster_arrive();  // phase 1: "this CTA runs"; waited for before the first remote store
  if (tid == 0) {
    mbar_init(reinterpret_cast<uint64_t*>(&S->bar_ctc), 1);
    mbar_init(reinterpret_cast<uint64_t*>(&S->bar_cand), 1);
    S->pend_cnt[0] = S->pend_cnt[1] = 0;
    S->overflow[0] = S->overflow[1] = 0;
    mbar_fence_init();
  }
  // ---- state of the previous position: four 16-byte records ------------------------------------------------------
  auto load_meta = [&]() {
    if (tid == 0) S->m_u = d.urec[pp][b];
    if (tid == 1) S->m_s = d.ustat[b];
    if (tid == 2) S->m_h0 = d.hrec[pp][2 * row];
    if (tid == 3) S->m_h1 = d.hrec[pp][2 * row + 1];
    if (CL && tid >= 4 && tid < 4 + d.NS) S->m_sc[tid - 4] = d.scores[pp][(long long)row * d.NS + (tid - 4)];
  };
  // TMA of the CTC inputs (one thread): parent r_sum (both rows for the initial hypothesis), token row, blank row
  auto issue_ctc = [&]() {
    if (!d.use_ctc) return;
    const int4 rm = S->m_h0;
    const int has_par = rm.x >= 0;
    const float* r_par = d.r[rp] + ((long long)(b * W + (has_par ? rm.x : h)) * 2) * Tp;
    const uint32_t rb = (uint32_t)Tp * 4u;
    uint64_t* bar = reinterpret_cast<uint64_t*>(&S->bar_ctc);
    mbar_expect_tx(bar, has_par ? 3u * rb : 2u * rb);
    if (has_par) {
      bulk_g2s(bR1, r_par + Tp, rb, bar);
      bulk_g2s(bA, xT_b + (long long)rm.y * Tp, rb, bar);
      bulk_g2s(bC, xT_b + (long long)d.blank * Tp, rb, bar);
    } else {
      bulk_g2s(bR0, r_par, rb, bar);
      bulk_g2s(bR1, r_par + Tp, rb, bar);
    }
  };
  bool idle = false;
  if (!pdl) load_meta();  // stepping launch: the previous position is complete
  __syncthreads();        // header, mbarriers (and, stepping, the state records) are in place
  if (!pdl) {
    idle = S->m_u.x || h >= S->m_u.y;  // idle CTAs leave before touching the rows
    if (!idle && tid == 0) issue_ctc();
  }
  int n_top = 0, cur = 0;
  if (!idle) {
    LGASR_T2(0, clock64());
    row_topk<NTH, GG, NT>(d, tm, sl, row, rowoff, n_sel, S, reinterpret_cast<unsigned long long*>(arena), tid, n_top, cur,
                          TR ? d.trace + (long long)row * 32 : nullptr);
    LGASR_T2(1, clock64());
    // candidate rows of the first group: requested now, consumed after the CTC scans
    if (mode == kModePreBeam && warp == 0)
      issue_cand_rows2(d, S->keys[cur], 0, n_top < sl.rows_per_group ? n_top : sl.rows_per_group, xT_b, arena,
                       reinterpret_cast<uint64_t*>(&S->bar_cand), lane);
  }
  if (pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    load_meta();
    __syncthreads();
    idle = S->m_u.x || h >= S->m_u.y;
    if (!idle && tid == 0) issue_ctc();
  }
  LGASR_T2(16 + 14, gtimer());
  const int n_run = S->m_u.y;
  if (idle) {
    if (h == 0 && tid == 0 && (S->m_u.x || n_run <= 0)) utt_carry_done(d, b, pp);
    // copies already in flight target this CTA's shared memory: let them land before leaving
    if (pdl && mode == kModePreBeam) mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand), 0);
    if (CL) {  // an idle CTA still takes part in both phases of the cluster barrier
      __syncwarp();
      cluster_wait();
      cluster_arrive();
    }
    return;
  }
  const int T = S->m_s.x, cur_len = S->m_u.z;
  const int last = S->m_h0.y;
  const int has_parent = S->m_h0.x >= 0;
  const int step = cur_len - d.P;  // decode position of THIS utterance
  const int olen = cur_len - 1;

  // ======================= phase 3: CTC state of this hypothesis ===============================================
  const float* p0 = bR0;  // phi arrays of phase 4 (frame-shifted): P0[u] = r_sum[u-1], P1[u] = r^b[u-1]
  const float* p1 = bR1;
  if (d.use_ctc) {
    float* r_out = d.r[rp ^ 1] + ((long long)row * 2) * Tp;
    mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_ctc), 0);
    LGASR_T2(2, clock64());
    if (has_parent) {
      const int olen_par = olen - 1;
      const int start = olen_par > 1 ? olen_par : 1;              // first frame of the recursion that makes this state
      const int start_par = olen_par - 1 > 1 ? olen_par - 1 : 1;  // the parent's own recursion start
      const bool rep = (last == S->m_h0.z) && (step != 1);        // repeated token: phi of the parent is its r^b (:131-140)
      const float rn0 = (olen_par == 0) ? bA[0] : LGASR_LOGZERO;  // r[start-1, 0]  (:128-129)
      const int n = T - start;
      // three frames per thread while they cover the utterance (an odd length keeps the lanes on distinct banks)
      int L = (n + NTH - 1) / NTH;
      L = L < 3 ? 3 : (L | 1);
      const int nwa = ((n + L - 1) / L + 31) >> 5;  // warps that own frames
      int s0 = start + tid * L, s1 = s0 + L;
      if (s0 > T) s0 = T;
      if (s1 > T) s1 = T;
      if (tid == 0 && n < 1) S->psi_eos = (T - 1 == start - 1) ? rn0 : LGASR_LOGZERO;  // no frame left to scan
      // ---- scan 1: r^n.  a = x_tok, b[t] = phi_parent[t-1] + x_tok[t] ----------------------------------------
      float c1 = LGASR_NEG, A1 = 0.f;
      if (!rep) {
#pragma unroll 1
        for (int t = s0; t < s1; ++t) {
          const float at = bA[t];
          c1 = lse2(c1 + at, bR1[t - 1] + at);
          bN[t] = c1;
          A1 += at;
        }
      } else {
#pragma unroll 1
        for (int t = s0; t < s1; ++t) {
          const int u = t - 1;
          const float phi = (u >= start_par) ? bR1[u - 1] + bC[u] : LGASR_LOGZERO;
          const float at = bA[t];
          c1 = lse2(c1 + at, phi + at);
          bN[t] = c1;
          A1 += at;
        }
      }
      const float yin1 = scan_prefix(c1, A1, rn0, tid, nwa, S->ws[0]);
      // scan 2 pass 1 rides on scan 1's fix-up: r_sum.  a = x_blank, b[t] = r^n[t]
      float run = 0.f, c2 = LGASR_NEG, A2 = 0.f;
#pragma unroll 1
      for (int t = s0; t < s1; ++t) {
        run += bA[t];
        const float rn = lse2(yin1 + run, bN[t]);
        r_out[t] = rn;
        const float ct = bC[t];
        c2 = lse2(c2 + ct, rn);
        bS[t] = c2;
        A2 += ct;
      }
      // entries of the state and of the phi arrays before the recursion / past the utterance (the parent's rows were
      // last read before the barrier of scan 1)
      for (int u = tid; u <= start; u += NTH) {
        bR0[u] = (u == start) ? rn0 : LGASR_NEG;
        bR1[u] = LGASR_NEG;
        if (u < start) {
          const float rv = (u == start - 1) ? rn0 : LGASR_LOGZERO;
          r_out[u] = rv;
          r_out[Tp + u] = rv;
        }
      }
      for (int u = T + 1 + tid; u < Tp; u += NTH) {
        bR0[u] = LGASR_NEG;
        bR1[u] = LGASR_NEG;
      }
      const float yin2 = scan_prefix(c2, A2, rn0, tid, nwa, S->ws[1]);
      run = 0.f;
      float prev = yin2;
#pragma unroll 1
      for (int t = s0; t < s1; ++t) {
        const float ct = bC[t];
        run += ct;
        const float rs = lse2(yin2 + run, bS[t]);
        const float rb = prev + ct;  // r^b[t] = r_sum[t-1] + x_blank[t]
        r_out[Tp + t] = rs;
        bR0[t + 1] = rs;
        bR1[t + 1] = rb;
        if (t + 1 >= T) S->psi_eos = rs;  // t == T - 1: r_sum at the end frame (:180-181)
        prev = rs;
      }
    } else {
      // initial hypothesis (position 0): the state is given; phi arrays staged in bN / bS
      const int pstart = olen > 1 ? olen : 1;  // a primer makes the first prefix longer than [sos]
      for (int u = tid; u < Tp; u += NTH) {
        const bool in = (u >= pstart && u < T);
        const float rs = in ? bR1[u - 1] : LGASR_NEG;
        bN[u] = rs;
        bS[u] = rs;  // r^b == r_sum while r^n is log-zero (:82-93)
        if (u < T) {
          r_out[u] = bR0[u];
          r_out[Tp + u] = bR1[u];
        }
      }
      if (tid == 0) S->psi_eos = bR1[T - 1];
      p0 = bN;
      p1 = bS;
    }
    __syncthreads();  // state + phi arrays complete
    LGASR_T2(3, clock64());
  } else {
    __syncthreads();
  }

  // ======================= phase 4: candidate records ==========================================================
  const float s_prev = __int_as_float(S->m_h1.x), hyp_score = __int_as_float(S->m_h1.y);
  const unsigned long long* tk = S->keys[cur];
  const int R = d.R;
  float* selbuf = arena + sl.arena_floats;
  const uint32_t sel0 = CL ? map_rank(selbuf, 0) : 0u;
  if (CL) {
    cluster_wait();  // phase 1: every CTA of the cluster is running, its shared memory may be written
    const int WR = W * R;
    if (tid == 0) {
      st_cluster(sel0 + 4u * (3 * WR + 2 * W + h), (unsigned)S->m_h0.w);
      st_cluster(sel0 + 4u * (3 * WR + 3 * W + h), (unsigned)last);
      st_cluster(sel0 + 4u * (3 * WR + 4 * W + h), __float_as_uint(s_prev));
    }
    if (tid < d.NS) st_cluster(sel0 + 4u * (3 * WR + 5 * W + h * d.NS + tid), __float_as_uint(S->m_sc[tid]));
  }
  auto put_rec = [&](int jr, float total, int tok, float psi) {
    if (CL) {
      const int e = h * R + jr;
      unsigned u = __float_as_uint(total + 0.0f);
      u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
      const unsigned long long ke = tok >= 0 ? (((unsigned long long)u << 32) | (unsigned)(~(h * d.V + tok))) : 0ull;
      st_cluster(sel0 + 8u * e, ke);
      st_cluster(sel0 + 4u * (2 * W * R + e), __float_as_uint(psi));
    } else {
      const long long o = (long long)row * R + jr;
      d.cand_total[o] = total;
      d.cand_tok[o] = tok;
      d.cand_psi[o] = psi;
    }
  };

  if (mode == kModeNoCtc) {
    for (int jr = tid; jr < d.W; jr += NTH) {
      float total = -INFINITY;
      int c = -1;
      if (jr < n_top) {
        c = key_idx(tk[jr]);
        total = __fadd_rn(key_val(tk[jr]), hyp_score);  // the key is always the full weighted sum here (lgasr_create)
      }
      put_rec(jr, total, c, 0.f);
    }
  } else {
    const float psi_eos = S->psi_eos;
    const int RPG = sl.rows_per_group;
    const int nq4 = Tp >> 2;
    constexpr float kL2E = 1.4426950408889634f;
    // one candidate per half warp: candidate i of a group belongs to half warp i % 16, round i / 16
    const int hw = tid >> 4, hl = lane & 15;
    int grp = 0;
    for (int g0 = 0; g0 < n_top; g0 += RPG, ++grp) {
      const int g1 = (g0 + RPG < n_top) ? g0 + RPG : n_top;
      if (g0 > 0) {
        __syncthreads();  // the arena rows of the previous group have been consumed
        if (warp == 0)
          issue_cand_rows2(d, tk, g0, g1, xT_b, arena, reinterpret_cast<uint64_t*>(&S->bar_cand), lane);
      }
      mbar_wait(reinterpret_cast<uint64_t*>(&S->bar_cand), (uint32_t)(grp & 1));
      if (grp == 0) LGASR_T2(4, clock64());
      for (int i0 = 0; i0 < g1 - g0; i0 += 2 * NW) {
        const int i = i0 + hw;
        const bool have = i < g1 - g0;
        const int jr = g0 + (have ? i : 0);
        const int c = key_idx(tk[jr]);
        const bool special = (c == d.eos) || (d.eos != d.blank && c == d.blank);
        // log psi = logsumexp({phi[t-1] + x[t] : start <= t < T} U {r^n[start-1]})  (:163-171), the way torch.logsumexp
        // does it (max, sum of exp, log): a one-sweep sum relative to max(phi) is 2-3 % faster but its roundings are
        // independent of the reference's (max |dscore| against the reference 9e-5 instead of 3e-5 after 64 positions).
        // phi is pre-shifted and log-zero outside [start, T), so the sums run over whole float4 rows.  (A special or
        // missing candidate runs the same code on whatever its arena slot holds and discards the result.)
        const float4* x4 = reinterpret_cast<const float4*>(arena + (jr - g0) * Tp);
        const float4* p4 = reinterpret_cast<const float4*>((c == last) ? p1 : p0);
        const float rn0 = (olen == 0) ? arena[(jr - g0) * Tp] : LGASR_LOGZERO;
        float m = rn0;
#pragma unroll 2
        for (int q = hl; q < nq4; q += 16) {
          const float4 xa = x4[q], pa = p4[q];
          m = fmaxf(fmaxf(fmaxf(m, pa.x + xa.x), fmaxf(pa.y + xa.y, pa.z + xa.z)), pa.w + xa.w);
        }
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULLMASK, m, o));
        // (e - m) first: the subtraction is exact near the maximum, where the terms matter
        float sum = (hl == 0) ? ex2_approx((rn0 - m) * kL2E) : 0.f;
#pragma unroll 2
        for (int q = hl; q < nq4; q += 16) {
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
          const float wfull = (d.key < 0) ? key_val(tk[jr]) : fuse_at2(tm, row, rowoff, c);
          put_rec(jr, __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi, s_prev))), hyp_score), c, psi);
        }
      }
    }
    if (warp == NW - 1) {
      for (int jr = n_top + lane; jr < n_sel; jr += 32) put_rec(jr, -INFINITY, -1, 0.f);  // V < S: unused slots
      // eos is always scored by the CTC scorer (:180-181); skip if it already is a candidate
      bool dup = false;
      for (int e = lane; e < n_top; e += 32) dup |= (key_idx(tk[e]) == d.eos);
      dup = __any_sync(FULLMASK, dup);
      if (lane == 0) {
        float total = -INFINITY;
        if (!dup) {
          const float wfull = (d.key < 0) ? S->w_eos : fuse_at2(tm, row, rowoff, d.eos);
          total = __fadd_rn(__fadd_rn(wfull, __fmul_rn(d.w_ctc, __fsub_rn(psi_eos, s_prev))), hyp_score);
        }
        put_rec(n_sel, total, dup ? -1 : d.eos, psi_eos);
      }
    }
  }
  LGASR_T2(5, clock64());

  // ======================= phase 5: exchange + selection ========================================================
  if (CL) {
    // phase 2 of the cluster barrier: every thread's remote stores are released by its own arrive; only rank 0 waits
    __syncwarp();
    cluster_arrive();
    LGASR_T2(6, clock64());
    LGASR_T2(16 + 15, gtimer());
    if (h != 0) return;
    cluster_wait();
    utt_select<true>(d, tm, b, pp, step, selbuf, n_run, cur_len, S->m_s.y, S->m_s.z, TR ? d.trace + (long long)row * 32 : nullptr);
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
    LGASR_T2(6, clock64());
    LGASR_T2(16 + 15, gtimer());
    if (!S->last_flag) return;
    utt_select<false>(d, tm, b, pp, step, arena, n_run, cur_len, S->m_s.y, S->m_s.z, TR ? d.trace + (long long)row * 32 : nullptr);
  }
  LGASR_T2(7, clock64());
  LGASR_T2(16 + 15, gtimer());
#undef LGASR_T2
}

}  // namespace lgasr
