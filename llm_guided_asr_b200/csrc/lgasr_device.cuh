// Device-side building blocks of the fused beam-search scoring step (sm_100a).
//
//  * log-semiring helpers (fp32, log-zero = -1e10 as espnet/nets/ctc_prefix_score.py:32)
//  * WarpTop<NL>: a sorted top-(32*NL) list distributed over the lanes of a warp (pre-beam / beam
//    selection without a torch.topk launch; batch_beam_search.py:101-128,302-308)
//  * warp_affine_scan: the CTC forward recursion r_t = logaddexp(r_{t-1} + a_t, b_t)
//    (ctc_prefix_score.py:155-160) as a warp-level segmented scan over the frames
//  * warp_prefix_lse: log psi = logsumexp_t(phi_{t-1} + x_t[c]) (+) r^n_{start-1}
//    (ctc_prefix_score.py:163-171) as a warp-shuffle log-sum-exp
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define LGASR_LOGZERO (-1.0e10f)
#define LGASR_NEG (-1.0e30f)  // identity of the scan algebra ("probability 0"), finite to avoid inf-inf
#define FULLMASK 0xffffffffu

namespace lgasr {

// m + log(exp(a-m) + exp(b-m)), the two-operand torch.logsumexp (one of the exps is exp(0) = 1).
// Default: MUFU ex2/lg2.  For this composite the hardware approximations are as good as the library calls:
// the log argument 1 + e lies in [1, 2], where lg2.approx has an ABSOLUTE error <= 2^-22 (1.6e-7 after the
// ln 2 scale), and the 2-ulp error of ex2.approx enters through d/de log(1+e) <= 1/2 (<= 1.2e-7); expf + logf
// give ~1.8e-7 worst case for 4x the instructions.  -DLGASR_ACCURATE_LSE selects the library calls.
__device__ __forceinline__ float lse2(float a, float b) {
  const float m = fmaxf(a, b);
  const float d = fminf(a, b) - m;  // <= 0
#ifdef LGASR_ACCURATE_LSE
  return m + logf(1.0f + expf(d));
#else
  float e, l;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(d * 1.4426950408889634f));
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(1.0f + e));
  return fmaf(l, 0.6931471805599453f, m);
#endif
}

// total order used by every selection: larger value first, then smaller index
__device__ __forceinline__ bool better(float va, int ia, float vb, int ib) {
  return va > vb || (va == vb && ia < ib);
}

// ------------------------------------------------------------------------------------------
// Sorted top list of capacity 32*NL held by one warp: rank r lives in register r/32 of lane r%32.
template <int NL>
struct WarpTop {
  float v[NL];
  int i[NL];
  float tv;  // threshold = entry at rank (cap-1)
  int ti;
  int cap;

  __device__ __forceinline__ void init(int capacity) {
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      v[j] = -INFINITY;
      i[j] = 0x7fffffff;
    }
    tv = -INFINITY;
    ti = 0x7fffffff;
    cap = capacity;
  }

  __device__ __forceinline__ bool passes(float val, int idx) const { return better(val, idx, tv, ti); }

  // insert a warp-uniform (val, idx); every lane of the warp must call this
  __device__ __forceinline__ void insert(float val, int idx, int lane) {
    float cv = 0.f;
    int ci = 0;
    int cc = 0;  // carry: the old entry just before this register's first rank, and its compare bit
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      int cmp = better(val, idx, v[j], i[j]) ? 1 : 0;
      float pv = __shfl_up_sync(FULLMASK, v[j], 1);
      int pi = __shfl_up_sync(FULLMASK, i[j], 1);
      int pc = __shfl_up_sync(FULLMASK, cmp, 1);
      if (lane == 0) {
        pv = cv;
        pi = ci;
        pc = cc;
      }
      cv = __shfl_sync(FULLMASK, v[j], 31);
      ci = __shfl_sync(FULLMASK, i[j], 31);
      cc = __shfl_sync(FULLMASK, cmp, 31);
      if (cmp) {
        if (pc) {
          v[j] = pv;
          i[j] = pi;
        } else {
          v[j] = val;
          i[j] = idx;
        }
      }
    }
    const int r = cap - 1;
    float tvv = v[0];
    int tii = i[0];
#pragma unroll
    for (int j = 1; j < NL; ++j)
      if ((r >> 5) == j) {
        tvv = v[j];
        tii = i[j];
      }
    tv = __shfl_sync(FULLMASK, tvv, r & 31);
    ti = __shfl_sync(FULLMASK, tii, r & 31);
  }

  // offer one per-lane element (valid or not); warp-collective
  __device__ __forceinline__ void offer(float val, int idx, bool valid, int lane) {
    unsigned any = __ballot_sync(FULLMASK, valid && passes(val, idx));
    while (any) {
      int src = __ffs(any) - 1;
      float bv = __shfl_sync(FULLMASK, val, src);
      int bi = __shfl_sync(FULLMASK, idx, src);
      if (passes(bv, bi)) insert(bv, bi, lane);
      any &= any - 1;
    }
  }

  // entry at rank r (warp-uniform r) broadcast to all lanes
  __device__ __forceinline__ void get(int r, float& val, int& idx) const {
    float vv = v[0];
    int ii = i[0];
#pragma unroll
    for (int j = 1; j < NL; ++j)
      if ((r >> 5) == j) {
        vv = v[j];
        ii = i[j];
      }
    val = __shfl_sync(FULLMASK, vv, r & 31);
    idx = __shfl_sync(FULLMASK, ii, r & 31);
  }
};

// ------------------------------------------------------------------------------------------
// Warp-level segmented scan of y[t] = lse2(y[t-1] + a[t], b[t]) for t in [start, T), y[start-1] = y0.
// a: shared-memory array; y: shared-memory array holding b[t] on entry, y[t] on exit.
__device__ __forceinline__ void warp_affine_scan(const float* a, float* y, int start, int T, float y0, int lane) {
  const int n = T - start;
  if (n <= 0) return;
  int L = (n + 31) >> 5;
  L |= 1;  // odd segment length: lanes hit distinct shared-memory banks
  int s0 = start + lane * L;
  int s1 = s0 + L;
  if (s0 > T) s0 = T;
  if (s1 > T) s1 = T;
  // pass 1: local recursion from the scan identity
  float cur = LGASR_NEG, A = 0.f;
  for (int t = s0; t < s1; ++t) {
    float at = a[t];
    cur = lse2(cur + at, y[t]);
    y[t] = cur;
    A += at;
  }
  // inclusive scan of the segment summaries (A, B): y_out = lse2(y_in + A, B)
  float SA = A, SB = cur;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    float pa = __shfl_up_sync(FULLMASK, SA, d);
    float pb = __shfl_up_sync(FULLMASK, SB, d);
    if (lane >= d) {
      SB = lse2(pb + SA, SB);
      SA = pa + SA;
    }
  }
  float ea = __shfl_up_sync(FULLMASK, SA, 1);
  float eb = __shfl_up_sync(FULLMASK, SB, 1);
  float yin = (lane == 0) ? y0 : lse2(y0 + ea, eb);
  // pass 2: fix-up (independent iterations)
  float run = 0.f;
  for (int t = s0; t < s1; ++t) {
    run += a[t];
    y[t] = lse2(yin + run, y[t]);
  }
}

// online log-sum-exp accumulator
struct Lse {
  float m, s;
  __device__ __forceinline__ void init() {
    m = LGASR_NEG;
    s = 0.f;
  }
  __device__ __forceinline__ void add(float x) {
    if (x > m) {
      s = s * expf(m - x) + 1.0f;
      m = x;
    } else {
      s += expf(x - m);
    }
  }
  __device__ __forceinline__ void merge(float om, float os) {
    float nm = fmaxf(m, om);
    s = s * expf(m - nm) + os * expf(om - nm);
    m = nm;
  }
  __device__ __forceinline__ float value() const { return m + logf(s); }
};

// log psi of one candidate token by one warp:
//   logsumexp( {phi[t-1] + x[t] : start <= t < T} U {rn0} )
// phi: shared memory (r_sum or r^b of the hypothesis), x: the token's row of xT in global memory.
__device__ __forceinline__ float warp_prefix_lse(const float* phi, const float* __restrict__ x, int start, int T,
                                                 float rn0, int lane) {
  Lse acc;
  acc.init();
  if (lane == 0) acc.add(rn0);
  for (int t = start + lane; t < T; t += 32) acc.add(phi[t - 1] + __ldg(x + t));
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    float om = __shfl_xor_sync(FULLMASK, acc.m, d);
    float os = __shfl_xor_sync(FULLMASK, acc.s, d);
    acc.merge(om, os);
  }
  // lanes may differ in the last bit (FMA contraction is not symmetric): make the result warp-uniform
  return __shfl_sync(FULLMASK, acc.value(), 0);
}

}  // namespace lgasr
