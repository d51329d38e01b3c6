// Stand-alone CTC prefix scorer kernels: the (n, V)-matrix contract of
// CTCPrefixScorer.batch_score_partial / select_state (espnet/nets/scorers/ctc.py:40-63,101-126)
// for callers that drive the scorer through the plug-in API instead of the fused step.
#pragma once
#include "lgasr_kernels.cuh"

namespace lgasr {

// one CTA per hypothesis
__global__ void __launch_bounds__(256) k_ctc_prefix_score(const float* __restrict__ xT, int T, int Tp, int V, int blank,
                                                          int eos, const float* __restrict__ r_prev,
                                                          const float* __restrict__ s_prev_arr,
                                                          const int* __restrict__ last_arr, int olen,
                                                          const long long* __restrict__ ids, int S,
                                                          float* __restrict__ scores, float* __restrict__ log_psi,
                                                          int win_s, int win_e) {
  extern __shared__ float smem[];
  float* phi_sum = smem;
  float* rbv = smem + Tp;
  const int h = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* rp = r_prev + (long long)h * 2 * Tp;
  const float s_prev = s_prev_arr ? s_prev_arr[h] : 0.f;
  const int last = last_arr[h];
  float* sc = scores + (long long)h * V;
  float* lp = log_psi + (long long)h * V;
  for (int t = tid; t < T; t += blockDim.x) {
    float rn = rp[t], rb = rp[Tp + t];
    phi_sum[t] = lse2(rn, rb);
    rbv[t] = rb;
  }
  if (S > 0) {
    const float fillv = __fsub_rn(LGASR_LOGZERO, s_prev);
    for (int v = tid; v < V; v += blockDim.x) {
      sc[v] = fillv;
      lp[v] = LGASR_LOGZERO;
    }
  }
  __syncthreads();
  // attention window (ctc_prefix_score.py:143-148): the sum runs over frames [win_s, win_e) (0, 0: [max(olen, 1), T))
  const int start = win_s > 0 ? win_s : (olen > 1 ? olen : 1);
  const int Tend = (win_e > 0 && win_e < T) ? win_e : T;
  const int ncand = S > 0 ? S : V;
  for (int j = warp; j < ncand; j += blockDim.x / 32) {
    const int c = S > 0 ? (int)ids[(long long)h * S + j] : j;
    if (c < 0 || c >= V) continue;
    const float* phi = (c == last) ? rbv : phi_sum;
    const float* xr = xT + (long long)c * Tp;
    const float rn0 = (olen == 0 && start == 1) ? __ldg(xr) : LGASR_LOGZERO;  // r[start-1, 0]
    const float psi = warp_prefix_lse(phi, xr, start, Tend, rn0, lane);
    if (lane == 0) {
      lp[c] = psi;
      sc[c] = __fsub_rn(psi, s_prev);
    }
  }
  __syncthreads();
  if (tid == 0) {
    const float pe = phi_sum[T - 1];  // eos is always scored (ctc_prefix_score.py:180-181)
    lp[eos] = pe;
    sc[eos] = __fsub_rn(pe, s_prev);
    if (eos != blank) {  // :183-185
      lp[blank] = LGASR_LOGZERO;
      sc[blank] = __fsub_rn(LGASR_LOGZERO, s_prev);
    }
  }
}

// one warp per selected (hypothesis, token)
__global__ void __launch_bounds__(32) k_ctc_select_state(const float* __restrict__ xT, int T, int Tp, int blank,
                                                         const float* __restrict__ r_prev,
                                                         const int* __restrict__ last_arr,
                                                         const int* __restrict__ sel_hyp,
                                                         const int* __restrict__ sel_tok, int olen,
                                                         float* __restrict__ r_new, int win_s, int win_e) {
  extern __shared__ float smem[];
  const int i = blockIdx.x, lane = threadIdx.x;
  const int h = sel_hyp[i], c = sel_tok[i];
  warp_ctc_state(xT, r_prev + (long long)h * 2 * Tp, r_new + (long long)i * 2 * Tp, 1, c, last_arr[h], olen, T, Tp, blank,
                 smem, smem + Tp, smem + 2 * Tp, smem + 3 * Tp, lane, win_s, win_e);
}

// CTCPrefixScoreTH.extend_state (ctc_prefix_score.py:246-269): one block per survivor state
__global__ void k_ctc_extend_state(const float* __restrict__ xT, int T_old, int T_new, int Tp_old, int Tp_new, int blank,
                                   const float* __restrict__ r_old, float* __restrict__ r_new) {
  extern __shared__ float xb[];  // blank posteriors of the new frames
  const int i = blockIdx.x;
  const float* ro = r_old + (long long)i * 2 * Tp_old;
  float* rn = r_new + (long long)i * 2 * Tp_new;
  const int start = T_old > 1 ? T_old : 1;
  for (int t = threadIdx.x; t < Tp_new; t += blockDim.x) {
    rn[t] = (t < start) ? ro[t] : LGASR_LOGZERO;
    if (t < start) rn[Tp_new + t] = ro[Tp_old + t];
    if (t >= T_new) rn[Tp_new + t] = LGASR_LOGZERO;
    if (t >= start && t < T_new) xb[t - start] = xT[(long long)blank * Tp_new + t];
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // sequential on purpose: the fp32 roundings of the reference's frame-by-frame loop
    float acc = ro[Tp_old + start - 1];
    for (int t = start; t < T_new; ++t) {
      acc = __fadd_rn(acc, xb[t - start]);
      rn[Tp_new + t] = acc;
    }
  }
}

}  // namespace lgasr
