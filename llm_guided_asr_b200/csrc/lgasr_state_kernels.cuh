// Kernels either side of the fused step: the full scorers' per-hypothesis caches
// re-ordered by the parents the step chose, and the (rows, V) log-softmax hand-off from raw logits.
//
//   k_reorder_rows      replaces the per-hypothesis Python state shuffling of the reference
//                       (espnet/nets/batch_beam_search.py:347-355 select_state per survivor, then
//                       espnet2/asr/decoder/transformer_decoder.py:268-272 torch.stack per layer and
//                       :285-286 the unstack) by ONE parent-indexed gather that reads the parents where
//                       k_step left them in HBM: no device->host round trip before the next decoder call.
//   k_log_softmax_rows  replaces torch.log_softmax over the (n, V) logits (transformer_decoder.py:212,
//                       espnet2/asr/llm/llama.py:260) for fp32 or bf16 logits; one row (or a row slice per
//                       CTA of a thread-block cluster for 50k-128k vocabularies) is staged in shared memory
//                       once, so HBM sees one read and one write of the row.
#pragma once
#include <cuda_bf16.h>

#include "lgasr_step.cuh"

namespace lgasr {

constexpr int kCopyThreads = 256;
constexpr int kCopyUnroll = 4;

// dst[o][row][0:row_bytes) = src[o][b*W + parent[row]][0:row_bytes) for row = b*W + j, j < n_run[b].
// VEC = uint4 (16-byte rows/strides/pointers) or unsigned (4-byte).  grid (chunks, B*W, n_outer).
template <typename VEC>
__global__ __launch_bounds__(kCopyThreads) void k_reorder_rows(const int* __restrict__ parent, const int* __restrict__ n_run,
                                                               int W, char* __restrict__ dst, const char* __restrict__ src,
                                                               long long dst_outer, long long src_outer, long long dst_row,
                                                               long long src_row, long long row_bytes) {
  const int row = blockIdx.y;
  const int b = row / W, j = row - b * W;
  if (j >= n_run[b]) return;
  const int p = parent[row];
  if (p < 0 || p >= W) return;
  const VEC* s = reinterpret_cast<const VEC*>(src + blockIdx.z * src_outer + (long long)(b * W + p) * src_row);
  VEC* t = reinterpret_cast<VEC*>(dst + blockIdx.z * dst_outer + (long long)row * dst_row);
  const long long n = row_bytes / (long long)sizeof(VEC);
  const long long stride = (long long)gridDim.x * kCopyThreads;
  long long i = (long long)blockIdx.x * kCopyThreads + threadIdx.x;
  for (; i + (kCopyUnroll - 1) * stride < n; i += kCopyUnroll * stride) {
    VEC v[kCopyUnroll];
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) v[u] = __ldcs(s + i + u * stride);
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) __stcs(t + i + u * stride, v[u]);
  }
  for (; i < n; i += stride) __stcs(t + i, __ldcs(s + i));
}

// ---- row log-softmax -------------------------------------------------------------------------------------
constexpr int kLsmThreads = 256;

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ unsigned cluster_nctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}

constexpr int kLsmChunks = 4;  // TMA chunks per slice (separate mbarriers: the maximum pass starts on chunk 0)

template <bool BF16>
struct LsmIn {
  using T = float;
  static constexpr int kVec = 4;  // elements per 16 bytes
  __device__ static __forceinline__ float at(const T* p, int i) { return p[i]; }
  __device__ static __forceinline__ void vec(const T* p, int i, float* f) {
    const float4 q = reinterpret_cast<const float4*>(p)[i];
    f[0] = q.x, f[1] = q.y, f[2] = q.z, f[3] = q.w;
  }
};
template <>
struct LsmIn<true> {
  using T = __nv_bfloat16;
  static constexpr int kVec = 8;
  __device__ static __forceinline__ float at(const T* p, int i) { return __bfloat162float(p[i]); }
  __device__ static __forceinline__ void vec(const T* p, int i, float* f) {
    const uint4 q = reinterpret_cast<const uint4*>(p)[i];
    const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // bf16 -> fp32 is a 16-bit shift
      f[2 * k] = __uint_as_float(w[k] << 16);
      f[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u);
    }
  }
};

// out[r][v] = x[r][v] - (m + log(sum_v exp(x[r][v] - m))), m = max_v x[r][v], all in fp32.
// One row per cluster (gridDim.x = cluster size CS along x, blockIdx.y = row); CTA c owns the contiguous
// slice [c*SL, min(V, (c+1)*SL)), staged ONCE in dynamic shared memory in its input type (SL elements).
// BF16: logits are bf16.  VECIO: V, SL, both leading dimensions and both pointers allow 16-byte accesses;
// the slice then arrives by kLsmChunks 1-D TMA bulk copies (cp.async.bulk + mbarrier), else by plain loads.
template <bool BF16, bool VECIO>
// out may be NULL (normaliser only); lse_out (may be NULL) receives the row's logsumexp: lgasr_step_ex then reads the
// LOGITS once, in their own type, and subtracts the normaliser in registers.
__global__ __launch_bounds__(kLsmThreads) void k_log_softmax_rows(const void* __restrict__ logits, long long ld_in,
                                                                  float* __restrict__ out, long long ld_out, int V, int SL,
                                                                  float* __restrict__ lse_out) {
  using In = LsmIn<BF16>;
  using T = typename In::T;
  constexpr int NV = In::kVec;
  extern __shared__ __align__(16) unsigned char lsm_raw[];
  T* row_s = reinterpret_cast<T*>(lsm_raw);
  __shared__ __align__(8) unsigned long long bar[kLsmChunks];
  __shared__ float red[kLsmThreads / 32];
  __shared__ float xm[16], xs[16];  // (max, sum) of every CTA of the cluster, written by the peers
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // a step kernel queued behind this one (programmatic launch) may set up its CTAs now; it reads the rows only after
  // its griddepcontrol.wait, i.e. when this grid has completed
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const unsigned CS = cluster_nctarank(), c = cluster_ctarank();
  const long long r = blockIdx.y;
  const int v0 = (int)c * SL;
  const int n = (V - v0 < SL) ? (V - v0 > 0 ? V - v0 : 0) : SL;
  const T* x = reinterpret_cast<const T*>(logits) + r * ld_in + v0;
  if (CS > 1) cluster_arrive();  // phase 1: every CTA of the cluster runs (its xm/xs may be written)

  float m = -INFINITY;
  const int nv = n / NV;  // VECIO: n % NV == 0
  // chunk k covers vectors [k*cv, min(nv, (k+1)*cv))
  const int cv = (nv + kLsmChunks - 1) / kLsmChunks;
  if (VECIO) {
    if (tid == 0) {
#pragma unroll
      for (int k = 0; k < kLsmChunks; ++k) mbar_init(reinterpret_cast<uint64_t*>(&bar[k]), 1);
      mbar_fence_init();
#pragma unroll
      for (int k = 0; k < kLsmChunks; ++k) {
        const int a = k * cv, b = (a + cv < nv) ? a + cv : nv;
        const uint32_t bytes = b > a ? (uint32_t)(b - a) * 16u : 0u;
        mbar_expect_tx(reinterpret_cast<uint64_t*>(&bar[k]), bytes);
        if (bytes) bulk_g2s(row_s + (long long)a * NV, x + (long long)a * NV, bytes, reinterpret_cast<uint64_t*>(&bar[k]));
      }
    }
    __syncthreads();
#pragma unroll 1
    for (int k = 0; k < kLsmChunks; ++k) {
      mbar_wait(reinterpret_cast<uint64_t*>(&bar[k]), 0);
      const int a = k * cv, b = (a + cv < nv) ? a + cv : nv;
      for (int i = a + tid; i < b; i += kLsmThreads) {
        float f[NV];
        In::vec(row_s, i, f);
#pragma unroll
        for (int e = 0; e < NV; ++e) m = fmaxf(m, f[e]);
      }
    }
  } else {
    for (int i = tid; i < n; i += kLsmThreads) {
      const T t = x[i];
      row_s[i] = t;
      m = fmaxf(m, In::at(&t, 0));
    }
  }
  m = warp_max(m);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < kLsmThreads / 32; ++w) m = fmaxf(m, red[w]);
  __syncthreads();
  // sum of exp(x - m) over the slice
  float s = 0.f;
  if (VECIO) {
    for (int i = tid; i < nv; i += kLsmThreads) {
      float f[NV];
      In::vec(row_s, i, f);
#pragma unroll
      for (int e = 0; e < NV; e += 2) s += __expf(f[e] - m) + __expf(f[e + 1] - m);
    }
  } else {
    for (int i = tid; i < n; i += kLsmThreads) s += __expf(In::at(row_s, i) - m);
  }
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  s = 0.f;
#pragma unroll
  for (int w = 0; w < kLsmThreads / 32; ++w) s += red[w];

  if (CS > 1) {
    // every CTA hands its (m, s) to every CTA of the cluster; all combine in rank order -> identical result
    cluster_wait();
    if (tid < (int)CS) {
      st_cluster(map_rank(&xm[c], (uint32_t)tid), __float_as_uint(m));
      st_cluster(map_rank(&xs[c], (uint32_t)tid), __float_as_uint(s));
    }
    cluster_arrive();
    cluster_wait();
    float M = xm[0];
    for (unsigned k = 1; k < CS; ++k) M = fmaxf(M, xm[k]);
    float S = 0.f;
    for (unsigned k = 0; k < CS; ++k) S += xs[k] * __expf(xm[k] - M);
    m = M;
    s = S;
  }
  const float lse = m + logf(s);
  if (lse_out && tid == 0 && c == 0) lse_out[r] = lse;
  if (!out) return;
  float* o = out + r * ld_out + v0;
  if (VECIO) {
    for (int i = tid; i < nv; i += kLsmThreads) {
      float f[NV];
      In::vec(row_s, i, f);
#pragma unroll
      for (int e = 0; e < NV; e += 4)
        __stcs(reinterpret_cast<float4*>(o) + (long long)i * (NV / 4) + e / 4,
               make_float4(f[e] - lse, f[e + 1] - lse, f[e + 2] - lse, f[e + 3] - lse));
    }
  } else {
    for (int i = tid; i < n; i += kLsmThreads) o[i] = In::at(row_s, i) - lse;
  }
}

// ---- timestamp-rule scorer of the OWSM recipe --------------------------------------------------------------
// out[r][v] = 0 or -inf by the rules of ScoreFilter.score (espnet2/bin/s2t_inference.py:88-116) applied to the
// prefix y = yseq[r][0:L).  One CTA per row: the prefix is reduced to five scalars (contains <notimestamps>,
// number of timestamp tokens, the last timestamp, y[-1], y[-3]), then the row is written with 16-byte stores.
// The reference evaluates the rules per hypothesis in Python (batch_score loops over score, :140-146).
constexpr int kFilterThreads = 128;

__global__ __launch_bounds__(kFilterThreads) void k_score_filter(const long long* __restrict__ yseq, long long ld_y, int L,
                                                                 int V, int notime, int first, int last, int sos, int eos,
                                                                 float* __restrict__ out, long long ld_out, int vec_ok) {
  __shared__ int s_has, s_cnt, s_lastpos;
  const int tid = threadIdx.x;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // see k_log_softmax_rows
  const long long* y = yseq + blockIdx.x * ld_y;
  if (tid == 0) {
    s_has = 0;
    s_cnt = 0;
    s_lastpos = -1;
  }
  __syncthreads();
  int has = 0, cnt = 0, lastpos = -1;
  for (int i = tid; i < L; i += kFilterThreads) {
    const long long t = y[i];
    has |= (t == notime);
    if (t >= first && t <= last) {
      ++cnt;
      lastpos = i;
    }
  }
  if (has) atomicOr(&s_has, 1);
  if (cnt) {
    atomicAdd(&s_cnt, cnt);
    atomicMax(&s_lastpos, lastpos);
  }
  __syncthreads();
  // rule -> (lo, hi, inside, eos_override): entry v is -inf iff ((lo <= v <= hi) == inside), except eos when
  // eos_override is 0 (forced 0.0) or 1 (forced -inf)
  int lo, hi, inside = 1, eos_ov = -1;
  const long long y1 = L >= 1 ? y[L - 1] : -1, y3 = L >= 3 ? y[L - 3] : (long long)sos + 1;
  if (s_has) {  // :93-95
    lo = first, hi = last;
  } else if (y3 == sos) {  // :96-99
    lo = first, hi = last, inside = 0;
  } else if (s_cnt & 1) {  // :102-106
    lo = first, hi = (int)y[s_lastpos];
    eos_ov = 1;
  } else if (y1 >= first && y1 <= last) {  // :109-113
    lo = (int)y1, hi = last, inside = 0;
    eos_ov = 0;
  } else {  // :114-116
    lo = 0, hi = -1, inside = 0;
  }
  const float NEG = -INFINITY;
  float* o = out + blockIdx.x * ld_out;
  auto val = [&](int v) -> float {
    if (v == eos && eos_ov >= 0) return eos_ov ? NEG : 0.f;
    return ((v >= lo && v <= hi) == (inside != 0)) ? NEG : 0.f;
  };
  if (vec_ok) {
    for (int i = tid; i < (V >> 2); i += kFilterThreads)
      reinterpret_cast<float4*>(o)[i] = make_float4(val(4 * i), val(4 * i + 1), val(4 * i + 2), val(4 * i + 3));
  } else {
    for (int v = tid; v < V; v += kFilterThreads) o[v] = val(v);
  }
}

}  // namespace lgasr
