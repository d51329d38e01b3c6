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

// out[r][v] = x[r][v] - (m + log(sum_v exp(x[r][v] - m))), m = max_v x[r][v], all in fp32.
// One row per cluster (gridDim.x = cluster size CS along x, blockIdx.y = row); CTA c owns the contiguous
// slice [c*SL, min(V, (c+1)*SL)) staged in dynamic shared memory (SL floats).  BF16: logits are bf16.
// VECIO: V, SL, both leading dimensions and both pointers allow 16-byte accesses.
template <bool BF16, bool VECIO>
__global__ __launch_bounds__(kLsmThreads) void k_log_softmax_rows(const void* __restrict__ logits, long long ld_in,
                                                                  float* __restrict__ out, long long ld_out, int V, int SL) {
  extern __shared__ __align__(16) float row_s[];
  __shared__ float red[kLsmThreads / 32];
  __shared__ float xm[16], xs[16];  // (max, sum) of every CTA of the cluster, written by the peers
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned CS = cluster_nctarank(), c = cluster_ctarank();
  const long long r = blockIdx.y;
  const int v0 = (int)c * SL;
  const int n = (V - v0 < SL) ? (V - v0 > 0 ? V - v0 : 0) : SL;
  if (CS > 1) cluster_arrive();  // phase 1: every CTA of the cluster runs (its xm/xs may be written)

  // stage the slice as fp32, running maximum on the way
  float m = -INFINITY;
  if (BF16) {
    const __nv_bfloat16* x = reinterpret_cast<const __nv_bfloat16*>(logits) + r * ld_in + v0;
    if (VECIO) {
      const uint4* x8 = reinterpret_cast<const uint4*>(x);
      for (int i = tid; i < (n >> 3); i += kLsmThreads) {
        const uint4 q = __ldcs(x8 + i);
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.y));
        const float2 cc = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.z));
        const float2 dd = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.w));
        reinterpret_cast<float4*>(row_s)[2 * i] = make_float4(a.x, a.y, b.x, b.y);
        reinterpret_cast<float4*>(row_s)[2 * i + 1] = make_float4(cc.x, cc.y, dd.x, dd.y);
        m = fmaxf(m, fmaxf(fmaxf(fmaxf(a.x, a.y), fmaxf(b.x, b.y)), fmaxf(fmaxf(cc.x, cc.y), fmaxf(dd.x, dd.y))));
      }
    } else {
      for (int i = tid; i < n; i += kLsmThreads) {
        const float f = __bfloat162float(x[i]);
        row_s[i] = f;
        m = fmaxf(m, f);
      }
    }
  } else {
    const float* x = reinterpret_cast<const float*>(logits) + r * ld_in + v0;
    if (VECIO) {
      const float4* x4 = reinterpret_cast<const float4*>(x);
      for (int i = tid; i < (n >> 2); i += kLsmThreads) {
        const float4 q = __ldcs(x4 + i);
        reinterpret_cast<float4*>(row_s)[i] = q;
        m = fmaxf(m, fmaxf(fmaxf(q.x, q.y), fmaxf(q.z, q.w)));
      }
    } else {
      for (int i = tid; i < n; i += kLsmThreads) {
        const float f = x[i];
        row_s[i] = f;
        m = fmaxf(m, f);
      }
    }
  }
  m = warp_max(m);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < kLsmThreads / 32; ++w) m = fmaxf(m, red[w]);
  __syncthreads();
  // sum of exp(x - m) over the slice (each thread re-reads what it staged itself or its neighbours did:
  // the barrier above orders the shared-memory writes)
  float s = 0.f;
  if (VECIO) {
    for (int i = tid; i < (n >> 2); i += kLsmThreads) {
      const float4 q = reinterpret_cast<const float4*>(row_s)[i];
      s += (__expf(q.x - m) + __expf(q.y - m)) + (__expf(q.z - m) + __expf(q.w - m));
    }
  } else {
    for (int i = tid; i < n; i += kLsmThreads) s += __expf(row_s[i] - m);
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
  const float lse = m + __logf(s);
  float* o = out + r * ld_out + v0;
  if (VECIO) {
    for (int i = tid; i < (n >> 2); i += kLsmThreads) {
      float4 q = reinterpret_cast<const float4*>(row_s)[i];
      q.x -= lse;
      q.y -= lse;
      q.z -= lse;
      q.w -= lse;
      __stcs(reinterpret_cast<float4*>(o) + i, q);
    }
  } else {
    for (int i = tid; i < n; i += kLsmThreads) o[i] = row_s[i] - lse;
  }
}

}  // namespace lgasr
