// Prediction-aware encoding (inter-CTC + Adapter "inter_league") helpers:
//   s2t_softmax_rows_fwd   dist = softmax(logit) over the vocabulary, from the row log-sum-exp the CTC loss of the
//                          same logits needs anyway (fairseq/modules/speech_to_text/adapter.py:215-217, quirk A13:
//                          softmax in the logits' dtype -> fp32 math, bf16 result)
//   s2t_softmax_rows_bwd   dlogit (+)= dist * (ddist - sum_v ddist * dist)
// The V -> C contraction soft = dist @ embed_adapter.weight (adapter.py:264-266) and x + soft (:296-297) run in the
// tcgen05 GEMM (residual epilogue); these two kernels are HBM-bound: one CTA per (b,t) row, 16-byte accesses.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

__device__ __forceinline__ void unpack8v(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8v(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]);
  v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}

__global__ void __launch_bounds__(256)
softmax_rows_fwd_kernel(const __nv_bfloat16* __restrict__ logits, long long ldv, const float* __restrict__ lse,
                        __nv_bfloat16* __restrict__ probs, long long ldp, int V) {
  pdl_entry();
  const long long row = blockIdx.x;
  const float l = lse[row];
  const __nv_bfloat16* x = logits + row * ldv;
  __nv_bfloat16* y = probs + row * ldp;
  const int nv = static_cast<int>(ldp >> 3);
  for (int i = threadIdx.x; i < nv; i += 256) {
    float f[8];
    unpack8v(*reinterpret_cast<const uint4*>(x + i * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = (i * 8 + j < V) ? __expf(f[j] - l) : 0.f;
    *reinterpret_cast<uint4*>(y + i * 8) = pack8v(f);
  }
}

__global__ void __launch_bounds__(256)
softmax_rows_bwd_kernel(const __nv_bfloat16* __restrict__ probs, const __nv_bfloat16* __restrict__ dprobs,
                        long long ldp, __nv_bfloat16* __restrict__ dlogits, long long ldv, int V, int accumulate) {
  pdl_entry();
  __shared__ float s_part[8];
  __shared__ float s_dot;
  const long long row = blockIdx.x;
  const __nv_bfloat16* p = probs + row * ldp;
  const __nv_bfloat16* g = dprobs + row * ldp;
  __nv_bfloat16* o = dlogits + row * ldv;
  const int nv = (V + 7) >> 3;
  float dot = 0.f;
  for (int i = threadIdx.x; i < nv; i += 256) {
    float a[8], b[8];
    unpack8v(*reinterpret_cast<const uint4*>(p + i * 8), a);
    unpack8v(*reinterpret_cast<const uint4*>(g + i * 8), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) dot += (i * 8 + j < V) ? a[j] * b[j] : 0.f;
  }
  dot = warp_sum(dot);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = dot;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += s_part[w];
    s_dot = t;
  }
  __syncthreads();
  dot = s_dot;
  for (int i = threadIdx.x; i < nv; i += 256) {
    float a[8], b[8], c[8];
    unpack8v(*reinterpret_cast<const uint4*>(p + i * 8), a);
    unpack8v(*reinterpret_cast<const uint4*>(g + i * 8), b);
    if (accumulate) unpack8v(*reinterpret_cast<const uint4*>(o + i * 8), c);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float v = (i * 8 + j < V) ? a[j] * (b[j] - dot) : 0.f;
      c[j] = accumulate ? c[j] + v : v;
    }
    *reinterpret_cast<uint4*>(o + i * 8) = pack8v(c);
  }
}


// Prediction-aware encoding with ground-truth mixing (curriculum, s2t_transformer.py:1904-1946 / 2061-2103 and
// adapter.py:244-262): rows whose mask is set take the (optionally smoothed) one-hot distribution of their aligned ground-truth token
// instead of the model's softmax: smooth -> 0.9 + 0.1 / V on the token, 0.1 / V elsewhere (pae-oracle-smooth).
__global__ void __launch_bounds__(256)
rows_truth_mix_kernel(__nv_bfloat16* __restrict__ probs, long long ldp, const int* __restrict__ truth_tok,
                       const int* __restrict__ mask, int V, int smooth) {
  pdl_entry();
  const long long row = blockIdx.x;
  if (mask[row] == 0) return;
  const int tok = truth_tok[row];
  const float on = smooth ? 0.9f + 0.1f / static_cast<float>(V) : 1.f;
  const float off = smooth ? 0.1f / static_cast<float>(V) : 0.f;
  __nv_bfloat16* y = probs + row * ldp;
  for (int v = threadIdx.x; v < static_cast<int>(ldp); v += 256)
    y[v] = __float2bfloat16(v < V ? (v == tok ? on : off) : 0.f);
}

// x[r][:] = 0 for rows whose mask is set (no gradient reaches the logits of a frame whose distribution was replaced)
__global__ void __launch_bounds__(256)
rows_zero_masked_kernel(__nv_bfloat16* __restrict__ x, long long ld, const int* __restrict__ mask, int ncols) {
  pdl_entry();
  const long long row = blockIdx.x;
  if (mask[row] == 0) return;
  __nv_bfloat16* y = x + row * ld;
  for (int v = threadIdx.x; v < ncols; v += 256) y[v] = __float2bfloat16(0.f);
}

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_softmax_rows_fwd(const void* logits, long long ldv, const float* lse, void* probs,
                                    long long ldp, int rows, int V, void* stream) {
  S2T_REQUIRE(logits && lse && probs, "s2t_softmax_rows_fwd: null pointer");
  S2T_REQUIRE(V > 0 && ldv % 8 == 0 && ldp % 8 == 0 && ldp >= V && ldv >= ldp,
              "s2t_softmax_rows_fwd: need V <= ldp <= ldv, both multiples of 8");
  if (rows <= 0) return 0;
  S2T_LAUNCH(softmax_rows_fwd_kernel, rows, 256, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(logits), ldv, lse, reinterpret_cast<__nv_bfloat16*>(probs), ldp, V);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_softmax_rows_bwd(const void* probs, const void* dprobs, long long ldp, void* dlogits,
                                    long long ldv, int rows, int V, int accumulate, void* stream) {
  S2T_REQUIRE(probs && dprobs && dlogits, "s2t_softmax_rows_bwd: null pointer");
  S2T_REQUIRE(V > 0 && ldv % 8 == 0 && ldp % 8 == 0 && ldp >= ((V + 7) & ~7) && ldv >= ((V + 7) & ~7),
              "s2t_softmax_rows_bwd: row strides must be multiples of 8 and cover round8(V)");
  if (rows <= 0) return 0;
  S2T_LAUNCH(softmax_rows_bwd_kernel, rows, 256, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(probs), reinterpret_cast<const __nv_bfloat16*>(dprobs), ldp, reinterpret_cast<__nv_bfloat16*>(dlogits), ldv, V, accumulate);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_rows_truth_mix(void* probs, long long ldp, const int* truth_tok, const int* mask, int rows, int V,
                                   int smooth, void* stream) {
  S2T_REQUIRE(probs && truth_tok && mask, "s2t_rows_truth_mix: null pointer");
  S2T_REQUIRE(V > 0 && ldp >= V, "s2t_rows_truth_mix: ldp must cover V");
  if (rows <= 0) return 0;
  S2T_LAUNCH(rows_truth_mix_kernel, rows, 256, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<__nv_bfloat16*>(probs), ldp, truth_tok, mask, V, smooth);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_rows_zero_masked(void* x, long long ld, const int* mask, int rows, int ncols, void* stream) {
  S2T_REQUIRE(x && mask, "s2t_rows_zero_masked: null pointer");
  S2T_REQUIRE(ncols >= 0 && ld >= ncols, "s2t_rows_zero_masked: ld must cover ncols");
  if (rows <= 0 || ncols == 0) return 0;
  S2T_LAUNCH(rows_zero_masked_kernel, rows, 256, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<__nv_bfloat16*>(x), ld, mask, ncols);
  S2T_LAUNCH_CHECK();
  return 0;
}
