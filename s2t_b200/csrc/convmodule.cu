// Conformer convolution-module kernels (fairseq/modules/convolution.py:8-120), channels-last:
//   s2t_dwconv1d_fwd / _bwd   depthwise Conv1d(C, C, k, padding=(k-1)/2, groups=C, bias=False) along time
//                             (convolution.py:48-56,96); the forward also emits per-tile channel sums for the
//                             BatchNorm that follows, the backward emits dx and per-tile weight-gradient partials
//   s2t_batchnorm_fwd / _bwd  nn.BatchNorm1d(C) over all B*T rows incl. padded frames in
//                             training mode (batch statistics, running-stat update) or eval mode (running
//                             statistics), fused with the activation that follows (ReLU / Swish, :99-100)
//   s2t_relpos_qbias_fwd/_bwd q + pos_bias_u, q + pos_bias_v (espnet_multihead_attention.py:343-347)
//   s2t_act_dropout           y = dropout(act(x)) (FeedForwardModule with a Swish activation keeps the
//                             pre-activation for the backward pass, s2t_transformer_layer.py:59-65)
// All are HBM/L2-bound: 4-byte bf16x2 accesses, a warp covers 64 consecutive channels of one row (128 B),
// time tiles staged in shared memory so every input element is read from HBM once.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kDwTile = 16;     // output time steps per block
constexpr int kDwThreads = 256;
constexpr int kDwMaxK = 31;

__device__ __forceinline__ float2 ld_bf2(const __nv_bfloat16* p) {
  const uint32_t v = *reinterpret_cast<const uint32_t*>(p);
  return make_float2(bf16_lo(v), bf16_hi(v));
}
__device__ __forceinline__ void st_bf2(__nv_bfloat16* p, float a, float b) {
  *reinterpret_cast<uint32_t*>(p) = pack_bf16(a, b);
}

// number of row groups a block splits its tile into: the largest power of two <= min(tile, threads / pairs)
__device__ __forceinline__ int row_groups(int pairs, int tile) {
  int g = 1;
  while (g * 2 <= tile && g * 2 * pairs <= static_cast<int>(blockDim.x)) g *= 2;
  return g;
}

// stage rows [t0 - halo, t0 + kDwTile + halo) of utterance b (zero outside [0, T)) into smem [rows][C]
__device__ __forceinline__ void stage_tile(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* s, int b, int t0,
                                           int T, int C, int halo) {
  const int rows = kDwTile + 2 * halo;
  const int cv = C >> 3;
  for (int i = threadIdx.x; i < rows * cv; i += blockDim.x) {
    const int r = i / cv, c8 = (i - r * cv) << 3;
    const int t = t0 - halo + r;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (t >= 0 && t < T) v = *reinterpret_cast<const uint4*>(x + (static_cast<long long>(b) * T + t) * C + c8);
    *reinterpret_cast<uint4*>(s + r * C + c8) = v;
  }
}

// y[b][t][c] = sum_j w[c][j] * x[b][t + j - pad][c]   (flip = 0: the forward conv)
// y[b][t][c] = sum_j w[c][j] * x[b][t - j + pad][c]   (flip = 1: the input gradient, x = dy)
// stats (optional): per-block fp32 [2][C] = sum_t y, sum_t y^2 over the tile's valid rows.
__global__ void __launch_bounds__(kDwThreads)
dwconv1d_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                __nv_bfloat16* __restrict__ y, float* __restrict__ stats, int T, int C, int K, int flip) {
  pdl_entry();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int pad = (K - 1) >> 1;
  const int tiles = (T + kDwTile - 1) / kDwTile;
  const int b = blockIdx.x / tiles, t0 = (blockIdx.x % tiles) * kDwTile;
  __nv_bfloat16* sx = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  float* sst = reinterpret_cast<float*>(smem_raw + static_cast<size_t>(kDwTile + 2 * pad) * C * 2);  // [G][2][C]
  stage_tile(x, sx, b, t0, T, C, pad);
  __syncthreads();
  const int pairs = C >> 1;
  const int G = row_groups(pairs, kDwTile);
  const int rpg = kDwTile / G;
  for (int item = threadIdx.x; item < pairs * G; item += blockDim.x) {
    const int cp = item % pairs, g = item / pairs, c = cp << 1;
    float w0[kDwMaxK], w1[kDwMaxK];
#pragma unroll
    for (int j = 0; j < kDwMaxK; ++j) {
      if (j < K) {
        const int jj = flip ? K - 1 - j : j;
        w0[j] = __bfloat162float(w[static_cast<long long>(c) * K + jj]);
        w1[j] = __bfloat162float(w[static_cast<long long>(c + 1) * K + jj]);
      }
    }
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
    for (int r = g * rpg; r < (g + 1) * rpg; ++r) {
      const int t = t0 + r;
      if (t >= T) break;
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int j = 0; j < kDwMaxK; ++j) {
        if (j < K) {
          const float2 v = ld_bf2(sx + (r + j) * C + c);
          a0 = fmaf(w0[j], v.x, a0);
          a1 = fmaf(w1[j], v.y, a1);
        }
      }
      st_bf2(y + (static_cast<long long>(b) * T + t) * C + c, a0, a1);
      s0 += a0; s1 += a1; q0 += a0 * a0; q1 += a1 * a1;
    }
    if (stats != nullptr) {
      sst[(g * 2 + 0) * C + c] = s0; sst[(g * 2 + 0) * C + c + 1] = s1;
      sst[(g * 2 + 1) * C + c] = q0; sst[(g * 2 + 1) * C + c + 1] = q1;
    }
  }
  if (stats != nullptr) {
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
      float a = 0.f;
      for (int g = 0; g < G; ++g) a += sst[g * 2 * C + i];
      stats[static_cast<long long>(blockIdx.x) * 2 * C + i] = a;
    }
  }
}

// weight-gradient partials: dwp[block][j][c] = sum_{t in tile} dy[b][t][c] * x[b][t + j - pad][c]
__global__ void __launch_bounds__(kDwThreads)
dwconv1d_wgrad_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                      float* __restrict__ dwp, int T, int C, int K) {
  pdl_entry();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int pad = (K - 1) >> 1;
  const int tiles = (T + kDwTile - 1) / kDwTile;
  const int b = blockIdx.x / tiles, t0 = (blockIdx.x % tiles) * kDwTile;
  __nv_bfloat16* sx = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* sd = sx + (kDwTile + 2 * pad) * C;
  stage_tile(x, sx, b, t0, T, C, pad);
  stage_tile(dy, sd, b, t0, T, C, 0);
  __syncthreads();
  const int pairs = C >> 1;
  float* out = dwp + static_cast<long long>(blockIdx.x) * K * C;
  for (int cp = threadIdx.x; cp < pairs; cp += blockDim.x) {
    const int c = cp << 1;
    float a0[kDwMaxK], a1[kDwMaxK];
#pragma unroll
    for (int j = 0; j < kDwMaxK; ++j) a0[j] = a1[j] = 0.f;
    for (int r = 0; r < kDwTile; ++r) {
      const float2 g = ld_bf2(sd + r * C + c);
#pragma unroll
      for (int j = 0; j < kDwMaxK; ++j) {
        if (j < K) {
          const float2 v = ld_bf2(sx + (r + j) * C + c);
          a0[j] = fmaf(g.x, v.x, a0[j]);
          a1[j] = fmaf(g.y, v.y, a1[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kDwMaxK; ++j) {
      if (j < K) *reinterpret_cast<float2*>(out + j * C + c) = make_float2(a0[j], a1[j]);
    }
  }
}

// dw[c][j] (bf16, the reference's (C,1,k) layout) = sum_blocks dwp[block][j][c]
__global__ void dwconv1d_wgrad_reduce_kernel(const float* __restrict__ dwp, int nblk, int C, int K,
                                             __nv_bfloat16* __restrict__ dw) {
  pdl_entry();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // i = j*C + c
  if (i >= K * C) return;
  float a = 0.f;
  for (int n = 0; n < nblk; ++n) a += dwp[static_cast<long long>(n) * K * C + i];
  const int j = i / C, c = i - j * C;
  dw[static_cast<long long>(c) * K + j] = __float2bfloat16(a);
}

// ---------------------------------------------------------------------------------------------
// BatchNorm1d + activation
// ---------------------------------------------------------------------------------------------
constexpr int kBnRows = 16;
constexpr int kBnThreads = 256;

__device__ __forceinline__ float act_apply(float z, int act) {
  if (act == S2T_ACT_RELU) return fmaxf(z, 0.f);
  if (act == S2T_ACT_SWISH) return z / (1.f + __expf(-z));
  return z;
}
__device__ __forceinline__ float act_grad(float z, int act) {
  if (act == S2T_ACT_RELU) return z > 0.f ? 1.f : 0.f;
  if (act == S2T_ACT_SWISH) {
    const float s = 1.f / (1.f + __expf(-z));
    return s * (1.f + z * (1.f - s));
  }
  return 1.f;
}

// mean / rstd of channels c, c+1: from the per-tile partial sums (training) or the running statistics (eval)
__device__ __forceinline__ void bn_channel_stats(const float* __restrict__ partials, int nparts, long long n,
                                                 const float* __restrict__ rmean, const float* __restrict__ rvar,
                                                 int training, float eps, int C, int c, float (&mean)[2],
                                                 float (&var)[2]) {
  if (training) {
    double s0 = 0, s1 = 0, q0 = 0, q1 = 0;
    for (int p = 0; p < nparts; ++p) {
      const float2 s = *reinterpret_cast<const float2*>(partials + (static_cast<long long>(p) * 2 + 0) * C + c);
      const float2 q = *reinterpret_cast<const float2*>(partials + (static_cast<long long>(p) * 2 + 1) * C + c);
      s0 += s.x; s1 += s.y; q0 += q.x; q1 += q.y;
    }
    const double m0 = s0 / n, m1 = s1 / n;
    mean[0] = static_cast<float>(m0); mean[1] = static_cast<float>(m1);
    var[0] = fmaxf(static_cast<float>(q0 / n - m0 * m0), 0.f);
    var[1] = fmaxf(static_cast<float>(q1 / n - m1 * m1), 0.f);
  } else {
    mean[0] = rmean[c]; mean[1] = rmean[c + 1];
    var[0] = rvar[c]; var[1] = rvar[c + 1];
  }
}

__global__ void __launch_bounds__(kBnThreads)
batchnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ partials, int nparts,
                     const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                     float* __restrict__ rmean, float* __restrict__ rvar, long long* __restrict__ nbt,
                     float* __restrict__ save, __nv_bfloat16* __restrict__ y, long long rows, int C, float eps,
                     float momentum, int act, int training) {
  pdl_entry();
  const int pairs = C >> 1;
  const int G = row_groups(pairs, kBnRows);
  const int rpg = kBnRows / G;
  const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
  for (int item = threadIdx.x; item < pairs * G; item += blockDim.x) {
    const int cp = item % pairs, g = item / pairs, c = cp << 1;
    float mean[2], var[2];
    bn_channel_stats(partials, nparts, rows, rmean, rvar, training, eps, C, c, mean, var);
    const float rs0 = rsqrtf(var[0] + eps), rs1 = rsqrtf(var[1] + eps);
    const float2 ga = ld_bf2(gamma + c), be = ld_bf2(beta + c);
    if (blockIdx.x == 0 && g == 0) {
      if (save != nullptr) {
        save[c] = mean[0]; save[c + 1] = mean[1];
        save[C + c] = rs0; save[C + c + 1] = rs1;
      }
      if (training && rmean != nullptr) {  // nn.BatchNorm1d running-stat update (unbiased variance)
        const float ub = rows > 1 ? static_cast<float>(rows) / static_cast<float>(rows - 1) : 1.f;
        rmean[c] = (1.f - momentum) * rmean[c] + momentum * mean[0];
        rmean[c + 1] = (1.f - momentum) * rmean[c + 1] + momentum * mean[1];
        rvar[c] = (1.f - momentum) * rvar[c] + momentum * var[0] * ub;
        rvar[c + 1] = (1.f - momentum) * rvar[c + 1] + momentum * var[1] * ub;
        if (c == 0 && nbt != nullptr) *nbt += 1;
      }
    }
    for (int r = g * rpg; r < (g + 1) * rpg; ++r) {
      const long long row = r0 + r;
      if (row >= rows) break;
      const float2 v = ld_bf2(x + row * C + c);
      const float z0 = (v.x - mean[0]) * rs0 * ga.x + be.x;
      const float z1 = (v.y - mean[1]) * rs1 * ga.y + be.y;
      st_bf2(y + row * C + c, act_apply(z0, act), act_apply(z1, act));
    }
  }
}

// partial sums of dz and dz * xhat over a row chunk; dz = da * act'(z), z = xhat * gamma + beta
__global__ void __launch_bounds__(kBnThreads)
batchnorm_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ x,
                            const float* __restrict__ save, const __nv_bfloat16* __restrict__ gamma,
                            const __nv_bfloat16* __restrict__ beta, float* __restrict__ partials,
                            long long rows, int C, int act) {
  pdl_entry();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* sst = reinterpret_cast<float*>(smem_raw);  // [G][2][C]
  const int pairs = C >> 1;
  const int G = row_groups(pairs, kBnRows);
  const int rpg = kBnRows / G;
  const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
  for (int item = threadIdx.x; item < pairs * G; item += blockDim.x) {
    const int cp = item % pairs, g = item / pairs, c = cp << 1;
    const float m0 = save[c], m1 = save[c + 1], rs0 = save[C + c], rs1 = save[C + c + 1];
    const float2 ga = ld_bf2(gamma + c), be = ld_bf2(beta + c);
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
    for (int r = g * rpg; r < (g + 1) * rpg; ++r) {
      const long long row = r0 + r;
      if (row >= rows) break;
      const float2 v = ld_bf2(x + row * C + c), d = ld_bf2(da + row * C + c);
      const float xh0 = (v.x - m0) * rs0, xh1 = (v.y - m1) * rs1;
      const float dz0 = d.x * act_grad(xh0 * ga.x + be.x, act), dz1 = d.y * act_grad(xh1 * ga.y + be.y, act);
      s0 += dz0; s1 += dz1; q0 += dz0 * xh0; q1 += dz1 * xh1;
    }
    sst[(g * 2 + 0) * C + c] = s0; sst[(g * 2 + 0) * C + c + 1] = s1;
    sst[(g * 2 + 1) * C + c] = q0; sst[(g * 2 + 1) * C + c + 1] = q1;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
    float a = 0.f;
    for (int g = 0; g < G; ++g) a += sst[g * 2 * C + i];
    partials[static_cast<long long>(blockIdx.x) * 2 * C + i] = a;
  }
}

// dx = gamma * rstd * (dz - mean(dz) - xhat * mean(dz * xhat))   (training)  |  gamma * rstd * dz   (eval)
__global__ void __launch_bounds__(kBnThreads)
batchnorm_bwd_apply_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ x,
                           const float* __restrict__ save, const __nv_bfloat16* __restrict__ gamma,
                           const __nv_bfloat16* __restrict__ beta, const float* __restrict__ partials, int nparts,
                           __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dgamma,
                           __nv_bfloat16* __restrict__ dbeta, long long rows, int C, int act, int training) {
  pdl_entry();
  const int pairs = C >> 1;
  const int G = row_groups(pairs, kBnRows);
  const int rpg = kBnRows / G;
  const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
  for (int item = threadIdx.x; item < pairs * G; item += blockDim.x) {
    const int cp = item % pairs, g = item / pairs, c = cp << 1;
    double s0 = 0, s1 = 0, q0 = 0, q1 = 0;
    for (int p = 0; p < nparts; ++p) {
      const float2 s = *reinterpret_cast<const float2*>(partials + (static_cast<long long>(p) * 2 + 0) * C + c);
      const float2 q = *reinterpret_cast<const float2*>(partials + (static_cast<long long>(p) * 2 + 1) * C + c);
      s0 += s.x; s1 += s.y; q0 += q.x; q1 += q.y;
    }
    if (blockIdx.x == 0 && g == 0) {
      st_bf2(dbeta + c, static_cast<float>(s0), static_cast<float>(s1));
      st_bf2(dgamma + c, static_cast<float>(q0), static_cast<float>(q1));
    }
    const float inv_n = training ? 1.f / static_cast<float>(rows) : 0.f;
    const float ms0 = static_cast<float>(s0) * inv_n, ms1 = static_cast<float>(s1) * inv_n;
    const float mq0 = static_cast<float>(q0) * inv_n, mq1 = static_cast<float>(q1) * inv_n;
    const float m0 = save[c], m1 = save[c + 1], rs0 = save[C + c], rs1 = save[C + c + 1];
    const float2 ga = ld_bf2(gamma + c), be = ld_bf2(beta + c);
    for (int r = g * rpg; r < (g + 1) * rpg; ++r) {
      const long long row = r0 + r;
      if (row >= rows) break;
      const float2 v = ld_bf2(x + row * C + c), d = ld_bf2(da + row * C + c);
      const float xh0 = (v.x - m0) * rs0, xh1 = (v.y - m1) * rs1;
      const float dz0 = d.x * act_grad(xh0 * ga.x + be.x, act), dz1 = d.y * act_grad(xh1 * ga.y + be.y, act);
      st_bf2(dx + row * C + c, ga.x * rs0 * (dz0 - ms0 - xh0 * mq0), ga.y * rs1 * (dz1 - ms1 - xh1 * mq1));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// relative-position query biases, activation + dropout
// ---------------------------------------------------------------------------------------------
// quv[r][0:C] = q[r] + u, quv[r][C:2C] = q[r] + v      (q = first C columns of rows with stride ldq)
__global__ void relpos_qbias_fwd_kernel(const __nv_bfloat16* __restrict__ q, long long ldq,
                                        const __nv_bfloat16* __restrict__ u, const __nv_bfloat16* __restrict__ v,
                                        __nv_bfloat16* __restrict__ quv, long long rows, int C) {
  pdl_entry();
  const int cv = C >> 3;
  const long long n = rows * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / cv;
    const int c = static_cast<int>(i - r * cv) << 3;
    const uint4 qv = *reinterpret_cast<const uint4*>(q + r * ldq + c);
    const uint4 uv = *reinterpret_cast<const uint4*>(u + c);
    const uint4 vv = *reinterpret_cast<const uint4*>(v + c);
    const uint32_t* qa = &qv.x; const uint32_t* ua = &uv.x; const uint32_t* va = &vv.x;
    uint4 ou, ov;
    uint32_t* oua = &ou.x; uint32_t* ova = &ov.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float a = bf16_lo(qa[k]), b = bf16_hi(qa[k]);
      oua[k] = pack_bf16(a + bf16_lo(ua[k]), b + bf16_hi(ua[k]));
      ova[k] = pack_bf16(a + bf16_lo(va[k]), b + bf16_hi(va[k]));
    }
    *reinterpret_cast<uint4*>(quv + r * 2 * C + c) = ou;
    *reinterpret_cast<uint4*>(quv + r * 2 * C + C + c) = ov;
  }
}

// dq[r] (row stride lddq) = dquv[r][0:C] + dquv[r][C:2C]
__global__ void relpos_qbias_bwd_kernel(const __nv_bfloat16* __restrict__ dquv, __nv_bfloat16* __restrict__ dq,
                                        long long lddq, long long rows, int C) {
  pdl_entry();
  const int cv = C >> 3;
  const long long n = rows * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / cv;
    const int c = static_cast<int>(i - r * cv) << 3;
    const uint4 a = *reinterpret_cast<const uint4*>(dquv + r * 2 * C + c);
    const uint4 b = *reinterpret_cast<const uint4*>(dquv + r * 2 * C + C + c);
    const uint32_t* aa = &a.x; const uint32_t* ba = &b.x;
    uint4 o;
    uint32_t* oa = &o.x;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      oa[k] = pack_bf16(bf16_lo(aa[k]) + bf16_lo(ba[k]), bf16_hi(aa[k]) + bf16_hi(ba[k]));
    *reinterpret_cast<uint4*>(dq + r * lddq + c) = o;
  }
}

// y = dropout(act(x)); dense bf16 [n], Philox group i <-> elements 8i..8i+7 (same numbering as s2t_dropout)
__global__ void act_dropout_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                   long long n_vec, int act, uint32_t thresh16, float inv_keep,
                                   unsigned long long seed, uint32_t stream_id,
                                   const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_vec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint4 v = reinterpret_cast<const uint4*>(x)[i];
    const uint32_t* va = &v.x;
    const uint32_t keep =
        thresh16 != 0 ? dropout_keep8(seed, stream_id, static_cast<unsigned long long>(i), thresh16) : 0xffu;
    uint4 o;
    uint32_t* oa = &o.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float a = act_apply(bf16_lo(va[k]), act), b = act_apply(bf16_hi(va[k]), act);
      a = ((keep >> (2 * k)) & 1u) ? a * inv_keep : 0.f;
      b = ((keep >> (2 * k + 1)) & 1u) ? b * inv_keep : 0.f;
      oa[k] = pack_bf16(a, b);
    }
    reinterpret_cast<uint4*>(y)[i] = o;
  }
}

static inline unsigned cm_grid(long long n, int threads) {
  long long g = (n + threads - 1) / threads;
  if (g > 148ll * 16) g = 148ll * 16;
  if (g < 1) g = 1;
  return static_cast<unsigned>(g);
}

}  // namespace s2t

using namespace s2t;
#define ST(s) reinterpret_cast<cudaStream_t>(s)
#define BF(p) reinterpret_cast<const __nv_bfloat16*>(p)
#define BFM(p) reinterpret_cast<__nv_bfloat16*>(p)

static int dw_check(const char* who, int batch, int T, int C, int K) {
  S2T_REQUIRE(batch > 0 && T > 0, "%s: empty batch", who);
  S2T_REQUIRE(C > 0 && C % 8 == 0, "%s: C must be a positive multiple of 8", who);
  S2T_REQUIRE(K >= 1 && K <= kDwMaxK && (K & 1), "%s: kernel size must be odd and <= %d", who, kDwMaxK);
  return 0;
}

extern "C" long long s2t_dwconv1d_tiles(int batch, int T) {
  return static_cast<long long>(batch) * ((T + kDwTile - 1) / kDwTile);
}

extern "C" int s2t_dwconv1d_fwd(const void* x, const void* w, void* y, float* bn_partials, int batch, int T,
                                int C, int K, void* stream) {
  S2T_REQUIRE(x && w && y, "s2t_dwconv1d_fwd: null pointer");
  if (int e = dw_check("s2t_dwconv1d_fwd", batch, T, C, K)) return e;
  const size_t smem = static_cast<size_t>(kDwTile + K - 1) * C * 2 + static_cast<size_t>(kDwTile) * 2 * C * 4;
  S2T_REQUIRE(smem <= 200 * 1024, "s2t_dwconv1d_fwd: C too large for the shared-memory tile");
  S2T_CHECK_CUDA(cudaFuncSetAttribute(dwconv1d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  S2T_LAUNCH(dwconv1d_kernel, static_cast<unsigned>(s2t_dwconv1d_tiles(batch, T)), kDwThreads, smem, ST(stream), BF(x), BF(w), BFM(y), bn_partials, T, C, K, 0);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_dwconv1d_bwd(const void* x, const void* w, const void* dy, void* dx, void* dw,
                                float* workspace, int batch, int T, int C, int K, void* stream) {
  S2T_REQUIRE(x && w && dy && dx && dw && workspace, "s2t_dwconv1d_bwd: null pointer");
  if (int e = dw_check("s2t_dwconv1d_bwd", batch, T, C, K)) return e;
  const unsigned tiles = static_cast<unsigned>(s2t_dwconv1d_tiles(batch, T));
  const size_t smem = static_cast<size_t>(kDwTile + K - 1) * C * 2 + static_cast<size_t>(kDwTile) * 2 * C * 4;
  S2T_REQUIRE(smem <= 200 * 1024, "s2t_dwconv1d_bwd: C too large for the shared-memory tile");
  S2T_CHECK_CUDA(cudaFuncSetAttribute(dwconv1d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  S2T_CHECK_CUDA(cudaFuncSetAttribute(dwconv1d_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  S2T_LAUNCH(dwconv1d_kernel, tiles, kDwThreads, smem, ST(stream), BF(dy), BF(w), BFM(dx), nullptr, T, C, K, 1);
  S2T_LAUNCH_CHECK();
  const size_t smem_w = static_cast<size_t>(2 * kDwTile + K - 1) * C * 2;
  S2T_LAUNCH(dwconv1d_wgrad_kernel, tiles, kDwThreads, smem_w, ST(stream), BF(x), BF(dy), workspace, T, C, K);
  S2T_LAUNCH_CHECK();
  S2T_LAUNCH(dwconv1d_wgrad_reduce_kernel, (K * C + 255) / 256, 256, 0, ST(stream), workspace, static_cast<int>(tiles), C, K, BFM(dw));
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_batchnorm_bwd_partials(long long rows) {
  return static_cast<int>((rows + kBnRows - 1) / kBnRows);
}

extern "C" int s2t_batchnorm_fwd(const void* x, const float* partials, int nparts, const void* gamma,
                                 const void* beta, float* running_mean, float* running_var,
                                 long long* num_batches_tracked, float* save_stats, void* y, long long rows, int C,
                                 float eps, float momentum, int act, int training, void* stream) {
  S2T_REQUIRE(x && gamma && beta && y, "s2t_batchnorm_fwd: null pointer");
  S2T_REQUIRE(C > 0 && C % 8 == 0 && rows > 0, "s2t_batchnorm_fwd: bad shape");
  S2T_REQUIRE(training ? (partials != nullptr && nparts > 0) : (running_mean && running_var),
              "s2t_batchnorm_fwd: training needs the channel-sum partials, eval needs the running statistics");
  S2T_LAUNCH(batchnorm_fwd_kernel, static_cast<unsigned>((rows + kBnRows - 1) / kBnRows), kBnThreads, 0, ST(stream), BF(x), partials, nparts, BF(gamma), BF(beta), running_mean, running_var, num_batches_tracked, save_stats, BFM(y), rows, C, eps, momentum, act, training);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_batchnorm_bwd(const void* da, const void* x, const float* save_stats, const void* gamma,
                                 const void* beta, float* partials, void* dx, void* dgamma, void* dbeta,
                                 long long rows, int C, int act, int training, void* stream) {
  S2T_REQUIRE(da && x && save_stats && gamma && beta && partials && dx && dgamma && dbeta,
              "s2t_batchnorm_bwd: null pointer");
  S2T_REQUIRE(C > 0 && C % 8 == 0 && rows > 0, "s2t_batchnorm_bwd: bad shape");
  const unsigned nblk = static_cast<unsigned>((rows + kBnRows - 1) / kBnRows);
  const size_t smem = static_cast<size_t>(kBnRows) * 2 * C * 4;
  S2T_REQUIRE(smem <= 200 * 1024, "s2t_batchnorm_bwd: C too large");
  S2T_CHECK_CUDA(cudaFuncSetAttribute(batchnorm_bwd_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  S2T_LAUNCH(batchnorm_bwd_reduce_kernel, nblk, kBnThreads, smem, ST(stream), BF(da), BF(x), save_stats, BF(gamma), BF(beta), partials, rows, C, act);
  S2T_LAUNCH_CHECK();
  S2T_LAUNCH(batchnorm_bwd_apply_kernel, nblk, kBnThreads, 0, ST(stream), BF(da), BF(x), save_stats, BF(gamma), BF(beta), partials, static_cast<int>(nblk), BFM(dx), BFM(dgamma), BFM(dbeta), rows, C, act, training);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_relpos_qbias_fwd(const void* q, long long ldq, const void* u, const void* v, void* quv,
                                    long long rows, int C, void* stream) {
  S2T_REQUIRE(q && u && v && quv, "s2t_relpos_qbias_fwd: null pointer");
  S2T_REQUIRE(C % 8 == 0 && ldq % 8 == 0, "s2t_relpos_qbias_fwd: C and ldq must be multiples of 8");
  if (rows <= 0) return 0;
  S2T_LAUNCH(relpos_qbias_fwd_kernel, cm_grid(rows * (C / 8), 256), 256, 0, ST(stream), BF(q), ldq, BF(u), BF(v), BFM(quv), rows, C);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_relpos_qbias_bwd(const void* dquv, void* dq, long long lddq, long long rows, int C,
                                    void* stream) {
  S2T_REQUIRE(dquv && dq, "s2t_relpos_qbias_bwd: null pointer");
  S2T_REQUIRE(C % 8 == 0 && lddq % 8 == 0, "s2t_relpos_qbias_bwd: C and lddq must be multiples of 8");
  if (rows <= 0) return 0;
  S2T_LAUNCH(relpos_qbias_bwd_kernel, cm_grid(rows * (C / 8), 256), 256, 0, ST(stream), BF(dquv), BFM(dq), lddq, rows, C);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_act_dropout(const void* x, void* y, long long n, int act, float p, unsigned long long seed,
                               unsigned int stream_id, const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(x && y, "s2t_act_dropout: null pointer");
  S2T_REQUIRE(n % 8 == 0, "s2t_act_dropout: n must be a multiple of 8");
  S2T_REQUIRE(p >= 0.f && p < 1.f, "s2t_act_dropout: p must be in [0,1)");
  if (n <= 0) return 0;
  S2T_LAUNCH(act_dropout_kernel, cm_grid(n / 8, 256), 256, 0, ST(stream), BF(x), BFM(y), n / 8, act, p > 0.f ? static_cast<uint32_t>(p * 65536.f + 0.5f) : 0u, 1.f / (1.f - p), seed, stream_id, seed_offset);
  S2T_LAUNCH_CHECK();
  return 0;
}
