// K3: LayerNorm forward / backward (fairseq/modules/layer_norm.py:30-35 -> nn.LayerNorm(C, 1e-5)).
// HBM-bound: one warp per row, 16-byte vector loads, the row lives in registers between the
// statistics pass and the normalisation pass, warp-shuffle reductions, fp32 math on bf16 storage.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kLnWarps = 8;
constexpr int kLnMaxVec = 4;  // C <= 32 lanes * 4 vectors * 8 elements = 1024

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]);
  v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}

template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                     const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ stats, int rows, int C, float eps) {
  pdl_entry();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kLnWarps + warp;
  if (row >= rows) return;
  const int nvec = C >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<long long>(row) * C);
  float v[NV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = i * 32 + lane;
    if (vi < nvec) {
      unpack8(xr[vi], v[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[i][j];
    }
  }
  const float mean = warp_sum(sum) / static_cast<float>(C);
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = i * 32 + lane;
    if (vi < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = v[i][j] - mean;
        sq += d * d;
      }
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / static_cast<float>(C) + eps);
  if (lane == 0 && stats != nullptr) {
    stats[2 * row] = mean;
    stats[2 * row + 1] = rstd;
  }
  uint4* yr = reinterpret_cast<uint4*>(y + static_cast<long long>(row) * C);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = i * 32 + lane;
    if (vi < nvec) {
      float g[8], b[8], o[8];
      unpack8(reinterpret_cast<const uint4*>(gamma)[vi], g);
      unpack8(reinterpret_cast<const uint4*>(beta)[vi], b);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
      yr[vi] = pack8(o);
    }
  }
}

// dx = rstd * (g - mean(g) - xhat * mean(g*xhat)), g = dy*gamma; per-CTA partial dgamma/dbeta.
template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ stats,
                     const __nv_bfloat16* __restrict__ dx_res, __nv_bfloat16* __restrict__ dx,
                     float* __restrict__ partials, int rows, int C, int rows_per_cta,
                     __nv_bfloat16* __restrict__ dx_drop, uint32_t thresh16, float drop_scale,
                     unsigned long long seed, uint32_t stream_id,
                     const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  extern __shared__ float s_part[];  // [kLnWarps][2][C]
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = C >> 3;
  float dg[NV][8], db[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = 0.f;

  float gm[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = i * 32 + lane;
    if (vi < nvec) unpack8(reinterpret_cast<const uint4*>(gamma)[vi], gm[i]);
  }

  const int row_begin = blockIdx.x * rows_per_cta;
  const int row_end = min(rows, row_begin + rows_per_cta);
  for (int row = row_begin + warp; row < row_end; row += kLnWarps) {
    const float mean = stats[2 * row], rstd = stats[2 * row + 1];
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<long long>(row) * C);
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + static_cast<long long>(row) * C);
    float xh[NV][8], g[NV][8];
    float s1 = 0.f, s2 = 0.f;
    // the residual gradient is only needed after the two row reductions: fetch it with the other operands so that its
    // latency is not paid a second time behind them
    const uint4* rr =
        dx_res ? reinterpret_cast<const uint4*>(dx_res + static_cast<long long>(row) * C) : nullptr;
    uint4 rpre[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = i * 32 + lane;
      rpre[i] = (rr != nullptr && vi < nvec) ? rr[vi] : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = i * 32 + lane;
      if (vi < nvec) {
        float xv[8], dv[8];
        unpack8(xr[vi], xv);
        unpack8(dyr[vi], dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mean) * rstd;
          g[i][j] = dv[j] * gm[i][j];
          s1 += g[i][j];
          s2 += g[i][j] * xh[i][j];
          dg[i][j] += dv[j] * xh[i][j];
          db[i][j] += dv[j];
        }
      }
    }
    s1 = warp_sum(s1) / static_cast<float>(C);
    s2 = warp_sum(s2) / static_cast<float>(C);
    uint4* dxr = reinterpret_cast<uint4*>(dx + static_cast<long long>(row) * C);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = i * 32 + lane;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - s1 - xh[i][j] * s2);
        if (rr) {
          float r[8];
          unpack8(rpre[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxr[vi] = pack8(o);
        if (dx_drop != nullptr) {  // second output: drop_scale * dropout(dx) for the next residual branch's backward
          const uint32_t keep = thresh16 != 0
              ? dropout_keep8(seed, stream_id, static_cast<unsigned long long>(row) * nvec + vi, thresh16) : 0xffu;
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = ((keep >> j) & 1u) ? o[j] * drop_scale : 0.f;
          reinterpret_cast<uint4*>(dx_drop + static_cast<long long>(row) * C)[vi] = pack8(o);
        }
      }
    }
  }
  // reduce the per-warp dgamma/dbeta partials across the CTA, one fp32 row per CTA to HBM
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = i * 32 + lane;
    if (vi < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s_part[(warp * 2 + 0) * C + vi * 8 + j] = dg[i][j];
        s_part[(warp * 2 + 1) * C + vi * 8 + j] = db[i][j];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    const int which = c / C, col = c % C;
    float acc = 0.f;
#pragma unroll
    for (int w = 0; w < kLnWarps; ++w) acc += s_part[(w * 2 + which) * C + col];
    partials[(static_cast<long long>(blockIdx.x) * 2 + which) * C + col] = acc;
  }
}

// dgamma / dbeta = column sums of the per-CTA partials.  Block = 32 columns x 8 partial slices, so the serial depth is
// n_partials / 8 coalesced loads (the reduce runs beside the dgrad chain on a side stream, but finishes in ~2 us).
__global__ void __launch_bounds__(256)
layernorm_bwd_reduce_kernel(const float* __restrict__ partials, int n_partials, int C,
                            __nv_bfloat16* __restrict__ dgamma, __nv_bfloat16* __restrict__ dbeta, int accumulate) {
  pdl_entry();
  __shared__ float s_acc[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;  // column of the [2][C] row
  float acc = 0.f;
  if (c < 2 * C)
    for (int p = slice; p < n_partials; p += 8) acc += partials[static_cast<long long>(p) * 2 * C + c];
  s_acc[slice][lane] = acc;
  __syncthreads();
  if (slice == 0 && c < 2 * C) {
#pragma unroll
    for (int k = 1; k < 8; ++k) acc += s_acc[k][lane];
    const int which = c / C, col = c % C;
    __nv_bfloat16* dst = which == 0 ? dgamma : dbeta;
    if (dst != nullptr) {
      if (accumulate) acc += __bfloat162float(dst[col]);
      dst[col] = __float2bfloat16(acc);
    }
  }
}

// The same reduction for up to kLnBatch LayerNorm instances in ONE launch (blockIdx.y = instance): the training step has
// 26-100 LayerNorms whose dgamma / dbeta are only needed when a gradient bucket is shipped (or at the end of backward),
// so their reductions are collected and run together instead of as one 5-10 us launch each.
constexpr int kLnBatch = 32;
struct LnReduceJob {
  const float* partials;
  __nv_bfloat16* dgamma;
  __nv_bfloat16* dbeta;
  int n_partials, C, accumulate;
};
struct LnReduceJobs {
  LnReduceJob j[kLnBatch];
};

__global__ void __launch_bounds__(256)
layernorm_bwd_reduce_batched_kernel(const __grid_constant__ LnReduceJobs jobs) {
  pdl_entry();
  const LnReduceJob& jb = jobs.j[blockIdx.y];
  __shared__ float s_acc[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const int C = jb.C;
  const int c = blockIdx.x * 32 + lane;
  if (blockIdx.x * 32 >= 2 * C) return;      // whole block beyond this instance's [2][C] row (block-uniform)
  float acc = 0.f;
  if (c < 2 * C)
    for (int p = slice; p < jb.n_partials; p += 8) acc += jb.partials[static_cast<long long>(p) * 2 * C + c];
  s_acc[slice][lane] = acc;
  __syncthreads();
  if (slice == 0 && c < 2 * C) {
#pragma unroll
    for (int k = 1; k < 8; ++k) acc += s_acc[k][lane];
    const int which = c / C, col = c % C;
    __nv_bfloat16* dst = which == 0 ? jb.dgamma : jb.dbeta;
    if (dst != nullptr) {
      if (jb.accumulate) acc += __bfloat162float(dst[col]);
      dst[col] = __float2bfloat16(acc);
    }
  }
}

static int ln_partials(int rows) {
  int n = (rows + kLnWarps - 1) / kLnWarps;            // one row per warp while the grid fits in one wave
  if (n > 296) n = 296;                                // 2 CTAs per SM on 148 SMs
  if (n < 1) n = 1;
  return n;
}

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y,
                                 float* stats, int rows, int C, float eps, void* stream) {
  S2T_REQUIRE(x && gamma && beta && y, "s2t_layernorm_fwd: null pointer");
  S2T_REQUIRE(C % 8 == 0 && C >= 8 && C <= 1024, "s2t_layernorm_fwd: C=%d must be a multiple of 8 in [8,1024]", C);
  if (rows <= 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = (rows + kLnWarps - 1) / kLnWarps;
  auto xp = reinterpret_cast<const __nv_bfloat16*>(x);
  auto gp = reinterpret_cast<const __nv_bfloat16*>(gamma);
  auto bp = reinterpret_cast<const __nv_bfloat16*>(beta);
  auto yp = reinterpret_cast<__nv_bfloat16*>(y);
  const int nv = (C / 8 + 31) / 32;
  switch (nv) {
    case 1: S2T_LAUNCH(layernorm_fwd_kernel<1>, grid, kLnWarps * 32, 0, st, xp, gp, bp, yp, stats, rows, C, eps); break;
    case 2: S2T_LAUNCH(layernorm_fwd_kernel<2>, grid, kLnWarps * 32, 0, st, xp, gp, bp, yp, stats, rows, C, eps); break;
    default: S2T_LAUNCH(layernorm_fwd_kernel<kLnMaxVec>, grid, kLnWarps * 32, 0, st, xp, gp, bp, yp, stats, rows, C, eps); break;
  }
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_layernorm_bwd_partials(int rows, int C) {
  (void)C;
  return ln_partials(rows);
}

extern "C" int s2t_layernorm_bwd(const void* dy, const void* x, const void* gamma,
                                 const float* stats, const void* dx_residual, void* dx,
                                 float* partials, int rows, int C, void* dx_drop, float drop_p,
                                 float drop_scale, unsigned long long seed, unsigned int stream_id,
                                 const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(dy && x && gamma && stats && dx && partials, "s2t_layernorm_bwd: null pointer");
  S2T_REQUIRE(C % 8 == 0 && C >= 8 && C <= 1024, "s2t_layernorm_bwd: C=%d must be a multiple of 8 in [8,1024]", C);
  S2T_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "s2t_layernorm_bwd: drop_p must be in [0,1)");
  if (rows <= 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int np = ln_partials(rows);
  const int rows_per_cta = (rows + np - 1) / np;
  const size_t smem = sizeof(float) * kLnWarps * 2 * C;
  auto dyp = reinterpret_cast<const __nv_bfloat16*>(dy);
  auto xp = reinterpret_cast<const __nv_bfloat16*>(x);
  auto gp = reinterpret_cast<const __nv_bfloat16*>(gamma);
  auto rp = reinterpret_cast<const __nv_bfloat16*>(dx_residual);
  auto dxp = reinterpret_cast<__nv_bfloat16*>(dx);
  auto ddp = reinterpret_cast<__nv_bfloat16*>(dx_drop);
  const uint32_t th = drop_p > 0.f ? static_cast<uint32_t>(drop_p * 65536.f + 0.5f) : 0u;
  const float sc = drop_scale / (1.f - drop_p);
  const int nv = (C / 8 + 31) / 32;
  if (smem > 48 * 1024) {
    S2T_CHECK_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<kLnMaxVec>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
#define S2T_LN_BWD(NV)                                                                                         \
  S2T_LAUNCH(layernorm_bwd_kernel<NV>, np, kLnWarps * 32, smem, st, dyp, xp, gp, stats, rp, dxp, partials, rows, C, rows_per_cta, ddp, th, sc, seed, stream_id, seed_offset)
  switch (nv) {
    case 1: S2T_LN_BWD(1); break;
    case 2: S2T_LN_BWD(2); break;
    default: S2T_LN_BWD(kLnMaxVec); break;
  }
#undef S2T_LN_BWD
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_layernorm_bwd_reduce(const float* partials, int rows, int C, void* dgamma, void* dbeta,
                                        int accumulate, void* stream) {
  S2T_REQUIRE(partials, "s2t_layernorm_bwd_reduce: null pointer");
  if (rows <= 0) return 0;
  S2T_LAUNCH(layernorm_bwd_reduce_kernel, (2 * C + 31) / 32, 256, 0, reinterpret_cast<cudaStream_t>(stream), partials, ln_partials(rows), C, reinterpret_cast<__nv_bfloat16*>(dgamma), reinterpret_cast<__nv_bfloat16*>(dbeta), accumulate);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_layernorm_bwd_reduce_batched(const s2t_ln_reduce_job* jobs, int n_jobs, void* stream) {
  S2T_REQUIRE(n_jobs >= 0 && (n_jobs == 0 || jobs != nullptr), "s2t_layernorm_bwd_reduce_batched: bad job list");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  for (int base = 0; base < n_jobs; base += kLnBatch) {
    const int n = n_jobs - base < kLnBatch ? n_jobs - base : kLnBatch;
    LnReduceJobs js;
    memset(&js, 0, sizeof(js));
    int cmax = 0;
    for (int i = 0; i < n; ++i) {
      const s2t_ln_reduce_job& in = jobs[base + i];
      S2T_REQUIRE(in.partials != nullptr && in.rows > 0 && in.C > 0, "s2t_layernorm_bwd_reduce_batched: bad job %d", base + i);
      js.j[i].partials = in.partials;
      js.j[i].dgamma = reinterpret_cast<__nv_bfloat16*>(in.dgamma);
      js.j[i].dbeta = reinterpret_cast<__nv_bfloat16*>(in.dbeta);
      js.j[i].n_partials = ln_partials(in.rows);
      js.j[i].C = in.C;
      js.j[i].accumulate = in.accumulate;
      if (in.C > cmax) cmax = in.C;
    }
    S2T_LAUNCH(layernorm_bwd_reduce_batched_kernel, dim3((2 * cmax + 31) / 32, n), 256, 0, st, js);
    S2T_LAUNCH_CHECK();
  }
  return 0;
}
