// HBM-bound elementwise / data-movement kernels of the S2T encoder path: GLU (subsampling.py:140-151),
// bias-gradient column sums, the Conv1d dgrad col2im, the embed block (s2t_transformer.py:1765-1794),
// padding masks and weight re-layouts.  All use 16-byte vector accesses on bf16 rows and grids sized
// from the element count.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

__device__ __forceinline__ void unpack8e(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8e(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]);
  v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

// ---------------------------------------------------------------------------------------------
// GLU: y = a * sigmoid(g), x = [a | g] along channels.  Optional padding mask by utterance length.
// ---------------------------------------------------------------------------------------------
__global__ void glu_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                               long long ldy, long long y_bs, long long rows, int C,
                               const int* __restrict__ lengths, int T) {
  pdl_entry();
  const int cv = C >> 3;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= rows * cv) return;
  const long long r = i / cv;
  const int c = static_cast<int>(i % cv) * 8;
  float o[8];
  bool valid = true;
  if (lengths != nullptr) valid = static_cast<int>(r % T) < lengths[r / T];
  if (valid) {
    float a[8], g[8];
    unpack8e(*reinterpret_cast<const uint4*>(x + r * 2 * C + c), a);
    unpack8e(*reinterpret_cast<const uint4*>(x + r * 2 * C + C + c), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = a[j] * sigmoidf_(g[j]);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = 0.f;
  }
  const long long yoff = y_bs != 0 ? (r / T) * y_bs + (r % T) * ldy : r * ldy;
  *reinterpret_cast<uint4*>(y + yoff + c) = pack8e(o);
}

__global__ void glu_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                               long long lddy, __nv_bfloat16* __restrict__ dx, long long rows, int C,
                               const int* __restrict__ lengths, int T) {
  pdl_entry();
  const int cv = C >> 3;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= rows * cv) return;
  const long long r = i / cv;
  const int c = static_cast<int>(i % cv) * 8;
  float da[8], dg[8];
  bool valid = true;
  if (lengths != nullptr) valid = static_cast<int>(r % T) < lengths[r / T];
  if (valid) {
    float a[8], g[8], d[8];
    unpack8e(*reinterpret_cast<const uint4*>(x + r * 2 * C + c), a);
    unpack8e(*reinterpret_cast<const uint4*>(x + r * 2 * C + C + c), g);
    unpack8e(*reinterpret_cast<const uint4*>(dy + r * lddy + c), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = sigmoidf_(g[j]);
      da[j] = d[j] * s;
      dg[j] = d[j] * a[j] * s * (1.f - s);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) da[j] = dg[j] = 0.f;
  }
  *reinterpret_cast<uint4*>(dx + r * 2 * C + c) = pack8e(da);
  *reinterpret_cast<uint4*>(dx + r * 2 * C + C + c) = pack8e(dg);
}

// ---------------------------------------------------------------------------------------------
// column sums (bias gradients): two stages, fp32 partials [slabs][C]
// ---------------------------------------------------------------------------------------------
constexpr int kColsumRowsPerSlab = 64;   // legacy workspace sizing (s2t_colsum_workspace_floats); no longer written
constexpr int kColsumCluster = 8;        // CTAs of one cluster split the rows of a 64-column strip

// One launch: a cluster of 8 CTAs owns a 64-column strip; CTA r sums rows r, r+8, ... (thread (x, y): 8 columns
// 8x..8x+7 of rows r + 8*(y + 32*i)), reduces its 32 row slices through shared memory, and the cluster leader
// gathers the 8 per-CTA strips over DSMEM and writes the bf16 result (optionally accumulating into it).
__global__ void __cluster_dims__(1, kColsumCluster, 1) __launch_bounds__(256)
colsum_cluster_kernel(const __nv_bfloat16* __restrict__ x, long long ld, int rows, int C,
                      __nv_bfloat16* __restrict__ out, int accumulate) {
  pdl_entry();
  __shared__ float s_acc[32][65];
  __shared__ __align__(16) float s_strip[64];
  const int lx = threadIdx.x & 7, ly = threadIdx.x >> 3;      // 8 column groups x 32 row slices
  const int col = blockIdx.x * 64 + lx * 8;
  const int rank = static_cast<int>(cluster_ctarank());
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (col < C) {
    const int stride = kColsumCluster * 32;
    int r = rank + kColsumCluster * ly;
    for (; r + 3 * stride < rows; r += 4 * stride) {            // four independent 16-byte loads in flight
      uint4 q[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        q[u] = *reinterpret_cast<const uint4*>(x + static_cast<long long>(r + u * stride) * ld + col);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float v[8];
        unpack8e(q[u], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    for (; r < rows; r += stride) {
      float v[8];
      unpack8e(*reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * ld + col), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) s_acc[ly][lx * 8 + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < 64) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) t += s_acc[k][threadIdx.x];
    s_strip[threadIdx.x] = t;
  }
  cluster_sync_all();
  if (rank == 0 && threadIdx.x < 64) {
    const int c = blockIdx.x * 64 + threadIdx.x;
    const uint32_t local = smem_u32(&s_strip[threadIdx.x]);
    float t = 0.f;
#pragma unroll
    for (int pr = 0; pr < kColsumCluster; ++pr) {
      float v;
      asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(dsmem_addr(local, pr)) : "memory");
      t += v;
    }
    if (c < C) {
      if (accumulate) t += __bfloat162float(out[c]);
      out[c] = __float2bfloat16(t);
    }
  }
  cluster_sync_all();   // the strips of ranks > 0 stay alive until the leader has read them
}

// ---------------------------------------------------------------------------------------------
// col2im for strided Conv1d dgrad
// ---------------------------------------------------------------------------------------------
__global__ void col2im1d_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx,
                                int batch, int t_in, int t_out, int C, int ksize, int stride, int pad) {
  pdl_entry();
  const int cv = C >> 3;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(batch) * t_in * cv) return;
  const int c = static_cast<int>(i % cv) * 8;
  const long long bt = i / cv;
  const int t = static_cast<int>(bt % t_in);
  const int b = static_cast<int>(bt / t_in);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int j = 0; j < ksize; ++j) {
    const int num = t + pad - j;
    if (num < 0 || num % stride != 0) continue;
    const int to = num / stride;
    if (to >= t_out) continue;
    float v[8];
    unpack8e(*reinterpret_cast<const uint4*>(
                 dcol + (static_cast<long long>(b) * t_out + to) * (static_cast<long long>(ksize) * C) +
                 static_cast<long long>(j) * C + c),
             v);
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] += v[q];
  }
  *reinterpret_cast<uint4*>(dx + (static_cast<long long>(b) * t_in + t) * C + c) = pack8e(acc);
}

// ---------------------------------------------------------------------------------------------
// embed block: y = dropout(scale * x + pos_table[t + 2]) for valid frames; padded frames take the
// zero positional row (s2t_transformer.py:1773-1794, sinusoidal_positional_embedding.py:60-105)
// ---------------------------------------------------------------------------------------------
__global__ void embed_fwd_kernel(const __nv_bfloat16* __restrict__ x, const int* __restrict__ lengths,
                                 const float* __restrict__ pos_table, int n_pos,
                                 __nv_bfloat16* __restrict__ y, int batch, int T, int C, float scale,
                                 uint32_t thresh16, float inv_keep, unsigned long long seed,
                                 uint32_t stream_id, const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const int cv = C >> 3;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(batch) * T * cv) return;
  const int c = static_cast<int>(i % cv) * 8;
  const long long bt = i / cv;
  const int t = static_cast<int>(bt % T);
  const int b = static_cast<int>(bt / T);
  float v[8];
  unpack8e(*reinterpret_cast<const uint4*>(x + bt * C + c), v);
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] *= scale;
  if (pos_table != nullptr && t < lengths[b]) {
    const int p = min(t + 2, n_pos - 1);
    const float4 p0 = *reinterpret_cast<const float4*>(pos_table + static_cast<long long>(p) * C + c);
    const float4 p1 = *reinterpret_cast<const float4*>(pos_table + static_cast<long long>(p) * C + c + 4);
    v[0] += p0.x; v[1] += p0.y; v[2] += p0.z; v[3] += p0.w;
    v[4] += p1.x; v[5] += p1.y; v[6] += p1.z; v[7] += p1.w;
  }
  if (thresh16 != 0) {
    const uint32_t keep = dropout_keep8(seed, stream_id, static_cast<unsigned long long>(i), thresh16);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * inv_keep : 0.f;
  }
  *reinterpret_cast<uint4*>(y + bt * C + c) = pack8e(v);
}

__global__ void embed_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
                                 long long n_vec, float scale, uint32_t thresh16, float inv_keep,
                                 unsigned long long seed, uint32_t stream_id,
                                 const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_vec) return;
  float v[8];
  unpack8e(reinterpret_cast<const uint4*>(dy)[i], v);
  if (thresh16 != 0) {
    const uint32_t keep = dropout_keep8(seed, stream_id, static_cast<unsigned long long>(i), thresh16);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * inv_keep : 0.f;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] *= scale;
  reinterpret_cast<uint4*>(dx)[i] = pack8e(v);
}

__global__ void mask_rows_kernel(__nv_bfloat16* __restrict__ x, const int* __restrict__ lengths,
                                 int batch, int T, int C, long long ld) {
  pdl_entry();
  const int cv = C >> 3;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(batch) * T * cv) return;
  const int c = static_cast<int>(i % cv) * 8;
  const long long bt = i / cv;
  const int t = static_cast<int>(bt % T);
  const int b = static_cast<int>(bt / T);
  if (t >= lengths[b]) *reinterpret_cast<uint4*>(x + bt * ld + c) = make_uint4(0, 0, 0, 0);
}

__global__ void permute_conv_weight_kernel(const __nv_bfloat16* __restrict__ w,
                                           __nv_bfloat16* __restrict__ out, int cout, int cin, int k,
                                           int inverse) {
  pdl_entry();
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n = static_cast<long long>(cout) * cin * k;
  if (i >= n) return;
  // forward: out[o][j][c] = w[o][c][j]; inverse: out[o][c][j] = w[o][j][c]
  const int o = static_cast<int>(i / (static_cast<long long>(cin) * k));
  const int rem = static_cast<int>(i % (static_cast<long long>(cin) * k));
  if (!inverse) {
    const int j = rem / cin, c = rem % cin;
    out[i] = w[(static_cast<long long>(o) * cin + c) * k + j];
  } else {
    const int c = rem / k, j = rem % k;
    out[i] = w[(static_cast<long long>(o) * k + j) * cin + c];
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                     long long n) {
  pdl_entry();
  const long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    const float4 v = *reinterpret_cast<const float4*>(x + i);
    uint2 o;
    o.x = pack_bf16(v.x, v.y);
    o.y = pack_bf16(v.z, v.w);
    *reinterpret_cast<uint2*>(y + i) = o;
  } else {
    for (long long j = i; j < n; ++j) y[j] = __float2bfloat16(x[j]);
  }
}
__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ y,
                                     long long n) {
  pdl_entry();
  const long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    const uint2 v = *reinterpret_cast<const uint2*>(x + i);
    *reinterpret_cast<float4*>(y + i) = make_float4(bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y));
  } else {
    for (long long j = i; j < n; ++j) y[j] = __bfloat162float(x[j]);
  }
}

static inline unsigned blocks_for(long long n, int threads) {
  return static_cast<unsigned>((n + threads - 1) / threads);
}

}  // namespace s2t

using namespace s2t;
#define ST(s) reinterpret_cast<cudaStream_t>(s)
#define BF(p) reinterpret_cast<__nv_bfloat16*>(p)
#define CBF(p) reinterpret_cast<const __nv_bfloat16*>(p)

extern "C" int s2t_glu_fwd(const void* x, void* y, long long ldy, long long y_bs, int rows, int C,
                           const int* lengths, int T, void* stream) {
  S2T_REQUIRE(x && y, "s2t_glu_fwd: null pointer");
  S2T_REQUIRE(C % 8 == 0 && ldy % 8 == 0 && y_bs % 8 == 0, "s2t_glu_fwd: C, ldy, y_bs must be multiples of 8");
  S2T_REQUIRE(y_bs == 0 || T > 0, "s2t_glu_fwd: y_bs needs T");
  if (rows <= 0) return 0;
  const long long n = static_cast<long long>(rows) * (C / 8);
  S2T_LAUNCH(glu_fwd_kernel, blocks_for(n, 256), 256, 0, ST(stream), CBF(x), BF(y), ldy, y_bs, rows, C, lengths, T > 0 ? T : 1);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_glu_bwd(const void* x, const void* dy, long long lddy, void* dx, int rows, int C,
                           const int* lengths, int T, void* stream) {
  S2T_REQUIRE(x && dy && dx, "s2t_glu_bwd: null pointer");
  S2T_REQUIRE(C % 8 == 0 && lddy % 8 == 0, "s2t_glu_bwd: C and lddy must be multiples of 8");
  if (rows <= 0) return 0;
  const long long n = static_cast<long long>(rows) * (C / 8);
  S2T_LAUNCH(glu_bwd_kernel, blocks_for(n, 256), 256, 0, ST(stream), CBF(x), CBF(dy), lddy, BF(dx), rows, C, lengths, T > 0 ? T : 1);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" long long s2t_colsum_workspace_floats(int rows, int C) {
  const long long slabs = (rows + kColsumRowsPerSlab - 1) / kColsumRowsPerSlab;
  return slabs * C;
}

extern "C" int s2t_colsum(const void* x, long long ld, int rows, int C, void* out, int accumulate,
                          float* workspace, long long workspace_floats, void* stream) {
  S2T_REQUIRE(x && out && workspace, "s2t_colsum: null pointer");
  S2T_REQUIRE(C % 8 == 0 && ld % 8 == 0, "s2t_colsum: C and ld must be multiples of 8");
  S2T_REQUIRE(workspace_floats >= s2t_colsum_workspace_floats(rows, C), "s2t_colsum: workspace too small");
  if (rows <= 0) return 0;
  (void)workspace_floats;
  S2T_LAUNCH(colsum_cluster_kernel, dim3((C + 63) / 64, kColsumCluster), 256, 0, ST(stream), CBF(x), ld, rows, C, BF(out), accumulate);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_col2im1d(const void* dcol, void* dx, int batch, int t_in, int t_out, int C,
                            int ksize, int stride, int pad, void* stream) {
  S2T_REQUIRE(dcol && dx, "s2t_col2im1d: null pointer");
  S2T_REQUIRE(C % 8 == 0 && stride > 0 && ksize > 0, "s2t_col2im1d: bad shape");
  const long long n = static_cast<long long>(batch) * t_in * (C / 8);
  if (n <= 0) return 0;
  S2T_LAUNCH(col2im1d_kernel, blocks_for(n, 256), 256, 0, ST(stream), CBF(dcol), BF(dx), batch, t_in, t_out, C, ksize, stride, pad);
  S2T_LAUNCH_CHECK();
  return 0;
}

static inline uint32_t thresh16(float p) { return static_cast<uint32_t>(p * 65536.f + 0.5f); }

extern "C" int s2t_embed_fwd(const void* x, const int* lengths, const float* pos_table, int n_pos,
                             void* y, int batch, int T, int C, float scale, float dropout_p,
                             unsigned long long seed, unsigned int stream_id,
                             const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(x && y && lengths, "s2t_embed_fwd: null pointer");
  S2T_REQUIRE(C % 8 == 0, "s2t_embed_fwd: C must be a multiple of 8");
  S2T_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "s2t_embed_fwd: bad dropout_p");
  const long long n = static_cast<long long>(batch) * T * (C / 8);
  if (n <= 0) return 0;
  S2T_LAUNCH(embed_fwd_kernel, blocks_for(n, 256), 256, 0, ST(stream), CBF(x), lengths, pos_table, n_pos, BF(y), batch, T, C, scale, dropout_p > 0.f ? thresh16(dropout_p) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_embed_bwd(const void* dy, void* dx, int batch, int T, int C, float scale,
                             float dropout_p, unsigned long long seed, unsigned int stream_id,
                             const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(dy && dx, "s2t_embed_bwd: null pointer");
  S2T_REQUIRE(C % 8 == 0, "s2t_embed_bwd: C must be a multiple of 8");
  const long long n = static_cast<long long>(batch) * T * (C / 8);
  if (n <= 0) return 0;
  S2T_LAUNCH(embed_bwd_kernel, blocks_for(n, 256), 256, 0, ST(stream), CBF(dy), BF(dx), n, scale, dropout_p > 0.f ? thresh16(dropout_p) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_mask_rows(void* x, const int* lengths, int batch, int T, int C, long long ld,
                             void* stream) {
  S2T_REQUIRE(x && lengths, "s2t_mask_rows: null pointer");
  S2T_REQUIRE(C % 8 == 0 && ld % 8 == 0, "s2t_mask_rows: C and ld must be multiples of 8");
  const long long n = static_cast<long long>(batch) * T * (C / 8);
  if (n <= 0) return 0;
  S2T_LAUNCH(mask_rows_kernel, blocks_for(n, 256), 256, 0, ST(stream), BF(x), lengths, batch, T, C, ld);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_permute_conv_weight(const void* w, void* out, int cout, int cin, int k,
                                       int inverse, void* stream) {
  S2T_REQUIRE(w && out, "s2t_permute_conv_weight: null pointer");
  const long long n = static_cast<long long>(cout) * cin * k;
  if (n <= 0) return 0;
  S2T_LAUNCH(permute_conv_weight_kernel, blocks_for(n, 256), 256, 0, ST(stream), CBF(w), BF(out), cout, cin, k, inverse);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream) {
  S2T_REQUIRE(x && y, "s2t_cast_f32_to_bf16: null pointer");
  if (n <= 0) return 0;
  S2T_LAUNCH(cast_f32_bf16_kernel, blocks_for((n + 3) / 4, 256), 256, 0, ST(stream), x, BF(y), n);
  S2T_LAUNCH_CHECK();
  return 0;
}
extern "C" int s2t_cast_bf16_to_f32(const void* x, float* y, long long n, void* stream) {
  S2T_REQUIRE(x && y, "s2t_cast_bf16_to_f32: null pointer");
  if (n <= 0) return 0;
  S2T_LAUNCH(cast_bf16_f32_kernel, blocks_for((n + 3) / 4, 256), 256, 0, ST(stream), CBF(x), y, n);
  S2T_LAUNCH_CHECK();
  return 0;
}
