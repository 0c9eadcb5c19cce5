// Small HBM-bound helpers the encoder engine needs around the GEMMs:
//   s2t_dropout        standalone (regenerable) dropout, e.g. CTC.forward's dropout on the encoder output
//                      (fairseq/modules/speech_to_text/ctc.py:59) and the backward of epilogue-fused dropouts
//   s2t_im2col1d       explicit im2col of the zero-framed Conv1d input for the weight gradient
//                      (fairseq/modules/speech_to_text/subsampling.py:133-151 backward)
//   s2t_counter_add    bumps the device-side dropout seed counter (one launch per captured step)
//   s2t_add_bf16       y += x (gradient accumulation of residual branches)
//   s2t_f32_sum        sums a short fp32 vector (per-utterance NLL -> batch loss, criterions/ctc.py:473)
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

__device__ __forceinline__ void unpack8a(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8a(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]);
  v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}

// element i*8..i*8+7 uses Philox group i: identical to the GEMM epilogue's numbering for a dense
// [rows][N] tensor with N % 8 == 0, so a mask drawn in an epilogue can be re-applied here.
__global__ void dropout_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                               long long n_vec, uint32_t thresh16, float scale,
                               unsigned long long seed, uint32_t stream_id,
                               const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_vec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v[8];
    unpack8a(reinterpret_cast<const uint4*>(x)[i], v);
    const uint32_t keep =
        thresh16 != 0 ? dropout_keep8(seed, stream_id, static_cast<unsigned long long>(i), thresh16) : 0xffu;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * scale : 0.f;
    reinterpret_cast<uint4*>(y)[i] = pack8a(v);
  }
}

// xcol[b*t_out + t][j*C + c] = xp[b*x_bs + (t*stride + j)*C + c]: each output row is one contiguous
// run of ksize*C elements of the framed input.
__global__ void im2col1d_kernel(const __nv_bfloat16* __restrict__ xp, long long x_bs,
                                __nv_bfloat16* __restrict__ xcol, int batch, int t_out, int C,
                                int ksize, int stride) {
  pdl_entry();
  const int rv = (ksize * C) >> 3;
  const long long n = static_cast<long long>(batch) * t_out * rv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int e = static_cast<int>(i % rv) * 8;
    const long long bt = i / rv;
    const int t = static_cast<int>(bt % t_out);
    const long long b = bt / t_out;
    const uint4 v = *reinterpret_cast<const uint4*>(xp + b * x_bs + static_cast<long long>(t) * stride * C + e);
    *reinterpret_cast<uint4*>(xcol + bt * (static_cast<long long>(ksize) * C) + e) = v;
  }
}

// y[b][pad + t][:] = t < lengths[b] ? x[b][t][:] : 0 — a stage output copied into the zero-framed input of the
// next Conv1d with the padding mask applied (Downsampling.forward, pdss2t_transformer.py:107-116)
__global__ void frame_rows_kernel(const __nv_bfloat16* __restrict__ x, const int* __restrict__ lengths,
                                  __nv_bfloat16* __restrict__ y, long long y_bs, int pad, int batch, int T, int C) {
  pdl_entry();
  const int cv = C >> 3;
  const long long n = static_cast<long long>(batch) * T * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cv) << 3;
    const long long bt = i / cv;
    const int t = static_cast<int>(bt % T);
    const int b = static_cast<int>(bt / T);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (lengths == nullptr || t < lengths[b]) v = *reinterpret_cast<const uint4*>(x + bt * C + c);
    *reinterpret_cast<uint4*>(y + b * y_bs + static_cast<long long>(pad + t) * C + c) = v;
  }
}

__global__ void counter_add_kernel(unsigned long long* c, unsigned long long inc) {
  pdl_entry();
  if (threadIdx.x == 0 && blockIdx.x == 0) *c += inc;
}

__global__ void add_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                long long n_vec) {
  pdl_entry();
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_vec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float a[8], b[8];
    unpack8a(reinterpret_cast<const uint4*>(x)[i], a);
    unpack8a(reinterpret_cast<const uint4*>(y)[i], b);
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] += a[j];
    reinterpret_cast<uint4*>(y)[i] = pack8a(b);
  }
}

__global__ void f32_sum_kernel(const float* __restrict__ x, int n, float scale, float* __restrict__ out) {
  pdl_entry();
  __shared__ float s[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += x[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) *out = v * scale;
  }
}

// out[0] = sum_r w[r] * sum_c x[r][c]
__global__ void f32_weighted_sum_kernel(const float* __restrict__ x, const float* __restrict__ w, int rows, int cols,
                                        float* __restrict__ out) {
  pdl_entry();
  __shared__ float s[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < rows * cols; i += blockDim.x) acc += w[i / cols] * x[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) *out = v;
  }
}

static inline unsigned grid_for(long long n, int threads) {
  long long g = (n + threads - 1) / threads;
  if (g > 148ll * 16) g = 148ll * 16;
  if (g < 1) g = 1;
  return static_cast<unsigned>(g);
}

}  // namespace s2t

using namespace s2t;
#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int s2t_dropout(const void* x, void* y, long long n, float p, float out_scale,
                           unsigned long long seed, unsigned int stream_id,
                           const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(x && y, "s2t_dropout: null pointer");
  S2T_REQUIRE(n % 8 == 0, "s2t_dropout: n must be a multiple of 8");
  S2T_REQUIRE(p >= 0.f && p < 1.f, "s2t_dropout: p must be in [0,1)");
  if (n <= 0) return 0;
  const long long nv = n / 8;
  S2T_LAUNCH(dropout_kernel, grid_for(nv, 256), 256, 0, ST(stream), reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), nv, p > 0.f ? static_cast<uint32_t>(p * 65536.f + 0.5f) : 0u, out_scale / (1.f - p), seed, stream_id, seed_offset);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_im2col1d(const void* xp, long long x_bs, void* xcol, int batch, int t_out, int C,
                            int ksize, int stride, void* stream) {
  S2T_REQUIRE(xp && xcol, "s2t_im2col1d: null pointer");
  S2T_REQUIRE(C % 8 == 0 && x_bs % 8 == 0 && ksize > 0 && stride > 0, "s2t_im2col1d: bad shape");
  const long long n = static_cast<long long>(batch) * t_out * (ksize * C / 8);
  if (n <= 0) return 0;
  S2T_LAUNCH(im2col1d_kernel, grid_for(n, 256), 256, 0, ST(stream), reinterpret_cast<const __nv_bfloat16*>(xp), x_bs, reinterpret_cast<__nv_bfloat16*>(xcol), batch, t_out, C, ksize, stride);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_frame_rows(const void* x, const int* lengths, void* y, long long y_bs, int pad, int batch,
                              int T, int C, void* stream) {
  S2T_REQUIRE(x && y, "s2t_frame_rows: null pointer");
  S2T_REQUIRE(C % 8 == 0 && y_bs % 8 == 0 && pad >= 0, "s2t_frame_rows: bad shape");
  const long long n = static_cast<long long>(batch) * T * (C / 8);
  if (n <= 0) return 0;
  S2T_LAUNCH(frame_rows_kernel, grid_for(n, 256), 256, 0, ST(stream), reinterpret_cast<const __nv_bfloat16*>(x), lengths, reinterpret_cast<__nv_bfloat16*>(y), y_bs, pad, batch, T, C);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_counter_add(unsigned long long* counter, unsigned long long inc, void* stream) {
  S2T_REQUIRE(counter, "s2t_counter_add: null pointer");
  S2T_LAUNCH(counter_add_kernel, 1, 32, 0, ST(stream), counter, inc);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_add_bf16(const void* x, void* y, long long n, void* stream) {
  S2T_REQUIRE(x && y, "s2t_add_bf16: null pointer");
  S2T_REQUIRE(n % 8 == 0, "s2t_add_bf16: n must be a multiple of 8");
  if (n <= 0) return 0;
  S2T_LAUNCH(add_bf16_kernel, grid_for(n / 8, 256), 256, 0, ST(stream), reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), n / 8);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_f32_weighted_sum(const float* x, const float* w, int rows, int cols, float* out, void* stream) {
  S2T_REQUIRE(x && w && out, "s2t_f32_weighted_sum: null pointer");
  S2T_REQUIRE(rows > 0 && cols > 0, "s2t_f32_weighted_sum: empty input");
  S2T_LAUNCH(f32_weighted_sum_kernel, 1, 256, 0, ST(stream), x, w, rows, cols, out);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_f32_sum(const float* x, int n, float scale, float* out, void* stream) {
  S2T_REQUIRE(x && out, "s2t_f32_sum: null pointer");
  S2T_LAUNCH(f32_sum_kernel, 1, 256, 0, ST(stream), x, n, scale, out);
  S2T_LAUNCH_CHECK();
  return 0;
}
