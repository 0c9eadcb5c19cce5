// Conv2dSubsampling support kernels (reference: fairseq/modules/speech_to_text/subsampling.py:162-228).
// The reference runs nn.Conv2d(k, stride, no padding) over a (B, C, D, T) tensor; here activations are channels-last
// [B][T][D][C] bf16 and every convolution is a tcgen05 GEMM (gemm.cu) over an explicit patch matrix:
//   col[(b, t', d')][(j*k + i)*Cin + c] = x[b][stride*t' + j][stride*d' + i][c]      (j: time tap, i: feature tap)
// padded with zero columns up to Kpad (16-byte rows).  All three kernels are HBM-bound copies / reductions.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {
namespace {

inline cudaStream_t ST2(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// one thread per (row, 8-wide column group) of the patch matrix
__global__ void __launch_bounds__(256)
im2col2d_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, int B, int T, int D, int Cin,
                int k, int stride, int T2, int D2, int Kpad) {
  pdl_entry();
  const int groups = Kpad >> 3;
  const long long n = static_cast<long long>(B) * T2 * D2 * groups;
  const int K = k * k * Cin;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < n;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(idx % groups);
    const long long row = idx / groups;
    const int d2 = static_cast<int>(row % D2);
    const int t2 = static_cast<int>((row / D2) % T2);
    const int b = static_cast<int>(row / (static_cast<long long>(D2) * T2));
    const int kk0 = g * 8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if ((Cin & 7) == 0) {
      if (kk0 < K) {      // 8 consecutive channels of one tap: one 16-byte load
        const int tap = kk0 / Cin, c = kk0 - tap * Cin;
        const int j = tap / k, i = tap - j * k;
        v = *reinterpret_cast<const uint4*>(
            x + ((static_cast<long long>(b) * T + stride * t2 + j) * D + stride * d2 + i) * Cin + c);
      }
    } else {
      __nv_bfloat16 e[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int kk = kk0 + u;
        float f = 0.f;
        if (kk < K) {
          const int tap = kk / Cin, c = kk - tap * Cin;
          const int j = tap / k, i = tap - j * k;
          f = __bfloat162float(x[((static_cast<long long>(b) * T + stride * t2 + j) * D + stride * d2 + i) * Cin + c]);
        }
        e[u] = __float2bfloat16(f);
      }
      v = *reinterpret_cast<const uint4*>(e);
    }
    *reinterpret_cast<uint4*>(col + row * Kpad + kk0) = v;
  }
}

// gather form of the transposed scatter: dx[b][t][d][c] = sum over taps (j, i) with t = stride*t' + j, d = stride*d' + i
__global__ void __launch_bounds__(256)
col2im2d_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, int B, int T, int D, int Cin,
                int k, int stride, int T2, int D2, int Kpad) {
  pdl_entry();
  const int cg = Cin >> 3;
  const long long n = static_cast<long long>(B) * T * D * cg;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < n;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(idx % cg) * 8;
    const long long pos = idx / cg;
    const int d = static_cast<int>(pos % D);
    const int t = static_cast<int>((pos / D) % T);
    const int b = static_cast<int>(pos / (static_cast<long long>(D) * T));
    float acc[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] = 0.f;
    for (int j = 0; j < k; ++j) {
      const int tt = t - j;
      if (tt < 0 || tt % stride != 0) continue;
      const int t2 = tt / stride;
      if (t2 >= T2) continue;
      for (int i = 0; i < k; ++i) {
        const int dd = d - i;
        if (dd < 0 || dd % stride != 0) continue;
        const int d2 = dd / stride;
        if (d2 >= D2) continue;
        const long long row = (static_cast<long long>(b) * T2 + t2) * D2 + d2;
        const uint4 q = *reinterpret_cast<const uint4*>(dcol + row * Kpad + (j * k + i) * Cin + c);
        acc[0] += bf16_lo(q.x); acc[1] += bf16_hi(q.x); acc[2] += bf16_lo(q.y); acc[3] += bf16_hi(q.y);
        acc[4] += bf16_lo(q.z); acc[5] += bf16_hi(q.z); acc[6] += bf16_lo(q.w); acc[7] += bf16_hi(q.w);
      }
    }
    uint4 o;
    o.x = pack_bf16(acc[0], acc[1]); o.y = pack_bf16(acc[2], acc[3]);
    o.z = pack_bf16(acc[4], acc[5]); o.w = pack_bf16(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(dx + pos * Cin + c) = o;
  }
}

constexpr int kStatParts = 296;   // two CTAs per SM

// partials[p][0][c] = sum_r x[r][c], partials[p][1][c] = sum_r x[r][c]^2 over the rows p, p + P, ... (BatchNorm
// batch statistics; the layout s2t_batchnorm_fwd consumes).  block (C/8 lanes-wide strips) x row slices.
__global__ void __launch_bounds__(256)
colstats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ partials, long long rows, int C) {
  pdl_entry();
  extern __shared__ float s_red[];   // [slices][2][C]
  const int cg = C >> 3;
  const int slices = blockDim.x / cg;
  const int lane_c = threadIdx.x % cg, sl = threadIdx.x / cg;
  float s[8], q[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) s[u] = q[u] = 0.f;
  if (sl < slices) {
    for (long long r = static_cast<long long>(blockIdx.x) * slices + sl; r < rows;
         r += static_cast<long long>(gridDim.x) * slices) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + r * C + lane_c * 8);
      const float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                          bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        s[u] += f[u];
        q[u] += f[u] * f[u];
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      s_red[(sl * 2 + 0) * C + lane_c * 8 + u] = s[u];
      s_red[(sl * 2 + 1) * C + lane_c * 8 + u] = q[u];
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
    float a = 0.f;
    for (int k2 = 0; k2 < slices; ++k2) a += s_red[k2 * 2 * C + i];
    partials[static_cast<long long>(blockIdx.x) * 2 * C + i] = a;
  }
}

}  // namespace
}  // namespace s2t

using namespace s2t;

extern "C" int s2t_im2col2d(const void* x, void* col, int batch, int T, int D, int Cin, int k, int stride, int Kpad,
                            void* stream) {
  S2T_REQUIRE(x && col, "s2t_im2col2d: null pointer");
  S2T_REQUIRE(k >= 1 && stride >= 1 && T >= k && D >= k, "s2t_im2col2d: input smaller than the kernel");
  S2T_REQUIRE(Kpad % 8 == 0 && Kpad >= k * k * Cin, "s2t_im2col2d: Kpad must be a multiple of 8 and >= k*k*Cin");
  const int T2 = (T - k) / stride + 1, D2 = (D - k) / stride + 1;
  const long long n = 1ll * batch * T2 * D2 * (Kpad / 8);
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  S2T_LAUNCH(im2col2d_kernel, static_cast<unsigned>(blocks), 256, 0, ST2(stream),
             reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(col), batch, T, D, Cin, k,
             stride, T2, D2, Kpad);
  return 0;
}

extern "C" int s2t_col2im2d(const void* dcol, void* dx, int batch, int T, int D, int Cin, int k, int stride, int Kpad,
                            void* stream) {
  S2T_REQUIRE(dcol && dx, "s2t_col2im2d: null pointer");
  S2T_REQUIRE(Cin % 8 == 0 && Kpad % 8 == 0 && Kpad >= k * k * Cin, "s2t_col2im2d: Cin and Kpad must be multiples of 8");
  const int T2 = (T - k) / stride + 1, D2 = (D - k) / stride + 1;
  const long long n = 1ll * batch * T * D * (Cin / 8);
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  S2T_LAUNCH(col2im2d_kernel, static_cast<unsigned>(blocks), 256, 0, ST2(stream),
             reinterpret_cast<const __nv_bfloat16*>(dcol), reinterpret_cast<__nv_bfloat16*>(dx), batch, T, D, Cin, k,
             stride, T2, D2, Kpad);
  return 0;
}

extern "C" int s2t_colstats_parts(long long rows) {
  (void)rows;
  return kStatParts;
}

extern "C" int s2t_colstats(const void* x, float* partials, long long rows, int C, void* stream) {
  S2T_REQUIRE(x && partials, "s2t_colstats: null pointer");
  S2T_REQUIRE(C % 8 == 0 && C > 0 && C / 8 <= 256 && rows > 0, "s2t_colstats: C must be a multiple of 8, <= 2048");
  const int cg = C / 8;
  const int slices = 256 / cg;
  const size_t smem = static_cast<size_t>(slices) * 2 * C * sizeof(float);
  S2T_REQUIRE(smem <= 48 * 1024, "s2t_colstats: C too large");
  S2T_LAUNCH(colstats_kernel, kStatParts, 256, smem, ST2(stream), reinterpret_cast<const __nv_bfloat16*>(x), partials,
             rows, C);
  return 0;
}
