// Attention-probability softmax (forward / backward) with key-padding mask, optional relative-position
// bias through the rel_shift index map, clamp and counter-based dropout.
//
// Reference semantics:
//   fairseq MHA   (multihead_attention.py:383-400): masked_fill(key_pad, -inf) -> softmax(fp32) -> dropout
//   ESPnet MHA    (espnet_multihead_attention.py:119-133): masked_fill(-inf) -> clamp(+-1e8) -> softmax(fp32)
//   rel_pos       (espnet_multihead_attention.py:292-311,343-352): scores = (AC + rel_shift(BD)) / sqrt(d_k),
//                 rel_shift(BD)[i][j] = BD[i][j - i + T - 1]   
// One warp per (batch*head, query) row; the row (<= a few KB of bf16) stays L1-resident across the
// max / sum / write passes; 16-byte vector accesses; fp32 math.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kSmWarps = 8;

struct RowCtx {
  const __nv_bfloat16* srow;
  const __nv_bfloat16* brow;
  int bias_mode, shift;  // shift = Tk - 1 - i for rel-shift
  float scale, clamp;
};

__device__ __forceinline__ float score_at(const RowCtx& c, int j) {
  float s = __bfloat162float(c.srow[j]);
  if (c.bias_mode == 1) s += __bfloat162float(c.brow[j]);
  else if (c.bias_mode == 2) s += __bfloat162float(c.brow[j + c.shift]);
  s *= c.scale;
  if (c.clamp > 0.f) s = fminf(fmaxf(s, -c.clamp), c.clamp);
  return s;
}

__global__ void __launch_bounds__(kSmWarps * 32)
attn_softmax_fwd_kernel(const __nv_bfloat16* __restrict__ scores, long long ld,
                        const __nv_bfloat16* __restrict__ bias, long long ld_bias, int bias_mode,
                        int bias_head_major, __nv_bfloat16* __restrict__ probs, __nv_bfloat16* __restrict__ probs_drop,
                        const int* __restrict__ klen, long long nrows, int heads, int Tq, int Tk,
                        float scale, float clamp, uint32_t thresh16, float inv_keep,
                        unsigned long long seed, uint32_t stream_id,
                        const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kSmWarps + warp;
  if (row >= nrows) return;
  const int i = static_cast<int>(row % Tq);
  const long long z = row / Tq;
  const int b = static_cast<int>(z / heads);
  int kl = klen ? klen[b] : Tk;
  kl = kl < Tk ? kl : Tk;
  RowCtx c;
  c.srow = scores + row * ld;
  // bias rows are ordered [batch][head][Tq] like the scores, or [head][batch][Tq] (bias_head_major: the
  // layout that lets the linear_pos gradient contract over batch*time in one GEMM)
  const long long brow_idx =
      bias_head_major ? ((z % heads) * (nrows / Tq / heads) + b) * Tq + i : row;
  c.brow = bias ? bias + brow_idx * ld_bias : nullptr;
  c.bias_mode = bias ? bias_mode : 0;
  c.shift = Tk - 1 - i;
  c.scale = scale;
  c.clamp = clamp;

  float m = -INFINITY;
  for (int j = lane; j < kl; j += 32) m = fmaxf(m, score_at(c, j));
  m = warp_max(m);
  float sum = 0.f;
  for (int j = lane; j < kl; j += 32) sum += __expf(score_at(c, j) - m);
  sum = warp_sum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;

  const int tk8 = (Tk + 7) & ~7;
  const unsigned long long grp_row = static_cast<unsigned long long>(row) * (tk8 >> 3);
  __nv_bfloat16* prow = probs + row * ld;
  __nv_bfloat16* drow = probs_drop ? probs_drop + row * ld : nullptr;
  for (int j0 = lane * 8; j0 < tk8; j0 += 256) {
    float p[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int j = j0 + q;
      p[q] = j < kl ? __expf(score_at(c, j) - m) * inv : 0.f;
    }
    uint4 pv;
    pv.x = pack_bf16(p[0], p[1]); pv.y = pack_bf16(p[2], p[3]);
    pv.z = pack_bf16(p[4], p[5]); pv.w = pack_bf16(p[6], p[7]);
    *reinterpret_cast<uint4*>(prow + j0) = pv;
    if (drow != nullptr) {
      if (thresh16 != 0) {
        const uint32_t keep = dropout_keep8(seed, stream_id, grp_row + (j0 >> 3), thresh16);
#pragma unroll
        for (int q = 0; q < 8; ++q) p[q] = ((keep >> q) & 1u) ? p[q] * inv_keep : 0.f;
        pv.x = pack_bf16(p[0], p[1]); pv.y = pack_bf16(p[2], p[3]);
        pv.z = pack_bf16(p[4], p[5]); pv.w = pack_bf16(p[6], p[7]);
      }
      *reinterpret_cast<uint4*>(drow + j0) = pv;
    }
  }
}

__global__ void __launch_bounds__(kSmWarps * 32)
attn_softmax_bwd_kernel(const __nv_bfloat16* __restrict__ probs, const __nv_bfloat16* __restrict__ dpd,
                        __nv_bfloat16* __restrict__ dscores, __nv_bfloat16* __restrict__ dbias,
                        long long ld_bias, int bias_head_major, const int* __restrict__ klen, long long nrows, int heads,
                        int Tq, int Tk, long long ld, float scale, uint32_t thresh16, float inv_keep,
                        unsigned long long seed, uint32_t stream_id,
                        const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kSmWarps + warp;
  if (row >= nrows) return;
  const int i = static_cast<int>(row % Tq);
  const long long z = row / Tq;
  const int b = static_cast<int>(z / heads);
  int kl = klen ? klen[b] : Tk;
  kl = kl < Tk ? kl : Tk;
  const int tk8 = (Tk + 7) & ~7;
  const unsigned long long grp_row = static_cast<unsigned long long>(row) * (tk8 >> 3);
  const __nv_bfloat16* prow = probs + row * ld;
  const __nv_bfloat16* grow = dpd + row * ld;

  // pass 1: dot = sum_j dP_j * P_j
  float dot = 0.f;
  for (int j0 = lane * 8; j0 < tk8; j0 += 256) {
    const uint4 pv = *reinterpret_cast<const uint4*>(prow + j0);
    const uint4 gv = *reinterpret_cast<const uint4*>(grow + j0);
    float p[8] = {bf16_lo(pv.x), bf16_hi(pv.x), bf16_lo(pv.y), bf16_hi(pv.y),
                  bf16_lo(pv.z), bf16_hi(pv.z), bf16_lo(pv.w), bf16_hi(pv.w)};
    float g[8] = {bf16_lo(gv.x), bf16_hi(gv.x), bf16_lo(gv.y), bf16_hi(gv.y),
                  bf16_lo(gv.z), bf16_hi(gv.z), bf16_lo(gv.w), bf16_hi(gv.w)};
    uint32_t keep = 0xffu;
    float ik = 1.f;
    if (thresh16 != 0) {
      keep = dropout_keep8(seed, stream_id, grp_row + (j0 >> 3), thresh16);
      ik = inv_keep;
    }
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if (j0 + q < kl && ((keep >> q) & 1u)) dot += g[q] * ik * p[q];
  }
  dot = warp_sum(dot);
  // pass 2: dS = scale * P * (dP - dot)
  __nv_bfloat16* orow = dscores + row * ld;
  for (int j0 = lane * 8; j0 < tk8; j0 += 256) {
    const uint4 pv = *reinterpret_cast<const uint4*>(prow + j0);
    const uint4 gv = *reinterpret_cast<const uint4*>(grow + j0);
    float p[8] = {bf16_lo(pv.x), bf16_hi(pv.x), bf16_lo(pv.y), bf16_hi(pv.y),
                  bf16_lo(pv.z), bf16_hi(pv.z), bf16_lo(pv.w), bf16_hi(pv.w)};
    float g[8] = {bf16_lo(gv.x), bf16_hi(gv.x), bf16_lo(gv.y), bf16_hi(gv.y),
                  bf16_lo(gv.z), bf16_hi(gv.z), bf16_lo(gv.w), bf16_hi(gv.w)};
    uint32_t keep = 0xffu;
    float ik = 1.f;
    if (thresh16 != 0) {
      keep = dropout_keep8(seed, stream_id, grp_row + (j0 >> 3), thresh16);
      ik = inv_keep;
    }
    float o[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float dp = ((keep >> q) & 1u) ? g[q] * ik : 0.f;
      o[q] = (j0 + q < kl) ? scale * p[q] * (dp - dot) : 0.f;
    }
    uint4 ov;
    ov.x = pack_bf16(o[0], o[1]); ov.y = pack_bf16(o[2], o[3]);
    ov.z = pack_bf16(o[4], o[5]); ov.w = pack_bf16(o[6], o[7]);
    *reinterpret_cast<uint4*>(orow + j0) = ov;
  }
  if (dbias != nullptr) {
    // un-shift: dBD[i][c] = dS[i][c + i - (Tk-1)] (zero outside), c in [0, 2Tk-1)
    __syncwarp();
    const long long brow_idx =
        bias_head_major ? ((z % heads) * (nrows / Tq / heads) + b) * Tq + i : row;
    __nv_bfloat16* brow = dbias + brow_idx * ld_bias;
    const int shift = Tk - 1 - i;
    const int nb = 2 * Tk - 1;
    const int nb8 = (nb + 7) & ~7;
    for (int cidx = lane; cidx < nb8; cidx += 32) {
      const int j = cidx - shift;
      __nv_bfloat16 v = __float2bfloat16(0.f);
      if (cidx < nb && j >= 0 && j < kl) v = orow[j];
      if (cidx < ld_bias) brow[cidx] = v;
    }
  }
}

// -----------------------------------------------------------------------------------------------------------------
// Register-resident variants for rows of at most 256*NV keys (every encoder configuration of the path: T' <= 1024):
// each lane owns NV 16-byte vectors of its row, loaded ONCE; max / sum / normalise / dropout run on registers, so the
// kernel is one read of the scores and one write of each output (the generic kernels above make three passes).
// -----------------------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kSmWarps * 32)
attn_softmax_fwd_reg_kernel(const __nv_bfloat16* __restrict__ scores, long long ld,
                            const __nv_bfloat16* __restrict__ bias, long long ld_bias, int bias_mode,
                            int bias_head_major, __nv_bfloat16* __restrict__ probs,
                            __nv_bfloat16* __restrict__ probs_drop, const int* __restrict__ klen, long long nrows,
                            int heads, int Tq, int Tk, float scale, float clamp, uint32_t thresh16, float inv_keep,
                            unsigned long long seed, uint32_t stream_id,
                            const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kSmWarps + warp;
  if (row >= nrows) return;
  const int i = static_cast<int>(row % Tq);
  const long long z = row / Tq;
  const int b = static_cast<int>(z / heads);
  int kl = klen ? klen[b] : Tk;
  kl = kl < Tk ? kl : Tk;
  const int tk8 = (Tk + 7) & ~7;
  const __nv_bfloat16* srow = scores + row * ld;
  const long long brow_idx = bias_head_major ? ((z % heads) * (nrows / Tq / heads) + b) * Tq + i : row;
  const __nv_bfloat16* brow = bias ? bias + brow_idx * ld_bias : nullptr;
  const int bm = bias ? bias_mode : 0;
  const int shift = Tk - 1 - i;
  float v[NV][8];
  float m = -INFINITY;
#pragma unroll
  for (int n = 0; n < NV; ++n) {
    const int j0 = n * 256 + lane * 8;
    if (j0 < tk8) {
      const uint4 q = *reinterpret_cast<const uint4*>(srow + j0);
      const float f[8] = {bf16_lo(q.x), bf16_hi(q.x), bf16_lo(q.y), bf16_hi(q.y),
                          bf16_lo(q.z), bf16_hi(q.z), bf16_lo(q.w), bf16_hi(q.w)};
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int j = j0 + e;
        float x = f[e];
        if (j < kl) {
          if (bm == 1) x += __bfloat162float(brow[j]);
          else if (bm == 2) x += __bfloat162float(brow[j + shift]);
          x *= scale;
          if (clamp > 0.f) x = fminf(fmaxf(x, -clamp), clamp);
        } else {
          x = -INFINITY;
        }
        v[n][e] = x;
        m = fmaxf(m, x);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[n][e] = -INFINITY;
    }
  }
  m = warp_max(m);
  float sum = 0.f;
#pragma unroll
  for (int n = 0; n < NV; ++n)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float x = v[n][e] == -INFINITY ? 0.f : __expf(v[n][e] - m);
      v[n][e] = x;
      sum += x;
    }
  sum = warp_sum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  const unsigned long long grp_row = static_cast<unsigned long long>(row) * (tk8 >> 3);
  __nv_bfloat16* prow = probs + row * ld;
  __nv_bfloat16* drow = probs_drop ? probs_drop + row * ld : nullptr;
#pragma unroll
  for (int n = 0; n < NV; ++n) {
    const int j0 = n * 256 + lane * 8;
    if (j0 < tk8) {
      float pr[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) pr[e] = v[n][e] * inv;
      uint4 pv;
      pv.x = pack_bf16(pr[0], pr[1]); pv.y = pack_bf16(pr[2], pr[3]);
      pv.z = pack_bf16(pr[4], pr[5]); pv.w = pack_bf16(pr[6], pr[7]);
      *reinterpret_cast<uint4*>(prow + j0) = pv;
      if (drow != nullptr) {
        if (thresh16 != 0) {
          const uint32_t keep = dropout_keep8(seed, stream_id, grp_row + (j0 >> 3), thresh16);
#pragma unroll
          for (int e = 0; e < 8; ++e) pr[e] = ((keep >> e) & 1u) ? pr[e] * inv_keep : 0.f;
          pv.x = pack_bf16(pr[0], pr[1]); pv.y = pack_bf16(pr[2], pr[3]);
          pv.z = pack_bf16(pr[4], pr[5]); pv.w = pack_bf16(pr[6], pr[7]);
        }
        *reinterpret_cast<uint4*>(drow + j0) = pv;
      }
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(kSmWarps * 32)
attn_softmax_bwd_reg_kernel(const __nv_bfloat16* __restrict__ probs, const __nv_bfloat16* __restrict__ dpd,
                            __nv_bfloat16* __restrict__ dscores, __nv_bfloat16* __restrict__ dbias, long long ld_bias,
                            int bias_head_major, const int* __restrict__ klen, long long nrows, int heads, int Tq,
                            int Tk, long long ld, float scale, uint32_t thresh16, float inv_keep,
                            unsigned long long seed, uint32_t stream_id,
                            const unsigned long long* __restrict__ seed_off) {
  pdl_entry();
  if (thresh16 != 0 && seed_off != nullptr) seed += *seed_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kSmWarps + warp;
  if (row >= nrows) return;
  const int i = static_cast<int>(row % Tq);
  const long long z = row / Tq;
  const int b = static_cast<int>(z / heads);
  int kl = klen ? klen[b] : Tk;
  kl = kl < Tk ? kl : Tk;
  const int tk8 = (Tk + 7) & ~7;
  const unsigned long long grp_row = static_cast<unsigned long long>(row) * (tk8 >> 3);
  const __nv_bfloat16* prow = probs + row * ld;
  const __nv_bfloat16* grow = dpd + row * ld;
  float pq[NV][8], dp[NV][8];
  float dot = 0.f;
#pragma unroll
  for (int n = 0; n < NV; ++n) {
    const int j0 = n * 256 + lane * 8;
    if (j0 < tk8) {
      const uint4 pv = *reinterpret_cast<const uint4*>(prow + j0);
      const uint4 gv = *reinterpret_cast<const uint4*>(grow + j0);
      const float pf[8] = {bf16_lo(pv.x), bf16_hi(pv.x), bf16_lo(pv.y), bf16_hi(pv.y),
                           bf16_lo(pv.z), bf16_hi(pv.z), bf16_lo(pv.w), bf16_hi(pv.w)};
      const float gf[8] = {bf16_lo(gv.x), bf16_hi(gv.x), bf16_lo(gv.y), bf16_hi(gv.y),
                           bf16_lo(gv.z), bf16_hi(gv.z), bf16_lo(gv.w), bf16_hi(gv.w)};
      uint32_t keep = 0xffu;
      float ik = 1.f;
      if (thresh16 != 0) {
        keep = dropout_keep8(seed, stream_id, grp_row + (j0 >> 3), thresh16);
        ik = inv_keep;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const bool live = j0 + e < kl;
        pq[n][e] = live ? pf[e] : 0.f;
        dp[n][e] = (live && ((keep >> e) & 1u)) ? gf[e] * ik : 0.f;
        dot += dp[n][e] * pq[n][e];
      }
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) pq[n][e] = dp[n][e] = 0.f;
    }
  }
  dot = warp_sum(dot);
  __nv_bfloat16* orow = dscores + row * ld;
#pragma unroll
  for (int n = 0; n < NV; ++n) {
    const int j0 = n * 256 + lane * 8;
    if (j0 < tk8) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = scale * pq[n][e] * (dp[n][e] - dot);
      uint4 ov;
      ov.x = pack_bf16(o[0], o[1]); ov.y = pack_bf16(o[2], o[3]);
      ov.z = pack_bf16(o[4], o[5]); ov.w = pack_bf16(o[6], o[7]);
      *reinterpret_cast<uint4*>(orow + j0) = ov;
    }
  }
  if (dbias != nullptr) {
    // un-shift: dBD[i][c] = dS[i][c + i - (Tk-1)] (zero outside), c in [0, 2Tk-1)
    __syncwarp();
    const long long brow_idx = bias_head_major ? ((z % heads) * (nrows / Tq / heads) + b) * Tq + i : row;
    __nv_bfloat16* brow = dbias + brow_idx * ld_bias;
    const int shift = Tk - 1 - i;
    const int nb = 2 * Tk - 1;
    const int nb8 = (nb + 7) & ~7;
    for (int cidx = lane; cidx < nb8; cidx += 32) {
      const int j = cidx - shift;
      __nv_bfloat16 val = __float2bfloat16(0.f);
      if (cidx < nb && j >= 0 && j < kl) val = orow[j];
      if (cidx < ld_bias) brow[cidx] = val;
    }
  }
}

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_attn_softmax_fwd(const void* scores, long long ld, const void* bias,
                                    long long ld_bias, int bias_mode, int bias_head_major, void* probs,
                                    void* probs_drop, const int* klen, int nbh, int heads, int Tq, int Tk, float scale,
                                    float clamp, float dropout_p, unsigned long long seed,
                                    unsigned int stream_id, const unsigned long long* seed_offset,
                                    void* stream) {
  S2T_REQUIRE(scores && probs, "s2t_attn_softmax_fwd: null pointer");
  S2T_REQUIRE(ld % 8 == 0 && ld >= ((Tk + 7) & ~7), "s2t_attn_softmax_fwd: ld must be a multiple of 8 and >= round8(Tk)");
  S2T_REQUIRE(heads > 0 && nbh % heads == 0, "s2t_attn_softmax_fwd: nbh must be a multiple of heads");
  S2T_REQUIRE(bias_mode != 2 || Tq == Tk, "s2t_attn_softmax_fwd: rel-shift bias needs Tq == Tk");
  S2T_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "s2t_attn_softmax_fwd: bad dropout_p");
  const long long nrows = static_cast<long long>(nbh) * Tq;
  if (nrows <= 0) return 0;
  const unsigned grid = static_cast<unsigned>((nrows + kSmWarps - 1) / kSmWarps);
  const int tk8 = (Tk + 7) & ~7;
  if (tk8 <= 256) {
    S2T_LAUNCH(attn_softmax_fwd_reg_kernel<1>, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(scores), ld, reinterpret_cast<const __nv_bfloat16*>(bias), ld_bias, bias_mode, bias_head_major, reinterpret_cast<__nv_bfloat16*>(probs), reinterpret_cast<__nv_bfloat16*>(probs_drop), klen, nrows, heads, Tq, Tk, scale, clamp, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  } else if (tk8 <= 512) {
    S2T_LAUNCH(attn_softmax_fwd_reg_kernel<2>, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(scores), ld, reinterpret_cast<const __nv_bfloat16*>(bias), ld_bias, bias_mode, bias_head_major, reinterpret_cast<__nv_bfloat16*>(probs), reinterpret_cast<__nv_bfloat16*>(probs_drop), klen, nrows, heads, Tq, Tk, scale, clamp, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  } else if (tk8 <= 1024) {
    S2T_LAUNCH(attn_softmax_fwd_reg_kernel<4>, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(scores), ld, reinterpret_cast<const __nv_bfloat16*>(bias), ld_bias, bias_mode, bias_head_major, reinterpret_cast<__nv_bfloat16*>(probs), reinterpret_cast<__nv_bfloat16*>(probs_drop), klen, nrows, heads, Tq, Tk, scale, clamp, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  } else {
    S2T_LAUNCH(attn_softmax_fwd_kernel, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(scores), ld, reinterpret_cast<const __nv_bfloat16*>(bias), ld_bias, bias_mode, bias_head_major, reinterpret_cast<__nv_bfloat16*>(probs), reinterpret_cast<__nv_bfloat16*>(probs_drop), klen, nrows, heads, Tq, Tk, scale, clamp, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  }
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_attn_softmax_bwd(const void* probs, const void* dprobs_drop, void* dscores,
                                    void* dbias, long long ld_bias, int bias_head_major, const int* klen,
                                    int nbh,
                                    int heads, int Tq, int Tk, long long ld, float scale,
                                    float dropout_p, unsigned long long seed, unsigned int stream_id,
                                    const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(probs && dprobs_drop && dscores, "s2t_attn_softmax_bwd: null pointer");
  S2T_REQUIRE(ld % 8 == 0 && ld >= ((Tk + 7) & ~7), "s2t_attn_softmax_bwd: ld must be a multiple of 8 and >= round8(Tk)");
  S2T_REQUIRE(heads > 0 && nbh % heads == 0, "s2t_attn_softmax_bwd: nbh must be a multiple of heads");
  S2T_REQUIRE(dbias == nullptr || Tq == Tk, "s2t_attn_softmax_bwd: rel-shift bias needs Tq == Tk");
  const long long nrows = static_cast<long long>(nbh) * Tq;
  if (nrows <= 0) return 0;
  const unsigned grid = static_cast<unsigned>((nrows + kSmWarps - 1) / kSmWarps);
  const int tk8 = (Tk + 7) & ~7;
  if (tk8 <= 256) {
    S2T_LAUNCH(attn_softmax_bwd_reg_kernel<1>, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(probs), reinterpret_cast<const __nv_bfloat16*>(dprobs_drop), reinterpret_cast<__nv_bfloat16*>(dscores), reinterpret_cast<__nv_bfloat16*>(dbias), ld_bias, bias_head_major, klen, nrows, heads, Tq, Tk, ld, scale, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  } else if (tk8 <= 512) {
    S2T_LAUNCH(attn_softmax_bwd_reg_kernel<2>, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(probs), reinterpret_cast<const __nv_bfloat16*>(dprobs_drop), reinterpret_cast<__nv_bfloat16*>(dscores), reinterpret_cast<__nv_bfloat16*>(dbias), ld_bias, bias_head_major, klen, nrows, heads, Tq, Tk, ld, scale, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  } else if (tk8 <= 1024) {
    S2T_LAUNCH(attn_softmax_bwd_reg_kernel<4>, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(probs), reinterpret_cast<const __nv_bfloat16*>(dprobs_drop), reinterpret_cast<__nv_bfloat16*>(dscores), reinterpret_cast<__nv_bfloat16*>(dbias), ld_bias, bias_head_major, klen, nrows, heads, Tq, Tk, ld, scale, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  } else {
    S2T_LAUNCH(attn_softmax_bwd_kernel, grid, kSmWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(probs), reinterpret_cast<const __nv_bfloat16*>(dprobs_drop), reinterpret_cast<__nv_bfloat16*>(dscores), reinterpret_cast<__nv_bfloat16*>(dbias), ld_bias, bias_head_major, klen, nrows, heads, Tq, Tk, ld, scale, dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u, 1.f / (1.f - dropout_p), seed, stream_id, seed_offset);
  }
  S2T_LAUNCH_CHECK();
  return 0;
}
