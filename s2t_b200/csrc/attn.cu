// Fused self-attention forward for one (utterance, head, 128-query tile): S = Q K^T on tcgen05 into TMEM, masked fp32
// softmax straight out of TMEM, dropout, P V on tcgen05, context written once.  Replaces three launches
// (scores GEMM, softmax, PV GEMM) and the bf16 round trip of the scores through HBM/L2.
//
// Reference semantics (fairseq MultiheadAttention, modules/multihead_attention.py:246-265, 372-416): q is scaled by
// d_k^-0.5, key padding masked with -inf, softmax in fp32, dropout on the probabilities, bmm with v.
// The probabilities (and their dropped copy) are still written out: the backward pass consumes them
// (softmax.cu attn_softmax_bwd regenerates the same Philox mask from (row, column group)).
//
// Shape limits: d_k == 64 (one 128-byte swizzle row), T <= 256 keys (one key tile: the whole row of scores lives in 256
// TMEM columns), no relative-position bias (the rel_pos path keeps the three-kernel form).
//
// Warp roles (576 threads): warp 0 lane 0 = TMA loads + both MMA issues; warp 1 = TMEM allocation; warps 2..17 = softmax:
// TMEM lane quarter = warp % 4 (one query row per lane), the four warps of a quarter each own 64 of the 256 key columns
// (exactly one k-block of the P operand tile), read them from TMEM ONCE into registers, and exchange the row maximum and
// the row sum through shared memory (one named barrier per quarter and pass): one exp per score.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {
namespace {

constexpr int kAttnThreads = 576;
constexpr int kDk = 64;
constexpr int kTq = 128;    // queries per CTA
constexpr int kTk = 256;    // keys per CTA (all of them)
constexpr int kSmQ = kTq * kDk * 2;          // 16 KB
constexpr int kSmK = kTk * kDk * 2;          // 32 KB
constexpr int kSmV = kTk * kDk * 2;          // 32 KB, 4 k-blocks of [64 keys][64 d_k]
constexpr int kSmP = kTq * kTk * 2;          // 64 KB, 4 k-blocks of [128 rows][64 keys], K-major, 128B swizzle
constexpr int kSmRed = 2 * 4 * kTq * 4;          // row max / row sum partials [2][4 parts][128 rows] fp32
constexpr int kAttnSmem = kSmQ + kSmK + kSmV + kSmP + kSmRed + 1024 + 256;

struct AttnParams {
  __nv_bfloat16* probs;        // [B*H][T][ld]
  __nv_bfloat16* probs_drop;   // same, after dropout (nullptr when dropout is off)
  long long ld;
  __nv_bfloat16* ctx;          // [B*T][ldctx], head h at columns h*64
  long long ldctx;
  const int* klen;             // [B] valid keys (nullptr: T)
  int B, H, T;
  float scale;
  uint32_t thresh16;
  float inv_keep;
  unsigned long long seed;
  const unsigned long long* seed_off;
  uint32_t stream_id;
};

__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fused_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kSmQ;
  uint8_t* sV = sK + kSmK;
  uint8_t* sP = sV + kSmV;
  float* s_red = reinterpret_cast<float*>(sP + kSmP);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kSmP + kSmRed);   // 0: Q+K landed, 1: V landed, 2: S ready, 3: P written, 4: O ready
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kTq, h = blockIdx.y, b = blockIdx.z;

  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    mbar_init(&bars[3], 512);
    mbar_init(&bars[4], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);     // S: columns [0, 256), O: columns [256, 320)
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(&bars[0], kSmQ + kSmK);
      tma_load_4d(sQ, &tmQ, &bars[0], 0, m0, h, b);
      tma_load_4d(sK, &tmK, &bars[0], 0, 0, h, b);
      mbar_expect_tx(&bars[1], kSmV);
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) tma_load_4d(sV + kb * 8192, &tmV, &bars[1], 0, kb * 64, h, b);
      // S = Q K^T : M = 128 queries, N = 256 keys, K = 64
      mbar_wait(&bars[0], 0);
      tc_fence_after();
      {
        constexpr uint32_t idesc = umma_idesc_bf16(kTq, kTk, 0, 0);
        const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tmem_base, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024), idesc,
                    k != 0 ? 1u : 0u);
        umma_commit(&bars[2]);
      }
      // O = P V : M = 128, N = 64 (d_k), K = 256 keys; P from shared memory (written by the softmax warps)
      mbar_wait(&bars[1], 0);
      mbar_wait(&bars[3], 0);
      tc_fence_after();
      {
        constexpr uint32_t idesc = umma_idesc_bf16(kTq, kDk, 0, 1);
        const uint32_t pa = smem_u32(sP), va = smem_u32(sV);
#pragma unroll
        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base + 256, umma_smem_desc(pa + kb * 16384 + k * 32, 16, 1024),
                      umma_smem_desc(va + kb * 8192 + k * 2048, 64 * 128, 1024), idesc, (kb | k) != 0 ? 1u : 0u);
        umma_commit(&bars[4]);
      }
    }
  } else if (warp >= 2) {
    const int q = warp & 3;
    const int part = (warp - 2) >> 2;        // which 64-column quarter of the key axis (= k-block of the P tile)
    const int row = q * 32 + lane;          // TMEM lane = query row of the tile
    const int i = m0 + row;
    const bool row_ok = i < p.T;
    int kl = p.klen ? p.klen[b] : p.T;
    kl = kl < p.T ? kl : p.T;
    unsigned long long seed = p.seed;
    if (p.thresh16 != 0 && p.seed_off != nullptr) seed += *p.seed_off;
    const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const long long grow = (static_cast<long long>(b) * p.H + h) * p.T + i;       // row of probs
    const int tk8 = (p.T + 7) & ~7;
    const unsigned long long grp_row = static_cast<unsigned long long>(grow) * (tk8 >> 3);
    const int col0 = part * 64;
    mbar_wait(&bars[2], 0);
    tc_fence_after();
    // my 64 scores -> registers (scaled; masked keys = -inf)
    float x[64];
    {
      uint32_t r0[32], r1[32];
      tmem_ld_32x32(t_addr + col0, r0);
      tmem_ld_32x32(t_addr + col0 + 32, r1);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        x[e] = col0 + e < kl ? __uint_as_float(r0[e]) * p.scale : -INFINITY;
        x[32 + e] = col0 + 32 + e < kl ? __uint_as_float(r1[e]) * p.scale : -INFINITY;
      }
    }
    float m = -INFINITY;
#pragma unroll
    for (int e = 0; e < 64; ++e) m = fmaxf(m, x[e]);
    s_red[part * kTq + row] = m;
    named_bar_sync(1 + q, 128);
    m = fmaxf(fmaxf(s_red[row], s_red[kTq + row]), fmaxf(s_red[2 * kTq + row], s_red[3 * kTq + row]));
    float sum = 0.f;
#pragma unroll
    for (int e = 0; e < 64; ++e) {
      x[e] = x[e] == -INFINITY ? 0.f : __expf(x[e] - m);
      sum += x[e];
    }
    s_red[(4 + part) * kTq + row] = sum;
    named_bar_sync(1 + q, 128);
    sum = (s_red[4 * kTq + row] + s_red[5 * kTq + row]) + (s_red[6 * kTq + row] + s_red[7 * kTq + row]);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    // probabilities -> global (for backward), dropped probabilities -> global and the PV operand tile
    __nv_bfloat16* prow = p.probs + grow * p.ld;
    __nv_bfloat16* drow = p.probs_drop ? p.probs_drop + grow * p.ld : nullptr;
    uint8_t* pdst = sP + part * 16384 + row * 128;
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      const int j0 = col0 + g * 8;
      float pr[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) pr[e] = x[g * 8 + e] * inv;
      uint4 pv;
      pv.x = pack_bf16(pr[0], pr[1]); pv.y = pack_bf16(pr[2], pr[3]);
      pv.z = pack_bf16(pr[4], pr[5]); pv.w = pack_bf16(pr[6], pr[7]);
      if (row_ok && j0 < tk8) *reinterpret_cast<uint4*>(prow + j0) = pv;
      if (p.thresh16 != 0) {
        const uint32_t keep = dropout_keep8(seed, p.stream_id, grp_row + (j0 >> 3), p.thresh16);
#pragma unroll
        for (int e = 0; e < 8; ++e) pr[e] = ((keep >> e) & 1u) ? pr[e] * p.inv_keep : 0.f;
        pv.x = pack_bf16(pr[0], pr[1]); pv.y = pack_bf16(pr[2], pr[3]);
        pv.z = pack_bf16(pr[4], pr[5]); pv.w = pack_bf16(pr[6], pr[7]);
      }
      if (drow != nullptr && row_ok && j0 < tk8) *reinterpret_cast<uint4*>(drow + j0) = pv;
      // A operand of the second MMA: K-major, 128-byte rows, 16-byte chunk index XOR (row & 7)
      *reinterpret_cast<uint4*>(pdst + ((g ^ (row & 7)) << 4)) = row_ok ? pv : make_uint4(0u, 0u, 0u, 0u);
    }
    fence_proxy_async();
    tc_fence_before();
    mbar_arrive(&bars[3]);
    // epilogue: context rows (two 32-column chunks: parts 0 and 1)
    if (part < 2) {
      mbar_wait(&bars[4], 0);
      tc_fence_after();
      uint32_t rg[32];
      tmem_ld_32x32(t_addr + 256 + part * 32, rg);
      tmem_ld_wait();
      if (row_ok) {
        __nv_bfloat16* o = p.ctx + (static_cast<long long>(b) * p.T + i) * p.ldctx + h * kDk + part * 32;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 ov;
          ov.x = pack_bf16(__uint_as_float(rg[g * 8 + 0]), __uint_as_float(rg[g * 8 + 1]));
          ov.y = pack_bf16(__uint_as_float(rg[g * 8 + 2]), __uint_as_float(rg[g * 8 + 3]));
          ov.z = pack_bf16(__uint_as_float(rg[g * 8 + 4]), __uint_as_float(rg[g * 8 + 5]));
          ov.w = pack_bf16(__uint_as_float(rg[g * 8 + 6]), __uint_as_float(rg[g * 8 + 7]));
          *reinterpret_cast<uint4*>(o + g * 8) = ov;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace
}  // namespace s2t

using namespace s2t;

extern "C" int s2t_attn_fused_fwd(const void* qkv, long long ldqkv, int C, void* probs, void* probs_drop, long long ld,
                                  void* ctx, long long ldctx, const int* klen, int batch, int heads, int T, int dk,
                                  float scale, float dropout_p, unsigned long long seed, unsigned int stream_id,
                                  const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(qkv && probs && ctx, "s2t_attn_fused_fwd: null pointer");
  S2T_REQUIRE(dk == kDk && T >= 1 && T <= kTk, "s2t_attn_fused_fwd: needs d_k == 64 and T <= 256 (got d_k=%d, T=%d)", dk, T);
  S2T_REQUIRE(heads * dk <= C && ldqkv >= 3ll * C && ldqkv % 8 == 0 && C % 8 == 0, "s2t_attn_fused_fwd: bad qkv layout");
  S2T_REQUIRE(ld % 8 == 0 && ld >= ((T + 7) & ~7) && ldctx % 8 == 0, "s2t_attn_fused_fwd: ld / ldctx must be multiples of 8");
  S2T_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "s2t_attn_fused_fwd: bad dropout_p");
  S2T_REQUIRE(dropout_p == 0.f || probs_drop != nullptr, "s2t_attn_fused_fwd: dropout needs the probs_drop output");
  if (batch <= 0) return 0;
  const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(qkv);
  CUtensorMap tq, tk, tv;
  int rc = make_operand_map(&tq, base, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTq);
  if (rc) return rc;
  rc = make_operand_map(&tk, base + C, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTk);
  if (rc) return rc;
  rc = make_operand_map(&tv, base + 2 * C, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, 64);
  if (rc) return rc;
  AttnParams p;
  memset(&p, 0, sizeof(p));
  p.probs = reinterpret_cast<__nv_bfloat16*>(probs);
  p.probs_drop = dropout_p > 0.f ? reinterpret_cast<__nv_bfloat16*>(probs_drop) : nullptr;
  p.ld = ld;
  p.ctx = reinterpret_cast<__nv_bfloat16*>(ctx);
  p.ldctx = ldctx;
  p.klen = klen;
  p.B = batch; p.H = heads; p.T = T;
  p.scale = scale;
  p.thresh16 = dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u;
  p.inv_keep = 1.f / (1.f - dropout_p);
  p.seed = seed;
  p.seed_off = seed_offset;
  p.stream_id = stream_id;
  static bool configured = false;
  if (!configured) {
    S2T_CHECK_CUDA(cudaFuncSetAttribute(attn_fused_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem));
    configured = true;
  }
  S2T_LAUNCH(attn_fused_fwd_kernel, dim3((T + kTq - 1) / kTq, heads, batch), kAttnThreads, kAttnSmem,
             reinterpret_cast<cudaStream_t>(stream), tq, tk, tv, p);
  return 0;
}
