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
  __nv_bfloat16* probs;        // [B*H][T][ld], or nullptr (the fused backward recomputes them from lse)
  __nv_bfloat16* probs_drop;   // same, after dropout (nullptr when dropout is off or probs is nullptr)
  float* lse;                  // [B*H][T] log-sum-exp of the scaled, masked scores (nullptr: not wanted)
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
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                      const AttnParams p) {
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
    tma_prefetch_desc(&tmO);
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
    if (p.lse != nullptr && part == 0 && row_ok) p.lse[grow] = sum > 0.f ? m + __logf(sum) : 0.f;
    // probabilities -> global (when the caller keeps them), dropped probabilities -> global and the PV operand tile
    __nv_bfloat16* prow = p.probs ? p.probs + grow * p.ld : nullptr;
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
      if (prow != nullptr && row_ok && j0 < tk8) *reinterpret_cast<uint4*>(prow + j0) = pv;
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
    // epilogue: context rows (parts 0 and 1: 32 columns each) -> bf16 tile in the retired Q buffer (swizzled 128-byte rows)
    // -> ONE TMA store (coalesced, clipped at T by the tensor map) instead of 32 row-strided 16-byte stores per warp
    if (part < 2) {
      mbar_wait(&bars[4], 0);
      tc_fence_after();
      uint32_t rg[32];
      tmem_ld_32x32(t_addr + 256 + part * 32, rg);
      tmem_ld_wait();
      uint8_t* orow = sQ + row * 128;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 ov;
        ov.x = pack_bf16(__uint_as_float(rg[g * 8 + 0]), __uint_as_float(rg[g * 8 + 1]));
        ov.y = pack_bf16(__uint_as_float(rg[g * 8 + 2]), __uint_as_float(rg[g * 8 + 3]));
        ov.z = pack_bf16(__uint_as_float(rg[g * 8 + 4]), __uint_as_float(rg[g * 8 + 5]));
        ov.w = pack_bf16(__uint_as_float(rg[g * 8 + 6]), __uint_as_float(rg[g * 8 + 7]));
        *reinterpret_cast<uint4*>(orow + (((part * 4 + g) ^ (row & 7)) << 4)) = ov;
      }
      fence_proxy_async();
      named_bar_sync(5, 256);
      if (warp == 2 && lane == 0) {
        tma_store_4d(&tmO, sQ, 0, m0, h, b);
        tma_store_commit();
        tma_store_wait_read<0>();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}


// ---------------------------------------------------------------------------------------------
// Fused self-attention BACKWARD for one (utterance, head): a cluster of two CTAs, one per 128-query tile.
//   S  = Q_i K^T, dP = dO_i V^T                      (tcgen05, 2 x 256 TMEM columns)
//   P  = exp(scale * S - lse) (key mask), Pd = dropout(P)            } 16 softmax warps, straight out of TMEM;
//   dS = P * (dropout'(dP) - delta), delta_i = dO_i . O_i            } no row reductions: lse and delta are per-row scalars
//   dQ_i = scale * dS K ;  dK += scale * dS^T Q_i ;  dV += Pd^T dO_i  (tcgen05; Pd / dS tiles in shared memory, read as the
//                                                    K-major A operand for dQ and as the MN-major A operand for dK / dV)
// dK / dV are sums over the two query tiles: rank r owns keys [128 r, 128 r + 128) and receives the peer's fp32 partial of
// that half over DSMEM (push, 64 KB each way into the retired Pd tile), adds its own from TMEM and stores bf16.
// Replaces five launches (dV, dP, softmax', dK, dQ) and the round trips of P, dP and dS through HBM/L2; the probabilities are
// recomputed from the saved row log-sum-exp, so the forward pass no longer writes them.
// ---------------------------------------------------------------------------------------------
constexpr int kSmDO = kTq * kDk * 2;                                   // 16 KB
constexpr int kSmT = kTq * kTk * 2;                                    // 64 KB per [128][256] bf16 tile
constexpr int kBwdSmem = kSmQ + kSmK + kSmV + kSmDO + 2 * kSmT + 1024 + 256;   // 225.25 KB of the 227 KB a CTA may own

struct AttnBwdParams {
  const __nv_bfloat16* o;      // forward context [B*T][ldo] (for delta)
  const __nv_bfloat16* d_o;    // its gradient, same layout
  long long ldo;
  const float* lse;            // [B*H][T]
  __nv_bfloat16* dqkv;         // [B*T][ldg]: dq | dk | dv at columns 0 | C | 2C, head h at h*64
  long long ldg;
  int C;
  const int* klen;
  int B, H, T;
  float scale;
  uint32_t thresh16;
  float inv_keep;
  unsigned long long seed;
  const unsigned long long* seed_off;
  uint32_t stream_id;
  unsigned long long* trace;   // debug phase stamps of CTA (0,0,0) (S2T_GEMM_TRACE=1)
};

__device__ __forceinline__ void attn_stamp(unsigned long long* trace, int slot) {
  if (trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[slot] = t;
  }
}

__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fused_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                      const __grid_constant__ CUtensorMap tmDQ, const __grid_constant__ CUtensorMap tmDK,
                      const __grid_constant__ CUtensorMap tmDV, const AttnBwdParams p) {
  pdl_launch_dependents();
  if (threadIdx.x == 0) attn_stamp(p.trace, 0);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kSmQ;
  uint8_t* sV = sK + kSmK;
  uint8_t* sDO = sV + kSmV;
  uint8_t* sPd = sDO + kSmDO;      // dropped probabilities; later the receive buffer of the dK / dV exchange
  uint8_t* sDS = sPd + kSmT;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + kSmT);            // 0: operands landed, 1: S + dP ready, 2: tiles written, 3: MMAs done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = static_cast<int>(cluster_ctarank());               // query tile of this CTA == key half it finalises
  const int m0 = rank * kTq, h = blockIdx.y, b = blockIdx.z;

  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 512);
    mbar_init(&bars[3], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ);
    tma_prefetch_desc(&tmDK);
    tma_prefetch_desc(&tmDV);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);     // S: [0, 256), dP: [256, 512); then dQ [0, 64), dV halves [64, 192), dK halves [192, 320)
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) attn_stamp(p.trace, 1);
  pdl_wait();
  if (threadIdx.x == 0) attn_stamp(p.trace, 2);

  const int q = warp & 3;
  const int part = warp >= 2 ? (warp - 2) >> 2 : 0;
  const int row = q * 32 + lane;
  const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(&bars[0], kSmQ + kSmK + kSmV + kSmDO);
      tma_load_4d(sQ, &tmQ, &bars[0], 0, m0, h, b);
      tma_load_4d(sDO, &tmDO, &bars[0], 0, m0, h, b);
      tma_load_4d(sK, &tmK, &bars[0], 0, 0, h, b);
      tma_load_4d(sV, &tmV, &bars[0], 0, 0, h, b);
      mbar_wait(&bars[0], 0);
      attn_stamp(p.trace, 3);
      tc_fence_after();
      const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK), va = smem_u32(sV), doa = smem_u32(sDO);
      {
        constexpr uint32_t idesc = umma_idesc_bf16(kTq, kTk, 0, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)      // S = Q K^T
          umma_bf16(tmem_base, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024), idesc,
                    k != 0 ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k)      // dP = dO V^T
          umma_bf16(tmem_base + 256, umma_smem_desc(doa + k * 32, 16, 1024), umma_smem_desc(va + k * 32, 16, 1024), idesc,
                    k != 0 ? 1u : 0u);
        umma_commit(&bars[1]);
      }
      mbar_wait(&bars[2], 0);            // Pd and dS tiles are in shared memory, S / dP have been read out of TMEM
      attn_stamp(p.trace, 5);
      tc_fence_after();
      const uint32_t pa = smem_u32(sPd), sa = smem_u32(sDS);
      {
        // dQ[i][d] = sum_j dS[i][j] K[j][d]: A = dS K-major (4 k-blocks of 64 keys), B = K MN-major
        constexpr uint32_t idesc = umma_idesc_bf16(kTq, kDk, 0, 1);
#pragma unroll
        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base, umma_smem_desc(sa + kb * 16384 + k * 32, 16, 1024),
                      umma_smem_desc(ka + kb * 8192 + k * 2048, 64 * 128, 1024), idesc, (kb | k) != 0 ? 1u : 0u);
      }
      {
        // dV[j][d] = sum_i Pd[i][j] dO[i][d], dK[j][d] = sum_i dS[i][j] Q[i][d]: A = the tile read MN-major (M = 128 keys of
        // half hj = two 64-key chunks 16 KB apart, K = 128 queries in steps of 16 rows = 2 KB), B = dO / Q MN-major
        constexpr uint32_t idesc = umma_idesc_bf16(128, kDk, 1, 1);
#pragma unroll
        for (int hj = 0; hj < 2; ++hj) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_bf16(tmem_base + 64 + hj * 64, umma_smem_desc(pa + hj * 32768 + ks * 2048, 16384, 1024),
                      umma_smem_desc(doa + ks * 2048, 64 * 128, 1024), idesc, ks != 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_bf16(tmem_base + 192 + hj * 64, umma_smem_desc(sa + hj * 32768 + ks * 2048, 16384, 1024),
                      umma_smem_desc(qa + ks * 2048, 64 * 128, 1024), idesc, ks != 0 ? 1u : 0u);
        }
      }
      umma_commit(&bars[3]);
    }
  } else if (warp >= 2) {
    const int i = m0 + row;
    const bool row_ok = i < p.T;
    int kl = p.klen ? p.klen[b] : p.T;
    kl = kl < p.T ? kl : p.T;
    unsigned long long seed = p.seed;
    if (p.thresh16 != 0 && p.seed_off != nullptr) seed += *p.seed_off;
    const long long grow = (static_cast<long long>(b) * p.H + h) * p.T + i;
    const int tk8 = (p.T + 7) & ~7;
    const unsigned long long grp_row = static_cast<unsigned long long>(grow) * (tk8 >> 3);
    const int col0 = part * 64;
    // delta_i = dO_i . O_i for the 32 rows of this warp, with coalesced loads: one load instruction covers four whole
    // 128-byte rows (8 lanes x 16 bytes each); the 8-lane partial dots are folded by shuffles and row L is handed to lane L
    // (a thread-per-row walk costs 32 cache lines per load instruction: 4 us of LSU time per CTA)
    float delta = 0.f, lse = 0.f;
    {
      const int sub = lane >> 3, l8 = lane & 7;
#pragma unroll
      for (int itr = 0; itr < 8; ++itr) {
        const int gi = m0 + q * 32 + itr * 4 + sub;
        float d = 0.f;
        if (gi < p.T) {
          const long long off = (static_cast<long long>(b) * p.T + gi) * p.ldo + h * kDk + l8 * 8;
          const uint4 ov = *reinterpret_cast<const uint4*>(p.o + off);
          const uint4 gv = *reinterpret_cast<const uint4*>(p.d_o + off);
          d = bf16_lo(ov.x) * bf16_lo(gv.x) + bf16_hi(ov.x) * bf16_hi(gv.x) + bf16_lo(ov.y) * bf16_lo(gv.y) +
              bf16_hi(ov.y) * bf16_hi(gv.y) + bf16_lo(ov.z) * bf16_lo(gv.z) + bf16_hi(ov.z) * bf16_hi(gv.z) +
              bf16_lo(ov.w) * bf16_lo(gv.w) + bf16_hi(ov.w) * bf16_hi(gv.w);
        }
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        d += __shfl_xor_sync(0xffffffffu, d, 4);
        const float v = __shfl_sync(0xffffffffu, d, (lane & 3) * 8);     // row itr*4 + (lane & 3)
        if ((lane >> 2) == itr) delta = v;
      }
      if (row_ok) lse = p.lse[grow];
    }
    mbar_wait(&bars[1], 0);
    if (threadIdx.x == 64) attn_stamp(p.trace, 4);
    tc_fence_after();
    uint8_t* pd_dst = sPd + part * 16384 + row * 128;
    uint8_t* ds_dst = sDS + part * 16384 + row * 128;
#pragma unroll 1
    for (int hf = 0; hf < 2; ++hf) {
      uint32_t rs[32], rp[32];
      tmem_ld_32x32(t_addr + col0 + hf * 32, rs);
      tmem_ld_32x32(t_addr + 256 + col0 + hf * 32, rp);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int j0 = col0 + hf * 32 + g * 8;
        uint32_t keep = 0xffu;
        if (p.thresh16 != 0) keep = dropout_keep8(seed, p.stream_id, grp_row + (j0 >> 3), p.thresh16);
        float pd[8], ds[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const bool live = row_ok && (j0 + e) < kl;
          const float pr = live ? __expf(__uint_as_float(rs[g * 8 + e]) * p.scale - lse) : 0.f;
          const float kf = ((keep >> e) & 1u) ? p.inv_keep : 0.f;
          pd[e] = pr * kf;
          ds[e] = pr * (kf * __uint_as_float(rp[g * 8 + e]) - delta);
        }
        uint4 v;
        v.x = pack_bf16(pd[0], pd[1]); v.y = pack_bf16(pd[2], pd[3]);
        v.z = pack_bf16(pd[4], pd[5]); v.w = pack_bf16(pd[6], pd[7]);
        const int chunk = ((hf * 4 + g) ^ (row & 7)) << 4;
        *reinterpret_cast<uint4*>(pd_dst + chunk) = v;
        v.x = pack_bf16(ds[0], ds[1]); v.y = pack_bf16(ds[2], ds[3]);
        v.z = pack_bf16(ds[4], ds[5]); v.w = pack_bf16(ds[6], ds[7]);
        *reinterpret_cast<uint4*>(ds_dst + chunk) = v;
      }
    }
    fence_proxy_async();
    tc_fence_before();
    mbar_arrive(&bars[2]);
    mbar_wait(&bars[3], 0);
    if (threadIdx.x == 64) attn_stamp(p.trace, 6);
    tc_fence_after();
    // dQ rows of this tile (parts 0 and 1: 32 columns each) -> bf16 tile in the retired Q buffer -> one TMA store
    if (part < 2) {
      uint32_t rg[32];
      tmem_ld_32x32(t_addr + part * 32, rg);
      tmem_ld_wait();
      uint8_t* orow = sQ + row * 128;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 ov;
        ov.x = pack_bf16(p.scale * __uint_as_float(rg[g * 8 + 0]), p.scale * __uint_as_float(rg[g * 8 + 1]));
        ov.y = pack_bf16(p.scale * __uint_as_float(rg[g * 8 + 2]), p.scale * __uint_as_float(rg[g * 8 + 3]));
        ov.z = pack_bf16(p.scale * __uint_as_float(rg[g * 8 + 4]), p.scale * __uint_as_float(rg[g * 8 + 5]));
        ov.w = pack_bf16(p.scale * __uint_as_float(rg[g * 8 + 6]), p.scale * __uint_as_float(rg[g * 8 + 7]));
        *reinterpret_cast<uint4*>(orow + (((part * 4 + g) ^ (row & 7)) << 4)) = ov;
      }
      fence_proxy_async();
      named_bar_sync(5, 256);
      if (warp == 2 && lane == 0) {
        tma_store_4d(&tmDQ, sQ, 0, m0, h, b);
        tma_store_commit();
      }
    }
  }
  // ---- dK / dV: exchange the key halves between the two query-tile CTAs ---------------------------------------------
  // every MMA of this CTA has retired before it arrives here (bars[3]; warp 0 lane 0 issued the commit, the softmax warps
  // waited for it), so after the first cluster barrier the peer's Pd tile may be overwritten
  __syncwarp();
  if (warp < 2) {            // the two control warps never waited for bars[3]: do it now so that "arrived" means "MMAs retired"
    mbar_wait(&bars[3], 0);
    tc_fence_after();
  }
  tc_fence_before();
  __syncwarp();
  if (threadIdx.x == 64) attn_stamp(p.trace, 7);
  cluster_sync_all();
  if (threadIdx.x == 64) attn_stamp(p.trace, 8);
  // which accumulator a thread moves: parts 0 / 1 = dV columns [0, 32) / [32, 64), parts 2 / 3 = dK columns likewise
  const int which_k = part >> 1;                  // 0 = dV, 1 = dK
  const int chalf = part & 1;
  const uint32_t recv_local = smem_u32(sPd);
  if (warp >= 2) {
    const int peer = rank ^ 1;
    uint32_t rg[32];
    tmem_ld_32x32(t_addr + 64 + which_k * 128 + peer * 64 + chalf * 32, rg);   // my partial of the PEER's key half
    tmem_ld_wait();
    const uint32_t dst = dsmem_addr(recv_local, static_cast<uint32_t>(peer));
#pragma unroll
    for (int g4 = 0; g4 < 8; ++g4) {
      const uint32_t off = static_cast<uint32_t>(((part * 8 + g4) * 128 + row) << 4);
      dsmem_st_f4(dst + off, make_float4(__uint_as_float(rg[g4 * 4]), __uint_as_float(rg[g4 * 4 + 1]),
                                         __uint_as_float(rg[g4 * 4 + 2]), __uint_as_float(rg[g4 * 4 + 3])));
    }
  }
  tc_fence_before();
  __syncwarp();
  if (threadIdx.x == 64) attn_stamp(p.trace, 9);
  cluster_sync_all();        // partials of my key half have landed in sPd
  if (threadIdx.x == 64) attn_stamp(p.trace, 10);
  if (warp >= 2) {
    tc_fence_after();
    uint32_t rg[32];
    tmem_ld_32x32(t_addr + 64 + which_k * 128 + rank * 64 + chalf * 32, rg);   // my own partial of my key half
    tmem_ld_wait();
    // my key half, summed, as a bf16 tile in the retired V (dV) / K (dK) buffer -> one TMA store each (clipped at T)
    const float sc = which_k ? p.scale : 1.f;
    const float4* recv = reinterpret_cast<const float4*>(sPd);
    uint8_t* orow = (which_k ? sK : sV) + row * 128;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const float4 r0 = recv[(part * 8 + 2 * g) * 128 + row], r1 = recv[(part * 8 + 2 * g + 1) * 128 + row];
      uint4 ov;
      ov.x = pack_bf16(sc * (__uint_as_float(rg[g * 8 + 0]) + r0.x), sc * (__uint_as_float(rg[g * 8 + 1]) + r0.y));
      ov.y = pack_bf16(sc * (__uint_as_float(rg[g * 8 + 2]) + r0.z), sc * (__uint_as_float(rg[g * 8 + 3]) + r0.w));
      ov.z = pack_bf16(sc * (__uint_as_float(rg[g * 8 + 4]) + r1.x), sc * (__uint_as_float(rg[g * 8 + 5]) + r1.y));
      ov.w = pack_bf16(sc * (__uint_as_float(rg[g * 8 + 6]) + r1.z), sc * (__uint_as_float(rg[g * 8 + 7]) + r1.w));
      *reinterpret_cast<uint4*>(orow + (((chalf * 4 + g) ^ (row & 7)) << 4)) = ov;
    }
    fence_proxy_async();
    named_bar_sync(6 + which_k, 256);
    if ((warp == 2 || warp == 10) && lane == 0) {
      tma_store_4d(which_k ? &tmDK : &tmDV, which_k ? sK : sV, 0, rank * 128, h, b);
      tma_store_commit();
    }
    if (warp == 2 && lane == 0) tma_store_wait_read<0>();      // also covers the dQ store issued by this thread
    if (warp == 10 && lane == 0) tma_store_wait_read<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
  if (threadIdx.x == 0) attn_stamp(p.trace, 11);
}

}  // namespace
}  // namespace s2t

using namespace s2t;

extern "C" int s2t_attn_fused_fwd(const void* qkv, long long ldqkv, int C, void* probs, void* probs_drop, long long ld,
                                  float* lse, void* ctx, long long ldctx, const int* klen, int batch, int heads, int T, int dk,
                                  float scale, float dropout_p, unsigned long long seed, unsigned int stream_id,
                                  const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(qkv && ctx, "s2t_attn_fused_fwd: null pointer");
  S2T_REQUIRE(probs != nullptr || lse != nullptr, "s2t_attn_fused_fwd: keep either the probabilities or the row log-sum-exp");
  S2T_REQUIRE(dk == kDk && T >= 1 && T <= kTk, "s2t_attn_fused_fwd: needs d_k == 64 and T <= 256 (got d_k=%d, T=%d)", dk, T);
  S2T_REQUIRE(heads * dk <= C && ldqkv >= 3ll * C && ldqkv % 8 == 0 && C % 8 == 0, "s2t_attn_fused_fwd: bad qkv layout");
  S2T_REQUIRE(ldctx % 8 == 0 && (probs == nullptr || (ld % 8 == 0 && ld >= ((T + 7) & ~7))),
              "s2t_attn_fused_fwd: ld / ldctx must be multiples of 8");
  S2T_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "s2t_attn_fused_fwd: bad dropout_p");
  S2T_REQUIRE(dropout_p == 0.f || probs == nullptr || probs_drop != nullptr,
              "s2t_attn_fused_fwd: dropout with kept probabilities needs the probs_drop output");
  if (batch <= 0) return 0;
  const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(qkv);
  CUtensorMap tq, tk, tv;
  int rc = make_operand_map(&tq, base, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTq);
  if (rc) return rc;
  rc = make_operand_map(&tk, base + C, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTk);
  if (rc) return rc;
  rc = make_operand_map(&tv, base + 2 * C, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, 64);
  if (rc) return rc;
  CUtensorMap to;
  rc = make_operand_map(&to, ctx, kDk, T, ldctx, heads, kDk, batch, 1ll * T * ldctx, kTq);
  if (rc) return rc;
  AttnParams p;
  memset(&p, 0, sizeof(p));
  p.probs = reinterpret_cast<__nv_bfloat16*>(probs);
  p.probs_drop = (dropout_p > 0.f && probs != nullptr) ? reinterpret_cast<__nv_bfloat16*>(probs_drop) : nullptr;
  p.lse = lse;
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
  static std::atomic<unsigned long long> configured{0};   // per device, see once_per_device
  S2T_CHECK_CUDA(once_per_device(configured, [&] { return cudaFuncSetAttribute(attn_fused_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem); }));
  S2T_LAUNCH(attn_fused_fwd_kernel, dim3((T + kTq - 1) / kTq, heads, batch), kAttnThreads, kAttnSmem,
             reinterpret_cast<cudaStream_t>(stream), tq, tk, tv, to, p);
  return 0;
}

extern "C" int s2t_attn_fused_bwd(const void* qkv, long long ldqkv, int C, const void* o, const void* d_o, long long ldo,
                                  const float* lse, void* dqkv, long long ldg, const int* klen, int batch, int heads,
                                  int T, int dk, float scale, float dropout_p, unsigned long long seed,
                                  unsigned int stream_id, const unsigned long long* seed_offset, void* stream) {
  S2T_REQUIRE(qkv && o && d_o && lse && dqkv, "s2t_attn_fused_bwd: null pointer");
  S2T_REQUIRE(dk == kDk && T >= 1 && T <= kTk, "s2t_attn_fused_bwd: needs d_k == 64 and T <= 256 (got d_k=%d, T=%d)", dk, T);
  S2T_REQUIRE(heads * dk <= C && ldqkv >= 3ll * C && ldqkv % 8 == 0 && C % 8 == 0 && ldg >= 3ll * C && ldg % 8 == 0 &&
                  ldo % 8 == 0 && ldo >= heads * dk,
              "s2t_attn_fused_bwd: bad layout");
  S2T_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "s2t_attn_fused_bwd: bad dropout_p");
  S2T_REQUIRE((reinterpret_cast<uintptr_t>(o) & 15) == 0 && (reinterpret_cast<uintptr_t>(d_o) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(dqkv) & 15) == 0,
              "s2t_attn_fused_bwd: o / d_o / dqkv must be 16-byte aligned");
  if (batch <= 0) return 0;
  const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(qkv);
  CUtensorMap tq, tk, tv, tdo;
  int rc = make_operand_map(&tq, base, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTq);
  if (rc) return rc;
  rc = make_operand_map(&tk, base + C, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTk);
  if (rc) return rc;
  rc = make_operand_map(&tv, base + 2 * C, kDk, T, ldqkv, heads, kDk, batch, 1ll * T * ldqkv, kTk);
  if (rc) return rc;
  rc = make_operand_map(&tdo, d_o, kDk, T, ldo, heads, kDk, batch, 1ll * T * ldo, kTq);
  if (rc) return rc;
  CUtensorMap tdq, tdk, tdv;
  const __nv_bfloat16* gbase = reinterpret_cast<const __nv_bfloat16*>(dqkv);
  rc = make_operand_map(&tdq, gbase, kDk, T, ldg, heads, kDk, batch, 1ll * T * ldg, kTq);
  if (rc) return rc;
  rc = make_operand_map(&tdk, gbase + C, kDk, T, ldg, heads, kDk, batch, 1ll * T * ldg, 128);
  if (rc) return rc;
  rc = make_operand_map(&tdv, gbase + 2 * C, kDk, T, ldg, heads, kDk, batch, 1ll * T * ldg, 128);
  if (rc) return rc;
  AttnBwdParams p;
  memset(&p, 0, sizeof(p));
  p.o = reinterpret_cast<const __nv_bfloat16*>(o);
  p.d_o = reinterpret_cast<const __nv_bfloat16*>(d_o);
  p.ldo = ldo;
  p.lse = lse;
  p.dqkv = reinterpret_cast<__nv_bfloat16*>(dqkv);
  p.ldg = ldg;
  p.C = C;
  p.klen = klen;
  p.B = batch; p.H = heads; p.T = T;
  p.scale = scale;
  p.thresh16 = dropout_p > 0.f ? static_cast<uint32_t>(dropout_p * 65536.f + 0.5f) : 0u;
  p.inv_keep = 1.f / (1.f - dropout_p);
  p.seed = seed;
  p.seed_off = seed_offset;
  p.stream_id = stream_id;
  p.trace = trace_buffer();
  static std::atomic<unsigned long long> configured{0};   // per device, see once_per_device
  S2T_CHECK_CUDA(once_per_device(configured, [&] { return cudaFuncSetAttribute(attn_fused_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem); }));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2, heads, batch);
  cfg.blockDim = dim3(kAttnThreads, 1, 1);
  cfg.dynamicSmemBytes = kBwdSmem;
  cfg.stream = reinterpret_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  S2T_CHECK_CUDA(cudaLaunchKernelEx(&cfg, attn_fused_bwd_kernel, tq, tk, tv, tdo, tdq, tdk, tdv, p));
  return 0;
}
