// s2t_gemm_bf16: the dense-contraction kernel family of the S2T encoder hot path.
//
// Replaces every torch.nn.Linear / torch.matmul / torch.bmm / nn.Conv1d(-as-GEMM) call on the
// reference path (SURVEY.md §2.3 "K4", e.g. fairseq/modules/s2t_transformer_layer.py:49-50,
// fairseq/modules/multihead_attention.py:246-265,372-416, fairseq/modules/speech_to_text/ctc.py:59).
//
// Design (sm_100a only):
//   * one CTA computes a 128 x BN fp32 accumulator tile held in TMEM (tcgen05.mma, cta_group::1),
//   * operands are bf16, staged global -> shared by TMA (cp.async.bulk.tensor, SWIZZLE_128B) through
//     a multi-stage mbarrier ring; each operand can be K-major ([rows][K], the nn.Linear layout) or
//     MN-major ([K][rows]) so that dgrad / wgrad / P.V / dS^T.Q need no transposes in HBM,
//   * warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected lane),
//     warps 2..5 = epilogue (tcgen05.ld -> registers -> fused epilogue -> vectorised global stores),
//   * fused epilogue: alpha, bias, ReLU / Swish, activation-backward gating, counter-based dropout,
//     output scale, residual add, bf16 or fp32 output,
//   * two batch dimensions (e.g. batch x heads) map to gridDim.z through 4-D tensor maps.
#include "common.cuh"
#include <stdlib.h>

#include <mutex>
#include <string.h>
#include <unordered_map>

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kGemmThreads = 320;    // split-K kernel: TMA warp, MMA warp, 8 epilogue warps
constexpr int kEpiWarps = 8;
constexpr int kColsumWarps = 2;      // persistent kernel only: column sums of the A operand (bias gradients), see below;
                                     // two warps keep the CTA at 384 threads = 170 registers per thread (no epilogue spills)
constexpr int kGemmThreadsP = kGemmThreads + 32 * kColsumWarps;

struct GemmKernelParams {
  int M, N, K;
  int nb1;
  int a_zm1, a_zm2, b_zm1, b_zm2;  // 0 = operand is broadcast along that batch dim (stride 0)
  int nbatch;
  int split_k;  // cluster size of the split-K kernel (1 = persistent kernel)
  int max_ctas; // grid bound of the persistent kernel (0 = one CTA per SM)
  int gate_tma; // persistent kernel, compile-time gate epilogue: gate tiles arrive by TMA (tmG) instead of per-row loads
  int out_fp32;
  int bias_vec;  // bias base 16B aligned -> 16-byte loads
  int aux_vec;   // gate / residual strides multiples of 8 and 16B-aligned bases -> 16-byte loads
  const __nv_bfloat16* residual;
  long long ldr, r_bs1, r_bs2;
  const __nv_bfloat16* gate;
  long long ldg, g_bs1, g_bs2;
  const __nv_bfloat16* bias;
  float alpha, out_scale;
  int act;        // S2T_ACT_*
  int gate_mode;  // S2T_GATE_*
  uint32_t drop_thresh16;
  float drop_inv_keep;
  unsigned long long drop_seed;
  const unsigned long long* drop_seed_off;  // optional device counter added to the seed (CUDA-graph replay)
  uint32_t drop_stream;
  void* out;            // split-K kernel stores its slice directly (the persistent kernel uses the TMA map)
  long long ldc, c_bs1, c_bs2;
  int out_vec;          // out base 16B aligned, ldc and batch strides multiples of 8 (bf16) / 4 (fp32)
  unsigned long long* trace;  // debug: globaltimer stamps of CTA 0's phases (S2T_GEMM_TRACE=1), else nullptr
  // persistent kernel, MN-major A (weight gradients dW = dY^T X): colsum[m] (+)= sum_k A[k][m] — the bias gradient of the
  // same layer, summed from the dY tiles the main loop stages in shared memory anyway (no separate pass over dY)
  __nv_bfloat16* colsum;
  int colsum_accumulate;
  int late_trigger;   // release the dependent launch when this CTA has issued its last operand load, not at entry
};

// GATE = 1: the persistent kernel stages the epilogue's gate operand through shared memory by TMA (32 KB: 4 lane
// quarters x 2 buffers x [32 rows][128 B]); one operand stage less to make room.
template <int BN, int GATE = 0>
struct GemmCfg {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = ((BN >= 256) ? 3 : (BN >= 128 ? 5 : 6)) - GATE;  // 144-160 KB operand ring
  static constexpr int kStagingBytes = kEpiWarps * 2 /*double buffer*/ * 4096;
  static constexpr int kGateBytes = GATE ? 4 * 2 * 4096 : 0;
  static constexpr int kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;            // two accumulator buffers
  static constexpr int kSmemBytes =
      kStages * kStageBytes + kStagingBytes + kGateBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

// compile-time epilogue code with a gate operand -> the TMA-staged gate path of the persistent kernel
constexpr int epi_gate_tma(int epi) { return (epi >= 0 && ((epi >> 2) & 3) != 0) ? 1 : 0; }

__device__ __forceinline__ void trace_stamp(const GemmKernelParams& p, int slot) {
  if (p.trace != nullptr && blockIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[slot] = t;
  }
}

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == S2T_ACT_RELU) return fmaxf(v, 0.f);
  if (act == S2T_ACT_SWISH) return v / (1.f + __expf(-v));
  return v;
}

// Epilogue variants.  EPI >= 0 fixes the feature set at compile time (bits: 0-1 act, 2-3 gate mode, 4 dropout,
// 5 residual/out_scale); EPI = -1 reads everything from the runtime parameters (any combination, slower code).
constexpr int kEpiGeneric = -1;
constexpr int epi_code(int act, int gate, int drop, int res) { return act | (gate << 2) | (drop << 4) | (res << 5); }

// Gate / residual operands of one staged output unit (up to 8 groups of 8 columns of this thread's row), fetched
// BEFORE the accumulator is waited for so that their L2/HBM latency hides behind the MMA main loop instead of being
// paid once per 8-column group on the epilogue's critical path.
struct AuxUnit {
  uint4 g[4];
  uint4 r[4];
  uint4 b[4];   // bias (per column, the same for every row) — also a dependent global load per 8 columns otherwise
};

template <int EPI>
__device__ __forceinline__ void aux_prefetch(AuxUnit& ax, const GemmKernelParams& p, int ucol0, int ngroups,
                                             bool row_ok, long long grow, long long rrow) {
  const bool need_g = (EPI < 0 ? (p.gate_mode != S2T_GATE_NONE) : (((EPI >> 2) & 3) != 0)) && p.aux_vec && !p.gate_tma;
  const bool need_r = (EPI < 0 ? true : ((EPI >> 5) & 1)) && p.residual != nullptr && p.aux_vec;
  const bool need_b = p.bias != nullptr && p.bias_vec;
  if (!row_ok || !(need_g || need_r || need_b)) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int col = ucol0 + i * 8;
    if (i < ngroups && col + 8 <= p.N) {
      if (need_b) ax.b[i] = __ldg(reinterpret_cast<const uint4*>(p.bias + col));
      if (need_g) ax.g[i] = __ldg(reinterpret_cast<const uint4*>(p.gate + grow + col));
      if (need_r) ax.r[i] = __ldg(reinterpret_cast<const uint4*>(p.residual + rrow + col));
    }
  }
}

template <int EPI>
__device__ __forceinline__ void epilogue8(float (&v)[8], const GemmKernelParams& p, int col, int nvalid,
                                          long long grow, long long rrow, unsigned long long drop_row,
                                          unsigned long long drop_seed, const uint4& gpre, const uint4& rpre,
                                          const uint4& bpre) {
  const int act = EPI < 0 ? p.act : (EPI & 3);
  const int gate_mode = EPI < 0 ? p.gate_mode : ((EPI >> 2) & 3);
  const bool drop = EPI < 0 ? (p.drop_thresh16 != 0) : ((EPI >> 4) & 1);
  const bool res = EPI < 0 ? true : ((EPI >> 5) & 1);
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] *= p.alpha;
  if (p.bias != nullptr) {
    if (nvalid == 8 && p.bias_vec) {
      const uint4 bv = bpre;
      v[0] += bf16_lo(bv.x); v[1] += bf16_hi(bv.x); v[2] += bf16_lo(bv.y); v[3] += bf16_hi(bv.y);
      v[4] += bf16_lo(bv.z); v[5] += bf16_hi(bv.z); v[6] += bf16_lo(bv.w); v[7] += bf16_hi(bv.w);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < nvalid) v[j] += __bfloat162float(p.bias[col + j]);
    }
  }
  if (act == S2T_ACT_RELU) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
  } else if (act == S2T_ACT_SWISH) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = v[j] * __frcp_rn(1.f + __expf(-v[j]));
  }
  if (gate_mode != S2T_GATE_NONE) {
    float gt[8];
    if (nvalid == 8 && p.aux_vec) {
      const uint4 gv = gpre;
      gt[0] = bf16_lo(gv.x); gt[1] = bf16_hi(gv.x); gt[2] = bf16_lo(gv.y); gt[3] = bf16_hi(gv.y);
      gt[4] = bf16_lo(gv.z); gt[5] = bf16_hi(gv.z); gt[6] = bf16_lo(gv.w); gt[7] = bf16_hi(gv.w);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) gt[j] = j < nvalid ? __bfloat162float(p.gate[grow + col + j]) : 0.f;
    }
    if (gate_mode == S2T_GATE_POSITIVE) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = gt[j] > 0.f ? v[j] : 0.f;
    } else if (gate_mode == S2T_GATE_SWISH_GRAD) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float sg = __frcp_rn(1.f + __expf(-gt[j]));
        v[j] *= sg * (1.f + gt[j] * (1.f - sg));
      }
    } else {  // S2T_GATE_MUL
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] *= gt[j];
    }
  }
  if (drop) {
    const uint32_t keep = dropout_keep8(drop_seed, p.drop_stream, drop_row + (col >> 3), p.drop_thresh16);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * p.drop_inv_keep : 0.f;
  }
  if (res) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= p.out_scale;
    if (p.residual != nullptr) {
      if (nvalid == 8 && p.aux_vec) {
        const uint4 rv = rpre;
        v[0] += bf16_lo(rv.x); v[1] += bf16_hi(rv.x); v[2] += bf16_lo(rv.y); v[3] += bf16_hi(rv.y);
        v[4] += bf16_lo(rv.z); v[5] += bf16_hi(rv.z); v[6] += bf16_lo(rv.w); v[7] += bf16_hi(rv.w);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < nvalid) v[j] += __bfloat162float(p.residual[rrow + col + j]);
      }
    }
  }
}

// Persistent, warp-specialised GEMM.  One CTA per SM walks tiles blockIdx.x, +gridDim.x, ...;
//   warp 0    TMA producer  : operand ring (full/empty mbarriers) running continuously across tiles
//   warp 1    MMA issuer    : tcgen05.mma into one of TWO TMEM accumulator buffers (tmem_full/tmem_empty)
//   warps 2-5 epilogue      : tcgen05.ld -> fused epilogue -> swizzled smem staging -> TMA store (coalesced,
//                             edge-clipped by the tensor map); overlaps the next tile's loads and MMAs
template <int BN, int A_MN, int B_MN, int EPI>
__global__ void __launch_bounds__(kGemmThreadsP, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmG,
                 const GemmKernelParams p) {
  constexpr int kGateTma = epi_gate_tma(EPI);
  using Cfg = GemmCfg<BN, kGateTma>;
  // the next kernel's set-up may overlap this one (common.cuh, PDL) — except behind a side-stream GEMM that was given an
  // SM share (late_trigger): its successor's CTAs would sit on SMs the latency-critical chain needs while they wait
  if (!p.late_trigger) pdl_launch_dependents();
  if (threadIdx.x == 0) trace_stamp(p, 0);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;  // 1024-aligned (stage sizes are multiples of 1024)
  uint8_t* gate_stage = staging + Cfg::kStagingBytes;         // [4 quarters][2][4096] (GATE variants only)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(gate_stage + Cfg::kGateBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;   // [2]
  uint64_t* tempty_bar = tfull_bar + 2;             // [2]
  uint64_t* gate_bar = tempty_bar + 2;              // [4 quarters][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gate_bar + 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (p.K + kBK - 1) / kBK;
  const int tiles_m = (p.M + kBM - 1) / kBM;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tiles_per_z = tiles_m * tiles_n;
  const int num_tiles = tiles_per_z * p.nbatch;

  if (threadIdx.x == 0) {
    // with column sums a stage is free once the MMAs have retired AND the kColsumWarps column-sum warps have read it
    const uint32_t empty_count = (A_MN && p.colsum != nullptr) ? 1u + kColsumWarps : 1u;
    for (int i = 0; i < Cfg::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], empty_count);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], kEpiWarps);
    }
    for (int i = 0; i < 8; ++i) mbar_init(&gate_bar[i], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    if (kGateTma) tma_prefetch_desc(&tmG);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // barriers, TMEM and tensor maps are set up: from here on global memory written by earlier kernels is touched
  if (threadIdx.x == 0) trace_stamp(p, 1);
  pdl_wait();
  if (threadIdx.x == 0) trace_stamp(p, 2);

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int z = tile / tiles_per_z;
        const int r = tile - z * tiles_per_z;
        const int m0 = (r % tiles_m) * kBM, n0 = (r / tiles_m) * BN;
        const int z1 = z % p.nb1, z2 = z / p.nb1;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % Cfg::kStages;
          const uint32_t ph = (it / Cfg::kStages) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          mbar_expect_tx(&full_bar[s], Cfg::kStageBytes);
          uint8_t* a_dst = smem + s * Cfg::kStageBytes;
          uint8_t* b_dst = a_dst + Cfg::kABytes;
          if (A_MN == 0) {
            tma_load_4d(a_dst, &tmA, &full_bar[s], kb * kBK, m0, z1 * p.a_zm1, z2 * p.a_zm2);
          } else {
#pragma unroll
            for (int c = 0; c < kBM / 64; ++c)
              tma_load_4d(a_dst + c * (kBK * 128), &tmA, &full_bar[s], m0 + c * 64, kb * kBK, z1 * p.a_zm1,
                          z2 * p.a_zm2);
          }
          if (B_MN == 0) {
            tma_load_4d(b_dst, &tmB, &full_bar[s], kb * kBK, n0, z1 * p.b_zm1, z2 * p.b_zm2);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)
              tma_load_4d(b_dst + c * (kBK * 128), &tmB, &full_bar[s], n0 + c * 64, kb * kBK, z1 * p.b_zm1,
                          z2 * p.b_zm2);
          }
        }
        if (tile == static_cast<int>(blockIdx.x)) trace_stamp(p, 3);
      }
      // late trigger: every operand load of this CTA is in flight — what is left is the pipeline drain and the last
      // epilogue (~2-3 us), about the launch latency of the dependent kernel, whose CTAs then wait only briefly
      if (p.late_trigger) pdl_launch_dependents();
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, A_MN, B_MN);
      uint32_t it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++lt) {
        const uint32_t buf = lt & 1;
        mbar_wait(&tempty_bar[buf], ((lt >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        // tiles of the first column block (n0 == 0) are the ones whose A stages the column-sum warps read; for every other
        // tile this thread stands in for their arrivals
        const bool colsum_standin = A_MN && p.colsum != nullptr && ((tile % tiles_per_z) / tiles_m) != 0;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % Cfg::kStages;
          const uint32_t ph = (it / Cfg::kStages) & 1;
          mbar_wait(&full_bar[s], ph);
          if (it == 0) trace_stamp(p, 4);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + s * Cfg::kStageBytes);
          const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t adesc = A_MN ? umma_smem_desc(a_addr + k * 2048, kBK * 128, 1024)
                                        : umma_smem_desc(a_addr + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? umma_smem_desc(b_addr + k * 2048, kBK * 128, 1024)
                                        : umma_smem_desc(b_addr + k * 32, 16, 1024);
            umma_bf16(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);  // frees this smem stage once the MMAs above retire
          if (colsum_standin) {
#pragma unroll
            for (int w = 0; w < kColsumWarps; ++w) mbar_arrive(&empty_bar[s]);
          }
        }
        umma_commit(&tfull_bar[buf]);  // accumulator of this tile complete
        if (lt == 0) trace_stamp(p, 5);
      }
    }
  } else if (warp >= 2 + kEpiWarps) {
    // ------------------------------ column sums of A (bias gradient) ----------
    // A stage = A as [64 k-rows][64 m] x 2 chunks, 128-byte rows, 16-byte chunk index XORed by (k & 7).  Warp w owns the
    // 64-column chunk w: lane = cl * 8 + g reads the 16-byte pieces cl and cl + 4 of the rows k = 8 kk + g, so the eight
    // lanes of a quarter-warp phase hit eight different piece positions (conflict-free) and the swizzle term is a per-lane
    // constant; 16 LDS.128 + 128 adds per stage and lane, the 8 row groups are folded by shuffles at the tile end.
    if (A_MN && p.colsum != nullptr) {
      const int cwarp = warp - 2 - kEpiWarps;
      const int g = lane & 7, cl = lane >> 3;
      const uint32_t cbase = static_cast<uint32_t>(cwarp * (kBK * 128) + g * 128);
      const uint32_t off0 = cbase + ((cl ^ g) << 4), off1 = cbase + (((cl + 4) ^ g) << 4);
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int r = tile % tiles_per_z;
        if (r / tiles_m != 0) {
          it += num_kb;
          continue;
        }
        const int m0 = (r % tiles_m) * kBM;
        float acc[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = 0.f;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % Cfg::kStages;
          const uint32_t ph = (it / Cfg::kStages) & 1;
          mbar_wait(&full_bar[s], ph);
          const uint8_t* a_src = smem + s * Cfg::kStageBytes;
#pragma unroll
          for (int kk = 0; kk < kBK / 8; ++kk) {
            const uint4 v = *reinterpret_cast<const uint4*>(a_src + off0 + kk * 1024);
            const uint4 w = *reinterpret_cast<const uint4*>(a_src + off1 + kk * 1024);
            acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x); acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
            acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z); acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
            acc[8] += bf16_lo(w.x); acc[9] += bf16_hi(w.x); acc[10] += bf16_lo(w.y); acc[11] += bf16_hi(w.y);
            acc[12] += bf16_lo(w.z); acc[13] += bf16_hi(w.z); acc[14] += bf16_lo(w.w); acc[15] += bf16_hi(w.w);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty_bar[s]);
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
          acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
          acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 4);
        }
        if (g == 0) {
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int col = m0 + cwarp * 64 + (cl + 4 * hh) * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (col + j < p.M) {
                float v = acc[hh * 8 + j];
                if (p.colsum_accumulate) v += __bfloat162float(p.colsum[col + j]);
                p.colsum[col + j] = __float2bfloat16(v);
              }
            }
          }
        }
      }
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    // 8 warps: warp w may touch TMEM lanes 32*(w%4)..; the two warps of a lane quarter split the tile's
    // 128-byte output units (64 bf16 / 32 fp32 columns) between them.
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    uint8_t* stage_base = staging + ew * 8192;
    const unsigned long long drop_seed =
        p.drop_seed + ((p.drop_thresh16 != 0 && p.drop_seed_off != nullptr) ? *p.drop_seed_off : 0ull);
    const int units = BN / 32;   // fp32 output: one 32-column TMEM chunk (128-byte rows) per staged unit
    uint32_t lt = 0, sb = 0, gcount = 0;
    // gate operand (bf16 output, compile-time gate epilogue): one TMA box [32 rows][64 columns] per staged unit into a
    // swizzled buffer of this lane quarter (read back conflict-free, one 16-byte chunk per thread and column group).
    // The issuing lane walks the quarter's unit sequence two units ahead of the consumers, across tile boundaries.
    const bool gtma = kGateTma && p.gate_tma;
    int ptile = blockIdx.x, pu = 0;
    uint32_t pcount = 0;
    auto gate_issue_next = [&]() {
      if (ptile >= num_tiles) return;
      const int pz = ptile / tiles_per_z;
      const int pr = ptile - pz * tiles_per_z;
      const int pm0 = (pr % tiles_m) * kBM, pn0 = (pr / tiles_m) * BN;
      uint64_t* gbar = &gate_bar[q * 2 + (pcount & 1)];
      mbar_expect_tx(gbar, 4096);
      tma_load_4d(gate_stage + q * 8192 + (pcount & 1) * 4096, &tmG, gbar, pn0 + pu * 64, pm0 + q * 32,
                  p.g_bs1 != 0 ? pz % p.nb1 : 0, p.g_bs2 != 0 ? pz / p.nb1 : 0);
      ++pcount;
      ++pu;
      if (pu == BN / 64 || pn0 + pu * 64 >= p.N) {
        pu = 0;
        ptile += gridDim.x;
      }
    };
    if (gtma && half == 0 && lane == 0) {
      gate_issue_next();
      gate_issue_next();
    }
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++lt) {
      const int z = tile / tiles_per_z;
      const int r = tile - z * tiles_per_z;
      const int m0 = (r % tiles_m) * kBM, n0 = (r / tiles_m) * BN;
      const int z1 = z % p.nb1, z2 = z / p.nb1;
      const uint32_t buf = lt & 1;
      const int row = m0 + q * 32 + lane;
      const long long rrow = static_cast<long long>(row) * p.ldr + z1 * p.r_bs1 + z2 * p.r_bs2;
      const long long grow = static_cast<long long>(row) * p.ldg + z1 * p.g_bs1 + z2 * p.g_bs2;
      const unsigned long long drop_row =
          (static_cast<unsigned long long>(z) * p.M + row) * static_cast<unsigned long long>((p.N + 7) >> 3);
      const bool row_ok = row < p.M;
      const uint32_t t_addr = tmem_base + buf * BN + (static_cast<uint32_t>(q * 32) << 16);
      if (!p.out_fp32) {
        // bf16 output: a staged unit is 64 columns (128-byte rows).  BOTH warps of the lane quarter work on every unit,
        // warp `half` converting its 32-column half; they meet at a named barrier, then one lane hands the unit to the
        // TMA store.  Four staging buffers per quarter: the store of unit k-3 has been read before unit k+1 is written.
        constexpr int kUnits = BN / 64;
        uint8_t* stage_q = staging + q * 16384;
        AuxUnit ax;
        if (n0 + half * 32 < p.N) aux_prefetch<EPI>(ax, p, n0 + half * 32, 4, row_ok, grow, rrow);
        mbar_wait(&tfull_bar[buf], (lt >> 1) & 1);
        if (ew == 0 && lane == 0 && lt < 2) trace_stamp(p, 6 + 2 * lt);
        tc_fence_after();
#pragma unroll 1
        for (int u = 0; u < kUnits; ++u) {
          const int ucol0 = n0 + u * 64;
          if (ucol0 >= p.N) break;  // warp-uniform
          uint32_t rg[32];
          tmem_ld_32x32(t_addr + u * 64 + half * 32, rg);
          uint8_t* sbuf = stage_q + (sb & 3) * 4096;
          uint8_t* srow = sbuf + lane * 128;
          if (gtma) {
            mbar_wait(&gate_bar[q * 2 + (gcount & 1)], (gcount >> 1) & 1);
            const uint8_t* grow_s = gate_stage + q * 8192 + (gcount & 1) * 4096 + lane * 128;
#pragma unroll
            for (int g = 0; g < 4; ++g)
              ax.g[g] = *reinterpret_cast<const uint4*>(grow_s + (((half * 4 + g) ^ (lane & 7)) << 4));
          }
          tmem_ld_wait();
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = ucol0 + half * 32 + g * 8;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(rg[g * 8 + j]);
            if (row_ok && col < p.N)
              epilogue8<EPI>(v, p, col, min(8, p.N - col), grow, rrow, drop_row, drop_seed, ax.g[g], ax.r[g], ax.b[g]);
            uint4 ov;
            ov.x = pack_bf16(v[0], v[1]); ov.y = pack_bf16(v[2], v[3]);
            ov.z = pack_bf16(v[4], v[5]); ov.w = pack_bf16(v[6], v[7]);
            *reinterpret_cast<uint4*>(srow + (((half * 4 + g) ^ (lane & 7)) << 4)) = ov;
          }
          fence_proxy_async();
          named_bar_sync(1 + q, 64);
          if (half == 0 && lane == 0) {
            tma_store_4d(&tmC, sbuf, ucol0, m0 + q * 32, z1, z2);
            tma_store_commit();
            tma_store_wait_read<2>();
            // both warps have read this unit's gate buffer (barrier above): refill it for the unit after next
            if (gtma) gate_issue_next();
          }
          ++sb;
          ++gcount;
          if (u + 1 < kUnits && ucol0 + 64 + half * 32 < p.N)
            aux_prefetch<EPI>(ax, p, ucol0 + 64 + half * 32, 4, row_ok, grow, rrow);
        }
      } else {
        AuxUnit ax;
        if (half < units && n0 + half * 32 < p.N) aux_prefetch<EPI>(ax, p, n0 + half * 32, 4, row_ok, grow, rrow);
        mbar_wait(&tfull_bar[buf], (lt >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int u = half; u < units; u += 2) {
          const int ucol0 = n0 + u * 32;
          if (ucol0 >= p.N) break;  // warp-uniform
          uint32_t rg[32];
          tmem_ld_32x32(t_addr + u * 32, rg);
          uint8_t* sbuf = stage_base + (sb & 1) * 4096;
          // this staging buffer was handed to a TMA store two stores ago: wait until it has been read
          if (lane == 0) tma_store_wait_read<1>();
          tmem_ld_wait();
          __syncwarp();
          uint8_t* srow = sbuf + lane * 128;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = ucol0 + g * 8;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(rg[g * 8 + j]);
            if (row_ok && col < p.N)
              epilogue8<EPI>(v, p, col, min(8, p.N - col), grow, rrow, drop_row, drop_seed, ax.g[g], ax.r[g], ax.b[g]);
            *reinterpret_cast<float4*>(srow + (((2 * g) ^ (lane & 7)) << 4)) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(srow + (((2 * g + 1) ^ (lane & 7)) << 4)) = make_float4(v[4], v[5], v[6], v[7]);
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_4d(&tmC, sbuf, ucol0, m0 + q * 32, z1, z2);
            tma_store_commit();
          }
          ++sb;
          if (u + 2 < units && ucol0 + 64 < p.N) aux_prefetch<EPI>(ax, p, ucol0 + 64, 4, row_ok, grow, rrow);
        }
      }
      if (ew == 0 && lane == 0 && lt < 2) trace_stamp(p, 7 + 2 * lt);
      // accumulator drained: hand the TMEM buffer back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[buf]);
    }
    if (lane == 0) tma_store_wait_read<0>();
    if (ew == 0 && lane == 0) trace_stamp(p, 10);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
  if (threadIdx.x == 0) trace_stamp(p, 11);
}

// Split-K GEMM for K-heavy problems with few output tiles (weight gradients: K = B*T rows; ffn.w_2; the
// vocabulary dgrad).  A thread-block cluster of S CTAs shares one output tile: CTA r reduces k-blocks
// [r*kb_per, (r+1)*kb_per) into its own TMEM accumulator, ranks > 0 park their fp32 partial tile in their
// shared memory (the idle operand ring), the leader pulls them over DSMEM (ld.shared::cluster), adds them to its
// own accumulator and runs the fused epilogue + TMA store.  No global workspace, one launch.
template <int BN, int A_MN, int B_MN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_splitk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                        const __grid_constant__ CUtensorMap tmC, const GemmKernelParams p) {
  using Cfg = GemmCfg<BN>;
  pdl_launch_dependents();
  if (threadIdx.x == 0) trace_stamp(p, 0);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int S = p.split_k;
  const int rank = static_cast<int>(cluster_ctarank());
  const int tile = blockIdx.x / S;
  const int num_kb = (p.K + kBK - 1) / kBK;
  const int kb_per = (num_kb + S - 1) / S;
  const int kb0 = rank * kb_per;
  const int kb1 = min(num_kb, kb0 + kb_per);  // host guarantees kb0 < num_kb for every rank
  const int tiles_m = (p.M + kBM - 1) / kBM;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tiles_per_z = tiles_m * tiles_n;
  const int z = tile / tiles_per_z;
  const int r = tile - z * tiles_per_z;
  const int m0 = (r % tiles_m) * kBM, n0 = (r / tiles_m) * BN;
  const int z1 = z % p.nb1, z2 = z / p.nb1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tfull_bar, 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, BN < 32 ? 32 : BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) trace_stamp(p, 1);
  pdl_wait();
  if (threadIdx.x == 0) trace_stamp(p, 2);

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
        const int s = it % Cfg::kStages;
        const uint32_t ph = (it / Cfg::kStages) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        mbar_expect_tx(&full_bar[s], Cfg::kStageBytes);
        uint8_t* a_dst = smem + s * Cfg::kStageBytes;
        uint8_t* b_dst = a_dst + Cfg::kABytes;
        if (A_MN == 0) {
          tma_load_4d(a_dst, &tmA, &full_bar[s], kb * kBK, m0, z1 * p.a_zm1, z2 * p.a_zm2);
        } else {
#pragma unroll
          for (int c = 0; c < kBM / 64; ++c)
            tma_load_4d(a_dst + c * (kBK * 128), &tmA, &full_bar[s], m0 + c * 64, kb * kBK, z1 * p.a_zm1,
                        z2 * p.a_zm2);
        }
        if (B_MN == 0) {
          tma_load_4d(b_dst, &tmB, &full_bar[s], kb * kBK, n0, z1 * p.b_zm1, z2 * p.b_zm2);
        } else {
#pragma unroll
          for (int c = 0; c < BN / 64; ++c)
            tma_load_4d(b_dst + c * (kBK * 128), &tmB, &full_bar[s], n0 + c * 64, kb * kBK, z1 * p.b_zm1,
                        z2 * p.b_zm2);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, A_MN, B_MN);
      for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
        const int s = it % Cfg::kStages;
        const uint32_t ph = (it / Cfg::kStages) & 1;
        mbar_wait(&full_bar[s], ph);
        if (it == 0) trace_stamp(p, 4);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * Cfg::kStageBytes);
        const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k) {
          const uint64_t adesc = A_MN ? umma_smem_desc(a_addr + k * 2048, kBK * 128, 1024)
                                      : umma_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? umma_smem_desc(b_addr + k * 2048, kBK * 128, 1024)
                                      : umma_smem_desc(b_addr + k * 32, 16, 1024);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(tfull_bar);
      trace_stamp(p, 5);
    }
  }
  const int ew = warp - 2;
  const int q = warp & 3;
  const int half = ew >> 2;
  const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
  const int prow = q * 32 + lane;
  // Reduce-scatter over the cluster, push style: rank r owns the tile's columns [r*w, (r+1)*w), w = BN / S.  Every rank
  // scatters its fp32 accumulator straight from TMEM registers into the owners' shared memory (remote stores are
  // fire-and-forget: no DSMEM round trip on anybody's critical path); after one cluster barrier each owner sums the S
  // slices from its LOCAL shared memory, runs the fused epilogue and stores its slice to global memory.
  // Receive buffer = the epilogue staging area (not part of the operand ring, so a fast peer may write into it while
  // this CTA's MMAs are still reading operands); layout: float4 index ((src*w/4 + j)*128 + row).
  const int w = BN / S;
  float4* recv = reinterpret_cast<float4*>(BN <= 128 ? staging : smem);
  if (BN > 128) {           // 128 KB of partials only fit in the operand ring: wait until every rank's MMAs have retired
    if (warp >= 2) {
      mbar_wait(tfull_bar, 0);
      tc_fence_after();
    }
    tc_fence_before();
    __syncwarp();
    cluster_sync_all();
  }
  if (warp >= 2) {
    mbar_wait(tfull_bar, 0);  // all MMAs of this CTA retired: accumulator complete
    tc_fence_after();
    const uint32_t recv_local = smem_u32(recv);
    const int w4 = w >> 2;
#pragma unroll 1
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t rg[32];
      tmem_ld_32x32(t_addr + c * 32, rg);
      tmem_ld_wait();
#pragma unroll
      for (int g4 = 0; g4 < 8; ++g4) {
        const int j4 = c * 8 + g4;            // float4 column of the tile
        const int owner = j4 / w4;
        const uint32_t off = static_cast<uint32_t>(((rank * w4 + (j4 - owner * w4)) * 128 + prow) << 4);
        dsmem_st_f4(dsmem_addr(recv_local, owner) + off,
                    make_float4(__uint_as_float(rg[g4 * 4]), __uint_as_float(rg[g4 * 4 + 1]),
                                __uint_as_float(rg[g4 * 4 + 2]), __uint_as_float(rg[g4 * 4 + 3])));
      }
    }
  }
  // Owner side.  Thread = one row; the two warp halves of a lane quarter split the slice's 8-column groups.
  const int ngroups = w >> 3;
  const int ngh = (ngroups + 1) >> 1;
  const int g_lo = half * ngh;
  const int g_hi = min(ngroups, g_lo + ngh);
  const int tcol0 = rank * w;
  const int row = m0 + prow;
  const long long rrow = static_cast<long long>(row) * p.ldr + z1 * p.r_bs1 + z2 * p.r_bs2;
  const long long grow = static_cast<long long>(row) * p.ldg + z1 * p.g_bs1 + z2 * p.g_bs2;
  const bool row_ok = row < p.M;
  AuxUnit ax;
  // bias / gate / residual of the first groups travel while the other ranks finish their partial tiles
  if (warp >= 2 && g_lo < g_hi && n0 + tcol0 + g_lo * 8 < p.N)
    aux_prefetch<EPI>(ax, p, n0 + tcol0 + g_lo * 8, min(4, g_hi - g_lo), row_ok, grow, rrow);
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();  // all partial tiles are visible cluster-wide
  if (warp >= 2) {
    if (ew == 0 && lane == 0) trace_stamp(p, 6);
    const unsigned long long drop_seed =
        p.drop_seed + ((p.drop_thresh16 != 0 && p.drop_seed_off != nullptr) ? *p.drop_seed_off : 0ull);
    const unsigned long long drop_row =
        (static_cast<unsigned long long>(z) * p.M + row) * static_cast<unsigned long long>((p.N + 7) >> 3);
    const long long orow = static_cast<long long>(row) * p.ldc + z1 * p.c_bs1 + z2 * p.c_bs2;
#pragma unroll 1
    for (int gb = g_lo; gb < g_hi; gb += 4) {
      if (gb != g_lo && n0 + tcol0 + gb * 8 < p.N)
        aux_prefetch<EPI>(ax, p, n0 + tcol0 + gb * 8, min(4, g_hi - gb), row_ok, grow, rrow);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int g = gb + i;
        const int tcol = tcol0 + g * 8;
        const int col = n0 + tcol;
        if (g < g_hi && col < p.N) {   // warp-uniform
          float4 t[16];
#pragma unroll
          for (int pr = 0; pr < 8; ++pr) {
            if (pr < S) {
              const float4* src = recv + (pr * (w >> 2) + g * 2) * 128 + prow;
              t[2 * pr] = src[0];
              t[2 * pr + 1] = src[128];
            }
          }
          float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int pr = 0; pr < 8; ++pr) {
            if (pr < S) {
              v[0] += t[2 * pr].x; v[1] += t[2 * pr].y; v[2] += t[2 * pr].z; v[3] += t[2 * pr].w;
              v[4] += t[2 * pr + 1].x; v[5] += t[2 * pr + 1].y; v[6] += t[2 * pr + 1].z; v[7] += t[2 * pr + 1].w;
            }
          }
          if (p.trace != nullptr && ew == 0 && lane == 0 && g == g_lo && v[0] != 12345.678f) trace_stamp(p, 7);
          if (row_ok) {
            const int nvalid = min(8, p.N - col);
            epilogue8<EPI>(v, p, col, nvalid, grow, rrow, drop_row, drop_seed, ax.g[i], ax.r[i], ax.b[i]);
            if (p.out_fp32) {
              float* o = reinterpret_cast<float*>(p.out) + orow + col;
              if (nvalid == 8 && p.out_vec) {
                *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (j < nvalid) o[j] = v[j];
              }
            } else {
              __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + orow + col;
              if (nvalid == 8 && p.out_vec) {
                uint4 ov;
                ov.x = pack_bf16(v[0], v[1]); ov.y = pack_bf16(v[2], v[3]);
                ov.z = pack_bf16(v[4], v[5]); ov.w = pack_bf16(v[6], v[7]);
                *reinterpret_cast<uint4*>(o) = ov;
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (j < nvalid) o[j] = __float2bfloat16(v[j]);
              }
            }
          }
        }
      }
    }
    if (ew == 0 && lane == 0) trace_stamp(p, 9);
  }
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();  // every rank keeps its shared memory alive until all peers have read it
  if (warp == 1) tmem_dealloc(tmem_base, BN < 32 ? 32 : BN);
  if (threadIdx.x == 0) trace_stamp(p, 11);
}

// ---------------------------------------------------------------------------------------------
// host side: tensor-map construction (cached) and dispatch
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

struct MapKey {
  uint64_t v[10];
  bool operator==(const MapKey& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < 10; ++i) {
      h ^= k.v[i];
      h *= 1099511628211ull;
    }
    return static_cast<size_t>(h);
  }
};

// Build (or fetch) the 4-D bf16 tensor map of one operand.
//   inner/outer: extents of the contiguous and the strided matrix dimension,
//   ld: stride of the outer dimension (elements), box_outer: rows per TMA box.
int make_operand_map(CUtensorMap* out, const void* ptr, uint64_t inner, uint64_t outer,
                            uint64_t ld, uint64_t nb1, uint64_t bs1, uint64_t nb2, uint64_t bs2,
                            uint32_t box_outer, int fp32) {
  const uint64_t es = fp32 ? 4 : 2;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  static std::mutex mu;
  MapKey key{{reinterpret_cast<uint64_t>(ptr), inner, outer, ld, nb1, bs1, nb2, bs2, box_outer, (uint64_t)fp32}};
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  EncodeTiledFn enc = get_encode_fn();
  S2T_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
  S2T_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "GEMM operand base must be 16B aligned");
  S2T_REQUIRE((ld * es) % 16 == 0, "GEMM operand/output leading stride must be a multiple of 16 bytes (got %llu elements)",
              (unsigned long long)ld);
  // Unit batch dims still need a legal (16B-multiple, non-zero) stride.
  if (bs1 == 0) nb1 = 1;  // broadcast batch dims collapse; the kernel zeroes the coordinate
  if (bs2 == 0) nb2 = 1;
  uint64_t s1 = (nb1 > 1 ? bs1 : ld * outer) * es, s2 = (nb2 > 1 ? bs2 : ld * outer) * es;
  if (s1 == 0) s1 = 16;
  if (s2 == 0) s2 = 16;
  S2T_REQUIRE(s1 % 16 == 0 && s2 % 16 == 0, "GEMM batch strides must be multiples of 16 bytes");
  cuuint64_t dims[4] = {inner, outer, nb1, nb2};
  cuuint64_t strides[3] = {ld * es, s1, s2};
  cuuint32_t box[4] = {fp32 ? 32u : 64u, box_outer, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                   const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  S2T_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu ld=%llu nb=(%llu,%llu) bs=(%llu,%llu)",
              (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld,
              (unsigned long long)nb1, (unsigned long long)nb2, (unsigned long long)bs1,
              (unsigned long long)bs2);
  {
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 65536) cache.clear();
    cache.emplace(key, *out);
  }
  return 0;
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

template <int BN, int A_MN, int B_MN, int EPI>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                       const GemmKernelParams& kp_in, int nbatch, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, epi_gate_tma(EPI)>;
  GemmKernelParams kp = kp_in;
  CUtensorMap tg = tc;     // placeholder when the variant has no TMA-staged gate
  kp.gate_tma = 0;
  if (epi_gate_tma(EPI) && kp.gate != nullptr && kp.aux_vec && !kp.out_fp32) {
    const int rc = make_operand_map(&tg, kp.gate, kp.N, kp.M, kp.ldg, kp.nb1, kp.g_bs1, kp.nbatch / kp.nb1, kp.g_bs2, 32);
    if (rc) return rc;
    kp.gate_tma = 1;
  }
  static std::atomic<unsigned long long> configured{0};   // per device, see once_per_device
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, EPI>;
  S2T_CHECK_CUDA(once_per_device(configured, [&] { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes); }));
  const long long tiles = 1ll * ((kp.N + BN - 1) / BN) * ((kp.M + kBM - 1) / kBM) * nbatch;
  const int cap = (kp.max_ctas > 0 && kp.max_ctas < sm_count()) ? kp.max_ctas : sm_count();
  const int grid = static_cast<int>(tiles < cap ? tiles : cap);
  S2T_LAUNCH(kern, grid, kGemmThreadsP, Cfg::kSmemBytes, stream, ta, tb, tc, tg, kp);
  return 0;
}

template <int BN, int A_MN, int B_MN, int EPI>
static int launch_gemm_splitk(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                              const GemmKernelParams& kp, int nbatch, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static std::atomic<unsigned long long> configured{0};   // per device, see once_per_device
  auto kern = gemm_bf16_splitk_kernel<BN, A_MN, B_MN, EPI>;
  S2T_CHECK_CUDA(once_per_device(configured, [&] { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes); }));
  const long long tiles = 1ll * ((kp.N + BN - 1) / BN) * ((kp.M + kBM - 1) / kBM) * nbatch;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(static_cast<unsigned>(tiles * kp.split_k), 1, 1);
  cfg.blockDim = dim3(kGemmThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kp.split_k;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  S2T_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, kp));
  return 0;
}

template <int BN>
static int dispatch_splitk(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                           const CUtensorMap& tc, const GemmKernelParams& kp, int nbatch,
                           cudaStream_t stream) {
  const bool plain = kp.act == S2T_ACT_NONE && kp.gate_mode == S2T_GATE_NONE && kp.drop_thresh16 == 0 &&
                     kp.residual == nullptr && kp.out_scale == 1.f;
#define S2T_SK_CASE(AM, BM_)                                                                        \
  if (a_mn == AM && b_mn == BM_)                                                                    \
    return plain ? launch_gemm_splitk<BN, AM, BM_, 0>(ta, tb, tc, kp, nbatch, stream)               \
                 : launch_gemm_splitk<BN, AM, BM_, kEpiGeneric>(ta, tb, tc, kp, nbatch, stream);
  S2T_SK_CASE(0, 0) S2T_SK_CASE(0, 1) S2T_SK_CASE(1, 0) S2T_SK_CASE(1, 1)
#undef S2T_SK_CASE
  return -1;
}

template <int BN>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                          const CUtensorMap& tc, const GemmKernelParams& kp, int nbatch,
                          cudaStream_t stream) {
  // compile-time epilogue variants for the combinations the training step launches most; anything else
  // runs the generic (runtime-parameter) epilogue
  const int act = kp.act, gm = kp.gate_mode, dr = kp.drop_thresh16 != 0 ? 1 : 0;
  const int rs = (kp.residual != nullptr || kp.out_scale != 1.f) ? 1 : 0;
  const int code = epi_code(act, gm, dr, rs);
#define S2T_GEMM_CASE(AM, BM_, CODE)                                         \
  if (a_mn == AM && b_mn == BM_ && code == (CODE))                           \
    return launch_gemm<BN, AM, BM_, (CODE)>(ta, tb, tc, kp, nbatch, stream);
  S2T_GEMM_CASE(0, 0, 0) S2T_GEMM_CASE(0, 1, 0) S2T_GEMM_CASE(1, 0, 0) S2T_GEMM_CASE(1, 1, 0)
  S2T_GEMM_CASE(0, 0, epi_code(0, 0, 0, 1)) S2T_GEMM_CASE(0, 0, epi_code(0, 0, 1, 1))
  S2T_GEMM_CASE(0, 0, epi_code(S2T_ACT_RELU, 0, 0, 0)) S2T_GEMM_CASE(0, 0, epi_code(S2T_ACT_RELU, 0, 1, 0))
  S2T_GEMM_CASE(0, 1, epi_code(0, S2T_GATE_POSITIVE, 0, 0)) S2T_GEMM_CASE(0, 1, epi_code(0, S2T_GATE_POSITIVE, 1, 0))
#undef S2T_GEMM_CASE
  if (!a_mn && !b_mn) return launch_gemm<BN, 0, 0, kEpiGeneric>(ta, tb, tc, kp, nbatch, stream);
  if (!a_mn && b_mn) return launch_gemm<BN, 0, 1, kEpiGeneric>(ta, tb, tc, kp, nbatch, stream);
  if (a_mn && !b_mn) return launch_gemm<BN, 1, 0, kEpiGeneric>(ta, tb, tc, kp, nbatch, stream);
  return launch_gemm<BN, 1, 1, kEpiGeneric>(ta, tb, tc, kp, nbatch, stream);
}

}  // namespace s2t

using namespace s2t;

// Debug phase trace (S2T_GEMM_TRACE=1): 16 globaltimer stamps written by CTA 0 of the last traced launch.
namespace s2t {
unsigned long long* trace_buffer();
}
static unsigned long long* gemm_trace_buffer() { return s2t::trace_buffer(); }
unsigned long long* s2t::trace_buffer() {
  static unsigned long long* buf = nullptr;
  static int init = 0;
  if (!init) {
    init = 1;
    const char* e = getenv("S2T_GEMM_TRACE");
    if (e != nullptr && e[0] == '1') {
      if (cudaMalloc(&buf, 16 * sizeof(unsigned long long)) != cudaSuccess) buf = nullptr;
      else cudaMemset(buf, 0, 16 * sizeof(unsigned long long));
    }
  }
  return buf;
}

extern "C" int s2t_gemm_trace_read(unsigned long long* host_out16) {
  unsigned long long* buf = gemm_trace_buffer();
  S2T_REQUIRE(buf != nullptr && host_out16 != nullptr, "s2t_gemm_trace_read: tracing is off (set S2T_GEMM_TRACE=1)");
  S2T_CHECK_CUDA(cudaMemcpy(host_out16, buf, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return 0;
}

extern "C" int s2t_gemm_bf16(const s2t_gemm_args* g, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  S2T_REQUIRE(g != nullptr, "s2t_gemm_bf16: null args");
  S2T_REQUIRE(g->M > 0 && g->N > 0 && g->K > 0, "s2t_gemm_bf16: empty problem M=%d N=%d K=%d", g->M,
              g->N, g->K);
  const int nb1 = g->nb1 > 0 ? g->nb1 : 1, nb2 = g->nb2 > 0 ? g->nb2 : 1;
  S2T_REQUIRE(nb1 * nb2 <= 65535, "s2t_gemm_bf16: too many batches");
  S2T_REQUIRE(g->a && g->b && g->out, "s2t_gemm_bf16: null operand");

  // Tile width and split-K factor from a small cost model (cycles per SM, measured constants: ~44 B/clk of
  // operand traffic per SM, ~13 clk per output column of a 128-row tile in the 8-warp epilogue).
  const int num_kb = (g->K + kBK - 1) / kBK;
  int bn = g->tile_n, split = 1;
  {
    const long long mt = (g->M + kBM - 1) / kBM;
    const long long nbz = static_cast<long long>(nb1) * nb2;
    double best = -1.0;
    const int cta_cap = (g->max_ctas > 0 && g->max_ctas < sm_count()) ? g->max_ctas : sm_count();
    const int sk_cap = cta_cap < 128 ? cta_cap : 128;
    for (int cand : {64, 128, 256}) {
      if (g->tile_n == 64 || g->tile_n == 128 || g->tile_n == 256) {
        if (cand != g->tile_n) continue;
      } else if (cand > 64 && g->N <= cand / 2) {
        continue;  // mostly-empty tile
      }
      const long long tiles = mt * ((g->N + cand - 1) / cand) * nbz;
      const double kb_cyc = (kBM + cand) * 128.0 / 44.0;
      const double epi = cand * 13.0;
      const long long waves = (tiles + cta_cap - 1) / cta_cap;
      const double main1 = num_kb * kb_cyc;
      double cost = 2500.0 + (waves - 1) * (main1 > epi ? main1 : epi) + main1 + epi;
      int s_here = 1;
      if (g->split_k >= 0 && num_kb >= 12) {
        // power-of-two clusters only (odd sizes pack badly into GPCs), one wave, >= 4 k-blocks per rank
        for (int sk : {2, 4, 8}) {
          if (g->split_k > 1 && sk > g->split_k) break;
          if (tiles * sk > sk_cap) break;
          const int kb_per = (num_kb + sk - 1) / sk;
          if (kb_per < 4 || (sk - 1) * kb_per >= num_kb) continue;
          // reduce-scatter tail: park the partial tile (~600 clk), receive (sk-1)/sk of one fp32 tile over DSMEM at
          // ~21 B/clk, run the epilogue on 1/sk of the tile
          const double c2 = 4500.0 + kb_per * kb_cyc + 600.0 + epi / sk + cand * 512.0 * (sk - 1) / sk / 21.0;
          if (c2 < cost || (g->split_k > 1 && sk > s_here)) {  // split_k > 1 forces the largest feasible factor
            cost = c2;
            s_here = sk;
          }
        }
      }
      if (best < 0 || cost < best) {
        best = cost;
        bn = cand;
        split = s_here;
      }
    }
  }

  if (g->colsum != nullptr) {
    S2T_REQUIRE(g->a_mn_major && nb1 * nb2 == 1, "s2t_gemm_bf16: colsum needs an MN-major A operand and no batch dims");
    split = 1;   // the column sums are taken by the persistent kernel
  }
  CUtensorMap ta, tb, tc;
  int rc;
  if (g->a_mn_major)
    rc = make_operand_map(&ta, g->a, g->M, g->K, g->lda, nb1, g->a_bs1, nb2, g->a_bs2, kBK);
  else
    rc = make_operand_map(&ta, g->a, g->K, g->M, g->lda, nb1, g->a_bs1, nb2, g->a_bs2, kBM);
  if (rc) return rc;
  if (g->b_mn_major)
    rc = make_operand_map(&tb, g->b, g->N, g->K, g->ldb, nb1, g->b_bs1, nb2, g->b_bs2, kBK);
  else
    rc = make_operand_map(&tb, g->b, g->K, g->N, g->ldb, nb1, g->b_bs1, nb2, g->b_bs2, bn);
  if (rc) return rc;
  // output: written by TMA stores of 32-row x 128-byte boxes (clipped at M / N by the map)
  S2T_REQUIRE((nb1 == 1 || g->c_bs1 != 0) && (nb2 == 1 || g->c_bs2 != 0),
              "s2t_gemm_bf16: batched output needs non-zero batch strides");
  rc = make_operand_map(&tc, g->out, g->N, g->M, g->ldc, nb1, g->c_bs1, nb2, g->c_bs2, 32, g->out_fp32 ? 1 : 0);
  if (rc) return rc;

  GemmKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  kp.M = g->M; kp.N = g->N; kp.K = g->K; kp.nb1 = nb1;
  kp.nbatch = nb1 * nb2;
  kp.split_k = split;
  kp.max_ctas = g->max_ctas;
  kp.a_zm1 = g->a_bs1 != 0; kp.a_zm2 = g->a_bs2 != 0;
  kp.b_zm1 = g->b_bs1 != 0; kp.b_zm2 = g->b_bs2 != 0;
  kp.out_fp32 = g->out_fp32 ? 1 : 0;
  kp.residual = reinterpret_cast<const __nv_bfloat16*>(g->residual);
  kp.ldr = g->ldr ? g->ldr : g->ldc; kp.r_bs1 = g->r_bs1; kp.r_bs2 = g->r_bs2;
  kp.gate = reinterpret_cast<const __nv_bfloat16*>(g->gate);
  kp.ldg = g->ldg ? g->ldg : g->ldc; kp.g_bs1 = g->g_bs1; kp.g_bs2 = g->g_bs2;
  kp.bias = reinterpret_cast<const __nv_bfloat16*>(g->bias);
  kp.alpha = g->alpha; kp.out_scale = g->out_scale;
  kp.drop_inv_keep = 1.f;
  kp.act = g->act; kp.gate_mode = g->gate ? g->gate_mode : S2T_GATE_NONE;
  if (g->dropout_p > 0.f) {
    S2T_REQUIRE(g->dropout_p < 1.f, "dropout_p must be < 1");
    kp.drop_thresh16 = static_cast<uint32_t>(g->dropout_p * 65536.f + 0.5f);
    kp.drop_inv_keep = 1.f / (1.f - g->dropout_p);
    kp.drop_seed = g->dropout_seed;
    kp.drop_seed_off = g->dropout_seed_offset;
    kp.drop_stream = g->dropout_stream;
  }
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  kp.bias_vec = kp.bias && al16(kp.bias) ? 1 : 0;
  bool vec = true;
  if (kp.residual)
    vec = vec && (kp.ldr % 8 == 0) && (kp.r_bs1 % 8 == 0) && (kp.r_bs2 % 8 == 0) && al16(kp.residual);
  if (kp.gate)
    vec = vec && (kp.ldg % 8 == 0) && (kp.g_bs1 % 8 == 0) && (kp.g_bs2 % 8 == 0) && al16(kp.gate);
  kp.aux_vec = vec ? 1 : 0;
  kp.trace = gemm_trace_buffer();
  kp.colsum = reinterpret_cast<__nv_bfloat16*>(g->colsum);
  kp.colsum_accumulate = g->colsum_accumulate;
  kp.late_trigger = g->late_trigger;
  kp.out = g->out;
  kp.ldc = g->ldc; kp.c_bs1 = nb1 > 1 ? g->c_bs1 : 0; kp.c_bs2 = nb2 > 1 ? g->c_bs2 : 0;
  {
    const long long m = g->out_fp32 ? 4 : 8;
    kp.out_vec = ((reinterpret_cast<uintptr_t>(g->out) & 15) == 0 && g->ldc % m == 0 && kp.c_bs1 % m == 0 &&
                  kp.c_bs2 % m == 0) ? 1 : 0;
  }

  const int nbatch = nb1 * nb2;
  if (split > 1) {
    switch (bn) {
      case 256: return dispatch_splitk<256>(g->a_mn_major, g->b_mn_major, ta, tb, tc, kp, nbatch, stream);
      case 128: return dispatch_splitk<128>(g->a_mn_major, g->b_mn_major, ta, tb, tc, kp, nbatch, stream);
      default: return dispatch_splitk<64>(g->a_mn_major, g->b_mn_major, ta, tb, tc, kp, nbatch, stream);
    }
  }
  switch (bn) {
    case 256: return dispatch_major<256>(g->a_mn_major, g->b_mn_major, ta, tb, tc, kp, nbatch, stream);
    case 128: return dispatch_major<128>(g->a_mn_major, g->b_mn_major, ta, tb, tc, kp, nbatch, stream);
    default: return dispatch_major<64>(g->a_mn_major, g->b_mn_major, ta, tb, tc, kp, nbatch, stream);
  }
}
