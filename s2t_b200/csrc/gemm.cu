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

#include <mutex>
#include <string.h>
#include <unordered_map>

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kGemmThreads = 192;

struct GemmKernelParams {
  int M, N, K;
  int nb1;
  int a_zm1, a_zm2, b_zm1, b_zm2;  // 0 = operand is broadcast along that batch dim (stride 0)
  void* out;
  int out_fp32;
  int vec_ok;  // ldc/ldr/ldg multiples of 8 and 16B-aligned bases -> 16-byte accesses allowed
  long long ldc, c_bs1, c_bs2;
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
};

template <int BN>
struct GemmCfg {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 2 : (BN >= 128 ? 3 : 4);  // <= 96 KB: two CTAs per SM
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == S2T_ACT_RELU) return fmaxf(v, 0.f);
  if (act == S2T_ACT_SWISH) return v / (1.f + __expf(-v));
  return v;
}

template <int BN, int A_MN, int B_MN>
__global__ void __launch_bounds__(kGemmThreads, 2)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmKernelParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* accum_bar = empty_bar + Cfg::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * kBM;
  const int n0 = blockIdx.x * BN;
  const int z = blockIdx.z;
  const int z1 = z % p.nb1;
  const int z2 = z / p.nb1;
  const int num_kb = (p.K + kBK - 1) / kBK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(accum_bar, 1);
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

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % Cfg::kStages;
        const uint32_t ph = (kb / Cfg::kStages) & 1;
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
    // ------------------------------ MMA issuer --------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, A_MN, B_MN);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % Cfg::kStages;
        const uint32_t ph = (kb / Cfg::kStages) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * Cfg::kStageBytes);
        const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k) {
          const uint64_t adesc = A_MN ? umma_smem_desc(a_addr + k * 2048, kBK * 128, 1024)
                                      : umma_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? umma_smem_desc(b_addr + k * 2048, kBK * 128, 1024)
                                      : umma_smem_desc(b_addr + k * 32, 16, 1024);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);  // frees this smem stage once the MMAs above retire
      }
      umma_commit(accum_bar);  // accumulator complete
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < p.M;
    mbar_wait(accum_bar, 0);
    tc_fence_after();

    const long long crow = static_cast<long long>(row) * p.ldc + z1 * p.c_bs1 + z2 * p.c_bs2;
    const long long rrow = static_cast<long long>(row) * p.ldr + z1 * p.r_bs1 + z2 * p.r_bs2;
    const long long grow = static_cast<long long>(row) * p.ldg + z1 * p.g_bs1 + z2 * p.g_bs2;
    // dropout counter: element (z, row, col) lives in Philox group (z*M+row)*ceil(N/8) + col/8, bit col%8
    const unsigned long long drop_row =
        (static_cast<unsigned long long>(z) * p.M + row) * static_cast<unsigned long long>((p.N + 7) >> 3);
    const unsigned long long drop_seed =
        p.drop_seed + ((p.drop_thresh16 != 0 && p.drop_seed_off != nullptr) ? *p.drop_seed_off : 0ull);

#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      const int col0 = n0 + c * 32;
      if (col0 >= p.N) break;  // warp-uniform
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
      tmem_ld_wait();
      if (!row_ok) continue;
#pragma unroll
      for (int g = 0; g < 4; ++g) {  // groups of 8 columns
        const int col = col0 + g * 8;
        if (col >= p.N) break;
        const int nvalid = min(8, p.N - col);
        const bool full8 = (nvalid == 8) && p.vec_ok;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]) * p.alpha;
        if (p.bias != nullptr) {
          if (nvalid == 8 && (col & 7) == 0 &&
              (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0) {
            const uint4 bv = *reinterpret_cast<const uint4*>(p.bias + col);
            v[0] += bf16_lo(bv.x); v[1] += bf16_hi(bv.x);
            v[2] += bf16_lo(bv.y); v[3] += bf16_hi(bv.y);
            v[4] += bf16_lo(bv.z); v[5] += bf16_hi(bv.z);
            v[6] += bf16_lo(bv.w); v[7] += bf16_hi(bv.w);
          } else {
            for (int j = 0; j < nvalid; ++j) v[j] += __bfloat162float(p.bias[col + j]);
          }
        }
        if (p.act != S2T_ACT_NONE) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = apply_act(v[j], p.act);
        }
        if (p.gate_mode != S2T_GATE_NONE) {
          float gt[8];
          if (full8) {
            const uint4 gv = *reinterpret_cast<const uint4*>(p.gate + grow + col);
            gt[0] = bf16_lo(gv.x); gt[1] = bf16_hi(gv.x);
            gt[2] = bf16_lo(gv.y); gt[3] = bf16_hi(gv.y);
            gt[4] = bf16_lo(gv.z); gt[5] = bf16_hi(gv.z);
            gt[6] = bf16_lo(gv.w); gt[7] = bf16_hi(gv.w);
          } else {
            for (int j = 0; j < 8; ++j)
              gt[j] = j < nvalid ? __bfloat162float(p.gate[grow + col + j]) : 0.f;
          }
          if (p.gate_mode == S2T_GATE_POSITIVE) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = gt[j] > 0.f ? v[j] : 0.f;
          } else if (p.gate_mode == S2T_GATE_SWISH_GRAD) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float s = 1.f / (1.f + __expf(-gt[j]));
              v[j] *= s * (1.f + gt[j] * (1.f - s));
            }
          } else {  // S2T_GATE_MUL
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] *= gt[j];
          }
        }
        if (p.drop_thresh16 != 0) {
          const uint32_t keep =
              dropout_keep8(drop_seed, p.drop_stream, drop_row + (col >> 3), p.drop_thresh16);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * p.drop_inv_keep : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= p.out_scale;
        if (p.residual != nullptr) {
          if (full8) {
            const uint4 rv = *reinterpret_cast<const uint4*>(p.residual + rrow + col);
            v[0] += bf16_lo(rv.x); v[1] += bf16_hi(rv.x);
            v[2] += bf16_lo(rv.y); v[3] += bf16_hi(rv.y);
            v[4] += bf16_lo(rv.z); v[5] += bf16_hi(rv.z);
            v[6] += bf16_lo(rv.w); v[7] += bf16_hi(rv.w);
          } else {
            for (int j = 0; j < nvalid; ++j) v[j] += __bfloat162float(p.residual[rrow + col + j]);
          }
        }
        if (p.out_fp32) {
          float* o = reinterpret_cast<float*>(p.out) + crow + col;
          if (full8) {
            *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
          } else {
            for (int j = 0; j < nvalid; ++j) o[j] = v[j];
          }
        } else {
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + crow + col;
          if (full8) {
            uint4 ov;
            ov.x = pack_bf16(v[0], v[1]);
            ov.y = pack_bf16(v[2], v[3]);
            ov.z = pack_bf16(v[4], v[5]);
            ov.w = pack_bf16(v[6], v[7]);
            *reinterpret_cast<uint4*>(o) = ov;
          } else {
            for (int j = 0; j < nvalid; ++j) o[j] = __float2bfloat16(v[j]);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN < 32 ? 32 : BN);
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
static int make_operand_map(CUtensorMap* out, const void* ptr, uint64_t inner, uint64_t outer,
                            uint64_t ld, uint64_t nb1, uint64_t bs1, uint64_t nb2, uint64_t bs2,
                            uint32_t box_outer) {
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  static std::mutex mu;
  MapKey key{{reinterpret_cast<uint64_t>(ptr), inner, outer, ld, nb1, bs1, nb2, bs2, box_outer, 0}};
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
  S2T_REQUIRE((ld * 2) % 16 == 0, "GEMM operand leading stride must be a multiple of 8 elements (got %llu)",
              (unsigned long long)ld);
  // Unit batch dims still need a legal (16B-multiple, non-zero) stride.
  if (bs1 == 0) nb1 = 1;  // broadcast batch dims collapse; the kernel zeroes the coordinate
  if (bs2 == 0) nb2 = 1;
  uint64_t s1 = (nb1 > 1 ? bs1 : ld * outer) * 2, s2 = (nb2 > 1 ? bs2 : ld * outer) * 2;
  if (s1 == 0) s1 = 16;
  if (s2 == 0) s2 = 16;
  S2T_REQUIRE(s1 % 16 == 0 && s2 % 16 == 0, "GEMM batch strides must be multiples of 8 elements");
  cuuint64_t dims[4] = {inner, outer, nb1, nb2};
  cuuint64_t strides[3] = {ld * 2, s1, s2};
  cuuint32_t box[4] = {64, box_outer, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides,
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

template <int BN, int A_MN, int B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmKernelParams& kp,
                       int nbatch, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool configured = false;
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN>;
  if (!configured) {
    S2T_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    configured = true;
  }
  dim3 grid((kp.N + BN - 1) / BN, (kp.M + kBM - 1) / kBM, nbatch);
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(ta, tb, kp);
  S2T_LAUNCH_CHECK();
  return 0;
}

template <int BN>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                          const GemmKernelParams& kp, int nbatch, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm<BN, 0, 0>(ta, tb, kp, nbatch, stream);
  if (!a_mn && b_mn) return launch_gemm<BN, 0, 1>(ta, tb, kp, nbatch, stream);
  if (a_mn && !b_mn) return launch_gemm<BN, 1, 0>(ta, tb, kp, nbatch, stream);
  return launch_gemm<BN, 1, 1>(ta, tb, kp, nbatch, stream);
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

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_gemm_bf16(const s2t_gemm_args* g, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  S2T_REQUIRE(g != nullptr, "s2t_gemm_bf16: null args");
  S2T_REQUIRE(g->M > 0 && g->N > 0 && g->K > 0, "s2t_gemm_bf16: empty problem M=%d N=%d K=%d", g->M,
              g->N, g->K);
  const int nb1 = g->nb1 > 0 ? g->nb1 : 1, nb2 = g->nb2 > 0 ? g->nb2 : 1;
  S2T_REQUIRE(nb1 * nb2 <= 65535, "s2t_gemm_bf16: too many batches");
  S2T_REQUIRE(g->a && g->b && g->out, "s2t_gemm_bf16: null operand");

  // tile width: widest BN that still gives every SM a CTA (small-M problems are latency bound)
  int bn = g->tile_n;
  if (bn != 64 && bn != 128 && bn != 256) {
    const long long mt = (g->M + kBM - 1) / kBM;
    const long long nbz = static_cast<long long>(nb1) * nb2;
    bn = 64;
    for (int cand : {256, 128}) {
      if (g->N < cand / 2 + 1 && cand > 64) continue;  // mostly-empty tile
      const long long ctas = mt * ((g->N + cand - 1) / cand) * nbz;
      if (ctas >= sm_count()) {
        bn = cand;
        break;
      }
    }
  }

  CUtensorMap ta, tb;
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

  GemmKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  kp.M = g->M; kp.N = g->N; kp.K = g->K; kp.nb1 = nb1;
  kp.a_zm1 = g->a_bs1 != 0; kp.a_zm2 = g->a_bs2 != 0;
  kp.b_zm1 = g->b_bs1 != 0; kp.b_zm2 = g->b_bs2 != 0;
  kp.out = g->out; kp.out_fp32 = g->out_fp32;
  kp.ldc = g->ldc; kp.c_bs1 = g->c_bs1; kp.c_bs2 = g->c_bs2;
  kp.residual = reinterpret_cast<const __nv_bfloat16*>(g->residual);
  kp.ldr = g->ldr ? g->ldr : g->ldc; kp.r_bs1 = g->r_bs1; kp.r_bs2 = g->r_bs2;
  kp.gate = reinterpret_cast<const __nv_bfloat16*>(g->gate);
  kp.ldg = g->ldg ? g->ldg : g->ldc; kp.g_bs1 = g->g_bs1; kp.g_bs2 = g->g_bs2;
  kp.bias = reinterpret_cast<const __nv_bfloat16*>(g->bias);
  kp.alpha = g->alpha; kp.out_scale = g->out_scale;
  kp.act = g->act; kp.gate_mode = g->gate ? g->gate_mode : S2T_GATE_NONE;
  if (g->dropout_p > 0.f) {
    S2T_REQUIRE(g->dropout_p < 1.f, "dropout_p must be < 1");
    kp.drop_thresh16 = static_cast<uint32_t>(g->dropout_p * 65536.f + 0.5f);
    kp.drop_inv_keep = 1.f / (1.f - g->dropout_p);
    kp.drop_seed = g->dropout_seed;
    kp.drop_seed_off = g->dropout_seed_offset;
    kp.drop_stream = g->dropout_stream;
  }
  const int esz = g->out_fp32 ? 4 : 2;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  bool vec = (kp.ldc % 8 == 0) && (kp.c_bs1 % 8 == 0) && (kp.c_bs2 % 8 == 0) && al16(g->out);
  (void)esz;
  if (kp.residual)
    vec = vec && (kp.ldr % 8 == 0) && (kp.r_bs1 % 8 == 0) && (kp.r_bs2 % 8 == 0) && al16(kp.residual);
  if (kp.gate)
    vec = vec && (kp.ldg % 8 == 0) && (kp.g_bs1 % 8 == 0) && (kp.g_bs2 % 8 == 0) && al16(kp.gate);
  kp.vec_ok = vec ? 1 : 0;

  const int nbatch = nb1 * nb2;
  switch (bn) {
    case 256: return dispatch_major<256>(g->a_mn_major, g->b_mn_major, ta, tb, kp, nbatch, stream);
    case 128: return dispatch_major<128>(g->a_mn_major, g->b_mn_major, ta, tb, kp, nbatch, stream);
    default: return dispatch_major<64>(g->a_mn_major, g->b_mn_major, ta, tb, kp, nbatch, stream);
  }
}
