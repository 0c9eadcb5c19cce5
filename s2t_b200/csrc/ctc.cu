// K7/K8/K9: CTC head log-softmax statistics, CTC loss forward/backward, greedy CTC decode.
//
// Replaces, on the reference path,
//   utils.log_softmax(logits, dim=-1) in fp32                      (fairseq/utils.py:477-481)
//   torch.nn.CTCLoss(blank=0, reduction="none", zero_infinity=True) (fairseq/criterions/ctc.py:243-245,
//                                                                    435-474; ATen LossCTC alpha/beta)
//   CTCDecoder greedy: argmax -> pad->blank -> unique_consecutive -> drop blank
//                                                                   (fairseq/models/speech_to_text/s2t_ctc.py:312-347)
//
// The log-probabilities are never materialised: the loss works from the logits, one fp32 row-LSE
// vector and the (2S+1)-state lattice; the gradient w.r.t. the LOGITS, softmax - occupancy, is written
// once (algorithmic HBM traffic = read logits + write grad).
//
// Kernels:
//   row_lse            one CTA per (b,t) row, online max/sum, 16-byte loads
//   ctc_gather         lp_g[b][t][s] = logit[b][t][l'_s] - lse[b][t]  (parallel gather, so the serial
//                      recursions below read contiguous, prefetchable fp32 rows)
//   ctc_alpha_beta     grid (B, 2): y=0 forward (alpha, nll), y=1 backward (beta) recursion, lattice
//                      row in shared memory, one __syncthreads per frame, lp rows prefetched 4 frames ahead
//   ctc_dense_grad     dlogits = scale * softmax(logits) (0 for padded frames / infeasible utterances)
//   ctc_occupancy      dlogits[b][t][l'_s] -= scale * gamma_t(s), duplicates of a label merged in smem
//   argmax / collapse  greedy decode
#include "common.cuh"

#include <mutex>
#include <stdlib.h>

#include <math.h>

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr float kNegInf = -INFINITY;

__device__ __forceinline__ float load_logit(const void* p, int is_bf16, long long idx) {
  return is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[idx])
                 : reinterpret_cast<const float*>(p)[idx];
}

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  if (m2 == kNegInf) return;
  if (m == kNegInf) {
    m = m2;
    s = s2;
    return;
  }
  const float mm = fmaxf(m, m2);
  s = s * __expf(m - mm) + s2 * __expf(m2 - mm);
  m = mm;
}

// ---------------------------------------------------------------------------------------------
// row LSE
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
row_lse_kernel(const void* __restrict__ logits, int is_bf16, long long ldv, int V,
               float* __restrict__ lse) {
  pdl_entry();
  __shared__ float s_m[8], s_s[8];
  const long long row = blockIdx.x;
  float m = kNegInf, s = 0.f;
  if (is_bf16) {
    const __nv_bfloat16* x = reinterpret_cast<const __nv_bfloat16*>(logits) + row * ldv;
    const int nv = V >> 3;
    for (int i = threadIdx.x; i < nv; i += 256) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + i * 8);
      const float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                          bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
      float lm = f[0];
#pragma unroll
      for (int j = 1; j < 8; ++j) lm = fmaxf(lm, f[j]);
      float ls = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) ls += __expf(f[j] - lm);
      online_merge(m, s, lm, ls);
    }
    for (int i = (nv << 3) + threadIdx.x; i < V; i += 256) online_merge(m, s, __bfloat162float(x[i]), 1.f);
  } else {
    const float* x = reinterpret_cast<const float*>(logits) + row * ldv;
    const int nv = V >> 2;
    for (int i = threadIdx.x; i < nv; i += 256) {
      const float4 v = *reinterpret_cast<const float4*>(x + i * 4);
      const float lm = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
      const float ls = __expf(v.x - lm) + __expf(v.y - lm) + __expf(v.z - lm) + __expf(v.w - lm);
      online_merge(m, s, lm, ls);
    }
    for (int i = (nv << 2) + threadIdx.x; i < V; i += 256) online_merge(m, s, x[i], 1.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
    const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
    online_merge(m, s, m2, s2);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_m[warp] = m;
    s_s[warp] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float M = s_m[0], S = s_s[0];
    for (int w = 1; w < 8; ++w) online_merge(M, S, s_m[w], s_s[w]);
    lse[row] = M + logf(S);
  }
}

// ---------------------------------------------------------------------------------------------
// workspace layout (all fp32 / int32, sizes in elements):
//   lp_g  [B][T][Sp]   alpha [B][T][Sp]   beta [B][T][Sp]   leader int [B][Sp]   nll_raw [B]
// ---------------------------------------------------------------------------------------------
struct CtcWs {
  float* lp_g;
  float* alpha;
  float* beta;
  int* leader;
  float* nll_raw;
};

__host__ __device__ inline long long ctc_ws_floats(int B, int T, int Sp) {
  return 3ll * B * T * Sp + 1ll * B * Sp + B;
}

__device__ __forceinline__ int ext_label(const int* tgt, int s, int blank) {
  return (s & 1) ? tgt[s >> 1] : blank;
}

__global__ void ctc_gather_kernel(const void* __restrict__ logits, int is_bf16, long long ldv,
                                  const float* __restrict__ lse, const int* __restrict__ targets,
                                  long long tgt_ld, const int* __restrict__ in_len,
                                  const int* __restrict__ tgt_len, int T, int Sp, int V, int blank,
                                  CtcWs ws) {
  pdl_entry();
  const int b = blockIdx.y;
  const int Tb = min(in_len[b], T);
  const int S = tgt_len[b];
  const int Se = 2 * S + 1;
  const int* tgt = targets + b * tgt_ld;
  // leaders (first extended position carrying the same label); blank positions lead to 0
  if (blockIdx.x == 0) {
    for (int s = threadIdx.x; s < Sp; s += blockDim.x) {
      int lead = s;
      if (s < Se) {
        if ((s & 1) == 0) lead = 0;
        else {
          const int l = tgt[s >> 1];
          for (int k = 0; k < (s >> 1); ++k)
            if (tgt[k] == l) {
              lead = 2 * k + 1;
              break;
            }
        }
      }
      ws.leader[b * Sp + s] = lead;
    }
  }
  const long long n = static_cast<long long>(Tb) * Se;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / Se), s = static_cast<int>(i % Se);
    int l = ext_label(tgt, s, blank);
    l = min(max(l, 0), V - 1);
    const long long row = static_cast<long long>(b) * T + t;
    ws.lp_g[row * Sp + s] = load_logit(logits, is_bf16, row * ldv + l) - lse[row];
  }
}

__device__ __forceinline__ float lse3(float a, float b, float c) {
  const float m = fmaxf(fmaxf(a, b), c);
  if (m == kNegInf) return kNegInf;
  return m + logf(__expf(a - m) + __expf(b - m) + __expf(c - m));
}

// branch-free log(e^a + e^b + e^c) on the SFU (ex2.approx / lg2.approx): the sum lies in [1, 3], where lg2.approx is
// accurate to ~2^-21 absolute; an all -inf triple gives log(0) = -inf.
__device__ __forceinline__ float lse3_fast(float a, float b, float c) {
  const float m = fmaxf(fmaxf(a, b), c);
  const float ms = (m == kNegInf) ? 0.f : m;
  return ms + __logf(__expf(a - ms) + __expf(b - ms) + __expf(c - ms));
}

template <int N>
__device__ __forceinline__ void vec_load(const float* __restrict__ p, float (&v)[N]) {
  if constexpr (N == 1) {
    v[0] = *p;
  } else if constexpr (N == 2) {
    const float2 t = *reinterpret_cast<const float2*>(p);
    v[0] = t.x; v[1] = t.y;
  } else {
#pragma unroll
    for (int i = 0; i < N; i += 4) {
      const float4 t = *reinterpret_cast<const float4*>(p + i);
      v[i] = t.x; v[i + 1] = t.y; v[i + 2] = t.z; v[i + 3] = t.w;
    }
  }
}

template <int N>
__device__ __forceinline__ void vec_store(float* __restrict__ p, const float (&v)[N]) {
  if constexpr (N == 1) {
    *p = v[0];
  } else if constexpr (N == 2) {
    *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
  } else {
#pragma unroll
    for (int i = 0; i < N; i += 4)
      *reinterpret_cast<float4*>(p + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
  }
}

constexpr int kCtcMaxPerThread = 4;  // Sp <= 4 * blockDim

// MASKED: the "imputer" lattice (torch_imputer/imputer.cu:57-225, 364-520) — force_emit[b][t] = index of the extended-label
// state that must be occupied at frame t, or -1; every other state of that frame is removed (alpha = beta = -inf).
template <bool MASKED>
__global__ void ctc_alpha_beta_kernel(const int* __restrict__ targets, long long tgt_ld,
                                      const int* __restrict__ in_len, const int* __restrict__ tgt_len,
                                      int T, int Sp, int blank, CtcWs ws,
                                      const int* __restrict__ force_emit, long long fe_ld) {
  pdl_entry();
  extern __shared__ float s_lat[];  // [2][Sp + 4]
  const int b = blockIdx.x;
  const bool fwd = blockIdx.y == 0;
  const int Tb = min(in_len[b], T);
  const int S = tgt_len[b];
  const int Se = 2 * S + 1;
  const int* tgt = targets + b * tgt_ld;
  const int stride = Sp + 4;
  float* buf0 = s_lat + 2;           // two guard cells on each side hold -inf
  float* buf1 = s_lat + stride + 2;
  for (int i = threadIdx.x; i < 2 * stride; i += blockDim.x) s_lat[i] = kNegInf;

  // can this state take the skip transition (from s-2 forward / to s+2 backward)?
  bool skip[kCtcMaxPerThread];
#pragma unroll
  for (int k = 0; k < kCtcMaxPerThread; ++k) {
    const int s = threadIdx.x + k * blockDim.x;
    skip[k] = false;
    if (s < Se && (s & 1)) {
      if (fwd) skip[k] = s >= 3 && tgt[s >> 1] != tgt[(s >> 1) - 1];
      else skip[k] = (s + 2 < Se) && tgt[s >> 1] != tgt[(s >> 1) + 1];
    }
  }
  __syncthreads();
  if (Tb <= 0) {
    if (fwd && threadIdx.x == 0) ws.nll_raw[b] = (S == 0) ? 0.f : INFINITY;
    return;
  }
  const float* lp = ws.lp_g + static_cast<long long>(b) * T * Sp;
  float* lat = (fwd ? ws.alpha : ws.beta) + static_cast<long long>(b) * T * Sp;

  float* prev = buf0;
  float* cur = buf1;
  // ---- boundary frame ----------------------------------------------------------------------------
  {
    const int t = fwd ? 0 : Tb - 1;
#pragma unroll
    for (int k = 0; k < kCtcMaxPerThread; ++k) {
      const int s = threadIdx.x + k * blockDim.x;
      if (s < Se) {
        float v = kNegInf;
        if (fwd) {
          if (s <= 1) v = lp[static_cast<long long>(t) * Sp + s];
        } else {
          if (s >= Se - 2) v = lp[static_cast<long long>(t) * Sp + s];
        }
        if (MASKED) {
          const int fe = force_emit[b * fe_ld + t];
          if (fe > -1 && fe != s) v = kNegInf;
        }
        prev[s] = v;
        lat[static_cast<long long>(t) * Sp + s] = v;
      }
    }
  }
  __syncthreads();
  // ---- recursion, lp rows prefetched kPf frames ahead ----------------------------------------------
  constexpr int kPf = 4;
  float pf[kPf][kCtcMaxPerThread];
#pragma unroll
  for (int d = 0; d < kPf; ++d) {
    const int step = 1 + d;
    const int t = fwd ? step : Tb - 1 - step;
#pragma unroll
    for (int k = 0; k < kCtcMaxPerThread; ++k) {
      const int s = threadIdx.x + k * blockDim.x;
      pf[d][k] = (step < Tb && s < Se) ? lp[static_cast<long long>(t) * Sp + s] : 0.f;
    }
  }
  for (int step0 = 1; step0 < Tb; step0 += kPf) {
#pragma unroll
    for (int d = 0; d < kPf; ++d) {
      const int step = step0 + d;
      if (step < Tb) {  // block-uniform
        const int t = fwd ? step : Tb - 1 - step;
        int fe = -1;
        if (MASKED) fe = force_emit[b * fe_ld + t];
#pragma unroll
        for (int k = 0; k < kCtcMaxPerThread; ++k) {
          const int s = threadIdx.x + k * blockDim.x;
          if (s < Se) {
            float a, bb, c;
            if (fwd) {
              a = prev[s];
              bb = prev[s - 1];
              c = skip[k] ? prev[s - 2] : kNegInf;
            } else {
              a = prev[s];
              bb = (s + 1 < Se) ? prev[s + 1] : kNegInf;
              c = skip[k] ? prev[s + 2] : kNegInf;
            }
            float v = lse3(a, bb, c) + pf[d][k];
            if (MASKED && fe > -1 && fe != s) v = kNegInf;
            cur[s] = v;
            lat[static_cast<long long>(t) * Sp + s] = v;
          }
          // refill this prefetch slot with the row kPf frames further on
          const int nstep = step + kPf;
          const int nt = fwd ? nstep : Tb - 1 - nstep;
          pf[d][k] = (nstep < Tb && s < Se) ? lp[static_cast<long long>(nt) * Sp + s] : 0.f;
        }
        __syncthreads();
        float* tmp = prev;
        prev = cur;
        cur = tmp;
      }
    }
  }
  if (fwd && threadIdx.x == 0) {
    const float a = prev[Se - 1];
    const float c = Se >= 2 ? prev[Se - 2] : kNegInf;
    const float m = fmaxf(a, c);
    const float ll = (m == kNegInf) ? kNegInf : m + logf(__expf(a - m) + __expf(c - m));
    ws.nll_raw[b] = -ll;
  }
}

// Warp-resident lattice: one warp per (utterance, direction), lane l owns the NS consecutive states
// [l*NS, (l+1)*NS) in registers, neighbours' edge states arrive by two warp shuffles per frame — no shared
// memory and no block barrier on the T-step critical path (the recursion is latency-bound: T dependent steps).
// log(e^a + e^b) with one ex2 and one lg2; (-inf, -inf) -> -inf
__device__ __forceinline__ float lse2_fast(float a, float b) {
  const float m = fmaxf(a, b), n = fminf(a, b);
  const float r = m + __logf(1.f + __expf(n - m));
  return m == kNegInf ? kNegInf : r;
}
// log(e^a + e^b + e^c) with two ex2 and one lg2 (the largest term is exactly 1); all -inf -> -inf
__device__ __forceinline__ float lse3_sorted(float a, float b, float c) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  const float m = fmaxf(hi, c), mid = fmaxf(lo, fminf(hi, c)), mn = fminf(lo, c);
  const float r = m + __logf(1.f + __expf(mid - m) + __expf(mn - m));
  return m == kNegInf ? kNegInf : r;
}

// One warp walks one (utterance, direction) lattice.  The gathered emission log-probs lp_g[b] (T x Sp fp32) are staged
// in shared memory by bulk async copies (whole utterance at once when it fits, else two alternating time chunks), so
// the T-step dependent chain only ever waits on shared-memory loads, warp shuffles and the SFU: global-load latency
// (and the scoreboard slots in-flight global loads would share with the shuffles) is off the critical path.
// Blank states have no skip transition: log-sum-exp of two terms (1 ex2 + 1 lg2); label states: three terms sorted so
// that the largest contributes exactly 1 (2 ex2 + 1 lg2).
template <int NS, bool fwd>
__device__ __forceinline__ void ctc_lattice_warp(const int* __restrict__ targets, long long tgt_ld,
                                                 const int* __restrict__ in_len, const int* __restrict__ tgt_len,
                                                 int T, int Sp, int Tc, const CtcWs& ws, float* s_fin, float* s_lp,
                                                 uint64_t* bars) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const int Tb = min(in_len[b], T);
  const int S = tgt_len[b];
  const int Se = 2 * S + 1;
  const int* tgt = targets + b * tgt_ld;
  if (Tb <= 0) {
    if (fwd && lane == 0) ws.nll_raw[b] = (S == 0) ? 0.f : INFINITY;
    return;
  }
  const float* lp = ws.lp_g + static_cast<long long>(b) * T * Sp;
  float* lat = (fwd ? ws.alpha : ws.beta) + static_cast<long long>(b) * T * Sp;
  const int n_chunks = (Tb + Tc - 1) / Tc;
  // chunk c = steps [c*Tc, min(Tb, (c+1)*Tc)); step -> frame t = step (fwd) or Tb-1-step (bwd): a contiguous row range
  auto issue = [&](int c) {
    const int st0 = c * Tc, st1 = min(Tb, st0 + Tc);
    const int t_lo = fwd ? st0 : Tb - st1;
    const uint32_t bytes = static_cast<uint32_t>(st1 - st0) * Sp * 4u;
    mbar_expect_tx(&bars[c & 1], bytes);
    bulk_copy_g2s(s_lp + static_cast<size_t>(c & 1) * Tc * Sp, lp + static_cast<long long>(t_lo) * Sp, bytes,
                  &bars[c & 1]);
  };
  if (lane == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
    issue(0);
    if (n_chunks > 1) issue(1);
  }
  __syncwarp();
  const int s0 = lane * NS;
  // Sp is a multiple of 8 (s2t_ctc_loss), so this lane's NS states are one aligned vector inside the row or all outside
  const bool in_row = s0 < Sp;
  bool skip[NS], live[NS];
  float prev[NS];
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    const int s = s0 + k;
    skip[k] = false;
    live[k] = s < Se;
    if (s < Se && (s & 1)) {
      if (fwd) skip[k] = s >= 3 && tgt[s >> 1] != tgt[(s >> 1) - 1];
      else skip[k] = (s + 2 < Se) && tgt[s >> 1] != tgt[(s >> 1) + 1];
    }
  }
  const long long row_step = fwd ? Sp : -static_cast<long long>(Sp);
  float* lat_row = lat + (fwd ? 0ll : static_cast<long long>(Tb - 1) * Sp) + s0;
  for (int c = 0; c < n_chunks; ++c) {
    const int st0 = c * Tc, st1 = min(Tb, st0 + Tc);
    mbar_wait(&bars[c & 1], (c >> 1) & 1);
    // row of `step` inside the chunk buffer: step - st0 (fwd) / st1 - 1 - step (bwd)
    const float* buf = s_lp + static_cast<size_t>(c & 1) * Tc * Sp + s0;
    const float* row = buf + (fwd ? 0 : (st1 - 1 - st0)) * Sp;
    const int rstep = fwd ? Sp : -Sp;
    int step = st0;
    if (c == 0) {
      float v0[NS] = {};
      if (in_row) vec_load<NS>(row, v0);
#pragma unroll
      for (int k = 0; k < NS; ++k) {
        const int s = s0 + k;
        prev[k] = (live[k] && (fwd ? (s <= 1) : (s >= Se - 2))) ? v0[k] : kNegInf;
      }
      if (in_row) vec_store<NS>(lat_row, prev);
      lat_row += row_step;
      row += rstep;
      step = 1;
    }
    float nxt[NS] = {};
    if (step < st1 && in_row) vec_load<NS>(row, nxt);
#pragma unroll 4
    for (; step < st1; ++step) {
      float e[NS];
#pragma unroll
      for (int k = 0; k < NS; ++k) e[k] = nxt[k];
      row += rstep;
      if (step + 1 < st1 && in_row) vec_load<NS>(row, nxt);
      // n1 / n2: the states one and two positions outside this lane's range (towards s-1 fwd, s+1 bwd)
      float n1, n2;
      if (fwd) {
        n1 = __shfl_up_sync(0xffffffffu, prev[NS - 1], 1);
        n2 = NS >= 2 ? __shfl_up_sync(0xffffffffu, prev[NS >= 2 ? NS - 2 : 0], 1)
                     : __shfl_up_sync(0xffffffffu, prev[0], 2);
        if (lane == 0) n1 = n2 = kNegInf;
        if (NS == 1 && lane == 1) n2 = kNegInf;
      } else {
        n1 = __shfl_down_sync(0xffffffffu, prev[0], 1);
        n2 = NS >= 2 ? __shfl_down_sync(0xffffffffu, prev[NS >= 2 ? 1 : 0], 1)
                     : __shfl_down_sync(0xffffffffu, prev[0], 2);
        if (lane == 31) n1 = n2 = kNegInf;
        if (NS == 1 && lane == 30) n2 = kNegInf;
      }
      float cur[NS];
#pragma unroll
      for (int k = 0; k < NS; ++k) {
        float bb, cc;
        if (fwd) {
          bb = k >= 1 ? prev[k >= 1 ? k - 1 : 0] : n1;
          cc = k >= 2 ? prev[k >= 2 ? k - 2 : 0] : (k == 1 ? n1 : n2);
        } else {
          bb = k + 1 < NS ? prev[k + 1 < NS ? k + 1 : 0] : n1;
          cc = k + 2 < NS ? prev[k + 2 < NS ? k + 2 : 0] : (k + 2 == NS ? n1 : n2);
        }
        float v;
        if (NS >= 2 && (k & 1) == 0) {          // blank state (s0 is even): no skip transition
          v = lse2_fast(prev[k], bb);
        } else {
          if (!skip[k]) cc = kNegInf;
          v = lse3_sorted(prev[k], bb, cc);
        }
        cur[k] = live[k] ? v + e[k] : kNegInf;   // states >= Se stay -inf
      }
      if (in_row) vec_store<NS>(lat_row, cur);
      lat_row += row_step;
#pragma unroll
      for (int k = 0; k < NS; ++k) prev[k] = cur[k];
    }
    if (c + 2 < n_chunks) {   // this buffer is free again: refill it with the chunk after next
      __syncwarp();
      if (lane == 0) {
        fence_proxy_async();
        issue(c + 2);
      }
    }
  }
  if (fwd) {
#pragma unroll
    for (int k = 0; k < NS; ++k) s_fin[s0 + k] = prev[k];
    __syncwarp();
    if (lane == 0) {
      const float a = s_fin[Se - 1];
      const float c = Se >= 2 ? s_fin[Se - 2] : kNegInf;
      const float m = fmaxf(a, c);
      const float ll = (m == kNegInf) ? kNegInf : m + logf(__expf(a - m) + __expf(c - m));
      ws.nll_raw[b] = -ll;
    }
  }
}

template <int NS>
__global__ void __launch_bounds__(32)
ctc_alpha_beta_warp_kernel(const int* __restrict__ targets, long long tgt_ld, const int* __restrict__ in_len,
                           const int* __restrict__ tgt_len, int T, int Sp, int Tc, CtcWs ws) {
  pdl_entry();
  extern __shared__ __align__(128) uint8_t s_dyn[];   // [1 or 2][Tc][Sp] fp32 emission rows
  __shared__ float s_fin[32 * NS];
  __shared__ __align__(8) uint64_t s_bars[2];
  float* s_lp = reinterpret_cast<float*>(s_dyn);
  if (blockIdx.y == 0) ctc_lattice_warp<NS, true>(targets, tgt_ld, in_len, tgt_len, T, Sp, Tc, ws, s_fin, s_lp, s_bars);
  else ctc_lattice_warp<NS, false>(targets, tgt_ld, in_len, tgt_len, T, Sp, Tc, ws, s_fin, s_lp, s_bars);
}

// dlogits = scale * softmax(logits) for valid frames of feasible utterances, 0 elsewhere; also
// publishes nll with zero_infinity applied.
__global__ void __launch_bounds__(256)
ctc_dense_grad_kernel(const void* __restrict__ logits, int is_bf16, long long ldv,
                      const float* __restrict__ lse, const int* __restrict__ in_len, int T, int V,
                      const float* __restrict__ nll_raw, float* __restrict__ nll_out,
                      void* __restrict__ dlogits, float scale, const float* __restrict__ scale_mul) {
  pdl_entry();
  if (scale_mul != nullptr) scale *= *scale_mul;   // device-side batch weight (0 for a data-parallel dummy batch)
  const long long row = blockIdx.x;
  const int b = static_cast<int>(row / T), t = static_cast<int>(row % T);
  // nll_raw == nullptr: launched beside the lattice kernel (s2t_ctc_loss); the occupancy kernel then publishes nll and
  // clears the rows of infeasible utterances
  const float nr = nll_raw != nullptr ? nll_raw[b] : 0.f;
  const bool feasible = isfinite(nr);
  if (t == 0 && threadIdx.x == 0 && nll_out != nullptr) nll_out[b] = feasible ? nr : 0.f;
  if (dlogits == nullptr) return;
  const bool live = feasible && t < in_len[b];
  const float l = lse[row];
  if (is_bf16) {
    const __nv_bfloat16* x = reinterpret_cast<const __nv_bfloat16*>(logits) + row * ldv;
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(dlogits) + row * ldv;
    const int nv = V >> 3;
    for (int i = threadIdx.x; i < nv; i += 256) {
      uint4 ov = make_uint4(0, 0, 0, 0);
      if (live) {
        const uint4 v = *reinterpret_cast<const uint4*>(x + i * 8);
        ov.x = pack_bf16(scale * __expf(bf16_lo(v.x) - l), scale * __expf(bf16_hi(v.x) - l));
        ov.y = pack_bf16(scale * __expf(bf16_lo(v.y) - l), scale * __expf(bf16_hi(v.y) - l));
        ov.z = pack_bf16(scale * __expf(bf16_lo(v.z) - l), scale * __expf(bf16_hi(v.z) - l));
        ov.w = pack_bf16(scale * __expf(bf16_lo(v.w) - l), scale * __expf(bf16_hi(v.w) - l));
      }
      *reinterpret_cast<uint4*>(o + i * 8) = ov;
    }
    for (int i = (nv << 3) + threadIdx.x; i < V; i += 256)
      o[i] = __float2bfloat16(live ? scale * __expf(__bfloat162float(x[i]) - l) : 0.f);
  } else {
    const float* x = reinterpret_cast<const float*>(logits) + row * ldv;
    float* o = reinterpret_cast<float*>(dlogits) + row * ldv;
    for (int i = threadIdx.x; i < V; i += 256) o[i] = live ? scale * expf(x[i] - l) : 0.f;
  }
}

// one CTA per (t, b): gamma_t(s) = exp(alpha + beta - lp + nll); merged per label; RMW on dlogits
__global__ void ctc_occupancy_kernel(const int* __restrict__ targets, long long tgt_ld,
                                     const int* __restrict__ in_len, const int* __restrict__ tgt_len,
                                     int T, int Sp, int V, int blank, CtcWs ws, int is_bf16,
                                     long long ldv, void* __restrict__ dlogits, float scale,
                                     const float* __restrict__ scale_mul, float* __restrict__ nll_out) {
  pdl_entry();
  if (scale_mul != nullptr) scale *= *scale_mul;
  extern __shared__ float s_occ[];  // [Sp]
  const int t = blockIdx.x, b = blockIdx.y;
  const float nr = ws.nll_raw[b];
  if (nll_out != nullptr) {
    // the dense part ran without knowing which utterances are feasible (zero_infinity): settle that here
    if (t == 0 && threadIdx.x == 0) nll_out[b] = isfinite(nr) ? nr : 0.f;
    if (!isfinite(nr) && t < min(in_len[b], T)) {
      const long long r0 = (static_cast<long long>(b) * T + t) * ldv;
      if (is_bf16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(dlogits) + r0;
        for (int i = threadIdx.x; i < V; i += blockDim.x) o[i] = __float2bfloat16(0.f);
      } else {
        float* o = reinterpret_cast<float*>(dlogits) + r0;
        for (int i = threadIdx.x; i < V; i += blockDim.x) o[i] = 0.f;
      }
    }
  }
  if (!isfinite(nr) || t >= min(in_len[b], T)) return;
  const int S = tgt_len[b];
  const int Se = 2 * S + 1;
  const int* tgt = targets + b * tgt_ld;
  for (int s = threadIdx.x; s < Se; s += blockDim.x) s_occ[s] = 0.f;
  __syncthreads();
  const long long base = (static_cast<long long>(b) * T + t) * Sp;
  for (int s = threadIdx.x; s < Se; s += blockDim.x) {
    const float g = ws.alpha[base + s] + ws.beta[base + s] - ws.lp_g[base + s] + nr;
    const float occ = (g == kNegInf || !(g == g)) ? 0.f : __expf(g);
    if (occ != 0.f) atomicAdd(&s_occ[ws.leader[b * Sp + s]], occ);
  }
  __syncthreads();
  const long long row = static_cast<long long>(b) * T + t;
  for (int s = threadIdx.x; s < Se; s += blockDim.x) {
    if (ws.leader[b * Sp + s] != s) continue;
    const float occ = s_occ[s];
    if (occ == 0.f) continue;
    int l = ext_label(tgt, s, blank);
    l = min(max(l, 0), V - 1);
    const long long idx = row * ldv + l;
    // lp_g of the leader state = logit - lse of this label: recompute softmax - occupancy in fp32 and round ONCE
    // (a read-modify-write of the already rounded scale * softmax loses the difference where p and occ nearly cancel,
    // i.e. on the label frames of a trained model)
    const float pr = __expf(ws.lp_g[base + s]);
    if (is_bf16) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(dlogits);
      o[idx] = __float2bfloat16(scale * (pr - occ));
    } else {
      float* o = reinterpret_cast<float*>(dlogits);
      o[idx] = scale * (pr - occ);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// greedy decode
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void argmax_take(float v, int i, float& best, int& bi) {
  if (v > best || (v == best && i < bi)) {
    best = v;
    bi = i;
  }
}

// One CTA per (b,t) row: argmax over V (lowest index on ties; 16-byte loads for bf16).  Padded frames decode to
// blank (s2t_ctc.py:329-338).  With frame_lp != NULL every row (padded ones too, as the reference's topk runs before
// the padding mask, s2t_ctc.py:325-331) also gets frame_lp = argmax != blank ? max_logit - logsumexp : 0.
__global__ void __launch_bounds__(256)
ctc_argmax_kernel(const void* __restrict__ logits, int is_bf16, long long ldv,
                  const int* __restrict__ in_len, int T, int V, int blank,
                  int* __restrict__ frame_ids, float* __restrict__ frame_lp) {
  pdl_entry();
  __shared__ float s_v[8];
  __shared__ int s_i[8];
  __shared__ float s_best;
  const long long row = blockIdx.x;
  const int b = static_cast<int>(row / T), t = static_cast<int>(row % T);
  const bool padded = t >= in_len[b];
  if (padded && frame_lp == nullptr) {
    if (threadIdx.x == 0) frame_ids[row] = blank;
    return;
  }
  float best = kNegInf;
  int bi = 0x7fffffff;
  if (is_bf16) {
    const __nv_bfloat16* x = reinterpret_cast<const __nv_bfloat16*>(logits) + row * ldv;
    const int nv = V >> 3;
    for (int i = threadIdx.x; i < nv; i += 256) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + i * 8);
      const float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                          bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
#pragma unroll
      for (int j = 0; j < 8; ++j) argmax_take(f[j], i * 8 + j, best, bi);
    }
    for (int i = (nv << 3) + threadIdx.x; i < V; i += 256) argmax_take(__bfloat162float(x[i]), i, best, bi);
  } else {
    const float* x = reinterpret_cast<const float*>(logits) + row * ldv;
    for (int i = threadIdx.x; i < V; i += 256) argmax_take(x[i], i, best, bi);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float v2 = __shfl_xor_sync(0xffffffffu, best, o);
    const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
    argmax_take(v2, i2, best, bi);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_v[warp] = best;
    s_i[warp] = bi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) argmax_take(s_v[w], s_i[w], best, bi);
    if (bi == 0x7fffffff) bi = blank;
    frame_ids[row] = padded ? blank : bi;
    s_best = best;
    s_i[0] = bi;
  }
  if (frame_lp == nullptr) return;
  __syncthreads();
  best = s_best;
  bi = s_i[0];
  float sum = 0.f;
  for (int i = threadIdx.x; i < V; i += 256) sum += __expf(load_logit(logits, is_bf16, row * ldv + i) - best);
  sum = warp_sum(sum);
  __syncthreads();
  if (lane == 0) s_v[warp] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int w = 0; w < 8; ++w) tot += s_v[w];
    frame_lp[row] = (bi != blank) ? -logf(tot) : 0.f;   // max_logit - lse = -log(sum exp(x - max))
  }
}

// one warp per utterance: keep[t] = id != blank && (t == 0 || id != id[t-1]); ballot prefix compaction
__global__ void ctc_collapse_kernel(const int* __restrict__ frame_ids, int T, int blank,
                                    int* __restrict__ tokens, int* __restrict__ out_len,
                                    const float* __restrict__ frame_lp, float* __restrict__ scores) {
  pdl_entry();
  const int b = blockIdx.x;
  const int lane = threadIdx.x;
  const int* ids = frame_ids + static_cast<long long>(b) * T;
  int* out = tokens + static_cast<long long>(b) * T;
  int count = 0;
  for (int t0 = 0; t0 < T; t0 += 32) {
    const int t = t0 + lane;
    int id = blank, prev = -1;
    if (t < T) {
      id = ids[t];
      prev = t > 0 ? ids[t - 1] : -1;
    }
    const bool keep = (t < T) && id != blank && id != prev;
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (keep) out[count + __popc(m & ((1u << lane) - 1u))] = id;
    count += __popc(m);
  }
  for (int t = count + lane; t < T; t += 32) out[t] = -1;
  if (lane == 0) out_len[b] = count;
  if (scores != nullptr) {  // score = -sum_t [argmax_t != blank] max log-prob (s2t_ctc.py:330-332, quirk A14)
    float acc = 0.f;
    for (int t = lane; t < T; t += 32) acc += frame_lp[static_cast<long long>(b) * T + t];
    acc = warp_sum(acc);
    if (lane == 0) scores[b] = -acc;
  }
}

}  // namespace s2t

using namespace s2t;
#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int s2t_row_lse(const void* logits, int is_bf16, long long ldv, int rows, int V, float* lse,
                           void* stream) {
  S2T_REQUIRE(logits && lse, "s2t_row_lse: null pointer");
  S2T_REQUIRE(V > 0, "s2t_row_lse: V must be positive");
  S2T_REQUIRE(ldv % (is_bf16 ? 8 : 4) == 0, "s2t_row_lse: ldv alignment");
  if (rows <= 0) return 0;
  S2T_LAUNCH(row_lse_kernel, rows, 256, 0, ST(stream), logits, is_bf16, ldv, V, lse);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" long long s2t_ctc_workspace_bytes(int batch, int T, int max_target_len) {
  const int Sp = (2 * max_target_len + 1 + 7) & ~7;   // rows are whole 16-byte vectors
  return 4 * ctc_ws_floats(batch, T, Sp) + 256;
}

static int ctc_loss_impl(const void* logits, int is_bf16, long long ldv, const float* lse,
                         const int* targets, long long tgt_ld, const int* input_lengths,
                         const int* target_lengths, int batch, int T, int V, int max_target_len,
                         int blank, float* nll, void* dlogits, float grad_scale,
                         const float* grad_scale_dev, void* workspace, long long workspace_bytes,
                         void* stream, const int* force_emit, long long fe_ld) {
  S2T_REQUIRE(logits && lse && targets && input_lengths && target_lengths && workspace,
              "s2t_ctc_loss: null pointer");
  S2T_REQUIRE(nll != nullptr, "s2t_ctc_loss: nll output required");
  S2T_REQUIRE(max_target_len >= 0 && tgt_ld >= max_target_len, "s2t_ctc_loss: bad target shape");
  S2T_REQUIRE(blank >= 0 && blank < V, "s2t_ctc_loss: blank out of range");
  S2T_REQUIRE(workspace_bytes >= s2t_ctc_workspace_bytes(batch, T, max_target_len),
              "s2t_ctc_loss: workspace too small");
  S2T_REQUIRE(ldv % (is_bf16 ? 8 : 4) == 0, "s2t_ctc_loss: ldv alignment");
  if (batch <= 0 || T <= 0) return 0;
  const int Sp = (2 * max_target_len + 1 + 7) & ~7;   // rows are whole 16-byte vectors
  int threads = ((Sp + 31) / 32) * 32;
  if (threads > 1024) threads = 1024;
  if (threads < 64) threads = 64;
  S2T_REQUIRE(Sp <= kCtcMaxPerThread * threads, "s2t_ctc_loss: target too long (2S+1=%d > %d)", Sp,
              kCtcMaxPerThread * threads);
  S2T_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "s2t_ctc_loss: workspace must be 16-byte aligned");
  CtcWs ws;
  float* base = reinterpret_cast<float*>(workspace);
  const long long n = 1ll * batch * T * Sp;
  ws.lp_g = base;
  ws.alpha = base + n;
  ws.beta = base + 2 * n;
  ws.leader = reinterpret_cast<int*>(base + 3 * n);
  ws.nll_raw = base + 3 * n + 1ll * batch * Sp;
  cudaStream_t st = ST(stream);

  // With a gradient wanted, scale * softmax (one pass over the [B*T, V] logits, every SM) runs on an internal side
  // stream beside gather + lattice (a handful of warps, ~T dependent steps); the occupancy pass joins the two.
  // Event fork / join: also legal while `st` is being captured into a CUDA graph.
  // (one side stream + event pair per device, created on first use under a lock)
  struct Side { cudaStream_t side = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr; };
  static Side sides[64];
  static std::mutex side_mu;
  const bool overlap = dlogits != nullptr;
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  if (overlap) {
    int dev = 0;
    S2T_CHECK_CUDA(cudaGetDevice(&dev));
    S2T_REQUIRE(dev >= 0 && dev < 64, "s2t_ctc_loss: device index %d out of range", dev);
    {
      std::lock_guard<std::mutex> lk(side_mu);
      Side& sd = sides[dev];
      if (sd.side == nullptr) {
        S2T_CHECK_CUDA(cudaStreamCreateWithFlags(&sd.side, cudaStreamNonBlocking));
        S2T_CHECK_CUDA(cudaEventCreateWithFlags(&sd.ev_fork, cudaEventDisableTiming));
        S2T_CHECK_CUDA(cudaEventCreateWithFlags(&sd.ev_join, cudaEventDisableTiming));
      }
      side = sd.side; ev_fork = sd.ev_fork; ev_join = sd.ev_join;
    }
    S2T_CHECK_CUDA(cudaEventRecord(ev_fork, st));
    S2T_CHECK_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
    S2T_LAUNCH(ctc_dense_grad_kernel, batch * T, 256, 0, side, logits, is_bf16, ldv, lse, input_lengths, T, V,
               static_cast<const float*>(nullptr), static_cast<float*>(nullptr), dlogits, grad_scale, grad_scale_dev);
    S2T_LAUNCH_CHECK();
    S2T_CHECK_CUDA(cudaEventRecord(ev_join, side));
  }

  {
    int gx = static_cast<int>((1ll * T * Sp + 255) / 256);
    if (gx > 64) gx = 64;
    if (gx < 1) gx = 1;
    S2T_LAUNCH(ctc_gather_kernel, dim3(gx, batch), 256, 0, st, logits, is_bf16, ldv, lse, targets, tgt_ld, input_lengths, target_lengths, T, Sp, V, blank, ws);
    S2T_LAUNCH_CHECK();
  }
  {
    const size_t smem = sizeof(float) * 2 * (Sp + 4);
    const dim3 g(batch, 2);
    // shared-memory staging of the emission rows: whole utterance when T*Sp*4 <= 200 KB, else two 100 KB time chunks
    const size_t row_bytes = static_cast<size_t>(Sp) * 4;
    int Tc = T;
    size_t lat_smem = static_cast<size_t>(T) * row_bytes;
    if (lat_smem > 200 * 1024) {
      Tc = static_cast<int>((100 * 1024) / row_bytes);
      lat_smem = 2 * static_cast<size_t>(Tc) * row_bytes;
    }
    if (force_emit != nullptr) {
      // masked lattice: the block kernel (states in shared memory), every frame checks its forced state
      S2T_LAUNCH(ctc_alpha_beta_kernel<true>, g, threads, smem, st, targets, tgt_ld, input_lengths, target_lengths, T, Sp, blank, ws,
                 force_emit, fe_ld);
    } else
#define S2T_CTC_WARP(NS)                                                                                          \
  do {                                                                                                            \
    S2T_CHECK_CUDA(cudaFuncSetAttribute(ctc_alpha_beta_warp_kernel<NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                        210 * 1024));                                                             \
    S2T_LAUNCH(ctc_alpha_beta_warp_kernel<NS>, g, 32, lat_smem, st, targets, tgt_ld, input_lengths, target_lengths, T, \
               Sp, Tc, ws);                                                                                       \
  } while (0)
    if (Sp <= 32)
      S2T_CTC_WARP(1);
    else if (Sp <= 64)
      S2T_CTC_WARP(2);
    else if (Sp <= 128)
      S2T_CTC_WARP(4);
    else if (Sp <= 256)
      S2T_CTC_WARP(8);
#undef S2T_CTC_WARP
    else
      S2T_LAUNCH(ctc_alpha_beta_kernel<false>, g, threads, smem, st, targets, tgt_ld, input_lengths, target_lengths, T, Sp, blank, ws,
                 static_cast<const int*>(nullptr), 0ll);
    S2T_LAUNCH_CHECK();
  }
  if (!overlap) {   // loss only: publish nll (zero_infinity applied)
    S2T_LAUNCH(ctc_dense_grad_kernel, batch * T, 256, 0, st, logits, is_bf16, ldv, lse, input_lengths, T, V, ws.nll_raw, nll, dlogits, grad_scale, grad_scale_dev);
    S2T_LAUNCH_CHECK();
  } else {
    S2T_CHECK_CUDA(cudaStreamWaitEvent(st, ev_join, 0));
    int othreads = ((Sp + 31) / 32) * 32;
    if (othreads > 256) othreads = 256;
    S2T_LAUNCH(ctc_occupancy_kernel, dim3(T, batch), othreads, sizeof(float) * Sp, st, targets, tgt_ld, input_lengths, target_lengths, T, Sp, V, blank, ws, is_bf16, ldv, dlogits, grad_scale, grad_scale_dev, nll);
    S2T_LAUNCH_CHECK();
  }
  return 0;
}

extern "C" int s2t_ctc_loss(const void* logits, int is_bf16, long long ldv, const float* lse,
                            const int* targets, long long tgt_ld, const int* input_lengths,
                            const int* target_lengths, int batch, int T, int V, int max_target_len,
                            int blank, float* nll, void* dlogits, float grad_scale,
                            const float* grad_scale_dev, void* workspace, long long workspace_bytes,
                            void* stream) {
  return ctc_loss_impl(logits, is_bf16, ldv, lse, targets, tgt_ld, input_lengths, target_lengths, batch, T, V,
                       max_target_len, blank, nll, dlogits, grad_scale, grad_scale_dev, workspace, workspace_bytes, stream,
                       nullptr, 0);
}

extern "C" int s2t_ctc_loss_masked(const void* logits, int is_bf16, long long ldv, const float* lse,
                                   const int* targets, long long tgt_ld, const int* force_emit, long long fe_ld,
                                   const int* input_lengths, const int* target_lengths, int batch, int T, int V,
                                   int max_target_len, int blank, float* nll, void* dlogits, float grad_scale,
                                   const float* grad_scale_dev, void* workspace, long long workspace_bytes,
                                   void* stream) {
  S2T_REQUIRE(force_emit != nullptr && fe_ld >= T, "s2t_ctc_loss_masked: force_emit [B][fe_ld >= T] required");
  return ctc_loss_impl(logits, is_bf16, ldv, lse, targets, tgt_ld, input_lengths, target_lengths, batch, T, V,
                       max_target_len, blank, nll, dlogits, grad_scale, grad_scale_dev, workspace, workspace_bytes, stream,
                       force_emit, fe_ld);
}

extern "C" int s2t_ctc_greedy(const void* logits, int is_bf16, long long ldv, const int* input_lengths,
                              int batch, int T, int V, int blank, int* tokens, int* out_lengths,
                              int* frame_ids, float* scores, float* frame_lp, void* stream) {
  S2T_REQUIRE(logits && input_lengths && tokens && out_lengths && frame_ids,
              "s2t_ctc_greedy: null pointer (frame_ids is a required [B][T] scratch/output)");
  S2T_REQUIRE((scores == nullptr) == (frame_lp == nullptr),
              "s2t_ctc_greedy: scores [B] and frame_lp [B][T] go together");
  S2T_REQUIRE(ldv % (is_bf16 ? 8 : 1) == 0, "s2t_ctc_greedy: bf16 rows must be 16-byte aligned (ldv % 8 == 0)");
  if (batch <= 0 || T <= 0) return 0;
  S2T_LAUNCH(ctc_argmax_kernel, batch * T, 256, 0, ST(stream), logits, is_bf16, ldv, input_lengths, T, V, blank, frame_ids, frame_lp);
  S2T_LAUNCH_CHECK();
  S2T_LAUNCH(ctc_collapse_kernel, batch, 32, 0, ST(stream), frame_ids, T, blank, tokens, out_lengths, frame_lp, scores);
  S2T_LAUNCH_CHECK();
  return 0;
}
