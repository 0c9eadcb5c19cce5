// CTC best alignment (Viterbi) on device, including the backtrack the reference does on the host.
// Reference: fairseq/torch_imputer/best_alignment.cu:62-201 (max-recursion + argmax paths, one CUDA thread per lattice
// state, one __syncthreads per frame) and fairseq/torch_imputer/imputer.py:245-259,282-325 (log_alpha and paths are
// copied to the host and every utterance is backtracked in a Python loop).
//
// Here: one CTA per utterance, thread s owns lattice state s.  alpha lives in two shared-memory rows, the back-pointers
// (0, 1 or 2 states back) as bytes in shared memory ([T][Sp], up to ~200 KB; global scratch beyond that), the emission
// log-probabilities of the next 8 frames are fetched while the current 8 are consumed (a gather of one float per state
// and frame from the [T][V] matrix: the only HBM traffic), and thread 0 walks the back-pointers at the end, so only the
// T state indices per utterance ever leave the SM.  Same arithmetic as the reference (one fp32 add per state and frame,
// strict '>' in the order s, s-1, s-2), so the state sequence is bit-exact against the CPU restatement.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {
namespace {

constexpr int kVitAhead = 8;   // frames of emissions in flight per thread

__global__ void __launch_bounds__(1024)
ctc_viterbi_kernel(const float* __restrict__ lp, long long lp_ts, long long lp_bs, const int* __restrict__ targets,
                   long long tgt_ld, const int* __restrict__ in_len, const int* __restrict__ tgt_len, int T, int Sp,
                   int blank, uint8_t* __restrict__ bp_global, int* __restrict__ states, float* __restrict__ nll) {
  pdl_entry();
  extern __shared__ float s_vit[];               // alpha[2][Sp], then (optionally) back-pointers [T][Sp] bytes
  float* alpha0 = s_vit;
  float* alpha1 = s_vit + Sp;
  const int b = blockIdx.x, s = threadIdx.x;
  uint8_t* bp = bp_global != nullptr ? bp_global + static_cast<long long>(b) * T * Sp
                                     : reinterpret_cast<uint8_t*>(s_vit + 2 * Sp);
  const int S = tgt_len[b];
  const int Se = 2 * S + 1;
  int Tn = in_len[b];
  Tn = Tn < T ? Tn : T;
  int* out = states + static_cast<long long>(b) * T;
  for (int t = (Tn > 0 ? Tn : 0) + s; t < T; t += blockDim.x) out[t] = -1;      // padding frames
  if (Tn <= 0) {
    if (s == 0 && nll != nullptr) nll[b] = INFINITY;
    return;
  }
  const int* tgt = targets + b * tgt_ld;
  const bool live = s < Se;
  const int lab = (live && (s & 1)) ? tgt[(s - 1) >> 1] : blank;
  const bool three = live && s > 1 && (s & 1) && tgt[(s - 3) >> 1] != lab;      // l_{s-2} != l_s (both blank when s is even)
  const float* col = lp + b * lp_bs + lab;
  // t = 0
  float a0 = -INFINITY;
  if (s == 0) a0 = col[0];
  else if (s == 1 && S > 0) a0 = col[0];
  if (s < Sp) alpha0[s] = a0;
  float* prev = alpha0;
  float* cur = alpha1;
  float em[kVitAhead], nx[kVitAhead];
#pragma unroll
  for (int j = 0; j < kVitAhead; ++j) nx[j] = (live && 1 + j < Tn) ? col[(1 + j) * lp_ts] : 0.f;
  for (int t0 = 1; t0 < Tn; t0 += kVitAhead) {
#pragma unroll
    for (int j = 0; j < kVitAhead; ++j) em[j] = nx[j];
#pragma unroll
    for (int j = 0; j < kVitAhead; ++j) {
      const int tn = t0 + kVitAhead + j;
      nx[j] = (live && tn < Tn) ? col[static_cast<long long>(tn) * lp_ts] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < kVitAhead; ++j) {
      const int t = t0 + j;
      if (t >= Tn) break;                        // uniform across the CTA
      __syncthreads();
      if (live) {
        float best = prev[s];
        int back = 0;
        if (s > 0) {
          const float a = prev[s - 1];
          if (a > best) { best = a; back = 1; }
        }
        if (three) {
          const float a = prev[s - 2];
          if (a > best) { best = a; back = 2; }
        }
        cur[s] = best + em[j];
        bp[static_cast<long long>(t) * Sp + s] = static_cast<uint8_t>(back);
      }
      float* tmp = prev; prev = cur; cur = tmp;
    }
  }
  __syncthreads();
  if (s == 0) {
    const float l1 = prev[2 * S];
    const float l2 = S > 0 ? prev[2 * S - 1] : -INFINITY;
    if (nll != nullptr) {
      float m = l1 > l2 ? l1 : l2;
      m = (m == -INFINITY) ? 0.f : m;
      nll[b] = -(logf(expf(l1 - m) + expf(l2 - m)) + m);
    }
    // numpy argmax over (alpha[2S-1], alpha[2S]): the first maximum wins
    int st = Se == 1 ? 0 : (l2 >= l1 ? 2 * S - 1 : 2 * S);
    out[Tn - 1] = st;
    for (int t = Tn - 1; t >= 1; --t) {
      st -= bp[static_cast<long long>(t) * Sp + st];
      out[t - 1] = st;
    }
  }
}

}  // namespace
}  // namespace s2t

using namespace s2t;

extern "C" long long s2t_ctc_viterbi_workspace_bytes(int batch, int T, int max_target_len) {
  const long long Sp = (2ll * max_target_len + 1 + 31) & ~31ll;
  const long long smem = 2 * Sp * 4 + static_cast<long long>(T) * Sp;
  return smem <= 200 * 1024 ? 0 : static_cast<long long>(batch) * T * Sp;
}

extern "C" int s2t_ctc_viterbi(const float* log_probs, long long lp_t_stride, long long lp_b_stride, const int* targets,
                               long long tgt_ld, const int* input_lengths, const int* target_lengths, int batch, int T,
                               int V, int max_target_len, int blank, int* states, float* nll, void* workspace,
                               long long workspace_bytes, void* stream) {
  S2T_REQUIRE(log_probs && targets && input_lengths && target_lengths && states, "s2t_ctc_viterbi: null pointer");
  S2T_REQUIRE(blank >= 0 && blank < V, "s2t_ctc_viterbi: blank out of range");
  S2T_REQUIRE(max_target_len >= 0 && tgt_ld >= max_target_len, "s2t_ctc_viterbi: bad target shape");
  if (batch <= 0 || T <= 0) return 0;
  const int Sp = (2 * max_target_len + 1 + 31) & ~31;
  S2T_REQUIRE(Sp <= 1024, "s2t_ctc_viterbi: target too long (2S+1 = %d > 1024 lattice states)", 2 * max_target_len + 1);
  const long long need = s2t_ctc_viterbi_workspace_bytes(batch, T, max_target_len);
  S2T_REQUIRE(need == 0 || (workspace != nullptr && workspace_bytes >= need), "s2t_ctc_viterbi: workspace too small");
  size_t smem = static_cast<size_t>(2) * Sp * 4 + (need == 0 ? static_cast<size_t>(T) * Sp : 0);
  static std::atomic<unsigned long long> configured{0};   // per device, see once_per_device
  S2T_CHECK_CUDA(once_per_device(configured, [&] { return cudaFuncSetAttribute(ctc_viterbi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024); }));
  S2T_LAUNCH(ctc_viterbi_kernel, batch, Sp, smem, reinterpret_cast<cudaStream_t>(stream), log_probs, lp_t_stride,
             lp_b_stride, targets, tgt_ld, input_lengths, target_lengths, T, Sp, blank,
             need == 0 ? static_cast<uint8_t*>(nullptr) : reinterpret_cast<uint8_t*>(workspace), states, nll);
  S2T_LAUNCH_CHECK();
  return 0;
}
