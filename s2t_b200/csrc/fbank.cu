// K1: Kaldi-compatible 80-dim log-mel filterbank + CMVN on sm_100a.
//
// Replaces torchaudio.compliance.kaldi.fbank as called by the reference at
// fairseq/data/audio/audio_utils.py:59-79 (num_mel_bins=80, sample_frequency=16000, everything else
// default: 25 ms povey window, 10 ms shift, snip_edges, remove_dc_offset, preemphasis 0.97, power
// spectrum, log floor = fp32 eps, dither 0) and UtteranceCMVN / GlobalCMVN
// (fairseq/data/audio/feature_transforms/utterance_cmvn.py:30-45, global_cmvn.py:26-29).
//
// One warp owns one frame: 400 samples -> DC removal -> pre-emphasis -> window -> 512-point radix-2
// FFT in shared memory (warp-synchronous, no block barriers) -> |X|^2 -> sparse mel projection (each
// filter only touches its [lo,hi) bins) -> log.  Loads are coalesced (lane-strided), the waveform is
// read once from HBM (frame overlap is served by L1/L2) and features are written once.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kFrameLen = 400;
constexpr int kFrameShift = 160;
constexpr int kFft = 512;
constexpr int kBins = kFft / 2 + 1;  // 257
constexpr int kFbankWarps = 8;

__device__ __forceinline__ int bitrev9(int i) { return __brev(static_cast<unsigned>(i)) >> 23; }

__global__ void __launch_bounds__(kFbankWarps * 32)
fbank_kernel(const float* __restrict__ wave, long long wave_ld, const int* __restrict__ num_samples,
             const float* __restrict__ mel_w, const int* __restrict__ mel_range,
             const float* __restrict__ window, int n_mels, void* __restrict__ out, int out_bf16,
             int max_frames, long long out_ld, long long out_bs, int* __restrict__ out_frames) {
  pdl_entry();
  // per warp: 512 complex values in the padded transpose layout (16 * L + (L >> 1) + r, see below); reused for the
  // 257 power-spectrum bins once the FFT is back in registers
  __shared__ float2 s_fft[kFbankWarps][kFft + 16];
  // twiddles exp(-2*pi*i*k/512) laid out per FFT stage: entry (half - 1) + pos = k = pos << (8 - stage), so that the
  // lanes of a stage read consecutive words (the strided k of a single table is a 16-way bank conflict mid-FFT)
  __shared__ float2 s_tw[kFft];
  // non-zero mel weights packed filter after filter (each FFT bin belongs to at most two triangular filters)
  constexpr int kMelPacked = 2 * kBins + 32;
  __shared__ float s_mw[kMelPacked];
  __shared__ int s_moff[129], s_mlo[128], s_mhi[128];

  const int b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int frame = blockIdx.x * kFbankWarps + warp;
  const int ns = num_samples[b];
  const int nframes = ns >= kFrameLen ? 1 + (ns - kFrameLen) / kFrameShift : 0;
  if (blockIdx.x == 0 && threadIdx.x == 0 && out_frames != nullptr)
    out_frames[b] = nframes < max_frames ? nframes : max_frames;

  for (int e = threadIdx.x; e < kFft - 1; e += blockDim.x) {
    const int st = 31 - __clz(e + 1);                 // stage: half = 1 << st
    const int k = (e + 1 - (1 << st)) << (8 - st);
    float s, c;
    sincospif(-2.0f * static_cast<float>(k) / static_cast<float>(kFft), &s, &c);
    s_tw[e] = make_float2(c, s);
  }
  if (threadIdx.x < 128) {
    const int f = threadIdx.x;
    const int lo = f < n_mels ? mel_range[2 * f] : 0, hi = f < n_mels ? mel_range[2 * f + 1] : 0;
    s_mlo[f] = lo;
    s_mhi[f] = hi > lo ? hi : lo;
  }
  __syncthreads();
  if (warp == 0) {   // exclusive scan of the filter widths, four consecutive filters per lane
    int cnt[4], tot = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      cnt[j] = s_mhi[lane * 4 + j] - s_mlo[lane * 4 + j];
      tot += cnt[j];
    }
    int incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int base = incl - tot;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      s_moff[lane * 4 + j] = base;
      base += cnt[j];
    }
    if (lane == 31) s_moff[128] = incl;
  }
  __syncthreads();
  const bool mel_packed = s_moff[128] <= kMelPacked;
  if (mel_packed && threadIdx.x < n_mels) {
    const int f = threadIdx.x, lo = s_mlo[f], n = s_mhi[f] - lo;
    const float* wrow = mel_w + static_cast<long long>(f) * kBins + lo;
    float* dst = s_mw + s_moff[f];
    for (int k = 0; k < n; ++k) dst[k] = __ldg(wrow + k);
  }
  __syncthreads();
  if (frame >= max_frames) return;

  float* orow_f = reinterpret_cast<float*>(out) + b * out_bs + static_cast<long long>(frame) * out_ld;
  __nv_bfloat16* orow_h =
      reinterpret_cast<__nv_bfloat16*>(out) + b * out_bs + static_cast<long long>(frame) * out_ld;
  if (frame >= nframes) {  // zero-fill padding frames (collater zero-pads, speech_to_text_dataset.py:267-285)
    for (int f = lane; f < n_mels; f += 32) {
      if (out_bf16) orow_h[f] = __float2bfloat16(0.f);
      else orow_f[f] = 0.f;
    }
    return;
  }

  // ---- load the frame (coalesced), remove DC offset --------------------------------------------
  const float* src = wave + b * wave_ld + static_cast<long long>(frame) * kFrameShift;
  float x[13];  // 400 = 12*32 + 16
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 13; ++i) {
    const int idx = i * 32 + lane;
    x[i] = idx < kFrameLen ? __ldg(src + idx) : 0.f;
    sum += x[i];
  }
  const float mean = warp_sum(sum) / static_cast<float>(kFrameLen);
  // ---- pre-emphasis (replicate-padded first sample), povey window ------------------------------------------
  // 512-point radix-2 DIT FFT held in registers.  Position p of the bit-reversed input order holds sample
  // bitrev9(p); with p = 16 * L + r the samples of "virtual lane" L are 32 * rev4(r) + rev5(L): exactly the
  // samples 32 * i + lane this lane has loaded, with L = rev5(lane), r = rev4(i).  Stages 0-3 (pairs inside one
  // group of 16 positions) therefore need no data movement at all; one transpose through shared memory regroups
  // to i = 256 * b8 + 16 * m + low4 (lane = 16 * b8 + low4, register m) for stages 4-7; stage 8 pairs lanes l, l ^ 16.
  float2 y[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) y[i] = make_float2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < 13; ++i) {
    const int idx = i * 32 + lane;
    const float cur = x[i] - mean;
    // previous sample lives in lane-1 of the same i, or lane 31 of i-1
    float prev = __shfl_up_sync(0xffffffffu, cur, 1);
    const float prev_row = __shfl_sync(0xffffffffu, (i > 0 ? x[i > 0 ? i - 1 : 0] - mean : 0.f), 31);
    if (lane == 0) prev = (i == 0) ? cur : prev_row;
    float v = 0.f;
    if (idx < kFrameLen) v = (cur - 0.97f * prev) * __ldg(window + idx);
    // rev4(i) as a compile-time constant after unrolling
    y[((i & 1) << 3) | ((i & 2) << 1) | ((i & 4) >> 1) | ((i & 8) >> 3)].x = v;
  }
  // ---- stages 0-3: registers only; twiddles are lane-uniform (broadcast reads) ------------------
#pragma unroll
  for (int st = 0; st < 4; ++st) {
    const int half = 1 << st;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int pos = j & (half - 1);
      const int i0 = ((j >> st) << (st + 1)) + pos;
      const int i1 = i0 + half;
      const float2 w = s_tw[half - 1 + pos];
      const float2 a = y[i0], c = y[i1];
      const float tr = c.x * w.x - c.y * w.y;
      const float ti = c.x * w.y + c.y * w.x;
      y[i0] = make_float2(a.x + tr, a.y + ti);
      y[i1] = make_float2(a.x - tr, a.y - ti);
    }
  }
  // ---- transpose: position p = 16 * L + r lives at word p + (L >> 1) (conflict-free 64-bit accesses both ways) ----
  float2* fft = s_fft[warp];
  {
    const int L = static_cast<int>(__brev(static_cast<unsigned>(lane)) >> 27);
    float2* dst = fft + 16 * L + (L >> 1);
#pragma unroll
    for (int r = 0; r < 16; ++r) dst[r] = y[r];
  }
  __syncwarp();
  const int b8 = lane >> 4, low4 = lane & 15;
#pragma unroll
  for (int m = 0; m < 16; ++m) {
    const int Li = b8 * 16 + m;
    y[m] = fft[16 * Li + (Li >> 1) + low4];
  }
  __syncwarp();   // the buffer is free again (power spectrum below)
  // ---- stages 4-7: registers; the twiddle depends on low4 and the low bits of m ---------------------------
#pragma unroll
  for (int st = 4; st < 8; ++st) {
    const int half = 1 << st, sub = 1 << (st - 4);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int mlow = j & (sub - 1);
      const int m0 = ((j >> (st - 4)) << (st - 3)) + mlow;
      const int m1 = m0 + sub;
      const float2 w = s_tw[half - 1 + mlow * 16 + low4];
      const float2 a = y[m0], c = y[m1];
      const float tr = c.x * w.x - c.y * w.y;
      const float ti = c.x * w.y + c.y * w.x;
      y[m0] = make_float2(a.x + tr, a.y + ti);
      y[m1] = make_float2(a.x - tr, a.y - ti);
    }
  }
  // ---- stage 8 + power spectrum: bins 0..255 are the sums (lanes with b8 = 0), bin 256 the one difference at
  // position 0; torchaudio does rfft().abs().pow(2) ---------------------------------------------------------
  float* pw = reinterpret_cast<float*>(fft);
#pragma unroll
  for (int m = 0; m < 16; ++m) {
    const float cx = __shfl_down_sync(0xffffffffu, y[m].x, 16);
    const float cy = __shfl_down_sync(0xffffffffu, y[m].y, 16);
    if (b8 == 0) {
      const float2 w = s_tw[255 + m * 16 + low4];
      const float tr = cx * w.x - cy * w.y;
      const float ti = cx * w.y + cy * w.x;
      const float vx = y[m].x + tr, vy = y[m].y + ti;
      const float a = sqrtf(vx * vx + vy * vy);
      pw[m * 16 + low4] = a * a;
      if (m == 0 && lane == 0) {
        const float ux = y[m].x - tr, uy = y[m].y - ti;
        const float a2 = sqrtf(ux * ux + uy * uy);
        pw[256] = a2 * a2;
      }
    }
  }
  __syncwarp();
  // ---- mel projection + log ---------------------------------------------------------------------
  for (int f = lane; f < n_mels; f += 32) {
    const int lo = s_mlo[f], hi = s_mhi[f];
    float acc = 0.f;
    if (mel_packed) {
      const float* wp = s_mw + s_moff[f] - lo;
      for (int k = lo; k < hi; ++k) acc = fmaf(pw[k], wp[k], acc);
    } else {
      const float* wrow = mel_w + static_cast<long long>(f) * kBins;
      for (int k = lo; k < hi; ++k) acc = fmaf(pw[k], __ldg(wrow + k), acc);
    }
    const float v = logf(fmaxf(acc, 1.1920928955078125e-07f));
    if (out_bf16) orow_h[f] = __float2bfloat16(v);
    else orow_f[f] = v;
  }
}

// Per-utterance CMVN.  The reference (utterance_cmvn.py:35-43) computes E[x^2]-E[x]^2 in fp32 numpy:
// an axis-0 reduction of a C-contiguous (T, 80) array accumulates row after row, i.e. per bin a strictly
// SEQUENTIAL fp32 sum over frames, and the catastrophic cancellation of that formula is part of the
// oracle.  So: frames are staged chunk-wise in shared memory by the whole CTA (coalesced),
// and one thread per bin replays the exact sequential fp32 accumulation (no FMA contraction).
constexpr int kCmvnChunk = 128;

// stats[b][0][f] = mean (0 when !norm_means), stats[b][1][f] = std (1 when !norm_vars).  One CTA per utterance; the
// next chunk is prefetched into registers while the current one is accumulated from shared memory.  Threads 0..127
// carry the per-bin sum chain, threads 128..255 the per-bin sum-of-squares chain (each one dependent fp32 add per
// frame, in frame order: numpy's axis-0 reduction order, so the result is bit-exact); shared loads are issued eight
// frames ahead of the adds.
__global__ void __launch_bounds__(256)
cmvn_utt_stats_kernel(const float* __restrict__ in, long long in_ld, long long in_bs,
                      const int* __restrict__ num_frames, int max_frames, int n_mels, int norm_means,
                      int norm_vars, float* __restrict__ stats) {
  pdl_entry();
  extern __shared__ float s_chunk[];  // [2][kCmvnChunk][n_mels]
  __shared__ float s_sq[128];
  constexpr int kPer = (kCmvnChunk * 128 + 255) / 256;  // n_mels <= 128
  const int b = blockIdx.x;
  int T = num_frames[b];
  if (T > max_frames) T = max_frames;
  const float* src = in + b * in_bs;
  const int per = (kCmvnChunk * n_mels + 255) / 256;
  float reg[kPer];
  auto fetch = [&](int t0) {
    const int nt = min(kCmvnChunk, T - t0);
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = threadIdx.x + k * 256;
      if (k < per && i < nt * n_mels) {
        const int t = i / n_mels, f = i - t * n_mels;
        reg[k] = src[static_cast<long long>(t0 + t) * in_ld + f];
      }
    }
  };
  auto stash = [&](float* buf, int t0) {
    const int nt = min(kCmvnChunk, T - t0);
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = threadIdx.x + k * 256;
      if (k < per && i < nt * n_mels) buf[i] = reg[k];
    }
  };
  const int f = threadIdx.x & 127;
  const bool squares = threadIdx.x >= 128;
  float acc = 0.f;
  // contiguous, 16-byte aligned rows: chunks arrive by bulk async copy (one instruction per chunk, completion on an
  // mbarrier), so the accumulation loop waits on shared memory only; otherwise they are staged through registers.
  __shared__ __align__(8) uint64_t s_bar[2];
  const bool bulk = in_ld == n_mels && (n_mels & 3) == 0 && (in_bs & 3) == 0 &&
                    (reinterpret_cast<uintptr_t>(in) & 15) == 0;
  auto issue = [&](int chunk) {   // thread 0
    const int t0 = chunk * kCmvnChunk;
    const uint32_t bytes = static_cast<uint32_t>(min(kCmvnChunk, T - t0)) * n_mels * 4u;
    mbar_expect_tx(&s_bar[chunk & 1], bytes);
    bulk_copy_g2s(s_chunk + (chunk & 1) * kCmvnChunk * n_mels, src + static_cast<long long>(t0) * in_ld, bytes,
                  &s_bar[chunk & 1]);
  };
  if (bulk) {
    if (threadIdx.x == 0) {
      mbar_init(&s_bar[0], 1);
      mbar_init(&s_bar[1], 1);
      fence_mbar_init();
      if (T > 0) issue(0);
      if (T > kCmvnChunk) issue(1);
    }
  } else if (T > 0) {
    fetch(0);
    stash(s_chunk, 0);
  }
  __syncthreads();
  int cur = 0;
  for (int t0 = 0; t0 < T; t0 += kCmvnChunk) {
    const int nt = min(kCmvnChunk, T - t0);
    const int chunk = t0 / kCmvnChunk;
    const bool more = t0 + kCmvnChunk < T;
    if (bulk) mbar_wait(&s_bar[chunk & 1], (chunk >> 1) & 1);
    else if (more) fetch(t0 + kCmvnChunk);
    const float* col = s_chunk + cur * kCmvnChunk * n_mels + f;
    if (f < n_mels) {
      int t = 0;
      for (; t + 8 <= nt; t += 8) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = col[(t + j) * n_mels];
        if (squares) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc = __fadd_rn(acc, __fmul_rn(v[j], v[j]));
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc = __fadd_rn(acc, v[j]);
        }
      }
      for (; t < nt; ++t) {
        const float v = col[t * n_mels];
        acc = squares ? __fadd_rn(acc, __fmul_rn(v, v)) : __fadd_rn(acc, v);
      }
    }
    if (!bulk && more) stash(s_chunk + (cur ^ 1) * kCmvnChunk * n_mels, t0 + kCmvnChunk);
    __syncthreads();
    if (bulk && threadIdx.x == 0 && t0 + 2 * kCmvnChunk < T) {   // buffer `cur` has been consumed by every thread
      fence_proxy_async();
      issue(chunk + 2);
    }
    cur ^= 1;
  }
  if (squares) s_sq[f] = acc;
  __syncthreads();
  if (threadIdx.x < n_mels) {
    const float sum = acc, sq = s_sq[f];
    const float Tn = static_cast<float>(T > 0 ? T : 1);
    const float m = __fdiv_rn(sum, Tn);
    float sd = 1.f;
    if (norm_vars) {
      const float var = __fsub_rn(__fdiv_rn(sq, Tn), __fmul_rn(m, m));
      sd = __fsqrt_rn(fmaxf(var, 1e-10f));
    }
    stats[(static_cast<long long>(b) * 2 + 0) * n_mels + threadIdx.x] = norm_means ? m : 0.f;
    stats[(static_cast<long long>(b) * 2 + 1) * n_mels + threadIdx.x] = sd;
  }
}

// out[b][t][f] = t < T_b ? (in - mean) / std : 0 over the whole [max_frames][n_mels] plane (grid.y = utterance);
// stats == NULL: plain cast / copy with zero fill.
__global__ void __launch_bounds__(256)
cmvn_apply_kernel(const float* __restrict__ in, long long in_ld, long long in_bs, void* __restrict__ out,
                  int out_bf16, long long out_ld, long long out_bs, const int* __restrict__ num_frames,
                  int max_frames, int n_mels, int norm_vars, const float* __restrict__ stats) {
  pdl_entry();
  const int b = blockIdx.y;
  int T = num_frames[b];
  if (T > max_frames) T = max_frames;
  const long long total = static_cast<long long>(max_frames) * n_mels;
  const float* src = in + b * in_bs;
  const float* mean = stats ? stats + static_cast<long long>(b) * 2 * n_mels : nullptr;
  const float* sd = stats ? mean + n_mels : nullptr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / n_mels), f = static_cast<int>(i - static_cast<long long>(t) * n_mels);
    float v = 0.f;
    if (t < T) {
      v = src[static_cast<long long>(t) * in_ld + f];
      if (stats != nullptr) {
        v = __fsub_rn(v, mean[f]);
        if (norm_vars) v = __fdiv_rn(v, sd[f]);
      }
    }
    const long long o = b * out_bs + static_cast<long long>(t) * out_ld + f;
    if (out_bf16) reinterpret_cast<__nv_bfloat16*>(out)[o] = __float2bfloat16(v);
    else reinterpret_cast<float*>(out)[o] = v;
  }
}

__global__ void cmvn_global_kernel(const float* __restrict__ in, long long in_ld, long long in_bs,
                                   void* __restrict__ out, int out_bf16, long long out_ld,
                                   long long out_bs, const int* __restrict__ num_frames,
                                   int max_frames, int n_mels, const float* __restrict__ mean,
                                   const float* __restrict__ std_) {
  pdl_entry();
  const int b = blockIdx.y;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(max_frames) * n_mels) return;
  const int t = static_cast<int>(i / n_mels), f = static_cast<int>(i % n_mels);
  int T = num_frames[b];
  float v = 0.f;
  if (t < T) v = (in[b * in_bs + static_cast<long long>(t) * in_ld + f] - mean[f]) / std_[f];
  const long long o = b * out_bs + static_cast<long long>(t) * out_ld + f;
  if (out_bf16) reinterpret_cast<__nv_bfloat16*>(out)[o] = __float2bfloat16(v);
  else reinterpret_cast<float*>(out)[o] = v;
}

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_fbank_num_frames(long long n) {
  return n >= kFrameLen ? static_cast<int>(1 + (n - kFrameLen) / kFrameShift) : 0;
}

extern "C" int s2t_fbank_fwd(const float* wave, long long wave_ld, const int* num_samples, int batch,
                             const float* mel_w, const int* mel_range, const float* window,
                             int n_mels, void* out, int out_bf16, int max_frames, long long out_ld,
                             long long out_bs, int* out_frames, void* stream) {
  S2T_REQUIRE(wave && num_samples && mel_w && mel_range && window && out, "s2t_fbank_fwd: null pointer");
  S2T_REQUIRE(n_mels > 0 && n_mels <= 128, "s2t_fbank_fwd: n_mels=%d out of range", n_mels);
  if (batch <= 0 || max_frames <= 0) return 0;
  dim3 grid((max_frames + kFbankWarps - 1) / kFbankWarps, batch);
  S2T_LAUNCH(fbank_kernel, grid, kFbankWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream), wave, wave_ld, num_samples, mel_w, mel_range, window, n_mels, out, out_bf16, max_frames, out_ld, out_bs, out_frames);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_cmvn_utt(const float* feats, long long in_ld, long long in_bs, void* out,
                            int out_bf16, long long out_ld, long long out_bs, const int* num_frames,
                            int batch, int max_frames, int n_mels, int norm_means, int norm_vars,
                            float* stats, void* stream) {
  S2T_REQUIRE(feats && out && num_frames, "s2t_cmvn_utt: null pointer");
  S2T_REQUIRE(n_mels > 0 && n_mels <= 128, "s2t_cmvn_utt: n_mels=%d out of range", n_mels);
  const bool norm = norm_means || norm_vars;
  S2T_REQUIRE(!norm || stats, "s2t_cmvn_utt: stats workspace (fp32 [batch][2][n_mels]) is required when normalising");
  if (batch <= 0 || max_frames <= 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (norm) {
    const size_t smem = sizeof(float) * 2 * kCmvnChunk * n_mels;
    S2T_CHECK_CUDA(cudaFuncSetAttribute(cmvn_utt_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
    S2T_LAUNCH(cmvn_utt_stats_kernel, batch, 256, smem, st, feats, in_ld, in_bs, num_frames, max_frames, n_mels, norm_means, norm_vars, stats);
    S2T_LAUNCH_CHECK();
  }
  const long long total = static_cast<long long>(max_frames) * n_mels;
  long long gx = (total + 256 * 4 - 1) / (256 * 4);
  if (gx > 1024) gx = 1024;
  S2T_LAUNCH(cmvn_apply_kernel, dim3(static_cast<unsigned>(gx), batch), 256, 0, st, feats, in_ld, in_bs, out, out_bf16, out_ld, out_bs, num_frames, max_frames, n_mels, norm_vars, norm ? stats : nullptr);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_cmvn_global(const float* feats, long long in_ld, long long in_bs, void* out,
                               int out_bf16, long long out_ld, long long out_bs,
                               const int* num_frames, int batch, int max_frames, int n_mels,
                               const float* mean, const float* std_, void* stream) {
  S2T_REQUIRE(feats && out && num_frames && mean && std_, "s2t_cmvn_global: null pointer");
  if (batch <= 0 || max_frames <= 0) return 0;
  const long long n = static_cast<long long>(max_frames) * n_mels;
  dim3 grid(static_cast<unsigned>((n + 255) / 256), batch);
  S2T_LAUNCH(cmvn_global_kernel, grid, 256, 0, reinterpret_cast<cudaStream_t>(stream), feats, in_ld, in_bs, out, out_bf16, out_ld, out_bs, num_frames, max_frames, n_mels, mean, std_);
  S2T_LAUNCH_CHECK();
  return 0;
}
