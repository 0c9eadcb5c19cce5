// SpecAugment on device (reference: fairseq/data/audio/feature_transforms/specaugment.py:80-131, the CPU numpy
// transform applied per utterance after CMVN).  The random draws stay on the host (s2t_b200/specaugment.py replays the
// reference's numpy draw order); the device applies them to the whole padded batch in place:
//   value[b] = mask_value, or the mean of the utterance's (num_frames x n_freqs) features when use_mean (:88-89)
//   freq mask (f0, f): columns [f0, f0+f) of every valid frame := value[b]
//   time mask (t0, t): frames [t0, t0+t) := value[b]
// Both kernels are HBM-bound and tiny next to the encoder: one pass over the utterance for the mean (32 CTAs per
// utterance, fp32, fixed reduction order), then 8 CTAs per mask rectangle.
#include "common.cuh"

#include <mutex>

#include "../../include/s2t_b200.h"

namespace s2t {
namespace {

__device__ __forceinline__ float ld_feat(const void* p, int is_bf16, long long i) {
  return is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i])
                 : reinterpret_cast<const float*>(p)[i];
}

__device__ __forceinline__ void st_feat(void* p, int is_bf16, long long i, float v) {
  if (is_bf16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
  else reinterpret_cast<float*>(p)[i] = v;
}

constexpr int kMeanChunks = 32;   // CTAs per utterance in the mean pass
constexpr int kRowSlices = 8;     // CTAs per mask rectangle

// grid (kMeanChunks, batch): frames dealt round-robin to (CTA, warp), lanes over the bins; one fp32 partial per CTA
// (fixed order inside the CTA); the apply kernel adds the kMeanChunks partials in index order.
__global__ void __launch_bounds__(256)
specaug_mean_kernel(const void* __restrict__ feats, int is_bf16, long long ld, long long bs,
                    const int* __restrict__ num_frames, int max_frames, int n_freqs, float* __restrict__ partials) {
  pdl_entry();
  __shared__ float s_sum[8];
  const int b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int nf = num_frames[b];
  nf = nf < max_frames ? nf : max_frames;
  float acc = 0.f;
  for (int t = blockIdx.x * 8 + warp; t < nf; t += kMeanChunks * 8) {
    const long long row = b * bs + static_cast<long long>(t) * ld;
    for (int f = lane; f < n_freqs; f += 32) acc += ld_feat(feats, is_bf16, row + f);
  }
  acc = warp_sum(acc);
  if (lane == 0) s_sum[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int w = 0; w < 8; ++w) a += s_sum[w];
    partials[b * kMeanChunks + blockIdx.x] = a;
  }
}

// grid (n_freq_masks + n_time_masks, batch, kRowSlices): rows of one rectangle dealt to (slice, warp), lanes over columns
__global__ void __launch_bounds__(256)
specaug_apply_kernel(void* __restrict__ feats, int is_bf16, long long ld, long long bs,
                     const int* __restrict__ num_frames, int max_frames, int n_freqs,
                     const int* __restrict__ freq_masks, int n_freq_masks,
                     const int* __restrict__ time_masks, int n_time_masks, float mask_value,
                     const float* __restrict__ partials) {
  pdl_entry();
  const int b = blockIdx.y, k = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int nf = num_frames[b];
  nf = nf < max_frames ? nf : max_frames;
  float value = mask_value;
  if (partials != nullptr) {
    float a = 0.f;
    for (int i = 0; i < kMeanChunks; ++i) a += partials[b * kMeanChunks + i];
    const long long n = static_cast<long long>(nf) * n_freqs;
    value = n > 0 ? a / static_cast<float>(n) : 0.f;
  }
  int r0, r1, c0, c1;
  if (k < n_freq_masks) {
    const int f0 = freq_masks[(b * n_freq_masks + k) * 2], f = freq_masks[(b * n_freq_masks + k) * 2 + 1];
    r0 = 0; r1 = nf;
    c0 = max(f0, 0); c1 = min(f0 + f, n_freqs);
  } else {
    const int j = k - n_freq_masks;
    const int t0 = time_masks[(b * n_time_masks + j) * 2], t = time_masks[(b * n_time_masks + j) * 2 + 1];
    r0 = max(t0, 0); r1 = min(t0 + t, nf);
    c0 = 0; c1 = n_freqs;
  }
  if (c1 <= c0) return;
  for (int r = r0 + blockIdx.z * 8 + warp; r < r1; r += kRowSlices * 8) {
    const long long row = b * bs + static_cast<long long>(r) * ld;
    for (int c = c0 + lane; c < c1; c += 32) st_feat(feats, is_bf16, row + c, value);
  }
}

}  // namespace
}  // namespace s2t

using namespace s2t;

extern "C" int s2t_specaugment(void* feats, int is_bf16, long long ld, long long bs, const int* num_frames, int batch,
                               int max_frames, int n_freqs, const int* freq_masks, int n_freq_masks,
                               const int* time_masks, int n_time_masks, float mask_value, int use_mean,
                               float* means, void* stream) {
  S2T_REQUIRE(feats && num_frames, "s2t_specaugment: null pointer");
  S2T_REQUIRE(n_freq_masks >= 0 && n_time_masks >= 0, "s2t_specaugment: negative mask count");
  S2T_REQUIRE(n_freq_masks == 0 || freq_masks != nullptr, "s2t_specaugment: freq_masks missing");
  S2T_REQUIRE(n_time_masks == 0 || time_masks != nullptr, "s2t_specaugment: time_masks missing");
  S2T_REQUIRE(!use_mean || means != nullptr, "s2t_specaugment: use_mean needs a [batch * 32] fp32 scratch buffer");
  S2T_REQUIRE(ld >= n_freqs && n_freqs > 0, "s2t_specaugment: ld must be >= n_freqs");
  if (batch <= 0 || max_frames <= 0 || n_freq_masks + n_time_masks == 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (use_mean) {
    S2T_LAUNCH(specaug_mean_kernel, dim3(kMeanChunks, batch), 256, 0, st, static_cast<const void*>(feats), is_bf16, ld, bs,
               num_frames, max_frames, n_freqs, means);
    S2T_LAUNCH_CHECK();
  }
  S2T_LAUNCH(specaug_apply_kernel, dim3(n_freq_masks + n_time_masks, batch, kRowSlices), 256, 0, st, feats, is_bf16, ld, bs,
             num_frames, max_frames, n_freqs, freq_masks, n_freq_masks, time_masks, n_time_masks, mask_value,
             static_cast<const float*>(use_mean ? means : nullptr));
  S2T_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// s2t_specaugment_static: the same two kernels for callers that have no way to obtain device memory for the partial
// sums of the mean fill — a data-loader worker or a plugin that is handed the feature batch but runs where cudaMalloc is
// off limits (inside someone else's memory pool or capture region, in a process whose allocator it does not own, under a
// "no allocation after start-up" policy).  The partial sums live in a fixed array inside the library image instead.
//
// The array below is device memory that every GPU loading this library keeps until the process exits (128 KiB each).
// On the host its NAME is only a handle for the runtime: passing `g_specaug_static_partials` itself to a kernel would
// hand the kernel a host address.  The device address has to be asked for, per device, with cudaGetSymbolAddress —
// static_partials_ptr() does that the first time each device is used and remembers the answer.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_specaug_static_partials[S2T_SPECAUGMENT_STATIC_MAX_BATCH * s2t::kMeanChunks];

namespace {

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
float* g_static_ptr[kStaticMaxDevices] = {};   // device address of the array on each device, nullptr = not asked yet

// Device address of the scratch array on the CURRENT device (looked up once per device, then cached).
int static_partials_ptr(float** out) {
  int dev = -1;
  S2T_CHECK_CUDA(cudaGetDevice(&dev));
  S2T_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "s2t_specaugment_static: device ordinal %d not supported (max %d)", dev,
              kStaticMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    S2T_CHECK_CUDA(cudaGetSymbolAddress(&p, g_specaug_static_partials));
    S2T_REQUIRE(p != nullptr, "s2t_specaugment_static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_static_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int s2t_specaugment_static(void* feats, int is_bf16, long long ld, long long bs, const int* num_frames,
                                      int batch, int max_frames, int n_freqs, const int* freq_masks, int n_freq_masks,
                                      const int* time_masks, int n_time_masks, float mask_value, int use_mean,
                                      void* stream) {
  S2T_REQUIRE(feats && num_frames, "s2t_specaugment_static: null pointer");
  S2T_REQUIRE(n_freq_masks >= 0 && n_time_masks >= 0, "s2t_specaugment_static: negative mask count");
  S2T_REQUIRE(n_freq_masks == 0 || freq_masks != nullptr, "s2t_specaugment_static: freq_masks missing");
  S2T_REQUIRE(n_time_masks == 0 || time_masks != nullptr, "s2t_specaugment_static: time_masks missing");
  S2T_REQUIRE(ld >= n_freqs && n_freqs > 0, "s2t_specaugment_static: ld must be >= n_freqs");
  S2T_REQUIRE(batch <= S2T_SPECAUGMENT_STATIC_MAX_BATCH,
              "s2t_specaugment_static: batch %d exceeds the built-in scratch (at most %d utterances per call); split the "
              "batch or use s2t_specaugment with your own buffer", batch, S2T_SPECAUGMENT_STATIC_MAX_BATCH);
  if (batch <= 0 || max_frames <= 0 || n_freq_masks + n_time_masks == 0) return 0;
  // Looked up even when use_mean == 0, so that ANY first call on a device completes the one-time set-up.
  float* partials = nullptr;
  if (int rc = static_partials_ptr(&partials)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (use_mean) {
    // Every partial the apply kernel reads (batch * kMeanChunks of them) is written by this launch first, so nothing
    // left over from an earlier call is ever read and nothing has to be zeroed.
    S2T_LAUNCH(specaug_mean_kernel, dim3(kMeanChunks, batch), 256, 0, st, static_cast<const void*>(feats), is_bf16, ld, bs,
               num_frames, max_frames, n_freqs, partials);
    S2T_LAUNCH_CHECK();
  }
  S2T_LAUNCH(specaug_apply_kernel, dim3(n_freq_masks + n_time_masks, batch, kRowSlices), 256, 0, st, feats, is_bf16, ld, bs,
             num_frames, max_frames, n_freqs, freq_masks, n_freq_masks, time_masks, n_time_masks, mask_value,
             static_cast<const float*>(use_mean ? partials : nullptr));
  S2T_LAUNCH_CHECK();
  return 0;
}
