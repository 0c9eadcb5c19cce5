// s2t_collate_frames: batch collation on the device (second half).
//
// Replaces _collate_frames + the index_select by descending length of SpeechToTextDataset.collater
// (fairseq/data/audio/speech_to_text_dataset.py:267-285, 411-423): the host ships the utterances of a batch PACKED back to
// back (sum of lengths, no padding bytes over PCIe) with their offsets, lengths and the collater's order; this kernel
// writes the zero-padded [B][T][F] batch (features, or F = 1 for raw waveforms) in that order and the sorted lengths.
// Pure data movement: bit-exact.
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

// one CTA per (output row b, chunk of the row): 16-byte vector copies when F*4 and the offsets allow, else scalar
__global__ void __launch_bounds__(256)
collate_frames_kernel(const float* __restrict__ packed, const long long* __restrict__ offsets,
                      const int* __restrict__ lengths, const int* __restrict__ order, int T, int F,
                      float* __restrict__ out, long long out_bs, int* __restrict__ lengths_out) {
  pdl_entry();
  const int b = blockIdx.y;
  const int src = order != nullptr ? order[b] : b;
  const int len = min(lengths[src], T);
  if (blockIdx.x == 0 && threadIdx.x == 0 && lengths_out != nullptr) lengths_out[b] = len;
  const long long n_valid = static_cast<long long>(len) * F;
  const long long n_row = static_cast<long long>(T) * F;
  const float* in = packed + offsets[src];
  float* o = out + b * out_bs;
  const bool vec = (F % 4 == 0) && ((offsets[src] & 3) == 0) && ((reinterpret_cast<uintptr_t>(packed) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(o) & 15) == 0) && (out_bs % 4 == 0);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (vec) {
    const long long nv_valid = n_valid >> 2, nv_row = n_row >> 2;
    const float4* in4 = reinterpret_cast<const float4*>(in);
    float4* o4 = reinterpret_cast<float4*>(o);
    for (; i < nv_row; i += stride) o4[i] = i < nv_valid ? in4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
    for (; i < n_row; i += stride) o[i] = i < n_valid ? in[i] : 0.f;
  }
}

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_collate_frames(const float* packed, const long long* offsets, const int* lengths, const int* order,
                                  int batch, int T, int F, float* out, long long out_bs, int* lengths_out,
                                  void* stream) {
  S2T_REQUIRE(packed && offsets && lengths && out, "s2t_collate_frames: null pointer");
  S2T_REQUIRE(T > 0 && F > 0 && out_bs >= 1ll * T * F, "s2t_collate_frames: bad shape (T=%d F=%d out_bs=%lld)", T, F, out_bs);
  if (batch <= 0) return 0;
  const long long n_row = 1ll * T * F;
  int gx = static_cast<int>((n_row / 4 + 255) / 256);
  if (gx > 64) gx = 64;
  if (gx < 1) gx = 1;
  S2T_LAUNCH(collate_frames_kernel, dim3(gx, batch), 256, 0, reinterpret_cast<cudaStream_t>(stream), packed, offsets,
             lengths, order, T, F, out, out_bs, lengths_out);
  S2T_LAUNCH_CHECK();
  return 0;
}
