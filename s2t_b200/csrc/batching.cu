// s2t_batch_by_size: host-side length-bucketed batching of an ORDERED utterance list (no device work).
//
// Replaces fairseq/data/data_utils_fast.pyx:20-103 (batch_by_size_vec), the Cython routine behind
// FairseqDataset.batch_by_size for `--max-tokens N --required-batch-size-multiple M` (egs/mustc/asr/run.sh:74-75).
// One pass keeps the last complete batch [start, end) and a tail [end, pos]: the tail joins the batch whenever the
// union is legal (padded tokens = members x longest member <= max_tokens, members <= max_sentences, member count
// < mult or a multiple of it); when the union overflows the batch is closed and the tail starts the next one; a
// tail that overflows by itself is closed at pos as well.
#include "common.cuh"

#include "../../include/s2t_b200.h"

extern "C" long long s2t_batch_by_size(const long long* num_tokens, long long n, long long max_tokens,
                                       long long max_sentences, int bsz_mult, long long* batch_ends) {
  if (n < 0 || (n > 0 && (num_tokens == nullptr || batch_ends == nullptr)) || bsz_mult < 1) {
    s2t::set_last_error("s2t_batch_by_size: bad arguments (n=%lld, bsz_mult=%d)", n, bsz_mult);
    return -1;
  }
  if (n == 0) return 0;
  if (max_tokens > 0) {
    for (long long i = 0; i < n; ++i) {
      if (num_tokens[i] > max_tokens) {
        s2t::set_last_error("s2t_batch_by_size: utterance %lld has %lld tokens > max_tokens=%lld", i, num_tokens[i],
                            max_tokens);
        return -1;
      }
    }
  }
  // batch_ends[cur] = end of the complete part of the batch being grown (batches 0..cur-1 are closed)
  long long cur = 0, start = 0, tail_max = 0, batch_max = 0;
  batch_ends[0] = 0;
  for (long long pos = 0; pos < n; ++pos) {
    const long long t = num_tokens[pos];
    if (t > tail_max) tail_max = t;
    const long long new_end = pos + 1;
    long long new_max = batch_max > tail_max ? batch_max : tail_max;
    const long long members = new_end - start;
    const bool overflow = (max_sentences > 0 && members > max_sentences) ||
                          (max_tokens > 0 && members * new_max > max_tokens);
    const bool mult_ok = members < bsz_mult || members % bsz_mult == 0;
    if (overflow) {
      // close the complete part; a tail [batch_ends[cur], pos] that overflows on its own is closed at pos as well
      if (max_tokens > 0 && tail_max * (new_end - batch_ends[cur]) > max_tokens) {
        batch_ends[++cur] = pos;
        tail_max = t;
      }
      start = batch_ends[cur];
      ++cur;
      new_max = tail_max;
    }
    if (overflow || mult_ok) {
      batch_ends[cur] = new_end;
      batch_max = new_max;
      tail_max = 0;
    }
  }
  if (batch_ends[cur] != n) batch_ends[++cur] = n;   // leftover tail (member count not a multiple of bsz_mult)
  return cur + 1;
}
