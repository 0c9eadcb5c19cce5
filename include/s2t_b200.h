/*
 * s2t_b200 — C ABI of the B200-native S2T speech-encoder hot path.
 *
 * Drop-in boundary (last row): these are the entry points the reference's Python
 * plugin layer binds instead of the torch / torchaudio library calls it makes today.  Every function
 *   - takes raw DEVICE pointers (unless the name ends in `_host`), integer shapes and a CUDA stream
 *     (passed as void* == cudaStream_t), never a torch type;
 *   - never allocates, frees or keeps a pointer: the caller owns all buffers and workspaces;
 *   - is asynchronous on `stream`;
 *   - returns 0 on success, a negative status otherwise; s2t_last_error() gives the message
 *     (thread-local).
 *
 * All kernels are compiled for sm_100a only; there is no CPU or other-arch fallback.
 *
 * Layout conventions: activations are row-major [rows = utterance-major b*T + t][channels];
 * bf16 = __nv_bfloat16 storage, fp32 accumulate.  "ld*" = row stride in ELEMENTS.
 */
#ifndef S2T_B200_H_
#define S2T_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define S2T_ABI_VERSION 5

/* status / introspection ---------------------------------------------------------------------- */
const char* s2t_last_error(void);
int s2t_abi_version(void);
/* compute capability of the current device as major*10+minor (100 on B200); negative on error */
int s2t_device_arch(void);
/* Programmatic dependent launch between consecutive kernels of a stream (default on; S2T_PDL=0 in the environment
 * turns it off).  Returns the previous setting.  Affects launches made after the call. */
int s2t_set_pdl(int on);
/* Debug: with S2T_GEMM_TRACE=1 in the environment CTA 0 of every persistent GEMM launch records 16 %globaltimer
 * stamps of its phases; this copies the stamps of the last launch to host_out16.  Negative when tracing is off. */
int s2t_gemm_trace_read(unsigned long long* host_out16);

/* --------------------------------------------------------------------------------------------- */
/* K4: tcgen05/TMEM GEMM family.                                                                  */
/* Replaces torch.nn.Linear / torch.matmul / torch.bmm / nn.Conv1d on the reference path:         */
/*   fairseq/modules/s2t_transformer_layer.py:49-50,59-65 (FFN), multihead_attention.py:246-265,  */
/*   372-416 (q/k/v/out proj, QK^T, PV), espnet_multihead_attention.py:39-42,343-352,             */
/*   modules/speech_to_text/subsampling.py:133-151 (Conv1d as implicit GEMM through an            */
/*   overlapping-row tensor map), modules/speech_to_text/ctc.py:59 (vocabulary projection).       */
/*                                                                                               */
/*   out[z][m][n] = residual + out_scale * dropout( gate( act( alpha * sum_k A[z][m][k]*B[z][n][k] */
/*                                                             + bias[n] ) ) )                    */
/* --------------------------------------------------------------------------------------------- */
enum { S2T_ACT_NONE = 0, S2T_ACT_RELU = 1, S2T_ACT_SWISH = 2 };
enum {
  S2T_GATE_NONE = 0,
  S2T_GATE_POSITIVE = 1,   /* v = gate > 0 ? v : 0            (ReLU backward from saved output)   */
  S2T_GATE_SWISH_GRAD = 2, /* v *= d swish(gate) / d gate     (Swish backward from saved input)   */
  S2T_GATE_MUL = 3         /* v *= gate                                                           */
};

typedef struct s2t_gemm_args {
  int M, N, K;
  int nb1, nb2; /* two batch dims (0/1 = none); gridDim.z = nb1*nb2, z = z2*nb1 + z1              */
  /* A: logical [M][K].  a_mn_major=0: memory [M][lda] (K contiguous).  =1: memory [K][lda] (M contiguous) */
  const void* a;
  int a_mn_major;
  long long lda, a_bs1, a_bs2;
  /* B: logical [N][K].  b_mn_major=0: memory [N][ldb] (K contiguous, the nn.Linear weight layout).
   *                     =1: memory [K][ldb] (N contiguous)                                        */
  const void* b;
  int b_mn_major;
  long long ldb, b_bs1, b_bs2;
  /* output [M][ldc], bf16 or fp32 */
  void* out;
  int out_fp32;
  long long ldc, c_bs1, c_bs2;
  /* epilogue */
  float alpha;      /* scales the accumulator                                                      */
  const void* bias; /* bf16 [N] or NULL                                                            */
  int act;          /* S2T_ACT_*                                                                   */
  const void* gate; /* bf16 [M][ldg] or NULL                                                       */
  int gate_mode;    /* S2T_GATE_*                                                                  */
  long long ldg, g_bs1, g_bs2;
  float dropout_p;  /* 0 = off; mask = Philox(seed, stream, element index), regenerable in backward */
  unsigned long long dropout_seed;
  unsigned int dropout_stream;
  float out_scale;
  const void* residual; /* bf16 [M][ldr] or NULL (ldr=0 -> ldc)                                    */
  long long ldr, r_bs1, r_bs2;
  int tile_n; /* 0 = auto, else 64/128/256                                                        */
  /* optional DEVICE counter added to dropout_seed when the kernel runs, so a captured CUDA graph    */
  /* draws a fresh mask on every replay (the host bumps the counter between replays); NULL = none   */
  const unsigned long long* dropout_seed_offset;
  /* split-K over a thread-block cluster (partials reduced through distributed shared memory):           */
  /* 0 = automatic, -1 = never, n > 1 = forced: the largest feasible factor <= n                          */
  int split_k;
  /* upper bound on the CTAs (= SMs, one CTA per SM) the launch may occupy; 0 = the whole GPU.  A GEMM that runs on a  */
  /* side stream beside a latency-critical chain of kernels is given a small share so that the chain's CTAs always      */
  /* find free SMs; with a bound the split-K factor is also limited so that tiles * split_k <= max_ctas                 */
  int max_ctas;
  /* optional, MN-major A only, no batch dims: colsum[m] (+)= sum_k A[m][k] (bf16 [M]) — for a weight gradient       */
  /* dW = dY^T X this is the bias gradient of the same nn.Linear, summed from the dY tiles the kernel stages anyway   */
  void* colsum;
  int colsum_accumulate;
  /* 1 = let the NEXT kernel of the stream start (programmatic dependent launch) only once every CTA has issued its last  */
  /* operand load instead of at entry: for GEMMs on a side stream with an SM share, whose successor's waiting CTAs would  */
  /* otherwise occupy SMs of the critical chain for the whole duration of this kernel                                    */
  int late_trigger;
} s2t_gemm_args;

int s2t_gemm_bf16(const s2t_gemm_args* args, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* K1: Kaldi-compatible log-mel filterbank (+ utterance CMVN).                                    */
/* Replaces fairseq/data/audio/audio_utils.py:59-79,82-103 (torchaudio.compliance.kaldi.fbank)    */
/* and feature_transforms/utterance_cmvn.py:30-45.                                                */
/* --------------------------------------------------------------------------------------------- */
/* number of frames for n samples, snip_edges=True: 1 + (n - 400) / 160 (0 if n < 400)            */
int s2t_fbank_num_frames(long long num_samples);
/*
 * wave      fp32 [B][wave_ld] samples in int16 scale (reference does not normalise, audio_utils.py:92)
 * num_samples int32 [B] valid samples per utterance
 * mel_w     fp32 [n_mels][257] mel filter weights computed on the host with torchaudio's fp32 formulas
 *           (column 256 is zero) — shared with the oracle so both use the identical matrix
 * mel_range int32 [n_mels][2] = [lo, hi) non-zero bin range of each filter (the matrix is sparse)
 * window    fp32 [400] povey window
 * out       [B] x out_bs + [max_frames][out_ld], fp32 or bf16; frames >= the utterance's own count
 *           are zero filled (as the collater pads, speech_to_text_dataset.py:267-285)
 * out_frames int32 [B] (written; may be NULL)
 */
int s2t_fbank_fwd(const float* wave, long long wave_ld, const int* num_samples, int batch,
                  const float* mel_w, const int* mel_range, const float* window, int n_mels,
                  void* out, int out_bf16, int max_frames, long long out_ld, long long out_bs,
                  int* out_frames, void* stream);
/* per-utterance, per-bin mean/variance normalisation over the valid frames (E[x^2]-E[x]^2 variance
 * from pre-centring sums, floor 1e-10, utterance_cmvn.py:35-43).  feats fp32 [B] x in_bs +
 * [max_frames][in_ld]; out fp32 or bf16 [B] x out_bs + [max_frames][out_ld] (may alias feats when
 * fp32); frames >= num_frames[b] are written as zeros.  stats: fp32 [batch][2][n_mels] workspace that
 * receives (mean, std) per utterance and bin; may be NULL when norm_means == norm_vars == 0 (the call
 * is then a cast / copy with zero fill).                                                            */
int s2t_cmvn_utt(const float* feats, long long in_ld, long long in_bs, void* out, int out_bf16,
                 long long out_ld, long long out_bs, const int* num_frames, int batch,
                 int max_frames, int n_mels, int norm_means, int norm_vars, float* stats,
                 void* stream);
/* (x - mean) / std with global statistics (feature_transforms/global_cmvn.py:26-29)              */
int s2t_cmvn_global(const float* feats, long long in_ld, long long in_bs, void* out, int out_bf16,
                    long long out_ld, long long out_bs, const int* num_frames, int batch,
                    int max_frames, int n_mels, const float* mean, const float* std_, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* K3: LayerNorm (fairseq/modules/layer_norm.py:30-35 -> nn.LayerNorm(C, eps=1e-5, affine)).      */
/* x, y bf16 [rows][C]; gamma/beta bf16 [C]; stats fp32 [rows][2] = (mean, rstd) saved for bwd.   */
/* --------------------------------------------------------------------------------------------- */
int s2t_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* stats,
                      int rows, int C, float eps, void* stream);
/* dx = LN'(dy) (+ dx_residual if non-NULL); the per-CTA dgamma/dbeta partial sums go to `partials`,
 * fp32 [n_partials][2][C] (n_partials from s2t_layernorm_bwd_partials).  Optional second output
 * dx_drop = drop_scale * dropout(dx) with the regenerable Philox mask of (seed, stream_id): the
 * gradient entering the next residual branch x + drop_scale * dropout(branch), so that branch's
 * backward needs no separate dropout pass.                                                          */
int s2t_layernorm_bwd_partials(int rows, int C);
int s2t_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* stats,
                      const void* dx_residual, void* dx, float* partials, int rows, int C,
                      void* dx_drop, float drop_p, float drop_scale, unsigned long long seed,
                      unsigned int stream_id, const unsigned long long* seed_offset, void* stream);
/* dgamma / dbeta (bf16 [C], either may be NULL) = sums of the partials written by s2t_layernorm_bwd
 * for the same (rows, C); accumulate != 0 adds to the existing values.  A separate call so the
 * caller can run it beside the dgrad chain.                                                         */
int s2t_layernorm_bwd_reduce(const float* partials, int rows, int C, void* dgamma, void* dbeta,
                             int accumulate, void* stream);
/* The same reduction for a list of LayerNorm instances in one launch per 32 jobs (the step's 26-100 dgamma / dbeta
 * reductions are only needed where a gradient bucket is shipped, so they are collected).  rows selects the number of
 * partial rows exactly as s2t_layernorm_bwd_partials(rows, C).                                                      */
typedef struct s2t_ln_reduce_job {
  const float* partials;
  int rows, C;
  void* dgamma; /* bf16 [C] or NULL */
  void* dbeta;  /* bf16 [C] or NULL */
  int accumulate;
} s2t_ln_reduce_job;
int s2t_layernorm_bwd_reduce_batched(const s2t_ln_reduce_job* jobs, int n_jobs, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* Elementwise / data-movement kernels of the path                                                */
/* --------------------------------------------------------------------------------------------- */
/* GLU over channel halves (subsampling.py:140-151: nn.GLU(dim=1) after each Conv1d):
 * x bf16 [rows][2C] -> y[rows][ldy] = x[:, :C] * sigmoid(x[:, C:]).  With lengths != NULL the rows are
 * utterance-major (row = b*T + t) and rows with t >= lengths[b] produce 0 (the padding mask the
 * encoder applies after subsampling, s2t_transformer.py:1765); the backward zeroes their gradient. */
/* y_bs != 0: row r = b*T + t is written at y + b*y_bs + t*ldy (output inside a per-utterance padded
 * buffer, e.g. the zero-framed input of the next Conv1d); y_bs == 0: dense rows y + r*ldy.          */
int s2t_glu_fwd(const void* x, void* y, long long ldy, long long y_bs, int rows, int C,
                const int* lengths, int T, void* stream);
int s2t_glu_bwd(const void* x, const void* dy, long long lddy, void* dx, int rows, int C,
                const int* lengths, int T, void* stream);
/* column sums: out[c] = sum_r x[r][c]  (bias gradients).  x bf16 [rows][ld]; out bf16 [C].        */
int s2t_colsum(const void* x, long long ld, int rows, int C, void* out, int accumulate,
               float* workspace, long long workspace_floats, void* stream);
long long s2t_colsum_workspace_floats(int rows, int C);
/* col2im for the strided Conv1d dgrad: dcol bf16 [B][T_out][k*C] -> dx bf16 [B][T_in][C],
 * dx[b][t][c] = sum_{(t_out,j): stride*t_out + j - pad == t} dcol[b][t_out][j*C + c]            */
int s2t_col2im1d(const void* dcol, void* dx, int batch, int t_in, int t_out, int C, int ksize,
                 int stride, int pad, void* stream);
/* embed block (s2t_transformer.py:1765-1794): y = dropout(scale * x * valid + pos_table[pos_idx])
 * with pos_idx = t + 2 for valid frames (padding_idx 1 + cumsum, sinusoidal_positional_embedding
 * via utils.make_positions) and the zero row for padded frames.  x,y bf16 [B][T][C];
 * lengths int32 [B]; pos_table fp32 [n_pos][C] (NULL = no positional term).                       */
int s2t_embed_fwd(const void* x, const int* lengths, const float* pos_table, int n_pos, void* y,
                  int batch, int T, int C, float scale, float dropout_p,
                  unsigned long long seed, unsigned int stream_id,
                  const unsigned long long* seed_offset, void* stream);
int s2t_embed_bwd(const void* dy, void* dx, int batch, int T, int C, float scale,
                  float dropout_p, unsigned long long seed, unsigned int stream_id,
                  const unsigned long long* seed_offset, void* stream);
/* zero rows t >= lengths[b] of x bf16 [B][T][ld] in place (padding masks, s2t_transformer.py:1765,1828) */
int s2t_mask_rows(void* x, const int* lengths, int batch, int T, int C, long long ld, void* stream);
/* weight re-layout for the implicit-GEMM Conv1d: w bf16 [Cout][Cin][k] -> [Cout][k][Cin] (and back, summing nothing) */
int s2t_permute_conv_weight(const void* w, void* out, int cout, int cin, int k, int inverse,
                            void* stream);
/* standalone regenerable dropout: y = out_scale * mask(seed, stream_id, i/8) * x / (1-p) (p = 0: plain
 * scaling), x,y bf16 [n], n % 8 == 0.
 * Same Philox numbering as the GEMM epilogue on a dense [rows][N%8==0] tensor, so this re-applies an
 * epilogue-drawn mask in the backward pass; also CTC.forward's dropout (modules/speech_to_text/ctc.py:59) */
int s2t_dropout(const void* x, void* y, long long n, float p, float out_scale,
                unsigned long long seed, unsigned int stream_id,
                const unsigned long long* seed_offset, void* stream);
/* explicit im2col of a zero-framed channels-last Conv1d input (for the weight gradient):
 * xcol[b*t_out + t][j*C + c] = xp[b*x_bs + (t*stride + j)*C + c]                                   */
int s2t_im2col1d(const void* xp, long long x_bs, void* xcol, int batch, int t_out, int C, int ksize,
                 int stride, void* stream);
/* y[b][pad + t][:] = t < lengths[b] ? x[b][t][:] : 0 : x bf16 [B][T][C] into the zero-framed input
 * y bf16 [B] x y_bs + [T + 2 pad][C] of the next Conv1d with the padding mask applied
 * (Downsampling.forward, pdss2t_transformer.py:107-116); lengths may be NULL                       */
int s2t_frame_rows(const void* x, const int* lengths, void* y, long long y_bs, int pad, int batch,
                   int T, int C, void* stream);
/* Fused self-attention forward (fairseq MultiheadAttention, modules/multihead_attention.py:246-265,372-416) for
 * d_k == 64 and T <= 256: per (utterance, head, 128-query tile) S = scale * Q K^T on tcgen05 into TMEM, key-padding mask,
 * fp32 softmax from TMEM, dropout, P V on tcgen05.  qkv bf16 [B*T][ldqkv] with q | k | v at columns 0 | C | 2C, head h at
 * h*64; outputs: probs / probs_drop bf16 [B*H][T][ld] (for the unfused backward pass; both may be NULL when lse is
 * given), lse fp32 [B*H][T] = log-sum-exp of the scaled masked scores (for s2t_attn_fused_bwd; may be NULL when probs
 * is given), ctx bf16 [B*T][ldctx].  Same Philox numbering as s2t_attn_softmax_fwd.                    */
int s2t_attn_fused_fwd(const void* qkv, long long ldqkv, int C, void* probs, void* probs_drop, long long ld,
                       float* lse, void* ctx, long long ldctx, const int* klen, int batch, int heads, int T, int dk,
                       float scale, float dropout_p, unsigned long long seed, unsigned int stream_id,
                       const unsigned long long* seed_offset, void* stream);
/* Fused self-attention backward (same shapes): one cluster of two CTAs per (utterance, head) recomputes
 * P = exp(scale S - lse) from the row log-sum-exp the forward pass saved (probs == NULL there), regenerates the dropout
 * mask, and produces dq | dk | dv (bf16 [B*T][ldg], columns 0 | C | 2C, head h at h*64) from o (forward context) and d_o
 * (its gradient), both bf16 [B*T][ldo].  Replaces the dV / dP / softmax' / dK / dQ launches of the unfused path
 * (multihead_attention.py:372-416 under autograd).                                                              */
int s2t_attn_fused_bwd(const void* qkv, long long ldqkv, int C, const void* o, const void* d_o, long long ldo,
                       const float* lse, void* dqkv, long long ldg, const int* klen, int batch, int heads, int T,
                       int dk, float scale, float dropout_p, unsigned long long seed, unsigned int stream_id,
                       const unsigned long long* seed_offset, void* stream);
/* Conv2dSubsampling (fairseq/modules/speech_to_text/subsampling.py:162-228): nn.Conv2d(k, stride, no padding) over a
 * channels-last activation x bf16 [B][T][D][Cin] as a GEMM over the patch matrix
 *   col[(b*T2 + t2)*D2 + d2][(j*k + i)*Cin + c] = x[b][stride*t2 + j][stride*d2 + i][c],  zero columns up to Kpad,
 * T2 = (T-k)/stride + 1, D2 = (D-k)/stride + 1; Kpad % 8 == 0.                                      */
int s2t_im2col2d(const void* x, void* col, int batch, int T, int D, int Cin, int k, int stride, int Kpad,
                 void* stream);
/* transpose of s2t_im2col2d (input gradient): dx[b][t][d][c] = sum of the dcol entries that read it; Cin % 8 == 0 */
int s2t_col2im2d(const void* dcol, void* dx, int batch, int T, int D, int Cin, int k, int stride, int Kpad,
                 void* stream);
/* per-channel sum / sum of squares of x bf16 [rows][C] as fp32 partials [s2t_colstats_parts(rows)][2][C] — the
 * batch statistics s2t_batchnorm_fwd consumes (BatchNorm2d over (B, D', T') == BatchNorm1d over the rows)   */
int s2t_colstats_parts(long long rows);
int s2t_colstats(const void* x, float* partials, long long rows, int C, void* stream);
/* CTC best alignment (Viterbi): replaces torch_imputer.best_alignment (fairseq/torch_imputer/best_alignment.cu:62-201  *
 * + the host backtrack of imputer.py:245-259,282-325).  log_probs fp32, element (b, t, v) at b*lp_b_stride +           *
 * t*lp_t_stride + v; targets int32 [batch][tgt_ld]; states int32 [batch][T]: the lattice state (0..2S) of every frame  *
 * < input_lengths[b], -1 beyond; nll fp32 [batch] or NULL (-logsumexp of the two final states of the max-recursion, as  *
 * the reference kernel returns).  workspace: s2t_ctc_viterbi_workspace_bytes() bytes (0 when the back-pointers fit in   *
 * shared memory).                                                                                                     */
long long s2t_ctc_viterbi_workspace_bytes(int batch, int T, int max_target_len);
int s2t_ctc_viterbi(const float* log_probs, long long lp_t_stride, long long lp_b_stride, const int* targets,
                    long long tgt_ld, const int* input_lengths, const int* target_lengths, int batch, int T, int V,
                    int max_target_len, int blank, int* states, float* nll, void* workspace, long long workspace_bytes,
                    void* stream);
/* SpecAugment masks applied on device, in place, to a padded feature batch [batch][max_frames][ld] (fp32 or bf16).    *
 * Replaces the apply half of fairseq/data/audio/feature_transforms/specaugment.py:80-131 (CPU numpy per utterance);   *
 * the random draws stay on the host in the reference's numpy order (s2t_b200/specaugment.py).                         *
 * freq_masks int32 [batch][n_freq_masks][2] = (f0, f): columns [f0, f0+f) of every frame < num_frames[b];            *
 * time_masks int32 [batch][n_time_masks][2] = (t0, t): frames [t0, t0+t); width 0 = no mask.                          *
 * Fill value = mask_value, or with use_mean the mean of the utterance's valid (frames x n_freqs) block computed on   *
 * the device BEFORE masking (means = fp32 [batch * 32] scratch of partial sums; the reference's mask_value=None,     *
 * :88-89).                                                                                                           */
int s2t_specaugment(void* feats, int is_bf16, long long ld, long long bs, const int* num_frames, int batch,
                    int max_frames, int n_freqs, const int* freq_masks, int n_freq_masks, const int* time_masks,
                    int n_time_masks, float mask_value, int use_mean, float* means, void* stream);
/* s2t_specaugment for callers that cannot allocate device memory.  No counterpart in the reference: it exists for hosts *
 * of this library that are given the feature batch but may not call cudaMalloc (or torch's allocator) themselves — a   *
 * data-loader worker or plugin running inside another component's memory pool or capture region, or under a            *
 * "nothing is allocated after start-up" rule — and so have nowhere to put the `means` scratch of the mean fill.         *
 * Same arguments and results as s2t_specaugment without `means`: the partial sums go to a fixed 128 KiB array that is   *
 * part of the library image (each GPU that loads the library holds it until the process exits).                         *
 *   - batch <= S2T_SPECAUGMENT_STATIC_MAX_BATCH, otherwise -1 and a message (split the batch, or use s2t_specaugment); *
 *   - the first call on each device looks up the array's device address (one cudaGetSymbolAddress, with that device     *
 *     current) and caches it: make that first call outside any stream capture;                                          *
 *   - the call itself never calls cudaMalloc / cudaFree or any allocating routine;                                      *
 *   - all calls on a device share the one array.  Each call overwrites the part it reads before reading it, so stale    *
 *     contents are harmless and nothing needs clearing — but two calls whose kernels can run at the same time on the    *
 *     same GPU (different streams, different host threads) race on it: keep them ordered on one stream or by events.     */
#define S2T_SPECAUGMENT_STATIC_MAX_BATCH 1024
int s2t_specaugment_static(void* feats, int is_bf16, long long ld, long long bs, const int* num_frames, int batch,
                           int max_frames, int n_freqs, const int* freq_masks, int n_freq_masks, const int* time_masks,
                           int n_time_masks, float mask_value, int use_mean, void* stream);
/* *counter += inc on the device (dropout seed counter of a captured step)                         */
int s2t_counter_add(unsigned long long* counter, unsigned long long inc, void* stream);
/* y += x, bf16 [n], n % 8 == 0                                                                    */
int s2t_add_bf16(const void* x, void* y, long long n, void* stream);
/* out[0] = scale * sum_i x[i]  (per-utterance NLL -> batch loss, criterions/ctc.py:473)            */
int s2t_f32_sum(const float* x, int n, float scale, float* out, void* stream);
/* out[0] = sum_r w[r] * sum_c x[r][c]: the weighted sum of the per-utterance NLLs of every loss term
 * (ctc_w*ctc + xctc_w*xctc + inter_ctc_w*mean(inter) + ..., criterions/ctc.py:1004-1015)           */
int s2t_f32_weighted_sum(const float* x, const float* w, int rows, int cols, float* out, void* stream);
/* plain bf16 <-> fp32 casts and fills used by the host glue */
int s2t_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream);
int s2t_cast_bf16_to_f32(const void* x, float* y, long long n, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* Attention softmax (multihead_attention.py:383-400; espnet_multihead_attention.py:119-133).     */
/* scores bf16 [nbh][Tq][ld] (q.k^T) -> probs = softmax_j( clamp(scale*(scores + bias)) ) over keys  */
/* j < klen[b] in fp32, cast to bf16; keys >= klen[b] get probability 0.  b = z / heads.           */
/* bias_mode: 0 none; 1 bias bf16 [nbh][Tq][ld_bias] added element-wise; 2 relative-position bias  */
/* BD bf16 [nbh][T][ld_bias >= 2T-1] read through the rel_shift map BD[i][j - i + T - 1]           */
/* (espnet_multihead_attention.py:292-311).  clamp <= 0 disables the clamp.                        */
/* probs_drop (optional) = dropout(probs) with the regenerable Philox mask.                        */
/* --------------------------------------------------------------------------------------------- */
/* bias_head_major != 0: the bias / dbias rows are ordered [head][batch][T] instead of the scores'    */
/* [batch][head][T] (then every head's dBD is one dense [batch*T][ld_bias] matrix).                  */
int s2t_attn_softmax_fwd(const void* scores, long long ld, const void* bias, long long ld_bias,
                         int bias_mode, int bias_head_major, void* probs, void* probs_drop,
                         const int* klen, int nbh,
                         int heads, int Tq, int Tk, float scale, float clamp, float dropout_p,
                         unsigned long long seed, unsigned int stream_id,
                         const unsigned long long* seed_offset, void* stream);
/* dscores = scale * probs * (dp - sum_k dp*probs), dp = dropout_mask * dprobs_drop / (1-p);
 * dbias (optional, bias_mode 2 layout [nbh][T][ld_bias]) receives the same values un-shifted.     */
int s2t_attn_softmax_bwd(const void* probs, const void* dprobs_drop, void* dscores, void* dbias,
                         long long ld_bias, int bias_head_major, const int* klen, int nbh, int heads,
                         int Tq, int Tk,
                         long long ld, float scale, float dropout_p, unsigned long long seed,
                         unsigned int stream_id, const unsigned long long* seed_offset, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* K6: Conformer convolution module (fairseq/modules/convolution.py:8-120), channels-last.        */
/* x, y bf16 [B][T][C]; depthwise weight bf16 [C][K] (the reference's (C,1,K) Conv1d weight);     */
/* K odd, <= 31; zero padding (K-1)/2 at both ends of every utterance's T frames.                 */
/* --------------------------------------------------------------------------------------------- */
/* number of time tiles (= thread blocks = rows of the partial buffers below) for a [B][T] batch  */
long long s2t_dwconv1d_tiles(int batch, int T);
/* y = depthwise_conv(x).  bn_partials (optional) fp32 [tiles][2][C] receives per-tile sum(y),    */
/* sum(y^2) per channel: the batch statistics s2t_batchnorm_fwd reduces (convolution.py:96-99).   */
int s2t_dwconv1d_fwd(const void* x, const void* w, void* y, float* bn_partials, int batch, int T,
                     int C, int K, void* stream);
/* dx = conv_transpose(dy, w); dw[c][j] = sum_{b,t} dy[b][t][c] * x[b][t+j-pad][c].              */
/* workspace fp32 [tiles][K][C].                                                                  */
int s2t_dwconv1d_bwd(const void* x, const void* w, const void* dy, void* dx, void* dw,
                     float* workspace, int batch, int T, int C, int K, void* stream);
/* nn.BatchNorm1d(C) over `rows` = B*T rows (padded frames included, quirk A10) + activation      */
/* (S2T_ACT_*).  training != 0: batch statistics from `partials` fp32 [nparts][2][C] (biased      */
/* variance), running_mean/var fp32 [C] updated with `momentum` (unbiased variance) and           */
/* *num_batches_tracked += 1 (all three optional).  training == 0: running statistics.            */
/* save_stats fp32 [2][C] receives (mean, rstd) for the backward pass.                             */
int s2t_batchnorm_fwd(const void* x, const float* partials, int nparts, const void* gamma,
                      const void* beta, float* running_mean, float* running_var,
                      long long* num_batches_tracked, float* save_stats, void* y, long long rows,
                      int C, float eps, float momentum, int act, int training, void* stream);
/* rows of the fp32 [n][2][C] scratch s2t_batchnorm_bwd needs                                      */
int s2t_batchnorm_bwd_partials(long long rows);
/* da = gradient w.r.t. the activation output; x = the BatchNorm input; dx, dgamma, dbeta bf16.    */
int s2t_batchnorm_bwd(const void* da, const void* x, const float* save_stats, const void* gamma,
                      const void* beta, float* partials, void* dx, void* dgamma, void* dbeta,
                      long long rows, int C, int act, int training, void* stream);

/* rel-pos attention query biases (espnet_multihead_attention.py:343-347): quv bf16 [rows][2C] =  */
/* (q + pos_bias_u | q + pos_bias_v), q = C columns with row stride ldq; u, v bf16 [C] = (H, d_k). */
int s2t_relpos_qbias_fwd(const void* q, long long ldq, const void* u, const void* v, void* quv,
                         long long rows, int C, void* stream);
/* dq (row stride lddq) = dquv[:, :C] + dquv[:, C:]  (d pos_bias_u/v = column sums of dquv)        */
int s2t_relpos_qbias_bwd(const void* dquv, void* dq, long long lddq, long long rows, int C,
                         void* stream);
/* y = dropout(act(x)), dense bf16 [n], n % 8 == 0; Philox numbering of s2t_dropout (the mask is   */
/* re-applied by a GEMM epilogue in the backward pass).  FFN with Swish: s2t_transformer_layer.py:59-65 */
int s2t_act_dropout(const void* x, void* y, long long n, int act, float p, unsigned long long seed,
                    unsigned int stream_id, const unsigned long long* seed_offset, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* Prediction-aware encoding: inter-CTC + Adapter "inter_league" (s2t_transformer.py:1881-1946,   */
/* 2043-2103; modules/speech_to_text/adapter.py:189-217,264-266,296-297).                         */
/* probs[r][v] = exp(logits[r][v] - lse[r]) for v < V, 0 for V <= v < ldp; bf16 in / out.         */
/* --------------------------------------------------------------------------------------------- */
int s2t_softmax_rows_fwd(const void* logits, long long ldv, const float* lse, void* probs,
                         long long ldp, int rows, int V, void* stream);
/* dlogits[r][v] (+)= probs * (dprobs - sum_v dprobs * probs); accumulate != 0 adds to dlogits     */
/* (the CTC-loss gradient of the same logits written by s2t_ctc_loss).                             */
int s2t_softmax_rows_bwd(const void* probs, const void* dprobs, long long ldp, void* dlogits,
                         long long ldv, int rows, int V, int accumulate, void* stream);

/* Ground-truth mixing of the prediction-aware adapter (curriculum: s2t_transformer.py:1904-1946, 2061-2103;    */
/* adapter.py:244-262): rows r with mask[r] != 0 get the one-hot distribution of truth_tok[r] (smooth != 0: 0.9 +  */
/* 0.1 / V on the token, 0.1 / V elsewhere — pae-oracle-smooth) instead of the model's softmax; probs bf16 [rows][ldp].*/
int s2t_rows_truth_mix(void* probs, long long ldp, const int* truth_tok, const int* mask, int rows, int V,
                        int smooth, void* stream);
/* x[r][0:ncols] = 0 where mask[r] != 0 (bf16 [rows][ld]): no gradient reaches the logits of a replaced frame         */
int s2t_rows_zero_masked(void* x, long long ld, const int* mask, int rows, int ncols, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* K7/K8/K9: CTC head softmax, CTC loss forward/backward, greedy decode.                          */
/* Replaces utils.log_softmax (fairseq/utils.py:477-481) + torch.nn.CTCLoss(blank, 'none',        */
/* zero_infinity=True) (fairseq/criterions/ctc.py:243-245,435-474) and CTCDecoder greedy          */
/* (fairseq/models/speech_to_text/s2t_ctc.py:312-347).                                            */
/* logits bf16 or fp32, utterance-major [B][T][ldv] (V valid columns).                            */
/* --------------------------------------------------------------------------------------------- */
/* lse[b][t] = logsumexp_v logits[b][t][v] (fp32) */
int s2t_row_lse(const void* logits, int is_bf16, long long ldv, int rows, int V, float* lse,
                void* stream);
long long s2t_ctc_workspace_bytes(int batch, int T, int max_target_len);
/*
 * targets int32 [B][tgt_ld] (already stripped of pad/eos, criterions/ctc.py:516-540),
 * input_lengths / target_lengths int32 [B].
 * Outputs: nll fp32 [B] (0 where infeasible: zero_infinity), and — when dlogits != NULL —
 * dlogits[b][t][v] = grad_scale[b-independent] * (softmax(logits)[v] - occupancy[b][t][v]) for
 * t < input_length, 0 elsewhere and 0 everywhere for infeasible utterances (the gradient of
 * sum_b nll[b] through log_softmax, criterions/ctc.py:467-473).
 */
int s2t_ctc_loss(const void* logits, int is_bf16, long long ldv, const float* lse,
                 const int* targets, long long tgt_ld, const int* input_lengths,
                 const int* target_lengths, int batch, int T, int V, int max_target_len, int blank,
                 float* nll, void* dlogits /*same dtype and layout as logits, or NULL*/,
                 float grad_scale, const float* grad_scale_dev /*optional device scalar multiplied into
                 grad_scale (0 = zero-weight dummy batch, trainer.py:687-691), or NULL*/,
                 void* workspace, long long workspace_bytes, void* stream);
/*
 * Masked ("imputer") CTC loss: torch_imputer.imputer_loss (fairseq/torch_imputer/imputer.cu:57-225,364-640), the loss
 * CtcCriterion.get_ctc_loss switches to when ctc_masked_loss is set and the encoder supplies force_emit
 * (criterions/ctc.py:455-465).  force_emit int32 [B][fe_ld]: index of the extended-label state (0 .. 2 S) that must be
 * occupied at frame t, or -1 (free); every other state of a forced frame is removed from the lattice.  Everything
 * else as s2t_ctc_loss (with all entries -1 the two agree).
 */
int s2t_ctc_loss_masked(const void* logits, int is_bf16, long long ldv, const float* lse,
                        const int* targets, long long tgt_ld, const int* force_emit, long long fe_ld,
                        const int* input_lengths, const int* target_lengths, int batch, int T, int V,
                        int max_target_len, int blank, float* nll, void* dlogits, float grad_scale,
                        const float* grad_scale_dev, void* workspace, long long workspace_bytes, void* stream);
/*
 * Greedy CTC: argmax over V (lowest index on ties), padded frames -> blank, collapse repeats,
 * drop blank.  tokens int32 [B][T] (first out_lengths[b] valid, rest -1); frame_ids int32 [B][T]
 * (required) receives the per-frame argmax after pad->blank.  scores fp32 [B] (optional, with the
 * fp32 [B][T] scratch frame_lp) = -sum_t [argmax_t != blank] max_v log_softmax(logits[b][t])[v] over
 * ALL T frames: the hypothesis score of CTCDecoder.generate (s2t_ctc.py:325-332, quirk A14).
 */
int s2t_ctc_greedy(const void* logits, int is_bf16, long long ldv, const int* input_lengths,
                   int batch, int T, int V, int blank, int* tokens, int* out_lengths,
                   int* frame_ids, float* scores, float* frame_lp, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* Fused update over the flat bf16 parameter / gradient buffers.  Replaces, per update, the trainer's */
/* optimizer.multiply_grads -> clip_grad_norm -> Adam.step -> copy to the low-precision parameters    */
/* (fairseq/trainer.py:729-759, fairseq/utils.py:328-369, fairseq/optim/adam.py:145-231,              */
/* fairseq/optim/fp16_optimizer.py).                                                                  */
/* s2t_grad_norm: total_norm[0] = mult * ||grad||_2 (fp32 accumulation, fixed summation order);       */
/* partials: fp32 scratch of s2t_grad_norm_partials(n) elements.                                      */
/* s2t_fused_adam: g = grad * grad_mult * clip_coef, clip_coef = min(1, max_norm / (total_norm[0] +   */
/* 1e-6)) when total_norm != NULL and max_norm > 0; m, v, master fp32 [n]; param = bf16(master);      */
/* step counts from 1; lr_dev / step_dev: optional device scalars that override lr / step (a captured */
/* graph replays with fresh values).  beta1 / beta2 are doubles (ABI 5): the reference forms 1 - beta in  */
/* double before rounding to fp32, and 1.f - 0.999f is 1.3e-5 away from that.  The update is skipped on the device (nothing is written) when   */
/* total_norm is given and total_norm[0] is NaN or Inf (overflowed gradients: the reference trainer   */
/* raises FloatingPointError before optimizer.step, fairseq/trainer.py:751-753; the host can read the */
/* same scalar), and when step_dev[0] < 1.  Without total_norm there is no overflow guard.            */
/* --------------------------------------------------------------------------------------------- */
int s2t_grad_norm_partials(long long n);
int s2t_grad_norm(const void* grad, long long n, float mult, float* partials, float* total_norm, void* stream);
int s2t_fused_adam(void* param, float* master, const void* grad, float* exp_avg, float* exp_avg_sq, long long n,
                   float lr, double beta1, double beta2, float eps, float weight_decay, long long step,
                   float grad_mult, const float* total_norm, float max_norm, const float* lr_dev,
                   const long long* step_dev, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* Batch collation on the device.  Replaces _collate_frames + index_select(order) of                 */
/* SpeechToTextDataset.collater (fairseq/data/audio/speech_to_text_dataset.py:267-285, 411-423).     */
/* packed fp32: the utterances of a batch back to back (utterance i at element offsets[i], lengths[i] */
/* frames of F values; F = 1 for raw waveforms); order int32 [B] (the collater's descending-length    */
/* permutation) or NULL; out fp32 [B] x out_bs + [T][F], zero padded; lengths_out int32 [B] or NULL.  */
/* --------------------------------------------------------------------------------------------- */
int s2t_collate_frames(const float* packed, const long long* offsets, const int* lengths, const int* order,
                       int batch, int T, int F, float* out, long long out_bs, int* lengths_out, void* stream);

/* --------------------------------------------------------------------------------------------- */
/* Length-bucketed batching for the data-parallel path (host only, no device work).               */
/* Replaces fairseq/data/data_utils_fast.pyx:20-103 batch_by_size_vec (behind                      */
/* FairseqDataset.batch_by_size, --max-tokens / --required-batch-size-multiple).                   */
/* num_tokens[i] = length of the i-th utterance of the ORDERED list (speech_to_text_dataset.py:    */
/* 505-512); limits <= 0 are off.  Writes the exclusive end position of every batch to             */
/* batch_ends (room for n + 1 entries) and returns the number of batches (-1 on error).            */
/* --------------------------------------------------------------------------------------------- */
long long s2t_batch_by_size(const long long* num_tokens, long long n, long long max_tokens,
                            long long max_sentences, int bsz_mult, long long* batch_ends);

#ifdef __cplusplus
}
#endif
#endif /* S2T_B200_H_ */
