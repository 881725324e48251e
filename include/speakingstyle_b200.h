/*
 * speakingstyle_b200 - C ABI of the B200-native text -> mel -> wave inference path.
 *
 * The reference (lordzuko/SpeakingStyle) is pure Python/PyTorch and has no FFI of its own
 * (SURVEY.md section 8b); each entry point below replaces the ATen call sequence of the reference lines it
 * cites, so that a Python `nn.Module` with the reference's signature can be a thin shim over it
 * (speakingstyle_b200/ holds those shims; INTEGRATION.md shows the ctypes binding).
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *   - activations are channels-last, row-major: [B, S, C] with C contiguous (the reference's own
 *     layout between modules; its transposes around Conv1d disappear);
 *   - the library never allocates or frees device memory: the caller owns every input, output and scratch
 *     buffer (each entry point documents the buffers it needs).  Its only process-wide state is the launch
 *     counter below, per-device caches of kernel attributes / SM counts (atomics), and the built-in scratch
 *     array of ss_length_regulate_noalloc;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*), nothing synchronises
 *     unless stated;
 *   - return value: SS_OK (0) or a negative ss_status; ss_status_string() names it and
 *     ss_last_error() returns a per-thread message with the failing check.  No exceptions cross
 *     the boundary; there is no CPU fallback - without a CUDA device every compute entry point
 *     returns SS_ERR_CUDA.
 */
#ifndef SPEAKINGSTYLE_B200_H_
#define SPEAKINGSTYLE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SS_API __attribute__((visibility("default")))

typedef enum {
  SS_OK = 0,
  SS_ERR_INVALID = -1,     /* bad shape / null pointer / unsupported combination */
  SS_ERR_UNSUPPORTED = -2, /* valid request this build has no kernel for */
  SS_ERR_CUDA = -3,        /* a CUDA runtime call or launch failed */
  SS_ERR_ALIGNMENT = -4    /* pointer or stride not aligned as documented */
} ss_status;

/* activation codes shared by the epilogues */
enum { SS_ACT_NONE = 0, SS_ACT_RELU = 1, SS_ACT_TANH = 2, SS_ACT_LRELU = 3 };
/* element types of activations at the boundary */
enum { SS_F32 = 0, SS_BF16 = 1 };

SS_API const char* ss_status_string(int status);
SS_API const char* ss_last_error(void);
SS_API int ss_version(void);
/* Number of kernels this library has launched from the calling process (bench.py's
 * `gpu_launches`); ss_launch_count_reset() zeroes it. */
SS_API int64_t ss_launch_count(void);
SS_API void ss_launch_count_reset(void);

/* ------------------------------------------------------------------------------------------
 * Encoder input: out[b,l,:] = emb[texts[b,l],:] + pe[l,:]
 * replaces transformer/Models.py:89-91 (src_word_emb + position_enc; PAD row of emb is zero).
 * texts [B,L] int64, emb [V,D] f32, pe [>=L, D] f32, out [B,L,D] f32.  D % 4 == 0.
 * ------------------------------------------------------------------------------------------ */
SS_API int ss_embed_posenc(const int64_t* texts, const float* emb, const float* pe, float* out, int B, int L,
                           int D, int V, void* stream);

/* out[b,t,:] = x[b,t,:] + pe[t,:]  - the position add of transformer/Models.py:158-160 for a
 * caller that uses Decoder on its own (inside FastSpeech2 the LengthRegulator kernel fuses it). */
SS_API int ss_add_posenc(const float* x, const float* pe, float* out, int B, int S, int C, void* stream);

/* ------------------------------------------------------------------------------------------
 * 1-D convolution / linear layer as an implicit GEMM over channels-last activations, fp32 FMA
 * accumulation (the accuracy class of the reference; used wherever an integer output is decided).
 * replaces nn.Conv1d / nn.Linear / nn.ConvTranspose1d call sites:
 *   transformer/SubLayers.py:39-41,54 (w_qs/w_ks/w_vs/fc), :88 (w_1, w_2),
 *   model/modules.py:216-237 (predictor convs), :329-339 (reference-encoder conv stack),
 *   model/blocks.py:68 (LinearNorm), model/fastspeech2.py:105 (mel_linear),
 *   transformer/Layers.py:140-148 (PostNet, BatchNorm folded by the caller),
 *   hifigan/models.py:150 (conv_pre), :152-153 (ups, via the phase form below), :96-103 (ResBlock).
 *
 * For output row t of batch b and tap j the input row is  t + j*dilation - pad  (zero outside
 * [0, S_in)).  acc[n] = sum_{j,ci} in_act(x[b, row, ci]) * w[(j*Cin + ci)*Cout + n].
 * epilogue: v = act(acc + bias[n]); v += residual[...] (if given); v = 0 where t >= lens[b] (if
 * given); y[...] = v;  accum[...] = (accum_init ? 0 : accum[...]) + accum_scale * v (if given).
 * The output element (b, t, n) lives at  y + b*y_batch_stride + y_offset + t*y_row_stride + n  and is
 * written only if  0 <= y_offset + t*y_row_stride + n < y_batch_elems  (this window is what lets a
 * ConvTranspose1d run as a k/stride-tap convolution with Cout = stride*C_out whose rows interleave
 * into the up-sampled sequence); residual and accum use the same addressing.
 * Requirements: Cin % 4 == 0, Cout % 4 == 0, strides and offsets % 4 == 0, 16-byte aligned pointers.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const float* x;          /* [B, S_in, Cin] */
  int64_t x_batch_stride;  /* elements */
  int64_t x_row_stride;    /* elements, >= Cin */
  const float* w;          /* [k*Cin, Cout] packed by the caller */
  const float* bias;       /* [Cout] or NULL */
  const float* residual;   /* same addressing as y, or NULL */
  const int32_t* lens;     /* [B] valid rows, or NULL */
  float* y;
  float* accum;            /* or NULL */
  int64_t y_batch_stride, y_row_stride, y_offset, y_batch_elems;
  int32_t B, S_in, S_out, Cin, Cout, k, dilation, pad;
  int32_t in_act;          /* SS_ACT_NONE or SS_ACT_LRELU, applied to x on load */
  float in_slope;
  int32_t act;             /* SS_ACT_* applied after bias */
  float slope;
  float accum_scale;
  int32_t accum_init;
} ss_conv1d_f32_params;
SS_API int ss_conv1d_f32(const ss_conv1d_f32_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * The same convolution on the 5th-generation tensor cores: tcgen05.mma with bf16 operands and fp32
 * accumulation in TMEM, operands fetched by TMA (zero padding and the halo across utterances come
 * from TMA's out-of-bounds fill; no im2col).  This is the throughput kernel of the bf16 precision
 * mode (Decoder, PostNet, HiFi-GAN: north_star tolerance relative L2 <= 1e-2); call sites as for
 * ss_conv1d_f32.  x [B,S_in,Cin] bf16; w [k, Cout, Cin] bf16 (tap-major, Cin contiguous).
 * epilogue: v = act(acc + bias); v += residual (f32 or bf16); v = 0 where t >= lens[b];
 *   y = v (f32 or bf16, if given);
 *   if accum: accum = (accum_init ? 0 : accum) + accum_scale*v (f32) and v <- that new value;
 *   y_act = bf16(act2(v)) (if given) - the activated copy the NEXT convolution consumes as its A
 *   operand (HiFi-GAN's leaky_relu -> conv pairs cannot apply the activation between TMA and MMA).
 * Output addressing / window as in ss_conv1d_f32.  Cin % 8 == 0, Cout % 8 == 0, strides % 8 == 0.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* x;           /* bf16 */
  int64_t x_batch_stride;  /* elements */
  int64_t x_row_stride;
  const void* w;           /* bf16 [k, Cout, Cin] */
  const float* bias;       /* f32 [Cout] or NULL */
  const void* residual;    /* residual_dtype, or NULL */
  const int32_t* lens;
  void* y;                 /* y_dtype, or NULL */
  void* y_act;             /* bf16, or NULL */
  float* accum;            /* f32, or NULL */
  int64_t y_batch_stride, y_row_stride, y_offset, y_batch_elems;
  int32_t B, S_in, S_out, Cin, Cout, k, dilation, pad;
  int32_t act;
  float slope;
  int32_t act2;
  float slope2;
  int32_t y_dtype, residual_dtype; /* SS_F32 / SS_BF16 */
  float accum_scale;
  int32_t accum_init;
  const int32_t* rows;     /* [B] or NULL - ragged schedule, see ss_conv1d_tf32_params */
  int32_t pair_c;          /* ss_conv1d_bf16x3 only: channels per row of the pair tensor y_act (0 = Cout; the phase
                              form of a transposed convolution has Cout = stride*cout GEMM columns but cout channels) */
} ss_conv1d_tc_params;
SS_API int ss_conv1d_tc(const ss_conv1d_tc_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * fp32-ACCURATE convolution on the tensor cores: 3xTF32 error-compensated tcgen05.mma (kind::tf32),
 * fp32 activations in and out - a drop-in for ss_conv1d_f32 (same call sites, same epilogue, same
 * addressing) for the chain that decides integers and for the fp32 precision mode.  The fp32 input
 * tile is split on chip into hi = tf32(x), lo = tf32(x - hi) (after the optional leaky-ReLU on load);
 * the weights arrive pre-split: w_hi = tf32(w), w_lo = tf32(w - w_hi), both [k, Cout, Cin] fp32.
 * acc = lo*w_hi + hi*w_lo + hi*w_hi  in fp32.  y is fp32; y_act (optional) = bf16(act2(v)).
 * Cin % 4 == 0, Cout % 4 == 0, strides/offsets % 4 == 0, 16-byte aligned pointers.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const float* x;
  int64_t x_batch_stride;
  int64_t x_row_stride;
  const float* w_hi;       /* [k, Cout, Cin] */
  const float* w_lo;       /* [k, Cout, Cin] */
  const float* bias;
  const float* residual;   /* f32 or NULL */
  const int32_t* lens;
  float* y;
  void* y_act;             /* bf16 or NULL */
  float* accum;
  int64_t y_batch_stride, y_row_stride, y_offset, y_batch_elems;
  int32_t B, S_in, S_out, Cin, Cout, k, dilation, pad;
  int32_t in_act;
  float in_slope;
  int32_t act;
  float slope;
  int32_t act2;
  float slope2;
  float accum_scale;
  int32_t accum_init;
  int32_t k_group;         /* 32-channel chunks per accumulator restart (0 = default 8): tcgen05 accumulates with
                              truncation, so partial sums are re-added in registers every k_group*4 MMAs.  8 keeps
                              rms / max error below the FP32-pipe kernel's (scripts/tf32_group_error.py) */
  const int32_t* rows;     /* [B] or NULL.  Ragged schedule: only output rows t < rows[b] of utterance b are wanted
                              (clamped to [0, S_out]); whole 128-row tiles past that are neither computed nor
                              written - their output memory keeps whatever it held.  For callers that know which
                              frames can still reach a result: the decoder layers whose padded frames the next
                              LayerNorm zeroes without reading them, and the vocoder when its wave is trimmed to
                              mel_len*hop (rows[b] = (mel_len[b] + halo)*up covers the receptive field).  B <= 1024. */
} ss_conv1d_tf32_params;
SS_API int ss_conv1d_tf32x3(const ss_conv1d_tf32_params* p, void* stream);
/* Same contract; the split activations are kept in tensor memory (tcgen05.st) and the MMAs take their A
 * operand from there, which removes the shared-memory re-reads that bound the variant above. */
SS_API int ss_conv1d_tf32x3_ts(const ss_conv1d_tf32_params* p, void* stream);
/* Same contract; two CTAs of a cluster compute one 256 x 128 tile with tcgen05 cta_group::2, each holding half
 * of the weight tile, which halves the shared-memory traffic per SM (layers with Cout <= 64 go to the kernel above). */
SS_API int ss_conv1d_tf32x3_ts2(const ss_conv1d_tf32_params* p, void* stream);

/* Same contract, for the NARROW layers (HiFi-GAN's 64- and 32-channel MRF stages, hifigan/models.py:96-103): the
 * activations of a 128-row tile are fetched once as a panel of 128 + (k-1)*dilation rows (tap j of row r is panel row
 * r + j*dilation) and the weights of all taps stay resident in shared memory when they fit (32 channels: k <= 12,
 * 64 channels: k <= 3) or stream through a ring once per tile, instead of one TMA fetch of both per (tap, chunk) step;
 * hi*W_hi and hi*W_lo are one N = 2*Cout instruction.  Same products as ss_conv1d_tf32x3_ts, equal to ~1 ulp (the
 * correction terms are summed per accumulator-restart group).  Takes: S_in == S_out, k >= 2, Cin in {32, 64},
 * Cout <= 64, 128 + (k-1)*dilation <= 256: ss_conv1d_tf32x3_panel_supported() returns 1 (resident weights) or 2
 * (streamed); anything else is 0 / SS_ERR_UNSUPPORTED. */
SS_API int ss_conv1d_tf32x3_panel(const ss_conv1d_tf32_params* p, void* stream);
SS_API int ss_conv1d_tf32x3_panel_supported(const ss_conv1d_tf32_params* p);

/* out = bf16(act(x)) element-wise over n elements (n % 8 == 0): the entry into the bf16 path. */
SS_API int ss_cast_bf16(const float* x, void* out, int64_t n, int act, float slope, void* stream);

/* ------------------------------------------------------------------------------------------
 * Split-bf16 ("bf16x3") convolution: fp32-class accuracy from three bf16 tensor-core products per multiply
 * (x_hi*w_hi + x_hi*w_lo + x_lo*w_hi, fp32 accumulation; ~2^-16 relative per product).  Used by the fp32 mode of
 * hifigan.Generator for the MRF ResBlock convolutions (reference hifigan/models.py:96-103), where it replaces
 * ss_conv1d_tf32x3 at twice the tensor-pipe rate.  Same parameter block and epilogue as ss_conv1d_tc, except:
 *   x      PAIR rows [B, S_in, 2*Cin] bf16 = [hi(Cin) | lo(Cin)], hi = bf16(v), lo = bf16(v - hi)
 *          (x_row_stride >= 2*Cin); produced by ss_split_bf16 or by this kernel's own y_act
 *   w      weight tiles packed for the tile width BN = ss_conv1d_bf16x3_plan(Cin, Cout, k, dilation) (0 = layer not
 *          handled: Cin must be 32 with Cout 32, or a multiple of 64 with Cout 64 or a multiple of 128;
 *          128 + (k-1)*dilation <= 256).  With W_hi = bf16(W), W_lo = bf16(W - W_hi), W[j, co, ci]:
 *            Cin % 64 == 0: tile (nt, c, j), index (nt*(Cin/64) + c)*k + j, is [2*BN rows, 64] bf16 with
 *                           row r < BN = W_hi[j, nt*BN + r, 64c .. 64c+64), row BN + r = W_lo[same]
 *            Cin == 32:     tile (nt, q), index nt*ceil(k/2) + q, is [64 rows, 64]: row r < 32 =
 *                           [W_hi[2q, nt*32 + r, 0..32) | W_hi[2q+1, ..]], row 32 + r the same of W_lo (tap k: zeros)
 *   y_act  PAIR rows [.., 2*Cout] of act2(v) for the next convolution (row-major over the same rows as y)
 * ss_split_bf16: out[r, 0:C] | out[r, C:2C] = hi | lo of act(x[r, :]), rows x C fp32 in, C % 8 == 0.
 * ------------------------------------------------------------------------------------------ */
SS_API int ss_conv1d_bf16x3(const ss_conv1d_tc_params* p, void* stream);
SS_API int ss_conv1d_bf16x3_plan(int Cin, int Cout, int k, int dilation);
/* The same kernel with ONE bf16 number per value: ss_conv1d_tc's contract (x [B,S_in,Cin] bf16, y_act a plain bf16 copy)
 * for the layer shapes ss_conv1d_bf16x3_plan accepts, with the activations fetched once per tile instead of once per tap.
 * w: the W_hi half of the tiles above ([BN rows, 64] per tile, same order). */
SS_API int ss_conv1d_bf16_panel(const ss_conv1d_tc_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * One HiFi-GAN ResBlock unit in one launch (hifigan/models.py:96-103), split-bf16 arithmetic:
 *   t = leaky_relu(c1(x) + b1, slope)   x: pair rows [B,S,2*C] of leaky_relu(y); c1: k taps, `dilation`; t stays on chip
 *   v = c2(t) + b2 + residual           c2: k taps, dilation 1; residual = y [B,S,C] f32
 *   y = v (if given; may alias residual);  accum / accum_scale / accum_init and y_pair = pair rows of act2(v): as in
 *   ss_conv1d_bf16x3.  Bit-identical to two ss_conv1d_bf16x3 launches; 16 instead of 24 bytes of HBM traffic per
 *   element.  w1 / w2: ss_conv1d_bf16x3 weight tiles of the layer's channel count.  y_pair must not alias x.
 * ss_resunit_bf16x3_supported(C, k, dilation): 1 for C == 32 with k odd <= 17 and for C == 64 with k odd <= 9 whose
 * weights fit next to the panels (k = 3), 128 + (k-1)*dilation <= 256; else 0.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* x;           /* bf16 pair rows [B,S,2*C] */
  const float* residual;   /* f32 [B,S,C] or NULL */
  const void* w1;          /* bf16 tiles of c1 */
  const float* b1;         /* f32 [C] */
  const void* w2;          /* bf16 tiles of c2 */
  const float* b2;         /* f32 [C] */
  float* y;                /* f32 [B,S,C] or NULL */
  void* y_pair;            /* bf16 pair rows [B,S,2*C] or NULL */
  float* accum;            /* f32 [B,S,C] or NULL */
  const int32_t* rows;     /* [B] or NULL - ragged schedule */
  int32_t B, S, C, k, dilation;
  float slope;             /* leaky-ReLU between c1 and c2 */
  int32_t act2;
  float slope2;
  float accum_scale;
  int32_t accum_init;
} ss_resunit_params;
SS_API int ss_resunit_bf16x3(const ss_resunit_params* p, void* stream);
SS_API int ss_resunit_bf16x3_supported(int channels, int k, int dilation);
SS_API int ss_split_bf16(const float* x, void* out, int64_t rows, int C, int act, float slope, void* stream);

/* ------------------------------------------------------------------------------------------
 * Row-wise LayerNorm with the element-wise steps the reference applies around it, one pass:
 *   v = x (+ residual); v = relu(v) if pre_relu; y = LN(v)*ln_w + ln_b  (eps 1e-5, biased var);
 *   FiLM (model/blocks.py:60-62): y = (s_gamma*gamma[b,:] + 1)*y + s_beta*beta[b,:] if gamma given;
 *   y = 0 where t >= lens[b] (masked_fill, transformer/Layers.py:27-28,34-35; modules.py:370-371);
 *   y += pe[t,:] if pe given (modules.py:385-387).
 * replaces nn.LayerNorm at SubLayers.py:55,91; modules.py:231,243,334.  C in {256, 1024}.
 * s_gamma/s_beta are device scalars (the learnable FiLM scales).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const float* x;
  const float* residual; /* or NULL */
  const float* ln_w;
  const float* ln_b;
  const float* gamma;    /* [B, C] or NULL */
  const float* beta;     /* [B, C] */
  const float* s_gamma;  /* [1] */
  const float* s_beta;   /* [1] */
  const int32_t* lens;   /* [B] or NULL */
  const float* pe;       /* [>=S, C] or NULL */
  float* y;
  void* y_bf16;          /* optional bf16 copy of y (A operand of the next tensor-core convolution), or NULL */
  int32_t B, S, C, pre_relu;
  int32_t bf16_pair;     /* y_bf16 is [B,S,2*C]: split-bf16 pair rows [hi(C) | lo(C)] for ss_conv1d_bf16x3 */
} ss_layernorm_params;
SS_API int ss_layernorm(const ss_layernorm_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * Masked multi-head self-attention, softmax fused, scores never stored.
 * replaces transformer/SubLayers.py:42-52 + transformer/Modules.py:14-23:
 *   ctx[b,q,h*dk:(h+1)*dk] = softmax_k( Q_h[b,q,:].K_h[b,k,:] / sqrt(dk),  k < lens[b] ) . V_h[b,k,:]
 * qkv [B,S,3*D] f32 holds Q | K | V (the three projections computed by ONE ss_conv1d_f32 call with
 * concatenated weights); head h uses columns h*dk.. of each third.  ctx [B,S,D] f32.
 * Query rows >= lens[b] are written as zeros (the reference masks them right after the following
 * LayerNorm).  dk in {32, 128}; D = H*dk.
 * ------------------------------------------------------------------------------------------ */
SS_API int ss_attention_f32(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, int dk,
                            void* stream);

/* The same operation, same arguments and same fp32 accuracy class, on the tensor cores, for dk == 32 (the
 * ReferenceEncoder's 8 heads, modules.py:353) and dk == 128 (Encoder / Decoder, 2 heads); other head sizes return
 * SS_ERR_UNSUPPORTED (use ss_attention_f32).  S = Q.K^T and O = P.V are 3xTF32 tcgen05.mma with accumulators in TMEM,
 * K / V streamed by TMA up to the utterance's length, V transposed on chip by the split warps (the MMA wants it
 * K-major), Q and P held in TMEM as the A operands, softmax statistics and the output accumulation across key tiles in
 * fp32 registers. */
SS_API int ss_attention_tc(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, int dk,
                           void* stream);

/* ------------------------------------------------------------------------------------------
 * Variance predictor tail: LayerNorm_2 -> [FiLM] -> Linear(C->1) -> mask, one warp per token.
 * replaces model/modules.py:243-258 (h is the output of conv1d_2 + ReLU).
 * pred[b,l] = (lens && l >= lens[b]) ? 0 : dot(FiLM(LN(h[b,l,:])), lin_w) + lin_b
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const float* h;       /* [B, L, C] */
  const float* ln_w;
  const float* ln_b;
  const float* gamma;   /* [B, C] or NULL */
  const float* beta;
  const float* s_gamma;
  const float* s_beta;
  const float* lin_w;   /* [C] */
  const float* lin_b;   /* [1] */
  const int32_t* lens;
  float* pred;          /* [B, L] */
  int32_t B, L, C;
} ss_predictor_head_params;
SS_API int ss_predictor_head(const ss_predictor_head_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * pitch / energy quantisation + embedding add, in place:
 *   pred[b,l] *= control;  idx = #(edges < pred[b,l])  (torch.bucketize right=False; NaN -> n_edges);
 *   x[b,l,:] += table[idx,:];  idx_out[b,l] = idx (if idx_out given)
 * replaces model/modules.py:88-91,126 and :99-102,131.  edges ascending, n_edges <= 1024.
 * ------------------------------------------------------------------------------------------ */
SS_API int ss_bucket_embed_add(float* pred, float control, const float* edges, int n_edges, const float* table,
                               float* x, int32_t* idx_out, int B, int L, int C, void* stream);
/* The same with one control value PER TOKEN, control_vec [B, L] (the fine-control notebook multiplies the prediction by
 * a [B, L] tensor, notebooks/control.ipynb cells 17-23; model/modules.py:88,99 broadcasts either form). */
SS_API int ss_bucket_embed_add_v(float* pred, const float* control_vec, const float* edges, int n_edges,
                                 const float* table, float* x, int32_t* idx_out, int B, int L, int C, void* stream);

/* ------------------------------------------------------------------------------------------
 * Duration rounding + LengthRegulator, two calls around ONE host read of mel_len_expand:
 *
 * ss_duration_scan  (model/modules.py:137-140,143 and the int() / max(.,0) of :193-194)
 *   d_rounded[b,l] = clamp(round_half_even(exp(log_d[b,l]) - 1) * d_control, min 0)    (if log_d given;
 *                    otherwise `d_rounded` is read as the caller's durations, e.g. d_targets)
 *   reps = max(trunc(d_rounded), 0);  ends[b,l] = inclusive prefix sum of reps   (int32)
 *   mel_len_expand[b] = ends[b,L-1]                (what LengthRegulator.LR returns, modules.py:181)
 *   mel_len_sum[b]    = (int32) sum_l d_rounded[b,l]  in float32, index order   (modules.py:143)
 *
 * ss_length_regulate  (model/modules.py:174-197 + utils/tools.py:319-337, + transformer/Models.py:158-160)
 *   out[b,t,:] = x[b, i, :] (+ pe[t,:] if pe given)   with i = #{ j : ends[b,j] <= t }   for t < ends[b,L-1]
 *   out[b,t,:] = 0 (+ pe[t,:])                         for the padding rows up to T
 *   mel2phone[b,t] = i, or -1 on padding (if given).  C % 4 == 0.
 * ------------------------------------------------------------------------------------------ */
SS_API int ss_duration_scan(const float* log_d, float d_control, float* d_rounded, int32_t* ends,
                            int32_t* mel_len_expand, int32_t* mel_len_sum, int B, int L, void* stream);
/* ss_duration_scan with a per-token d_control [B, L] (model/modules.py:137-140 with a tensor control). */
SS_API int ss_duration_scan_v(const float* log_d, const float* d_control_vec, float* d_rounded, int32_t* ends,
                              int32_t* mel_len_expand, int32_t* mel_len_sum, int B, int L, void* stream);
SS_API int ss_length_regulate(const float* x, const int32_t* ends, const float* pe, float* out,
                              int32_t* mel2phone, int B, int L, int T, int C, void* stream);

/* ------------------------------------------------------------------------------------------
 * ss_length_regulate_noalloc - LengthRegulator in ONE call that needs no buffer beyond its inputs and
 * outputs.  (No counterpart in the reference: it exists for callers of this library, not of the model.)
 *
 * Who needs it: a caller that cannot obtain device memory at call time - a synthesis worker that was
 * given a fixed set of device buffers at start-up and must not call cudaMalloc on its request path
 * (it synchronises the whole device and can fail once memory is fragmented), or a plug-in living in
 * a host runtime that keeps its allocator to itself.  ss_duration_scan + ss_length_regulate make such
 * a caller supply a [B, L] int32 buffer for the prefix sums; here that buffer is an array inside the
 * library image (SS_LR_NOALLOC_MAX_TOKENS int32 = 256 KiB, resident on every GPU that loads the
 * library until the process exits).
 *
 *   reps = max(trunc(durations[b,l]), 0);  mel_len[b] = sum_l reps            (int32, written for every b)
 *   out[b,t,:] = x[b, i, :] (+ pe[t,:])  with i = #{ j : prefix[b,j] <= t }   for t < min(mel_len[b], T)
 *   out[b,t,:] = 0 (+ pe[t,:])                                                 for the remaining rows up to T
 * i.e. LengthRegulator.forward(x, duration, max_len=T) of model/modules.py:199-201; the caller fixes T
 * (its output buffer is [B, T, C]); an utterance longer than T is cut at T, as F.pad would not allow.
 *
 * Limits and rules:
 *   - B*L <= SS_LR_NOALLOC_MAX_TOKENS, L <= 12288 and B <= 65535, otherwise SS_ERR_UNSUPPORTED (use the two-call form);
 *   - no cudaMalloc / cudaFree / allocating library call is made, ever;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *     that device current) - make that first call outside any stream capture;
 *   - all calls on a device share the one array: each call overwrites the part it uses before reading
 *     it (nothing needs clearing), but two calls whose kernels can overlap on the GPU - different
 *     streams without an ordering between them - race on it.  Serialise them.
 * ------------------------------------------------------------------------------------------ */
#define SS_LR_NOALLOC_MAX_TOKENS 65536
SS_API int ss_length_regulate_noalloc(const float* x, const float* durations, const float* pe, float* out,
                                      int32_t* mel_len, int B, int L, int T, int C, void* stream);

/* mask[b,t] = (t >= lens[b])  as bytes (torch.bool storage) - utils/tools.py:110-118 */
SS_API int ss_length_mask(const int32_t* lens, uint8_t* mask, int B, int T, void* stream);

/* out[b,:] = (1/S) * sum_t x[b,t,:]   over the PADDED length S (model/modules.py:398). */
SS_API int ss_mean_pool(const float* x, float* out, int B, int S, int C, void* stream);

/* Batch-invariant mode (opt-in, deviates from the reference by design - see DESIGN.md):
 * out[b,:] = (1/n_b) * sum_{t<n_b} x[b,t,:], n_b = clamp(lens[b], 1, S): the mean over the utterance's own frames,
 * so that the style vector does not depend on how far the batch was padded. */
SS_API int ss_mean_pool_lens(const float* x, const int32_t* lens, float* out, int B, int S, int C, void* stream);
/* x[b,t,:] = 0 for t >= lens[b] (in place): closes the padding leaks the reference leaves open after the speaker /
 * pitch / energy embedding adds (fastspeech2.py:74-77, modules.py:126,131). */
SS_API int ss_mask_rows(float* x, const int32_t* lens, int B, int S, int C, void* stream);

/* x[b,l,:] += table[ids[b],:] for every l, padding included (model/fastspeech2.py:74-77). */
SS_API int ss_add_speaker(float* x, const int64_t* ids, const float* table, int B, int L, int C, int n_speakers,
                          void* stream);

/* ------------------------------------------------------------------------------------------
 * HiFi-GAN output head: wave = tanh(conv_post(leaky_relu(x, 0.01)))  (hifigan/models.py:161-163)
 * x [B,T,C] f32 (C <= 64), w [k*C] packed tap-major, one output channel.
 * wave [B,T] f32 (if given) and/or pcm [B,T] int16 = trunc(wave * pcm_scale) (utils/model.py:105-108).
 * ------------------------------------------------------------------------------------------ */
SS_API int ss_hifigan_post(const float* x, const float* w, const float* bias, float* wave, int16_t* pcm,
                           float pcm_scale, float in_slope, int B, int T, int C, int k, void* stream);

/* Ragged PCM epilogue: what utils/model.py:105-113 (x32768, astype int16, trim each wave to
 * mel_len*hop) and utils/tools.py:222-230 do on the host after a full fp32 D2H.  Same head as ss_hifigan_post,
 * but utterance b writes only its first n_samples[b] samples (clamped to [0,T]), contiguously from
 * pcm[out_offset[b]]; samples past an utterance's end are neither computed nor written.  n_samples [B] int32 and
 * out_offset [B] int64 are device arrays.  pcm may be device memory or page-locked host memory (under unified
 * addressing its host pointer is also its device pointer): then the trimmed int16 samples are the only bytes that
 * cross PCIe and no separate copy is needed - synchronise the stream before reading them on the host. */
SS_API int ss_hifigan_post_trim(const float* x, const float* w, const float* bias, const int32_t* n_samples,
                                const int64_t* out_offset, int16_t* pcm, float pcm_scale, float in_slope, int B, int T,
                                int C, int k, void* stream);

/* ------------------------------------------------------------------------------------------
 * Reference-mel front-end: audio/stft.py:51-81 (STFT.transform), :160-178 (TacotronSTFT.mel_spectrogram),
 * audio/audio_processing.py:85-91 (dynamic_range_compression).  The two contractions are ss_conv1d_tf32x3_ts calls
 * (Fourier basis as a k = n_fft/hop convolution over hop-sized rows; mel filterbank as a k = 1 layer); these are the
 * steps around them.
 * ------------------------------------------------------------------------------------------ */
/* out[b, i] = y[b, reflect(i - half)] for i < n_out <= T + 2*half   (F.pad(mode="reflect"), stft.py:59-64) */
SS_API int ss_reflect_pad(const float* y, float* out, int B, int T, int half, int n_out, void* stream);
/* spec [rows, spec_stride] = [re(n_bins) | im(n_bins) | pad] -> mag [rows, mag_stride] = sqrt(re^2+im^2), zero in the
 * padding columns (stft.py:76-79); energy [rows] = ||mag||_2 (stft.py:176), may be NULL */
SS_API int ss_stft_magnitude(const float* spec, float* mag, float* energy, int64_t rows, int n_bins, int spec_stride,
                             int mag_stride, void* stream);
/* *flag |= 1 if any y[i] is outside [lo, hi] or NaN (the `assert torch.min(y) >= -1 and torch.max(y) <= 1` of
 * stft.py:168-169 as one pass with one 4-byte read-back); the caller zeroes *flag first. */
SS_API int ss_range_flag(const float* y, int64_t n, float lo, float hi, int32_t* flag, void* stream);
/* x[i] = log(max(x[i], clip))  in place (audio_processing.py:85-91, C = 1) */
SS_API int ss_log_clamp(float* x, int64_t n, float clip, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SPEAKINGSTYLE_B200_H_ */
