/* se_b200.h -- C ABI of libse_b200.so: the B200 (sm_100a) STFT / complex-conv / mask hot path.
 *
 * The reference (shs2783/Speech-Enhancement) is pure Python and has no FFI of its own; the
 * boundary it fixes is the nn.Module surface of four classes.  Each
 * entry point below is what the Python shim under speech_enhancement_b200/ binds with ctypes
 * in place of the ATen call the reference makes at the cited line.  Plain pointers and
 * integers only: no torch types, no libtorch link dependency.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to fp32 (unless the name ends in _bf16), contiguous
 *     in the layout stated; the library never allocates, frees or synchronises.
 *   - `stream` is a cudaStream_t passed as void*; launches are asynchronous on it.
 *   - return value: SE_OK (0) or a negative SE_ERR_* code; se_last_error() gives the message
 *     of the last failure on the calling thread.
 */
#ifndef SE_B200_H_
#define SE_B200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define SE_API __attribute__((visibility("default")))
#else
#define SE_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define SE_OK                 0
#define SE_ERR_BAD_ARG       -1   /* null pointer, non-positive size, inconsistent dims      */
#define SE_ERR_UNSUPPORTED   -2   /* shape / option outside what the kernels implement       */
#define SE_ERR_CUDA          -3   /* a CUDA runtime or driver call failed (see last_error)    */
#define SE_ERR_WORKSPACE     -4   /* caller-provided workspace too small                      */

/* mask variants, se_mask_fwd `kind` */
#define SE_MASK_DCCRN_R        0  /* _2008_00264_dccrn.py:188-190 */
#define SE_MASK_DCCRN_C        1  /* _2008_00264_dccrn.py:191-193 */
#define SE_MASK_DCCRN_E        2  /* _2008_00264_dccrn.py:194-205 */
#define SE_MASK_DCUNET_UNB     3  /* _1903_03107_dcunet.py:159-161 */
#define SE_MASK_DCUNET_SIGMOID 4  /* _1903_03107_dcunet.py:163-165 */
#define SE_MASK_DCUNET_TANH    5  /* _1903_03107_dcunet.py:167-182 */
#define SE_MASK_CARN           6  /* _2104_05267_carn.py:165-166 (sign quirk kept) */
#define SE_MASK_FRCRN          7  /* _2206_07293_frcrn.py:141-152 */

/* conv engine selector, `engine` argument */
#define SE_ENGINE_AUTO         0  /* tcgen05 where the shape qualifies, else SIMT fp32       */
#define SE_ENGINE_SIMT         1  /* CUDA-core fp32 implicit GEMM                            */
#define SE_ENGINE_UMMA         2  /* tcgen05 bf16x3 split implicit GEMM (error if unsupported)*/

SE_API int         se_version(void);
SE_API const char* se_last_error(void);
/* 1 when the current device is compute capability 10.x, 0 otherwise, <0 on CUDA error */
SE_API int         se_device_ok(void);
/* number of kernels this library has launched since it was loaded (bench.py gpu_launches) */
SE_API int64_t     se_launch_count(void);

/* ---- ConvSTFT.forward (conv_stft.py:48-56): reflect pad + F.conv1d(stride=hop) -------------
 * x [B, L]; W [C2, win] (the module's `weight` buffer, [C2,1,win] contiguous); out [B, C2, T]
 * with T = (L + 2*pad*center - win)/hop + 1.                                                   */
SE_API int se_stft_fwd(const float* x, int64_t B, int64_t L, const float* W, int C2, int win, int hop,
                int pad, int center, float* out, int64_t T, void* stream);

/* ---- return_mag_phase branch (conv_stft.py:58-64): spec [B,2F,T] -> mags, phase [B,F,T] ---- */
SE_API int se_magphase_fwd(const float* spec, int64_t B, int F, int64_t T, float* mags, float* phase,
                    void* stream);

/* ---- ConviSTFT.forward (conv_stft.py:96-116) -------------------------------------------------
 * spec [B, C2, T]  (or mags [B, F, T] when phase != NULL, phase [B, F, T]);
 * A_packed [hop_pad, J*C2]: the module's inverse basis `weight` [C2, win] regrouped by
 * se_istft_pack_basis (J = ceil(win/hop));  window [win];
 * out [B, out_len] = (overlap-add / (coff + 1e-8))[crop_front : crop_front + out_len].         */
SE_API int se_istft_pack_basis(const float* Winv, int C2, int win, int hop, float* A_packed, void* stream);
SE_API int se_istft_fwd(const float* spec, const float* phase_or_null, const float* A_packed,
                 const float* window, int64_t B, int C2, int64_t T, int win, int hop,
                 int64_t crop_front, int64_t out_len, float* out, void* stream);

/* ---- FFT engine for ConvSTFT / ConviSTFT (HBM-bound; csrc/stft_fft.cu) --------------------------------
 * Valid when the module's basis IS the windowed real DFT of conv_stft.py:7-26 (`weight[c, 0, n]` =
 * window[n] * cos / -sin(2 pi f n / fft), inverse: window[n] * pinv(.)^T) -- the host shim verifies the
 * buffer against that closed form and otherwise stays on the dense engines above, so a checkpoint with a
 * different basis is still honoured.  window [win] is row 0 of the forward `weight` (forward) / the
 * module's `window` buffer (inverse).
 *   se_stft_fft_fwd    x [B, L] -> out [B, fft+2, T]: reflect pad (pad = 0 when center is False), window,
 *                      512-point real FFT per frame                                   (conv_stft.py:48-56)
 *   se_istft_fft_fwd   spec [B, fft+2, T] -> out [B, out_len]: inverse FFT + the rank-2 pseudo-inverse
 *                      term, window, overlap-add, * scale[n] (= 1 / (sum window^2 + 1e-8), [(T+J-1)*hop],
 *                      NULL = 1), crop [crop_front, crop_front + out_len)            (conv_stft.py:101-114)
 *   se_stft_fft_supported  1 when (win, hop, fft) is covered: fft == 512, win <= fft, hop <= 256,
 *                      ceil(win / hop) <= 8                                                              */
SE_API int se_stft_fft_supported(int win, int hop, int fft);
SE_API int se_stft_fft_fwd(const float* x, int64_t B, int64_t L, const float* window, int win, int hop, int fft,
                           int pad, float* out, int64_t T, void* stream);
SE_API int se_istft_fft_fwd(const float* spec, int64_t B, int64_t T, const float* window, int win, int hop, int fft,
                            const float* scale, int64_t crop_front, int64_t out_len, float* out, void* stream);

/* ---- PCM_S 16 wire format fused into the FFT engine -----------------------------------
 * The reference reads 16-bit PCM wav files (audio_dataset.py:12-21; torchaudio.load scales by 1 / 32768) and writes
 * them (mix_audio.py:144-146, trainer.py:255-262: torchaudio.save(encoding="PCM_S", bits_per_sample=16), i.e.
 * FFmpeg's av_clip_int16(lrintf(x * 2^15))).
 *   se_stft_fft_fwd_pcm16   se_stft_fft_fwd on int16 samples: the conversion happens in the stage load
 *   se_istft_fft_fwd_pcm16  se_istft_fft_fwd writing int16: clamp(-1, 1) (_2008_00264_dccrn.py:184) + round to
 *                           nearest + saturate in the overlap-add epilogue
 *   se_pcm16_to_f32 / se_pcm16_from_f32   the same two conversions as standalone passes over n samples (for the
 *                           framings the FFT engine does not cover)                                               */
SE_API int se_stft_fft_fwd_pcm16(const int16_t* x, int64_t B, int64_t L, const float* window, int win, int hop, int fft,
                                 int pad, float* out, int64_t T, void* stream);
SE_API int se_istft_fft_fwd_pcm16(const float* spec, int64_t B, int64_t T, const float* window, int win, int hop, int fft,
                                  const float* scale, int64_t crop_front, int64_t out_len, int16_t* out, void* stream);
SE_API int se_pcm16_to_f32(const int16_t* x, int64_t n, float* y, void* stream);
SE_API int se_pcm16_from_f32(const float* x, int64_t n, int16_t* y, void* stream);

/* ---- ComplexConv2d / ComplexConvTranspose2d (complex_nn.py:52-91) ----------------------------
 * se_cconv_pack_weight: Wr, Wi are the real_conv / imag_conv weights
 *   (conv: [Co, Ci, kh, kw]; transposed: [Ci, Co, kh, kw]); packed is the real GEMM operand
 *   [2Co, 2Ci*kh*kw] = [[Wr,-Wi],[Wi,Wr]] with K ordered (ci2, dh, dw); bias_packed [2Co] =
 *   [br-bi, br+bi] (only written when br, bi != NULL).
 * se_cconv2d_fwd / se_cconvT2d_fwd: x [B, 2Ci, H, W] -> y [B, 2Co, Ho, Wo], NCHW fp32.        */
SE_API int se_cconv_pack_weight(const float* Wr, const float* Wi, const float* br, const float* bi,
                         int Co, int Ci, int kh, int kw, int transposed,
                         float* packed, float* bias_packed, void* stream);
SE_API int se_cconv2d_fwd(const float* x, const float* packed, const float* bias_packed_or_null,
                   int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw,
                   int sh, int sw, int ph, int pw, float* y, int Ho, int Wo,
                   int engine, void* workspace, int64_t workspace_bytes, void* stream);
SE_API int se_cconvT2d_fwd(const float* x, const float* packed, const float* bias_packed_or_null,
                    int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw,
                    int sh, int sw, int ph, int pw, float* y, int Ho, int Wo,
                    int engine, void* workspace, int64_t workspace_bytes, void* stream);
/* bytes of scratch the UMMA engine needs for this problem (0 for SIMT) */
SE_API int64_t se_cconv_workspace_bytes(int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw,
                                 int sh, int sw, int transposed);

/* ---- allocation-free variant of se_cconv2d_fwd ---------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, layouts and result as se_cconv2d_fwd(engine = SE_ENGINE_UMMA) without the workspace
 * pair: the CL2 input + bf16 weight staging lives in a file-scope `static __device__` array of
 * SE_STATIC_WS_BYTES inside the library (every device that loads the library holds it until exit).
 * The call allocates and frees nothing.  Problems whose se_cconv_workspace_bytes() exceeds
 * SE_STATIC_WS_BYTES are rejected with SE_ERR_WORKSPACE; shapes outside the tensor engine
 * (Cin2 < 16, Cout2 > 256, kh*kw > 49) with SE_ERR_UNSUPPORTED.  All calls on a device share the one
 * array: calls that may overlap on the GPU (different streams or host threads) race -- serialise them. */
#define SE_STATIC_WS_BYTES (32 * 1024 * 1024)
SE_API int se_cconv2d_fwd_static(const float* x, const float* packed, const float* bias_packed_or_null,
                          int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw,
                          int sh, int sw, int ph, int pw, float* y, int Ho, int Wo, void* stream);

/* ---- tcgen05 conv engine, chainable form ------------------------------------------------------
 * Activations travel between layers in "CL2": channels-last bf16 hi|lo planes [B][2][H][W][Cp]
 * (value = hi + lo, Cp = channels padded to a multiple of 64, pad channels zero).
 *   se_cl2_from_nchw / se_cl2_to_nchw   convert the nn.Module layout (NCHW fp32) to / from CL2.
 *   se_cconv_prep_weight                packed fp32 [Cout2][Cin2][taps] (se_cconv_pack_weight) ->
 *                                       bf16 hi|lo GEMM operand [taps][2][npad][Cp_total]; chan_map[col]
 *                                       names the module input channel behind GEMM column `col`
 *                                       (-1 = padding, NULL = identity) so that one layer can read
 *                                       its K dimension from two CL2 tensors (decoder input + skip,
 *                                       complex_nn.py:4-16 complex_concat / torch.cat) without a copy.
 *   se_cconv_gemm                       the layer itself; replaces complex_nn.py:52-65 (+ the caller's
 *                                       causal F.pad, _2008_00264_dccrn.py:29, through pad_left).   */
typedef struct se_cconv_args {
  int transposed;              /* 0: ComplexConv2d, 1: ComplexConvTranspose2d                      */
  int64_t B;
  int nsrc;                    /* 1 or 2 CL2 sources concatenated along K                           */
  const void* src[2];          /* bf16 [B][2][Hpitch][Wpitch][Cp]                                   */
  int src_Cp[2];
  int src_Hpitch[2], src_Wpitch[2];   /* allocated rows / pixels per row (>= H, W)                  */
  int H, W;                    /* logical input extent read from every source                       */
  int pad_top, pad_left;       /* zero rows / columns the caller would F.pad before the layer (conv only) */
  int Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo;
  const void* wq;              /* from se_cconv_prep_weight, Cp_total = src_Cp[0] (+ src_Cp[1])     */
  const float* bias;           /* [Cout2] or NULL                                                   */
  float slope;                 /* epilogue activation v > 0 ? v : slope * v   (1 = none)            */
  float* y_nchw;               /* [B][Cout2][Ho][Wo] fp32 or NULL                                   */
  void* y_cl2;                 /* bf16 [B][2][Ho][Wo][y_Cp] or NULL                                 */
  int y_Cp;
  /* optional fp32 output addressing (all zero = NCHW): element (b, co, ho, wo) is written to
   * y_nchw[b*y_bs + ho*y_hs + pos + y_off] with pos = co*y_cs + wo*y_ws, skipped unless
   * 0 <= pos + y_off < y_lim (when y_lim > 0) and multiplied by y_scale[pos] (when non-NULL).
   * This is how the STFT writes [B, 2F, T] from a batch-as-rows image and how the iSTFT scatters
   * hop phases into the waveform with the window^2 normaliser and the centre crop fused
   * (conv_stft.py:101-114).                                                                        */
  int64_t y_bs, y_cs, y_hs, y_ws, y_off, y_lim;
  const float* y_scale;
  /* optional per-source extents (0 = H, W): a source with src_H < H or src_W < W reads as zero beyond its
   * own bottom / right edge -- the F.pad(x, ...) to the skip's size in front of the decoder layers
   * (_1903_03107_dcunet.py:89-92, _2206_07293_frcrn.py:97-98), done by TMA out-of-bounds fill.       */
  int src_H[2], src_W[2];
  /* optional real channel count of every source (0 = src_Cp): channels [src_C, src_Cp) are zero in the CL2
   * activations and in the prepared weights, so the engine skips 16-channel k-steps that are all padding. */
  int src_C[2];
  /* optional output folding (0 or 1 = off): the launch is a forward convolution over the QUOTIENT grid whose Cout2 =
   * fold_h * fold_w * fold_cg accumulator columns are fold_h x fold_w neighbouring output pixels side by side -- column
   * group ph = r * fold_w + s of quotient pixel (q, p) is output pixel (q * fold_h + r, p * fold_w + s), channel =
   * column - ph * fold_cg (< fold_cout2).  Ho, Wo are the real output extent; kh, kw, sh, sw, ph, pw describe the
   * convolution in quotient coordinates; `wq` / `bias` are built for the folded columns (zero weight blocks where a
   * pixel has no tap at an input offset).  Two uses: the sh * sw output phases of a ComplexConvTranspose2d as ONE stride-1
   * convolution (one wide MMA per input offset instead of one narrow MMA per kernel tap), and pixel pairs of a narrow
   * forward / stride-1 layer (k + s taps of 2 N columns instead of 2 k taps of N): complex_nn.py:67-91 layers with few
   * output channels, where every 128-row MMA is bound by its A-operand read, not by N.                              */
  int fold_h, fold_w, fold_cg, fold_cout2;
  /* with folding: the launch covers only the output pixels (out_h_add + i * out_h_mul, out_w_add + j * out_w_mul)
   * (0 = every pixel: mul 1, add 0) -- an axis whose phases are NOT folded is served by one launch per phase */
  int out_h_mul, out_h_add, out_w_mul, out_w_add;
} se_cconv_args;
SE_API int se_cl2_from_nchw(const float* x, int64_t B, int C, int H, int W, void* cl2, int Cp, void* stream);
/* fp32 x[b*xs_b + c*xs_c + h*xs_h + w] -> CL2 (same as se_cl2_from_nchw for arbitrary outer strides). */
SE_API int se_cl2_from_strided(const float* x, int64_t xs_b, int64_t xs_c, int64_t xs_h, int64_t B, int C, int H,
                        int W, void* cl2, int Cp, void* stream);
/* row-major fp32 x[M][ld] (first K columns) -> CL2 image of M pixels x K channels (one image row);
 * with se_cconv_gemm (1x1 kernel, y_cs = 1) this is a bf16x3 linear layer y = x W^T + b
 * (nn.Linear / the LSTM input projections of complex_nn.py:93-145).                              */
SE_API int se_cl2_from_rows(const float* x, int64_t M, int K, int64_t ld, void* cl2, int Cp, void* stream);
/* STFT framing for the tensor-core path (replaces the F.pad(reflect) + strided conv1d unfolding of
 * /root/reference/models/conv_stft.py:60-63): cl2[s][b][q][col] = reflect_pad(x)[b][q*hop + col],
 * bf16 hi|lo planes, col < hop (zero above), q < Q.  One image row per utterance.              */
SE_API int se_stft_frames_cl2(const float* x, int64_t B, int64_t L, int hop, int pad, int Q, int Cp, void* cl2,
                       void* stream);
SE_API int se_cl2_to_nchw(const void* cl2, int64_t B, int C, int H, int W, int Cp, float* y, void* stream);
/* rows [B*T][ld] fp32, column part*(C/2)*F + c*F + f = channel part*C/2 + c at frequency row f (the [B, T, C*F] output
 * of the LSTM block, _2008_00264_dccrn.py:161-165) -> CL2 [B][2][F][T][Cp]: the reference's permute + reshape back to
 * [B, C, F, T] fused with the layout conversion.                                                                   */
SE_API int se_cl2_from_lstm_rows(const float* rows, int64_t ld, int64_t B, int C, int F, int T, void* cl2, int Cp,
                                 void* stream);
SE_API int64_t se_cconv_wq_elems(int Cout2, int Cp_total, int taps);
SE_API int se_cconv_prep_weight(const float* packed, int Cout2, int Cin2, int taps, const int* chan_map,
                                int Cp_total, void* wq, void* stream);
SE_API int se_cconv_gemm(const se_cconv_args* args, void* stream);
/* First encoder layer (2 input channels, K = Cin2*kh*kw <= 20: HBM-bound): direct
 * fp32 kernel reading the input in place through strides (elements; w contiguous) -- e.g. the real
 * and imaginary row blocks of the STFT output with the DC bin skipped (_2008_00264_dccrn.py:152-156)
 * and the causal frame pad (:29) -- and writing NCHW fp32 and/or CL2.  Forward convolution only.   */
SE_API int se_cconv_small_cin(const float* x, int64_t xs_b, int64_t xs_c, int64_t xs_h, int64_t B, int Cin2,
                              int H, int W, int pad_top, int pad_left, const float* packed, const float* bias,
                              float slope, int Cout2, int kh, int kw, int sh, int sw, int ph, int pw,
                              float* y_nchw, void* y_cl2, int y_Cp, int Ho, int Wo, void* stream);

/* Gather step of the scatter-form transposed convolution for layers with 2 or 4 output channels (last
 * decoder layer of DCCRN / DCUNet, _2008_00264_dccrn.py:104-111, _1903_03107_dcunet.py:79-81):
 * P [B][H][W][Np] holds, per INPUT pixel, the products with every tap (column tap*Cout2 + co; produced by
 * se_cconv_gemm as a 1x1 convolution with taps*Cout2 output columns);
 * y[b,co,ho,wo] = act(bias[co] + sum of the P entries whose tap lands on (ho, wo)).                   */
SE_API int se_cconvT_col2im(const float* P, int64_t B, int H, int W, int Np, int Cout2, int kh, int kw, int sh,
                            int sw, int ph, int pw, int Ho, int Wo, const float* bias, float slope, float* y,
                            void* stream);

/* ---- weight gradient of the packed complex convolution (autograd of complex_nn.py:52-91) -------------
 * Activations are staged as NC2: NCHW bf16 hi|lo planes [B][2][C][H][Wp], Wp = W rounded up to 8.
 *   G[tap][cu][cv] += sum_{b,hu,wu} U[b,cu,hu,wu] * V[b,cv, hu*sh + dh - ph, wu*sw + dw - pw]
 * conv : U = dY, V = X  ->  d packed[co][ci][tap] = G[tap][co][ci]
 * convT: U = X,  V = dY ->  d packed[co][ci][tap] = G[tap][ci][co]
 * U is staged with se_nc2_from_nchw; V with se_nc2_gather_w, one copy per kernel column dw already sampled
 * on U's column grid (TMA needs a 16-byte aligned innermost start, so the column shift cannot be a box
 * coordinate).  G (fp32, [kh*kw][Cu][Cv]) must be zeroed by the caller; partial sums are added atomically. */
SE_API int64_t se_nc2_elems(int64_t B, int C, int H, int W);
SE_API int se_nc2_from_nchw(const float* x, int64_t B, int C, int H, int W, void* nc2, void* stream);
SE_API int64_t se_nc2_gather_elems(int64_t B, int C, int H, int Wu, int kw);
SE_API int se_nc2_gather_w(const float* x, int64_t B, int C, int H, int W, int Wu, int kw, int sw, int pw, void* out,
                           void* stream);
SE_API int se_cconv_wgrad(const void* u_nc2, int Cu, int Hu, int Wu, const void* v_gather, int Cv, int Hv,
                          int64_t B, int kh, int kw, int sh, int ph, float* G, void* stream);

/* ---- CCBAM attention on a CL2 skip activation (ccbam.py:28-106; SURVEY.md 8f side item) -----------------
 * x_cl2 [B][2][P = H*W][Cp] bf16 hi|lo, C channels = [C/2 real | C/2 imaginary].
 *   se_ccbam_channel_stats  partial [B][se_ccbam_num_partials(P)][2][Cp]: per-block sum and max over pixels
 *                           (the caller finishes the tiny reduction: mean = sum(partials)/P, max = max(partials))
 *   se_ccbam_spatial_pool   pooled [B][4][P] = mean, max over the real channels and mean, max over the
 *                           imaginary channels of x * ca[b, c]                              (ccbam.py:71-80)
 *   se_ccbam_apply          y_cl2 = x * ca[b, c] + sa[b, part(c)]  (sa [B][2][P])            (ccbam.py:98-106) */
SE_API int se_ccbam_num_partials(int64_t P);
SE_API int se_ccbam_channel_stats(const void* x_cl2, int64_t B, int64_t P, int C, int Cp, float* partial, void* stream);
SE_API int se_ccbam_spatial_pool(const void* x_cl2, const float* ca, int64_t B, int64_t P, int C, int Cp, float* pooled,
                                 void* stream);
SE_API int se_ccbam_apply(const void* x_cl2, const float* ca, const float* sa, int64_t B, int64_t P, int C, int Cp,
                          void* y_cl2, void* stream);
/*   se_ccbam_channel_attn   ca [B][C] = sigmoid(fc(mean) + fc(max)) from the partial statistics, fc(v) = w2 relu(w1 v)
 *                           with dense real matrices w1 [R2][C], w2 [C][R2] (the shared complex MLP of ccbam.py:39-42;
 *                           the reference's ComplexLinear is block diagonal, complex_nn.py:106-113)        (ccbam.py:45-63)
 *   se_ccbam_spatial_attn   sa [B][2][H*W] = sigmoid(relu(bias + 7x7 conv of pooled, zero padding 3)); packed [2][4*49],
 *                           bias [2] = the complex conv with its eval-mode complex BN folded in            (ccbam.py:65-86) */
SE_API int se_ccbam_channel_attn(const float* partial, int nblk, int64_t P, const float* w1, const float* w2, int64_t B,
                                 int C, int Cp, int R2, float* ca, void* stream);
SE_API int se_ccbam_spatial_attn(const float* pooled, const float* packed, const float* bias, int64_t B, int H, int W,
                                 float* sa, void* stream);

/* ---- train-mode ComplexBatchNorm2d (complex_nn.py:244-322; SURVEY.md 8f rank 1, the half that is not folded) ------
 * x, y, g, dx [B, 2C, H, W] NCHW fp32, channels [0, C) real and [C, 2C) imaginary.
 *   se_cbn_splits     S = number of partial-sum slots per channel for this shape
 *   se_cbn_stats      partial [C][S][5] (fp64): sum xr, xi, xr^2, xi^2, xr*xi -- the batch statistics of :266-281
 *   se_cbn_apply      y = Z x + c with coef [6][C] = Zrr, Zri, Zir, Zii, cr, ci (Z = W V^-1/2, c = B - Z M; :283-322)
 *   se_cbn_bwd_stats  partial [C][S][6] (fp64): sum gr, gi, gr*xr, gr*xi, gi*xr, gi*xi
 *   se_cbn_bwd_apply  dx = Z^T g + a0 + A x with coef [10][C] = Zrr, Zri, Zir, Zii, ar0, arr, ari, ai0, air, aii
 * The C-sized algebra between the passes (means, covariance, whitening, its derivative) stays on the host side.     */
SE_API int se_cbn_splits(int64_t B, int C, int H, int W);
SE_API int se_cbn_stats(const float* x, int64_t B, int C, int H, int W, double* partial, void* stream);
SE_API int se_cbn_apply(const float* x, const float* coef, int64_t B, int C, int H, int W, float* y, void* stream);
SE_API int se_cbn_bwd_stats(const float* x, const float* g, int64_t B, int C, int H, int W, double* partial, void* stream);
SE_API int se_cbn_bwd_apply(const float* x, const float* g, const float* coef, int64_t B, int C, int H, int W, float* dx,
                            void* stream);

/* ---- mask / magnitude-phase math (file:line per SE_MASK_* above) -----------------------------
 * Four input planes, each [B, F, T] addressed as base + b*bstride + f*fstride + t (elements),
 * so slices of the STFT output and of the decoder output are consumed in place.  For
 * SE_MASK_FRCRN the mask planes have F-2 rows.  out [B, 2F, T] contiguous (real | imag).       */
SE_API int se_mask_fwd(int kind,
                const float* nr, int64_t nr_bs, int64_t nr_fs,
                const float* ni, int64_t ni_bs, int64_t ni_fs,
                const float* mr, int64_t mr_bs, int64_t mr_fs,
                const float* mi, int64_t mi_bs, int64_t mi_fs,
                int64_t B, int F, int64_t T, float* out, void* stream);

/* backward of se_mask_fwd (what autograd derives from the reference's pointwise graph): grad_out
 * [B, 2F, T] -> d_nr, d_ni [B, F, T] and d_mr, d_mi [B, Fm, T] (Fm = F-2 for SE_MASK_FRCRN), contiguous. */
SE_API int se_mask_bwd(int kind,
                const float* nr, int64_t nr_bs, int64_t nr_fs,
                const float* ni, int64_t ni_bs, int64_t ni_fs,
                const float* mr, int64_t mr_bs, int64_t mr_fs,
                const float* mi, int64_t mi_bs, int64_t mi_fs,
                int64_t B, int F, int64_t T, const float* grad_out, float* d_nr, float* d_ni,
                float* d_mr, float* d_mi, void* stream);

/* polar -> rectangular (conv_stft.py:96-100): mags, phase [B,F,T] -> spec [B,2F,T] */
SE_API int se_polar_to_rect(const float* mags, const float* phase, int64_t B, int F, int64_t T,
                     float* spec, void* stream);

/* ---- recurrent half of an LSTM layer (replaces the time-step loop of the four
 * nn.LSTM passes of /root/reference/models/modules/complex_nn.py:131-142) -------------------------
 * pre_gates [N][T][pre_ld >= 4H] = W_ih x + b_ih + b_hh (gate order i|f|g|o), w_hh [4H][H], zero state;
 * writes h_out [N][T][H].  One persistent CTA per batch row(s), no inter-CTA synchronisation.     */
SE_API int se_lstm_rec_supported(int H);
SE_API int se_lstm_rec_fwd(const float* pre_gates, int64_t pre_ld, const float* w_hh, int64_t N, int64_t T, int H,
                           float* h_out, void* stream);

/* two LSTMs over the same input in one launch (the real_lstm / imag_lstm of a ComplexLSTM layer,
 * complex_nn.py:131-142): LSTM i reads pre-gate columns [i*4H, (i+1)*4H) and writes h_out<i> [N][T][H]. */
SE_API int se_lstm_rec_pair_fwd(const float* pre_gates, int64_t pre_ld, const float* w_hh0, const float* w_hh1,
                                int64_t N, int64_t T, int H, float* h_out0, float* h_out1, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SE_B200_H_ */
