/* ivit_b200 -- C ABI of the B200 (sm_100a) kernels behind the I-ViT `quantization_utils`
 * module API.
 *
 * The reference (lionnus/I-ViT) has no native code and no FFI: its boundary is the Python
 * module API of `models/quantization_utils`.  This header is the
 * seam *below* that API: the new Python `quantization_utils` modules call these entry
 * points through ctypes where the reference calls stock torch ops.  Each entry point names
 * the reference code it replaces (paths relative to /root/reference).
 *
 * Conventions: plain pointers and sizes only (device pointers unless noted), no torch
 * types; every call enqueues on `stream` (a cudaStream_t passed as void*), never
 * allocates, never synchronises, is re-entrant per stream, and returns 0 or a
 * cudaError_t / negative ivit error code.  All integer "codes" are the reference's
 * `round(x / scale)`; float artefacts of its float32 simulation are replayed on device.
 */
#ifndef IVIT_B200_H
#define IVIT_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IVIT_ERR_BAD_ARG (-1)
#define IVIT_ERR_UNSUPPORTED (-2)
#define IVIT_ERR_DRIVER (-3)

/* Library / device probe.  ivit_abi_version() needs no GPU.  2 = this header (round 2: fields appended to
 * ivit_attention_params_t, ivit_window_attention_params_t and ivit_gemm_epilogue_t; zero-initialised structs of version 1
 * callers keep their meaning). */
int ivit_abi_version(void);
/* Programmatic dependent launch of the hot kernels (GEMM, attention, LayerNorm, GELU table): their prologues
 * overlap the predecessor's tail inside a stream / CUDA graph.  Returns the previous setting.  Process-wide. */
int ivit_set_pdl(int on);
int ivit_device_sm(int* major, int* minor);           /* cudaGetDeviceProperties of current device */
const char* ivit_last_error(void);

/* ---- QuantAct on a float tensor (no predecessor scale) --------------------------------
 * quant_modules.py:374-377 -> quant_utils.py:13-49,79-97 (SymmetricQuantFunction):
 * q = clamp(round(fl32(1/s) * x), -128, 127).  inv_scale = fl32(1/s) computed by the caller. */
int ivit_quantize_f32_i8(const float* x, int8_t* out, float inv_scale, int64_t n, void* stream);
/* Same, fused with the im2col of a kernel=stride=P convolution (layers_quant.py:191-198 +
 * quant_modules.py:478-511): [B,Cin,H,W] fp32 -> [B*(H/P)*(W/P), Cin*P*P] int8. */
int ivit_quantize_patches_f32_i8(const float* x, int8_t* out, float inv_scale, int B, int Cin, int H, int W, int P,
                                 void* stream);

/* The same for uint8 images: torchvision's ToTensor (x / 255) and Normalize ((x - mean[c]) / std[c]) of the data
 * loader (scripts/inference.py builds them into its transform) followed by qact_input.  A pixel has 256 values
 * per channel, so the caller tabulates all three steps in float32, in that order, as lut int8 [Cin][256]
 * (device memory); the kernel gathers codes into the im2col layout.  x uint8 [B,Cin,H,W], 4-byte aligned. */
int ivit_quantize_patches_u8_i8(const uint8_t* x, int8_t* out, const int8_t* lut, int B, int Cin, int H, int W, int P,
                                void* stream);

/* ---- IntLayerNorm + following QuantAct --------------------------------------------------
 * kind 0: ivit_modules.py:30-65 (IVITIntLayerNorm), kind 1: ibert_modules.py:112-158
 * (IBERTIntLayerNorm; shift = its `shift` buffer, use_int_sqrt as the ctor flag), then
 * quant_modules.py:379-387 / quant_utils.py:193-253 (fixedpoint_mul, per-channel ratio).
 * x: int8 [rows, C] codes (row stride ldx bytes: the final `norm` runs on the class-token rows
 * only, vit_quant.py:302-303) with scalar scale s_in; bias_int[C], sf[C] (the LN's per-channel
 * output scale), mult[C] = m*2^-e of sf[c]/s_out as float64; out int8 [rows, C]. */
int ivit_layernorm_requant_i8(int kind, const int8_t* x, int64_t ldx, int8_t* out, int rows, int C, float s_in,
                              const float* bias_int, const float* sf, const double* mult, int qmin, int qmax,
                              float shift, int use_int_sqrt, void* stream);
/* Two kernels for kind 0: packed lanes (C % 128 == 0, default) and strided lanes (any supported C).
 * variant 0 = choose per call, 1 = always the second.  Returns the previous setting.  Process-wide.
 * The 16-bit entry points below follow the same switch: 0 = the restructured kernels (x/s as two fused operations on
 * 1/s, taken only for scales the library has checked against the IEEE quotient over every code), 1 = always the
 * division kernels (and ivit_gelu_lut_apply_i8 one row per warp at every width).  For A/B timing and tests. */
int ivit_layernorm_set_variant(int variant);
/* The check behind that choice, host arithmetic only (no GPU needed): 1 if, for scale s and EVERY code of the given
 * width (8 or 16 bits), c + fma(-c, s, r) * (1/s) with r = fl(q s), c = fl(r (1/s)) equals the IEEE quotient r / s --
 * the reference's `x / scaling_factor` on the float carrier (ivit_modules.py:36).  Cached per (s, bits). */
int ivit_xeff_fma_ok(float s, int bits);

/* ---- GELU (+ mlp.qact1) as a per-layer table --------------------------------------------
 * kind 0: ivit_modules.py:89-126 (ShiftGELU, n = 23, output_bit), kind 1:
 * ibert_modules.py:203-235.  Builds lut[256][256] int8, indexed [rowmax+128][code+128], with
 * the following QuantAct (scalar mult_out, clamp) applied.  s_in = scale of the int8 input. */
int ivit_gelu_lut_build(int kind, int8_t* lut, float s_in, int output_bit, double mult_out, int qmin, int qmax,
                        float* s_gelu_out_host, void* stream);
/* 65536 raw ShiftGELU integers (float32), for tests of the un-fused module. */
int ivit_gelu_ivit_raw_table(float* out65536, float s_in, int output_bit, void* stream);
int ivit_gelu_lut_apply_i8(const int8_t* x, int8_t* out, const int8_t* lut, int rows, int C, void* stream);

/* ---- GELU without any allocation ----------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * ivit_gelu_lut_build(kind, lut, s_in, output_bit, mult_out, qmin, qmax) followed by
 * ivit_gelu_lut_apply_i8(x, out, lut, rows, C), but the table and the per-row maximum codes live in a
 * file-scope device array of the library (320 KB per GPU that loads it, held until exit), so the call
 * allocates nothing: no cudaMalloc / cudaFree, no torch allocator.  rows > IVIT_GELU_NOALLOC_MAX_ROWS returns
 * IVIT_ERR_UNSUPPORTED (ivit_last_error() says why).  All calls share that array: it keeps the previous
 * call's contents (every entry a call reads is rewritten by that call first), and calls that overlap on
 * the GPU -- different streams or host threads -- race; serialise them on one stream. */
#define IVIT_GELU_NOALLOC_MAX_ROWS 65536
int ivit_gelu_i8_noalloc(int kind, const int8_t* x, int8_t* out, int rows, int C, float s_in, int output_bit,
                         double mult_out, int qmin, int qmax, float* s_gelu_out_host, void* stream);

/* ---- stand-alone QuantAct on integer codes (un-fused module path) -----------------------
 * quant_utils.py:193-253.  z int32 accumulators (after QuantLinear / QuantMatMul) or int8 codes
 * (after a previous QuantAct) [rows, C]; s_in/mult have C entries if per_channel else 1;
 * identity (may be NULL) int8 [id_rows, C] with scalar mult_id, broadcast over rows when
 * 0 < id_rows < rows (the positional embedding of vit_quant.py:295-296). */
int ivit_requant_i32_i8(const int32_t* z, const float* s_in, const double* mult, int per_channel,
                        const int8_t* identity, int64_t id_rows, double mult_id, int8_t* out, int64_t rows, int C,
                        int qmin, int qmax, void* stream);
int ivit_requant_i8_i8(const int8_t* z, const float* s_in, const double* mult, int per_channel,
                       const int8_t* identity, int64_t id_rows, double mult_id, int8_t* out, int64_t rows, int C,
                       int qmin, int qmax, void* stream);
int ivit_dequant_i8_f32(const int8_t* q, float scale, float* out, int64_t n, void* stream);
int ivit_fill_cls_rows_i8(int8_t* out, const int8_t* cls_row, int B, int N, int C, void* stream);

/* ---- 16-bit activation locations (vit_quant.py:180-187 `*_bw = 16`; scripts/inference.py:297 --bitwidth 16)
 * The residual stream then carries int16 codes.  Same LayerNorm + QuantAct as above on int16 rows
 * (row stride ldx in ELEMENTS; the int32 square of ivit_modules.py:41 wraps as in the reference),
 * the class-token row fill, and the stand-alone QuantAct over any of int32 / int16 / int8 inputs
 * (z_bits), int8 / int16 identity (id_bits) and int8 / int16 output (out_bits). */
int ivit_layernorm_requant_i16_i8(int kind, const int16_t* x, int64_t ldx, int8_t* out, int rows, int C, float s_in,
                                  const float* bias_int, const float* sf, const double* mult, int qmin, int qmax,
                                  float shift, int use_int_sqrt, void* stream);
/* QuantAct on a float tensor at any width up to 16 bits (quant_utils.py:49,93; e.g. qact_pos at 16 bit):
 * out int8 (out_bits 8) or int16 (out_bits 16) = clamp(round(inv_scale * x), qmin, qmax). */
int ivit_quantize_f32_codes(const float* x, void* out, int out_bits, float inv_scale, int64_t n, int qmin, int qmax,
                            void* stream);
/* The same LayerNorm + QuantAct when the reference's input tensor is CHANNEL-STRIDED (Swin stage 0: the
 * layout of `conv.flatten(2).transpose(1, 2)`, layers_quant.py:198, survives every elementwise op up to the
 * first PatchMerging): torch then sums the mean (ivit_modules.py:37) in its outer-reduction order, which
 * depends on the token's index within its image (tokens_per_image).  x int8 (x_bits 8) or int16 (16). */
int ivit_layernorm_requant_strided(int kind, const void* x, int x_bits, int64_t ldx, int8_t* out, int rows, int C,
                                   float s_in, const float* bias_int, const float* sf, const double* mult, int qmin,
                                   int qmax, float shift, int use_int_sqrt, int tokens_per_image, void* stream);
int ivit_fill_cls_rows_i16(int16_t* out, const int16_t* cls_row, int B, int N, int C, void* stream);
int ivit_requant_codes(const void* z, int z_bits, const float* s_in, const double* mult, int per_channel,
                       const void* identity, int id_bits, int64_t id_rows, double mult_id, void* out, int out_bits,
                       int64_t rows, int C, int qmin, int qmax, void* stream);

/* ---- QuantLinear / QuantConv2d(im2col) GEMM with fused QuantAct epilogues ---------------
 * quant_modules.py:186-226 (QuantLinear.forward: F.linear on integer-valued floats) and the
 * QuantAct(s) that follow it in vit_quant.py / layers_quant.py.  tcgen05 kind::i8 MMA, INT32
 * accumulators in TMEM, TMA-fed.  A int8 [M,K] (row stride lda bytes), W int8 [N,K]
 * (row-major, K contiguous), bias int32 [N] or NULL.
 *
 * epilogue:
 *   IVIT_EPI_ACC32    out int32 [M,N] = acc + bias                       (QuantLinear alone)
 *   IVIT_EPI_REQUANT  out int8 = clamp(rint((acc+bias)*mult[n]), qmin, qmax)
 *                     (QuantLinear -> QuantAct: attn.qkv/qact1, mlp.fc1/qact_gelu, ...)
 *   IVIT_EPI_RESIDUAL c1 = clamp(rint((acc+bias)*mult[n]), qmin, qmax);
 *                     out int8 = clamp(rint(c1*mult2) + rint(id*mult_id), qmin2, qmax2)
 *                     (attn.proj -> attn.qact3 -> Block.qact2(identity), vit_quant.py:86-87,147;
 *                      mlp.fc2 -> mlp.qact2 -> Block.qact4(identity), :151-153;
 *                      patch_embed.proj -> patch_embed.qact -> qact1(identity=pos), :289-296)
 *   IVIT_EPI_LOGITS   out float32 = float(acc+bias) * scale_out[n]       (head, vit_quant.py:311)
 * Row remap (patch embed writes tokens 1..N-1 of every image, identity = positional codes):
 *   out row  = (r / group) * out_group_stride + (r % group) + out_row_off
 *   id  row  = id_broadcast ? (r % group) + out_row_off : out row
 * Pass group = 0 for the identity mapping. */
enum { IVIT_EPI_ACC32 = 0, IVIT_EPI_REQUANT = 1, IVIT_EPI_RESIDUAL = 2, IVIT_EPI_LOGITS = 3 };
typedef struct {
  int epilogue;
  const int32_t* bias;       /* [N] or NULL */
  const double* mult;        /* [N]  (REQUANT, RESIDUAL) */
  const float* s_in;         /* [N]  per-channel scale of (acc+bias): slow path for |z| >= 2^22 */
  int qmin, qmax;
  double mult2;              /* RESIDUAL */
  double mult_id;            /* RESIDUAL */
  int qmin2, qmax2;          /* RESIDUAL */
  const int8_t* identity;    /* RESIDUAL: int8, row stride ld_id bytes */
  int64_t ld_id;
  const float* scale_out;    /* LOGITS: [N] */
  int group, out_group_stride, out_row_off, id_broadcast;
  int force_slow;            /* 1: always take the general requant path (|mult| >= 256 or huge codes) */
  int z_bound_ok;            /* 1: caller proved |acc + bias| < 2^22 for every output (weights-only bound):
                              * the epilogue skips its per-element range check */
  /* 16-bit activation locations (vit_quant.py:180-187 *_bw = 16; quant_utils.py:247-251 clamp for k = 16):
   * out_bits 16 -> REQUANT / RESIDUAL store int16 codes (ldo in int16 elements); id_bits 16 -> the
   * RESIDUAL identity is int16 (ld_id in int16 elements).  0 or 8 = int8.  RESIDUAL's first clamp
   * (qmin, qmax) may then be a 16-bit range too (attention_out_bw / mlp_out_bw / patch_embed_bw). */
  int out_bits, id_bits;
  /* REQUANT only, optional: the GELU that follows the QuantAct (layers_quant.py:145-154: mlp.fc1 -> qact_gelu ->
   * act -> qact1), fused into the launch.  gelu_lut = the layer's [256][256] table of ivit_gelu_lut_build
   * (ShiftGELU ivit_modules.py:89-126 or I-BERT GELU ibert_modules.py:203-235, + mlp.qact1); the QuantAct codes
   * still go to `out`, GELU(codes) to gelu_out (int8 [M, N], row stride ldo; may be `out` itself).  One kernel
   * when the activation-stationary GEMM applies (K <= 768, int8 codes, z_bound_ok), else the GEMM followed by
   * the table pass.  NULL: no GELU. */
  const int8_t* gelu_lut;
  int8_t* gelu_out;
} ivit_gemm_epilogue_t;
int ivit_gemm_i8(const int8_t* A, int64_t lda, const int8_t* W, int M, int N, int K, void* out, int64_t ldo,
                 const ivit_gemm_epilogue_t* epi, void* stream);
/* ivit_gemm_i8 has three tcgen05 kernels: tokens on TMEM lanes (every epilogue), channels on TMEM lanes
 * (REQUANT / RESIDUAL hot shapes), and the activation-stationary form of the latter (REQUANT, K <= 768: a CTA
 * keeps a token tile in shared memory for all channel blocks; carries the fused GELU).  layout 0 = choose per
 * call (default), 1 = always the first, 2 = never the third.  Returns the previous setting.  Process-wide; for
 * A/B timing and tests. */
int ivit_gemm_set_layout(int layout);
/* Straightforward CUDA-core int8 GEMM with the same epilogues: on-device cross-check for
 * the tcgen05 kernel at sizes the CPU oracle cannot reach.  Not used by the product path. */
int ivit_gemm_i8_simple(const int8_t* A, int64_t lda, const int8_t* W, int M, int N, int K, void* out, int64_t ldo,
                        const ivit_gemm_epilogue_t* epi, void* stream);

/* ---- fused attention ---------------------------------------------------------------------
 * vit_quant.py:65-85: QuantMatMul(q,k^T) (quant_modules.py:404-409) -> *scale -> qact_attn1
 * -> IntSoftmax (kind 0: ivit_modules.py:150-179 Shiftmax; kind 1: ibert_modules.py:275-319)
 * -> QuantMatMul(attn, v) -> transpose(1,2).reshape -> attn.qact2.  Scores never leave the SM.
 * qkv int8 [B, N, 3, H, hd] (the layout QuantLinear(qkv)+qact1 produce); out int8 [B, N, H*hd]. */
typedef struct {
  int kind;                  /* 0 ivit, 1 ibert, 2 ppoly (ppoly_modules.py:296-342: the exp term only from a caller-built
                              * exp_table; 2^32 numerator, denominator clamped to >= 1) */
  int output_bit;            /* softmax output bits: 8, or 16 (softmax_bw = 16: codes up to 32768, u16 operand
                              * split into two u8 planes, two MMAs) */
  double mult_attn;          /* qact_attn1: m*2^-e of (s_q*s_k*scale)/s_attn */
  int qmin_attn, qmax_attn;
  float s_attn;              /* scale of qact_attn1 output = softmax input */
  double mult_out;           /* attn.qact2: (s_softmax*s_v)/s_out */
  int qmin_out, qmax_out;
  /* ibert only: inner QuantAct(16) of IBERTIntSoftmax (ibert_modules.py:262,310) */
  double mult_exp16;
  float s_exp16;
  /* optional device table [256][256] float32 built by ivit_softmax_table_build from the SAME
   * params: exp term indexed [row-max code + 128][code + 128].  NULL: computed per element. */
  const float* exp_table;
  /* optional device table [256] float32 from ivit_softmax_table_collapse: the same term indexed
   * [row-max code - code]; only valid when that call reported the 2-D table collapsible.  Takes
   * precedence over exp_table (the kernel keeps 16 copies in shared memory instead of gathering). */
  const float* exp_table_1d;
  /* output_bit 16 only: float32 scale of the P.V accumulator (s_softmax * s_v, quant_modules.py:409);
   * |acc| reaches 8e8 there, so attn.qact2 replays the float32 recovery round(fl32(acc*s)/s)
   * (quant_utils.py:220) before the dyadic multiply. */
  float s_pv;
  /* optional, with exp_table / exp_table_1d: the table's saturation.  IntSoftmax clamps its argument
   * (ivit_modules.py:155 max(x, n * x0); ibert_modules.py:288), so every entry whose row-max code exceeds the
   * code by sat_k or more holds the same value sat_e.  The caller finds (sat_k, sat_e) by scanning the built
   * table (bit for bit); the kernel then skips the lookup for those scores.  sat_k = 0: not used. */
  int sat_k;
  float sat_e;
} ivit_attention_params_t;
/* The integer-exp term of IntSoftmax depends only on (code, row-max code) of qact_attn1's int8
 * output: tabulate it once per frozen layer with the kernel's own arithmetic. */
int ivit_softmax_table_build(const ivit_attention_params_t* p, float* table65536, void* stream);
/* table256[d] = table65536[127+128][127-d+128]; *collapsible (device int) = 1 iff every valid
 * entry satisfies table65536[m][c] == table256[m - c] bit for bit, i.e. the exp term depends on
 * the difference of the two codes only (x/s recovers every int8 code exactly), else 0. */
int ivit_softmax_table_collapse(const float* table65536, float* table256, int* collapsible, void* stream);
int ivit_attention_i8(const int8_t* qkv, int8_t* out, int B, int N, int H, int hd,
                      const ivit_attention_params_t* p, void* stream);
/* The kernel runs 2 or 4 threads per query row (8 or 16 warps per CTA; default 2).  Returns the
 * previous setting; other values are ignored.  Process-wide; for A/B timing and tests. */
int ivit_attention_set_threads_per_row(int tpr);
/* Two kernels for an 8-bit softmax output: thread = query row throughout (default; also the 16-bit kernel), or
 * the softmax walking one query row per warp-step (every lane of a table lookup reads the same table row; more
 * instructions, measured slower).  variant 0 = choose per call (the first), 1 = always the first, 2 = the second
 * where it applies.  Returns the previous setting.  Process-wide; for A/B timing and tests. */
int ivit_attention_set_variant(int variant);
/* Debug taps of the same kernel: softmax codes [B,H,N,N] (what int_softmax emits): uint8, or
 * uint16 when output_bit = 16 (`probs` then points to uint16 storage). */
int ivit_attention_probs_i8(const int8_t* qkv, uint8_t* probs, int B, int N, int H, int hd,
                            const ivit_attention_params_t* p, void* stream);

/* ---- Swin (swin_quant.py) ------------------------------------------------------------------
 * WindowAttention.forward (:121-169) for every (image, window, head): matmul_1 -> * scale -> qact_attn1
 * -> qact2 with the quantised relative-position bias as identity (:140-147) -> + shifted-window mask
 * (0 / -100 added to the float carrier, :149-153; built on the device from :223-247) -> IntSoftmax
 * (Shiftmax, ivit_modules.py:150-179) -> matmul_2 -> qact3.  Roll and window partition / reverse
 * (:18-50, :258-289) are index maps inside the kernel: qkv int8 [B, res*res, 3, nH, hd] and out int8
 * [B, res*res, nH*hd] are in image token order.  bias_codes int8 [nH, ws*ws, ws*ws] = qact_table codes
 * gathered by relative_position_index.  probs (may be NULL): softmax codes [B*nW, nH, N, N] u8. */
typedef struct {
  double mult_attn;          /* qact_attn1: (s_q*s_k*scale) / s_attn1 */
  int qmin_attn, qmax_attn;
  double mult2, mult_bias;   /* qact2: s_attn1 / s2 and s_table / s2 (identity branch) */
  int qmin2, qmax2;
  float s2;                  /* scale of qact2's output = softmax input */
  const int8_t* bias_codes;
  int output_bit;            /* softmax output bits (<= 8) */
  double mult_out;           /* qact3: (s_softmax*s_v) / s3 */
  int qmin_out, qmax_out;
  /* optional: the [256][256] float32 table of ivit_softmax_table_build for kind 0, s_attn = s2 (exp term
   * over [row-max code + 128][code + 128]).  With a shift mask it is used only if mask_clamps != 0: the
   * caller has checked, in float32, that the largest masked score (code qmax2, carrier - 100) minus the
   * smallest possible row maximum (code qmin2) is negative and maps below 15 * floor(-1/s2) under
   * x + floor(x/2) - floor(x/16), so that int_exp_shift (ivit_modules.py:151-155) clamps every masked entry
   * and returns -floor(-1/s2).  NULL: computed per score. */
  const float* exp_table;
  int mask_clamps;
  /* optional int32 [nH, ws*ws, ws*ws]: the identity branch of qact2 evaluated once per layer,
   * rint(float64(bias_codes) * mult_bias) (quant_utils.py:232-245; batch independent).  NULL: per score. */
  const int32_t* bias_terms;
  /* 0: qkv / out in image token order [B, res*res, .] (roll, partition and reverse are index maps in the kernel).
   * 1: qkv / out already in window order [B*nW, ws*ws, .], i.e. what WindowAttention.forward itself receives and
   *    returns (swin_quant.py:121-169) after the caller's torch.roll + window_partition (:258-272); the shift mask is
   *    still generated from (res, ws, shift).  The module-by-module frozen path uses this form. */
  int windowed;
  /* optional, with exp_table: every table entry whose code lies at least sat_k below its row-max code equals sat_e bit
   * for bit (int_exp_shift clamps its argument, ivit_modules.py:155), so the kernel need not gather it.  sat_k = 0:
   * gather every entry. */
  int sat_k;
  float sat_e;
} ivit_window_attention_params_t;
int ivit_window_attention_i8(const int8_t* qkv, int8_t* out, uint8_t* probs, int B, int res, int ws, int shift, int nH,
                             int hd, const ivit_window_attention_params_t* p, void* stream);
/* Three kernels: thread-per-row with two (window, head) units per CTA (7x7 windows, head_dim 32, tabulated exp term:
 * the default where it applies), CTA per unit with the products on tcgen05 (head_dim 32), CTA per unit on dp4a (any
 * head_dim % 4 == 0).  variant 0 = choose per call, 1 = always dp4a, 2 = never the thread-per-row kernel.  Returns the
 * previous setting.  Process-wide; for A/B timing and tests. */
int ivit_window_attention_set_variant(int variant);
/* PatchMerging.forward's strided slices + cat (:339-344): [B, res, res, C] -> [B, (res/2)^2, 4C]. */
int ivit_patch_merge_gather(const void* x, void* out, int elem_bits, int B, int res, int C, void* stream);
/* AdaptiveAvgPool1d(1) over the tokens of the float carrier + qact3 (:554-555), torch's summation order:
 * x int8 [B, L, C] with scale s_in -> out int8 [B, C] = clamp(rint(rint(mean / s_in) * mult)). */
int ivit_avgpool_requant_i8(const int8_t* x, int8_t* out, int B, int L, int C, float s_in, double mult, int qmin,
                            int qmax, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* IVIT_B200_H */
