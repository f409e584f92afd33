/*
 * libufm_b200 — C ABI of the B200-native UFM dense-correspondence engine.
 *
 * The reference (labrat97/UFM) is pure PyTorch and has no FFI of its own; its boundary for this path is the Python
 * method  UniFlowMatchModelsBase.predict_correspondences_batched()  (uniflowmatch/models/base.py:137-234) which runs
 * UniFlowMatchConfidence.forward (uniflowmatch/models/ufm.py:562-662) or
 * UniFlowMatchClassificationRefinement.forward (ufm.py:843-1009).  The entry points below are what a ctypes binding of
 * that method binds: plain pointers and sizes, int status returns (0 = ok), no exceptions across
 * the ABI, caller owns I/O buffers, the engine owns weights and workspace, one engine per GPU, calls on one engine are
 * stream-ordered and must not overlap.
 */
#ifndef UFM_B200_H
#define UFM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ufm_engine ufm_engine;

enum { UFM_F32 = 0, UFM_BF16 = 1, UFM_U8 = 2 };
enum { UFM_NCHW = 0, UFM_NHWC = 1 };

/* Architecture of one model instance (the reference's constructor kwargs, ufm.py:483-509 / 720-754, reduced to the
 * numbers the engine needs).  All spatial sizes are those of the inference resolution (base.py:86-100). */
typedef struct ufm_config {
  int img_h, img_w, patch;                 /* H, W (multiples of patch), patch size 14 */
  /* encoder: DINOv2 ViT (UC/models/encoders/dinov2.py) */
  int enc_dim, enc_depth, enc_heads;       /* depth after keep_first_n_layers */
  int enc_take_first, enc_take_last;       /* block indices whose normed output is returned (first only used by Refine) */
  /* info sharing (UC/models/info_sharing/{global,alternating}_attention_transformer.py) */
  int info_dim, info_depth, info_heads, info_hidden;
  int info_take0, info_take1;              /* indices[0], indices[1] */
  int info_alternating;                    /* 0: global attention over both views; 1: alternating (odd layers attend
                                            * within a view); 2: per-view cross-attention branches */
  int info_qk_norm, info_layerscale, info_has_proj;
  /* DPT heads (UC/models/prediction_heads/dpt.py) */
  int dpt_layer_dims[4];
  int dpt_feature_dim, dpt_hidden;
  int flow_out_ch;                         /* 2 */
  int unc_out_ch;                          /* 0 = no uncertainty head, else the sum of its adaptors' channels (<= 8) */
  float flow_scale[2], flow_shift[2];      /* FlowAdaptor: flow = raw*scale + shift (adaptors.py:77-120) */
  /* Refine (ufm.py:710-1178) */
  int refine;                              /* 0/1 */
  int cls_in, cls_hidden, cls_out, cls_patch, refine_P;
  float temperature;
  /* uncertainty-head adaptors in AdaptorMap order (ufm.py:553-556, UC/models/prediction_heads/adaptors.py:899-985).
   * unc_n_adaptors == 0 with unc_out_ch == 1 means the single 'non_occluded_mask' MaskAdaptor. */
  int unc_n_adaptors;
  int unc_kind[4];                         /* UFM_ADAPT_* */
  float unc_vmin[4], unc_vmax[4];          /* ConfidenceAdaptor vmin / vmax */
  /* UNet fine-feature branch of UFM-Refine (use_unet_feature, ufm.py:817-825, 916-918, 967-982; the UNet is
   * UNet(3, 16, [64,128,256,512]), unet_encoder.py:26-72): 0 = off, 1 = feature_combine_method "conv",
   * 2 = "modulate".  Needs refine = 1. */
  int unet;
  /* custom_positional_encoding="rope" of the global-attention info sharing (global_attention_transformer.py:110-113:
   * RoPE2D(freq=100, F0=1) on q and k of every block, positions = the patch grid of each view). 0/1. */
  int info_rope;
  /* FlowAdaptor.output_normalized_coordinate (adaptors.py:110-118): flow = raw * flow_scale + flow_shift - (x, y) of the
   * pixel, with flow_scale = flow_shift = (W/2, H/2).  0/1. */
  int flow_coord;
  /* norm_intermediate of the encoder (dinov2.py:259-261: applies to EVERY returned index, the last one included) and of
   * the info-sharing transformer (global_attention_transformer.py:549, alternating:666, cross:449: intermediates only,
   * the final features are always normed).  1 = apply the final LayerNorm (the default), 0 = return the raw residual
   * stream (a plain fp32 -> bf16 conversion pass). */
  int enc_norm_intermediate, info_norm_intermediate;
  /* (mean, std) of the encoder's data_norm_type (UC/models/encoders/image_normalizations.py:17-35) applied to uint8
   * inputs as (v / 255 - mean) / std (base.py:215-229).  All zero = the "dinov2" constants. */
  float img_mean[3], img_std[3];
  /* DPTFeature.hooks of the flow head and of the uncertainty head (dpt.py:47,198-219): input stage k of the head reads
   * entry hooks[k] of [encoder last, info intermediate 0, info intermediate 1, info final].  hooks_set = 0: [0,1,2,3]. */
  int hooks_set;
  int dpt_hooks[4], unc_hooks[4];
  /* register tokens of the hub dinov2_vit*14_reg models (with_registers=True, dinov2.py:24,49): 0 or 4 rows between the cls
   * row and the patch rows of every image, dropped with cls from the returned features. */
  int enc_registers;
} ufm_config;

/* adaptor kinds and their channel counts: mask 1 ('non_occluded_mask'), covariance 3 ('flow_cov'), confidence 1
 * ('keypoint_confidence').  Outputs other than the mask are read back with ufm_stage_copy: "flow_cov", "flow_cov_inv"
 * [B,3,H,W], "flow_cov_log_det" [B,1,H,W], "keypoint_confidence" [B,H,W]. */
enum { UFM_ADAPT_MASK = 1, UFM_ADAPT_COV_EXP_TANH = 2, UFM_ADAPT_CONF_EXP = 3, UFM_ADAPT_CONF_SIGMOID = 4,
       UFM_ADAPT_CONF_SOFTMAX = 5 };

/* Create an engine on CUDA device `device`.  Fails (non-zero) if no sm_100 GPU is present: there is no CPU fallback. */
int ufm_create(const ufm_config* cfg, int device, ufm_engine** out);
void ufm_destroy(ufm_engine* e);
const char* ufm_last_error(ufm_engine* e);

/* Hand one packed weight tensor to the engine (copied; the caller may free its buffer).  `name` is the engine's
 * packed-weight name (ufm_b200/packing.py maps the reference state_dict keys, SURVEY.md 8b, onto these names).
 * `ptr` may be a host or a device pointer. */
int ufm_set_weight(ufm_engine* e, const char* name, const void* ptr, int dtype, const int64_t* shape, int ndim);

/* Allocate workspace for up to `max_pairs` pairs per call and validate that every weight is present. */
int ufm_prepare(ufm_engine* e, int max_pairs);

/* One forward pass over `pairs` image pairs already resident in device memory.
 *   src, tgt : [pairs, H, W, 3] (UFM_NHWC) or [pairs, 3, H, W] (UFM_NCHW); UFM_U8 raw pixels (normalised inside as
 *              base.py:215-229) or UFM_F32 already normalised with the encoder's data_norm_type
 *   flow_out : [pairs, 2, H, W] fp32, x component first (Refine: after local refinement)
 *   covis_out: [pairs, H, W] fp32 in [0,1], or NULL (ignored when the model has no uncertainty head)
 * `stream` is a cudaStream_t; the call is asynchronous on it. */
int ufm_forward(ufm_engine* e, const void* src, const void* tgt, int pairs, int in_dtype, int in_layout,
                float* flow_out, float* covis_out, void* stream);

/* Same, with HOST buffers: pinned-or-pageable host inputs are copied H2D, results copied D2H, and the call returns
 * after the stream has drained.  This is the end-to-end entry measured as `e2e` by bench.py. */
int ufm_forward_host(ufm_engine* e, const void* src, const void* tgt, int pairs, int in_dtype, int in_layout,
                     float* flow_out, float* covis_out);

/* Ask the next forward calls to also keep optional Refine outputs (residual [B,2,H,W], log_softmax [B,H,W,P*P]). */
int ufm_set_keep_aux(ufm_engine* e, int keep);

/* Row of info_sharing.view_pos_table that the NON-reference view adds in the following passes (the reference view always
 * adds row 0).  Default 1.  With use_rand_idx_pe_for_non_reference_views the reference draws the row per forward with
 * torch.randint(1, max_num_views) (global_attention_transformer.py:520-524; alternating: :281-286); the host mirror draws
 * the same number from the same generator and hands it over here.  row == number of table rows selects the all-zero row
 * the engine appends (alternating attention without use_pe_for_non_reference_views). */
int ufm_set_view_index(ufm_engine* e, int row);

/* Parity hooks: describe / copy an internal stage buffer of the LAST forward (device-to-device or device-to-host).
 * Names: enc_first, enc_last, info_inter0, info_inter1, info_final, dpt1, dptu, head1_raw, headu_raw, flow_reg,
 * desc, residual, log_softmax.  Layouts are the engine's own (documented in DESIGN.md). */
int ufm_stage_desc(ufm_engine* e, const char* name, int* dtype, int64_t* shape, int* ndim);
int ufm_stage_copy(ufm_engine* e, const char* name, void* dst, int64_t bytes);

/* Number of kernels launched by the last ufm_forward call. */
int ufm_last_launch_count(ufm_engine* e);

/* Per-kernel-family device timing for bench.py's roofline (no reference counterpart; the reference's profiling aid is
 * UC/models/encoders/utils.py:11-86).  While profiling is on, every launch is bracketed by CUDA events on the forward's
 * stream; ufm_profile_read drains the stream and returns, per family, the summed kernel durations [ms], the algorithmic
 * FLOPs (tensor-bound families) or algorithmic HBM bytes (bandwidth-bound families) and the launch count accumulated
 * since the last read.  Arrays have UFM_NUM_CATS entries. */
enum { UFM_CAT_GEMM = 0,   /* gemm_kernel, taps == 1: Linear / 1x1 conv / folded ConvT */
       UFM_CAT_CONV = 1,   /* gemm_kernel, taps == 9: implicit-GEMM 3x3 conv */
       UFM_CAT_ATTN = 2,   /* attn_kernel */
       UFM_CAT_LN = 3,     /* layernorm_kernel, qk_norm_kernel */
       UFM_CAT_RESIZE = 4, /* resize_ac_kernel */
       UFM_CAT_MOVE = 5,   /* patch im2col, cls rows, pixel shuffles, stride-2 im2col */
       UFM_CAT_REFINE = 6, /* refine_kernel */
       UFM_NUM_CATS = 7 };
int ufm_set_profile(ufm_engine* e, int on);
int ufm_profile_read(ufm_engine* e, double* ms, double* flops, double* bytes, int64_t* launches);

/* Process-wide tuning switches for kernel experiments (no reference counterpart).  which = 0: GEMM epilogue variant
 * (0 = row-per-thread direct epilogue everywhere; 1 = TMA slab epilogues where the output geometry allows, including the
 * mixed second-residual / dual-output mode and the TMA reduce-add in-place residual (default); 2 = as 1 without the mixed mode;
 * | 4 = in-place residuals through a TMA-loaded residual slab instead of the reduce-add store); which = 1: CTA-pair (cta_group::2) GEMM
 * tiles, bit 0 = for the 256-column tiles, bit 1 = also for the 128-column tiles (default 3; -1 = default / UFM_PAIR); which = 2: attention, every n-th exp2 pair on the FMA pipe; which = 3: 1 = halo-resident A operand
 * for the 3x3 convolutions (default), 0 = nine shifted boxes; which = 4: 1 = programmatic dependent launches (default); which = 5: 1 = column slices for
 * the tiles of a GEMM's partly filled last wave (default); which = 6: bit 0 = 512 x 256 pair tiles (two row blocks per
 * CTA, one accumulator set) for the whole waves of the large row-matrix GEMMs, bit 1 = the next item's first row block
 * runs ahead while the second accumulator is drained, bit 2 = also for GELU epilogues, bit 3 = only K >= 2048 or (K >= 1024
 * and N >= 2048), bit 4 = only K >= 2048 (default 27; -1 = back to the default / UFM_ROWS2; the rows behind the last whole
 * wave of 512-row items run as column-sliced 256-row half items of the same launch); which = 7: warp-role
 * shift of the GEMM kernel (0 default; 4 = TMA / MMA warps above the epilogue warps); which = 8: 1 = passes of <= 4 pairs
 * run their independent branches (uncertainty head, classification head + UNet branch, the four input chains of each DPT
 * head) on side streams beside the flow head (default; bit-identical results), 0 = one stream; which = 9: 1 = GELU of the
 * bf16 TMA epilogue on packed f32x2 arithmetic (default; same bits as the scalar form), -1 = default / UFM_GELU_PACKED;
 * which = 10: 1 = L2 prefetch of the residual slabs of an in-place reduce-add tile (measured slower, default 0).
 * Returns non-zero for an unknown switch. */
int ufm_debug_flag(int which, int value);

/* ---- kernel-level entry points (unit parity tests; device pointers, asynchronous on `stream`) ---- */

/* out = epilogue(A (*) Wt): see ufm_b200/csrc/kernels.h GemmParams.  A is bf16 with logical extents (C, X, Y, Nimg) and
 * element strides (1, a_sx, a_sy, a_sn); Wt is bf16 [n_out, taps*cin_pad]. */
int ufm_op_gemm(const void* a, int C, int X, int Y, int Nimg, int64_t a_sx, int64_t a_sy, int64_t a_sn, int tw, int th,
                const void* wt, int n_out, int taps, int cin_pad, const float* bias, const float* scale,
                const float* res1, const float* res2, int64_t r2n, int64_t r2y, int64_t r2x, float* out_f32,
                void* out_bf16, int64_t sn, int64_t sy, int64_t sx, int64_t out_off, uint32_t flags,
                const float* tail_w2, const float* tail_b2, const float* tail_scale, const float* tail_shift,
                int tail_out, float* out2, void* stream, char* err, int errlen);
int ufm_op_attention(const void* qkv, int n_seq, int seq_len, int heads, float scale, void* out, void* stream,
                     char* err, int errlen);
int ufm_op_layernorm(const float* x, int n_seq, int src_seq_len, int src_skip, int tokens, int D, const float* w,
                     const float* b, float eps, int remap_B, void* out_bf16, int64_t ob_pitch, int64_t ob_col,
                     float* out_f32, int64_t of_pitch, int64_t of_col, void* stream);
int ufm_op_qk_norm(void* qkv, int64_t rows, int heads, const float* qw, const float* qb, const float* kw,
                   const float* kb, float eps, void* stream);
/* RoPE2D (UC/models/libs/croco/pos_embed.py:108-155, base 100) on the q and k columns of a fused qkv buffer, in place:
 * token (y, x) = ((row mod tokens_per_view) / grid_w, (row mod tokens_per_view) mod grid_w). */
int ufm_op_rope(void* qkv, int64_t rows, int heads, int tokens_per_view, int grid_w, void* stream);
int ufm_op_resize_ac(const float* in_f32, const void* in_bf16, int n, int h, int w, int C, int rh, int rw, int oh,
                     int ow, float* out_f32, void* out_bf16, void* stream);
int ufm_op_shuffle(const void* in, void* out, int is_f32, int n, int h, int w, int k, int C, void* stream);
int ufm_op_im2col_s2(const void* in, void* out, int n, int h, int w, int C, int oh, int ow, void* stream);
/* Non-native input sizes.  ufm_op_prep_resize replaces the normalisation of
 * predict_correspondences_batched (uniflowmatch/models/base.py:215-229) fused with ResizeToFixedManipulation.__call__
 * (uniflowmatch/utils/flow_resizing.py:276-354: bilinear, align_corners=False, antialias=True): `in` is [B,H0,W0,3]
 * (nhwc) or [B,3,H0,W0], uint8 (mean/std used) or fp32 (v*mul+add used; NULL = identity); `out` is normalised fp32
 * [B,3,H,W].  ufm_op_unmap replaces unmap_predicted_flow (flow_resizing.py:749-877) and unmap_predicted_channels
 * (flow_resizing.py:955-1010): regions are [top,bottom,left,right]; src_mul/tgt_mul are (x,y) region-size ratios.  flow /
 * flow_out may be NULL (channels only); plane b of cov_out is multiplied by cov_mul[b % cov_mul_n] (NULL = 1), which is the
 * covariance rescaling of base.py:295-319 when the three covariance planes are passed as 3B single-channel images. */
int ufm_op_prep_resize(const void* in, int is_u8, int nhwc, int B, int H0, int W0, int H, int W, const float* mean,
                       const float* std, const float* mul, const float* add, float* out, void* stream);
int ufm_op_unmap(const float* flow, const float* cov, int B, int H, int W, int H0, int W0, const int* rep, const int* src0,
                 const int* src1, const float* src_mul, const float* tgt_mul, float* flow_out, float* cov_out,
                 const float* cov_mul, int cov_mul_n, void* stream);
/* desc: bf16 descriptors, channel-group planar [2B images (b,view)][2][H][W][8] (16 channels) */
int ufm_op_refine(const void* desc, int B, int H, int W, const float* flow_in, const float* bias25,
                  float inv_temperature, float* flow_out, float* residual, float* log_softmax, void* stream);


/* 3x3 convolution, stride 2, padding 1 (the im2col_s2_kernel + gemm_kernel pair of the DPT head's last input stage) that
 * ALLOCATES NOTHING: no cudaMalloc / cudaFree and no allocating library routine on any call, so it may be used under CUDA
 * graph capture, from a serving loop or from a harness that forbids allocation.  It has no counterpart in the reference.
 * The im2col operand, which ufm_op_im2col_s2 + ufm_op_gemm keep in a
 * caller-allocated buffer, lives in a file-scope `static __device__` array of UFM_CONV3X3_S2_STATIC_MAX_BYTES that every
 * GPU loading the library holds until the process exits (with CUDA's lazy module loading the 16 MiB appear at the
 * first call on a device, which is also when the array's device address is taken; later calls change nothing).
 *   in      : bf16 [n, h, w, C] (NHWC), C a multiple of 64
 *   wt      : bf16 [n_out, 9 * C], taps row-major (ky, kx) then channels (the conv weight permuted to [n_out, 3, 3, C])
 *   bias    : fp32 [n_out] or NULL
 *   out_f32 : fp32 [n, (h-1)/2+1, (w-1)/2+1, n_out]
 * Works on the CURRENT device (all pointers must belong to it), asynchronous on `stream`.  Returns 2 with a message in
 * `err` when n * oh * ow * 9 * C * 2 bytes exceed the maximum, 1 on any other error.  All calls on one GPU share the
 * array: calls that overlap on the GPU (different streams, no ordering between them) race on it; calls on one stream
 * are safe.  Every element a call reads was written by that call, so nothing is carried over from the previous one. */
#define UFM_CONV3X3_S2_STATIC_MAX_BYTES (16u << 20)
int ufm_op_conv3x3_s2_static(const void* in, int n, int h, int w, int C, const void* wt, int n_out, const float* bias,
                             float* out_f32, void* stream, char* err, int errlen);

#ifdef __cplusplus
}
#endif
#endif /* UFM_B200_H */
