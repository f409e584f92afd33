// Host-side launch interface of the UFM engine kernels (all sm_100a, hand-written).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ufm {

// ------------------------------------------------------------------------------------------------
// Tensor maps (cuTensorMapEncodeTiled through cudaGetDriverEntryPoint; no link-time libcuda).
// dims/strides innermost first; strides in BYTES for dims 1..rank-1. bf16 elements, SWIZZLE_128B.
// ------------------------------------------------------------------------------------------------
bool make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box, char* err, int errlen);
// same for fp32 (is_f32) or bf16 elements
bool make_tmap(CUtensorMap* out, bool is_f32, const void* base, int rank, const uint64_t* dims,
               const uint64_t* strides_bytes, const uint32_t* box, char* err, int errlen);

// ------------------------------------------------------------------------------------------------
// tcgen05 GEMM / implicit-GEMM convolution.
//   out[n, y, x, :N] = epilogue( sum_{tap, c} A[n, y+dy(tap), x+dx(tap), c] * Wt[:, tap*cin_pad + c] )
// A is addressed through a 4-D tensor map (C, X, Y, Nimg) with box (64, tw, th, 1), tw*th == 128 (a plain
// row-major GEMM is X = M, Y = Nimg = 1, tw = 128).  Wt is a 2-D map (Ktotal, N) with box (64, BN).
// Epilogue per output element v (fp32 accumulator):
//   v += bias[col]; if GELU v = gelu_erf(v); v *= scale[col]; v += res1[...]; v += res2[...];
//   out_f32[...] = v;  out_bf16[...] = RELU_BF16 ? max(v,0) : v
// TAIL mode (DPT processor tail, dpt.py:273-277 + adaptors): v = relu(v + bias); y_o = sum_col v * w2[o][col] + b2[o];
//   y_o = affine / sigmoid adaptor; planar store out_f32[n][o][y][x] (and raw logits to out2 if given).
// ------------------------------------------------------------------------------------------------
enum : uint32_t {
  GEMM_GELU = 1u << 0,
  GEMM_RELU_BF16 = 1u << 1,
  GEMM_TAIL = 1u << 2,
  GEMM_TAIL_SIGMOID = 1u << 3,
  GEMM_TAIL_COORD = 1u << 4,   // TAIL: after the affine, subtract the pixel's (x, y) from channels (0, 1)
  // set by build_gemm only (EPI_TMA_F32, res1 == out_f32): the in-place residual add is done by the L2 through a TMA
  // reduce-add store (x += v) instead of loading the residual slab into the SM; one fp32 rounding either way
  GEMM_RED_INPLACE = 1u << 8,
  // set by build_gemm only (GEMM_GELU with the EPI_TMA_BF16 epilogue): GELU on packed f32x2 arithmetic, same bits
  GEMM_GELU_PACKED = 1u << 9,
  // set by build_gemm only (with GEMM_RED_INPLACE): the x slabs of a tile are prefetched into L2 under its main loop
  GEMM_RED_PREFETCH = 1u << 10,
};

// Epilogue paths of the GEMM kernel (chosen per launch by the host, engine.cu build_gemm):
//   EPI_DIRECT    one output row per thread, vector loads/stores straight to global (any combination of outputs/residuals,
//                 TAIL mode); costs 32 L1 wavefronts per warp instruction, so it paces short-K GEMMs
//   EPI_TMA_F32   fp32 output, optional fp32 residual with the same addressing (in-place residual stream): 32-column
//                 slabs staged in swizzled smem, residual slab in by TMA load, result out by TMA store
//   EPI_TMA_BF16  bf16 output only: 64-column slabs, out by TMA store
//   EPI_TMA_MIX   EPI_TMA_F32's slabs (res1 in by TMA load, fp32 result - if any - out by TMA store) plus a coalesced
//                 second pass over each slab for a res2 with its own strides and / or a bf16 copy of the result (the RCU
//                 convs of the DPT fusion blocks: two residuals, fp32 + ReLU'd bf16 outputs)
enum : int { EPI_DIRECT = 0, EPI_TMA_F32 = 1, EPI_TMA_BF16 = 2, EPI_TMA_MIX = 3 };

constexpr int kTailMax = 8;  // output channels of the fused TAIL epilogue
struct alignas(64) GemmParams {
  CUtensorMap tmA;
  CUtensorMap tmB;
  CUtensorMap tmOut;     // EPI_TMA_*: 4-D (N, X, Y, Nimg) over the output, box (32 fp32 | 64 bf16, bw, 32/bw, 1)
  CUtensorMap tmRes;     // EPI_TMA_F32 / EPI_TMA_MIX with res1: same geometry over the residual
  CUtensorMap tmBs;      // split_log2 > 0: tmB with a box of (BN / ctas) >> split_log2 rows (column slices of the last wave)
  int epi_mode;
  // tiles with index >= split_from (the last, partly filled wave of the static schedule) are computed as 2^split_log2
  // column slices each; split_log2 = 0: none
  int split_from, split_log2;
  // tile geometry
  int tw, th;            // A box extents, tw*th == 128, tw a power of two
  int tw_log2;
  int ctas;              // 1, or 2 = CTA-pair tiles (tmB box is then BN/2 rows)
  int halo;              // 1 = 3x3 conv with the halo-resident A operand: tw x th = 8 x 16, tmA box (64, 10, 18, 1)
  int rows2;             // 1 = every CTA of a pair owns TWO 128-row blocks (pair tile 512 x 256, one accumulator set):
                         //     the B half-tile in smem feeds both row blocks, -25 % L2 -> SM operand bytes per FLOP.
                         //     Plain GEMMs only (taps 1, ctas 2, BN 256, TMA epilogue without residual load)
  // rows2 launches: items [0, r2_items) are 512-row items (r2_items / tiles_n row units x tiles_n column tiles, whole waves of
  // the pair units); the rows behind them -- pair tiles (256 rows) r2_half_mt0 ... -- run as r2_half_tiles half items (row
  // block 0 / accumulator 0 only), each cut into 2^split_log2 column slices (tmBs), in the same launch
  int r2_items, r2_half_mt0, r2_half_tiles;
  int warp_shift;        // role of physical warp w = (w + warp_shift) % 12 (tuning: 4 puts TMA / MMA warps above the epilogue warps)
  int X, Y, Nimg;        // logical output extents per image
  int tiles_x, tiles_y;  // ceil(X/tw), ceil(Y/th)
  int N;                 // output channels (multiple of 32, or anything in TAIL mode with N <= 128)
  int tiles_n;           // ceil(N/BN)
  int taps;              // 1 or 9 (3x3, pad 1, stride 1)
  int cchunks;           // ceil(Cin/64)
  int cin_pad;           // K pitch of one tap in Wt (multiple of 64)
  uint32_t flags;
  // epilogue operands
  const float* bias;     // [N] or null
  const float* scale;    // [N] or null (LayerScale)
  const float* res1;     // same addressing as out (sN,sY,sX)
  const float* res2;     // own strides (r2N,r2Y,r2X); used for row-broadcast adds as well
  float* out_f32;
  __nv_bfloat16* out_bf16;
  long long sN, sY, sX;        // element strides of out / res1
  long long r2N, r2Y, r2X;     // element strides of res2
  long long out_off;           // element offset added to out / res1 addresses
  // TAIL
  const float* w2;       // [tail_out][N]
  float tail_b2[kTailMax];
  float tail_scale[kTailMax];
  float tail_shift[kTailMax];
  int tail_out;          // <= kTailMax
  float* out2;           // optional planar raw (logits) output
};

// BN in {128, 256}.  Returns cudaError from the launch.
cudaError_t launch_gemm(const GemmParams& p, int BN, int num_sms, cudaStream_t stream);
void set_gemm_epilogue_variant(int v);  // tuning hook: 1 = TMA-store epilogues where eligible (default), 0 = direct only
int gemm_epilogue_variant();
void set_gemm_pair_mode(int v);         // tuning hook: 1 = CTA-pair (cta_group::2) tiles where BN = 256 (default)
int gemm_pair_mode();
void set_gemm_halo_mode(int v);         // tuning hook: bit 0 = halo-resident A operand for the 3x3 convs (default on)
int gemm_halo_mode();
void set_gemm_split_mode(int v);        // tuning hook: 1 = column-slice the tiles of a partly filled last wave (default on)
int gemm_split_mode();
void set_gemm_rows2_mode(int v);        // tuning hook (ufm_debug_flag 6): bit 0 = 512 x 256 pair tiles for the large row-matrix GEMMs,
int gemm_rows2_mode();                  //   bit 1 = the next item's first row block starts before the second accumulator is drained
void set_gemm_gelu_packed(int v);       // tuning hook (ufm_debug_flag 9): 1 = packed f32x2 GELU in the bf16 TMA epilogue
int gemm_gelu_packed();
void set_gemm_red_prefetch(int v);      // tuning hook (ufm_debug_flag 10): 1 = L2 prefetch of the x slabs of an in-place reduce-add
int gemm_red_prefetch();
void set_gemm_warp_shift(int v);        // tuning hook (ufm_debug_flag 7): see GemmParams::warp_shift
int gemm_warp_shift();

// ------------------------------------------------------------------------------------------------
// Flash-style attention forward, head_dim 64, non-causal: O = softmax(Q K^T * scale) V per (sequence, head).
// qkv is the fused projection output [n_seq * seq_len, 3, heads, 64] bf16 (row pitch 3*heads*64);
// out is [n_seq * seq_len, heads * 64] bf16.  Optional per-head LayerNorm of q and k over head_dim
// (qk_norm, transformer_blocks.py:164) is applied by a separate kernel before the call.
// ------------------------------------------------------------------------------------------------
struct alignas(64) AttnParams {
  CUtensorMap tmQKV;     // 4-D (64, 3*heads, seq_len, n_seq), box (64, 1, 128, 1)
  int n_seq, seq_len, heads;
  float scale_log2;      // scale * log2(e)
  __nv_bfloat16* out;
  long long out_pitch;   // heads*64
  int n_qblk, n_items;   // filled by launch_attention: query tiles per sequence, work items = n_seq*heads*n_qblk
};
cudaError_t launch_attention(const AttnParams& p, cudaStream_t stream);
void set_attention_poly(int every);  // tuning switch (ufm_debug_flag 2): every n-th exp2 pair on the FMA pipe, 0 = none

// ------------------------------------------------------------------------------------------------
// SIMT kernels (bandwidth-bound stages)
// ------------------------------------------------------------------------------------------------
// LayerNorm over the last dim D (fp32 stats, eps) of x[rows, D] fp32 -> bf16 and/or fp32, with an optional row remap:
// source row r = seq*src_seq_len + src_skip + t (t < tokens), destination row = dst_row(seq, t) * dst_pitch + dst_col.
// dst sequence index = (seq % views_inner) * 0 ... see simt.cu for the exact mapping modes.
struct LNParams {
  const float* x; int D; float eps;
  const float* w; const float* b;
  int n_seq, src_seq_len, src_skip, tokens;   // rows processed: n_seq * tokens
  // destination sequence remap: dseq = (seq / swap_inner) + (seq % swap_inner) * swap_outer_stride ... (see kernel)
  int reverse;            // 1: process rows last-to-first (L2 reuse between consecutive kernels, see simt.cu)
  int remap_B;            // 0: dseq = seq.  B>0: seq = v*B + b  ->  dseq = b*2 + v   (encoder (view,b) -> info (b,view))
  __nv_bfloat16* out_bf16; long long ob_pitch; long long ob_col;   // element pitch/offset
  float* out_f32; long long of_pitch; long long of_col;
};
cudaError_t launch_layernorm(const LNParams& p, cudaStream_t stream);

// per-head LayerNorm of q and k (do_norm) and / or RoPE2D with base 100 (rope_n = tokens per view > 0, rope_w = patch
// grid width; pos_embed.py:108-155) inside the fused qkv buffer (in place, bf16), head_dim 64
cudaError_t launch_qk_norm(__nv_bfloat16* qkv, long long rows, int heads, const float* qw, const float* qb,
                           const float* kw, const float* kb, float eps, cudaStream_t stream, int do_norm = 1,
                           int rope_n = 0, int rope_w = 0);

// Patch im2col: image (uint8 or fp32, NHWC or NCHW) -> A[(img, py, px), Kpad] bf16 with
// k = c*P*P + ky*P + kx (conv weight flatten order), zero padded to Kpad.  uint8 pixels are normalised
// (x/255 - mean)/std (base.py:215-229) before the bf16 rounding, as under the reference's autocast.
struct PatchParams {
  const void* src; const void* tgt;   // each [B, ...]; image index = view*B + b
  int B, H, W, P, Kpad;
  int is_u8, is_nhwc;
  float mean[3], inv_std[3];
  __nv_bfloat16* out;
};
cudaError_t launch_patch_im2col(const PatchParams& p, cudaStream_t stream);

// x[img, 0, :] = cls + pos[0]   (fp32 residual stream, seq_len = 1 + tokens)
cudaError_t launch_cls_rows(float* x, const float* prefix, int n_prefix, int n_img, int seq_len, int D, cudaStream_t s);

// Pixel shuffle of a GEMM result: in[(n, y, x), (i, j, c)] -> out[n, y*k+i, x*k+j, c]   (ConvTranspose k==s and
// F.pixel_shuffle; the (i,j,c) column order is produced by host-side weight row permutation)
cudaError_t launch_shuffle_bf16(const __nv_bfloat16* in, __nv_bfloat16* out, int n, int h, int w, int k, int C, cudaStream_t s);
cudaError_t launch_shuffle_f32(const float* in, float* out, int n, int h, int w, int k, int C, cudaStream_t s);

// im2col for the 3x3 stride-2 pad-1 conv (dpt.py:163-169): in NHWC bf16 [n,h,w,C] -> out [(n,oy,ox), 9*C]
cudaError_t launch_im2col_s2(const __nv_bfloat16* in, __nv_bfloat16* out, int n, int h, int w, int C, int oh, int ow, cudaStream_t s);

// Bilinear resize, align_corners=True, NHWC: in (fp32 or bf16) [n,h,w,C] -> out [n,oh,ow,C] (crop to oh x ow of the
// resized (rh x rw) grid: scale = (h-1)/(rh-1)).
cudaError_t launch_resize_ac(const float* in_f32, const __nv_bfloat16* in_bf16, int n, int h, int w, int C, int rh, int rw,
                             int oh, int ow, float* out_f32, __nv_bfloat16* out_bf16, cudaStream_t s);

// Local classification refinement (ufm.py:1012-1178): desc NHWC fp32 [(b, view), H, W, C=16]; flow planar [B,2,H,W].
// desc: channel-group planar bf16 descriptors [2B images (b,view)][2][H][W][8]
struct RefineParams {
  const __nv_bfloat16* desc; int B, H, W;
  const float* flow_in;       // [B,2,H,W]
  const float* bias25; float inv_temperature;
  float* flow_out;            // [B,2,H,W] = flow_in + residual
  float* residual;            // optional [B,2,H,W]
  float* log_softmax;         // optional [B,H,W,25]
};
cudaError_t launch_refine(const RefineParams& p, cudaStream_t s);
// pixel_shuffle(k) of z[(img,ty,tx)][(i,j,c)] bf16 into desc[img][C/8][h*k][w*k][8]
cudaError_t launch_desc_shuffle(const __nv_bfloat16* z, __nv_bfloat16* desc, int n, int h, int w, int k, int C, cudaStream_t s);

// ------------------------------------------------------------------------------------------------
// Pre/post-processing for non-native input sizes (prepost.cu; flow_resizing.py:276-354, 749-877, 955-1010)
// ------------------------------------------------------------------------------------------------
struct PrepParams {
  const void* in; int is_u8, nhwc;     // [B,H0,W0,3] or [B,3,H0,W0], uint8 or fp32
  int B, H0, W0, H, W;
  float scale_y, scale_x;              // H0 / H, W0 / W as fp32 (ATen area_pixel_compute_scale)
  float mean[3], std[3];               // u8: (v/255 - mean) / std
  float mul[3], add[3];                // fp32: v * mul + add (identity or re-normalisation)
  float* out;                          // [B,3,H,W] normalised fp32
};
cudaError_t launch_prep_resize(const PrepParams& p, cudaStream_t s);
struct UnmapParams {
  const float* flow; const float* cov; // [B,2,H,W], [B,H,W] (may be null) at the inference resolution
  int B, H, W, H0, W0;
  int rep[4], src0[4], src1[4];        // [top, bottom, left, right]: represented region, source regions of image 0 / 1
  float src_mul[2], tgt_mul[2];        // (x, y): source / target region size over represented region size
  float* flow_out; float* cov_out;     // [B,2,H0,W0], [B,H0,W0] (may be null); flow/flow_out may be null (channels only)
  float cov_mul[4]; int cov_mul_n;     // cov_out plane b is scaled by cov_mul[b % cov_mul_n] (covariance ratios, base.py:311-315)
};
cudaError_t launch_unmap(const UnmapParams& p, cudaStream_t s);
// uncertainty-head adaptors (adaptors.py:899-985): raw [B,C,H,W] -> per-adaptor outputs
enum { ADAPT_MASK = 1, ADAPT_COV_EXP_TANH = 2, ADAPT_CONF_EXP = 3, ADAPT_CONF_SIGMOID = 4, ADAPT_CONF_SOFTMAX = 5 };
struct UncAdaptParams {
  const float* raw; int B, C, H, W;
  int n, kind[4], ch[4]; float vmin[4], vmax[4];
  float* mask; float* cov; float* cov_inv; float* cov_log_det; float* conf; float* conf_stats;
};
cudaError_t launch_unc_adaptors(const UncAdaptParams& p, cudaStream_t s);

// ------------------------------------------------------------------------------------------------
// UNet fine-feature branch of UFM-Refine (unet.cu; unet_encoder.py:26-72, ufm.py:817-825, 967-982)
// ------------------------------------------------------------------------------------------------
struct UnetStemParams {
  const void* src; const void* tgt;    // each [B, ...] uint8 / fp32, NHWC / NCHW; output image index = b*2 + view
  int B, H, W, is_u8, is_nhwc;
  float mean[3], inv_std[3];
  const float* w; const float* b;      // downs.0.conv.0: [64][3][3][3] fp32, [64]
  __nv_bfloat16* out;                  // [2B, H, W, 64] = relu(conv3x3(normalised image))
};
cudaError_t launch_unet_stem(const UnetStemParams& p, cudaStream_t s);
// 2x2/s2 max pool of the first C channels of in[n,H,W,in_pitch] -> out[n,H/2,W/2,C] (dense)
cudaError_t launch_unet_pool(const __nv_bfloat16* in, long long in_pitch, int n, int H, int W, int C, __nv_bfloat16* out,
                             cudaStream_t s);
// ConvTranspose(k=s=2) GEMM result g[(n,y,x)][(ky,kx,c)] -> out[n,oy,ox,out_off + c] (pixel pitch out_pitch), nearest
// resize from (2h, 2w) to (oh, ow) as F.interpolate does when the skip is larger (unet_encoder.py:66-67)
cudaError_t launch_unet_up(const __nv_bfloat16* g, int n, int h, int w, int C, __nv_bfloat16* out, long long out_pitch,
                           int out_off, int oh, int ow, cudaStream_t s);
struct UnetCombineParams {
  const __nv_bfloat16* x0;             // [n, H, W, 64] last DoubleConv output
  const __nv_bfloat16* desc_cls;       // classification features, planar [n][2][H][W][8]
  __nv_bfloat16* desc;                 // combined descriptors, same layout
  float* unet_feat;                    // optional [n][16][H][W] fp32 (final_conv output, parity stage)
  int n, H, W, method;                 // method 1 = "conv", 2 = "modulate"
  const float *wf, *bf;                // final_conv transposed [64][16], [16]
  const float *w1, *b1;                // conv1 transposed [32][32], [32] (method 1)
  const float *w2, *b2;                // conv2 transposed [32 | 16][16], [16]
};
cudaError_t launch_unet_combine(const UnetCombineParams& p, cudaStream_t s);

// NHWC fp32 [n,H,W,C] -> NCHW fp32 (stage dumps / descriptor export)
cudaError_t launch_nhwc_to_nchw(const float* in, float* out, int n, int H, int W, int C, cudaStream_t s);

// early programmatic-launch trigger (launch.cuh), one constant per translation unit
cudaError_t ufm_set_pdl_early_gemm(int v, cudaStream_t s);
cudaError_t ufm_set_pdl_early_attention(int v, cudaStream_t s);
cudaError_t ufm_set_pdl_early_simt(int v, cudaStream_t s);
cudaError_t ufm_set_pdl_early_prepost(int v, cudaStream_t s);
cudaError_t ufm_set_pdl_early_unet(int v, cudaStream_t s);

}  // namespace ufm
