// Bandwidth-bound SIMT kernels of the UFM engine (sm_100a): LayerNorm, patch im2col, pixel shuffles, strided
// im2col, align_corners bilinear resampling, the fused local classification refinement, layout export.
// All are coalesced along the channel / pixel axis with 128-bit accesses; none has data reuse worth a TMA stage
// except the refinement gather, which relies on L1 for the overlapping 8x8 windows of neighbouring pixels.
#include <math.h>
#include <type_traits>

#include <stdlib.h>

#include "kernels.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace ufm {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------------ LayerNorm
// One warp per row, the row held in registers (D <= 1024, D % 128 == 0): fp32 two-pass statistics as nn.LayerNorm
// (transformer_blocks.py norm1/norm2, final `norm` on returned layers: global_attention_transformer.py:546-550,586).
constexpr int LN_MAXV = 8;
constexpr int LN_BLOCKS_PER_SM = 2;
struct LNRow {
  float4 v[LN_MAXV];
};

__device__ __forceinline__ void ln_load(const LNParams& p, long long rows, long long r, int nv, int lane, LNRow& a) {
  // rows are walked from the last to the first: the kernel before this one (a residual GEMM) wrote x in ascending row
  // order and x (fp32) is larger than L2, so the rows written last are the ones still resident; and the GEMM after this
  // one reads the normalised rows in ascending order, i.e. starts with what was written last here
  const long long row = p.reverse ? rows - 1 - r : r;
  const int seq = (int)(row / p.tokens);
  const int t = (int)(row - (long long)seq * p.tokens);
  const float4* xr = reinterpret_cast<const float4*>(p.x + ((long long)seq * p.src_seq_len + p.src_skip + t) * p.D) + lane;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (i < nv) a.v[i] = __ldg(xr + i * 32);
}

__device__ __forceinline__ void ln_row(const LNParams& p, long long rows, long long r, int nv, int lane, const LNRow& a) {
  const long long row = p.reverse ? rows - 1 - r : r;
  const int seq = (int)(row / p.tokens);
  const int t = (int)(row - (long long)seq * p.tokens);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (i < nv) s += a.v[i].x + a.v[i].y + a.v[i].z + a.v[i].w;
  const float mean = warp_sum(s) / p.D;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (i < nv) {
      const float e = a.v[i].x - mean, f = a.v[i].y - mean, g = a.v[i].z - mean, h = a.v[i].w - mean;
      q += e * e + f * f + g * g + h * h;
    }
  const float rstd = rsqrtf(warp_sum(q) / p.D + p.eps);
  int dseq = seq;
  if (p.remap_B > 0) {
    const int vv = seq / p.remap_B, bb = seq - vv * p.remap_B;
    dseq = bb * 2 + vv;
  }
  const long long drow = (long long)dseq * p.tokens + t;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (i < nv) {
      const int col = (i * 32 + lane) * 4;
      float4 y = a.v[i];  // p.w == nullptr: nn.Identity in place of the norm (a dtype / layout conversion only)
      if (p.w) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(p.w + col));
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.b + col));
        y.x = (a.v[i].x - mean) * rstd * w.x + b.x;
        y.y = (a.v[i].y - mean) * rstd * w.y + b.y;
        y.z = (a.v[i].z - mean) * rstd * w.z + b.z;
        y.w = (a.v[i].w - mean) * rstd * w.w + b.w;
      }
      if (p.out_bf16) {
        uint2 o = make_uint2(pack_bf16(y.x, y.y), pack_bf16(y.z, y.w));
        *reinterpret_cast<uint2*>(p.out_bf16 + drow * p.ob_pitch + p.ob_col + col) = o;
      }
      if (p.out_f32) *reinterpret_cast<float4*>(p.out_f32 + drow * p.of_pitch + p.of_col + col) = y;
    }
}

// Persistent: LN_BLOCKS_PER_SM blocks per SM, every warp strides over the rows with the next row's loads in flight
// while the current one is reduced and stored (two register row buffers, loop unrolled by two so they never move).
// 38432 x 1024 fp32 -> bf16 in the ncu launch list: 42.1-42.8 us (5.5 TB/s, 84 % of the measured copy bandwidth)
// against 43.2-44.4 us for the earlier one-row-per-warp grid of 4804 short-lived blocks (profiles/r01_v5 / r01_v6
// launches): a small gain, the kernel was already close to the HBM roofline; what the in-step figure (4.4-4.5 TB/s)
// adds is the launch gap around a 40 us kernel.  Three blocks per SM (85 registers, spills) and one block per SM are
// both slower (61 / 66 us event-timed against 49).
__global__ void __launch_bounds__(256, LN_BLOCKS_PER_SM) layernorm_kernel(const LNParams p) {
  pdl_wait();
  const long long rows = (long long)p.n_seq * p.tokens;
  const long long nw = (long long)gridDim.x * (blockDim.x >> 5);
  long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nv = p.D >> 7;
  LNRow a, b;
  ln_load(p, rows, r, nv, lane, a);
  for (;;) {
    long long rn = r + nw;
    if (rn < rows) ln_load(p, rows, rn, nv, lane, b);
    ln_row(p, rows, r, nv, lane, a);
    if (rn >= rows) break;
    r = rn;
    rn = r + nw;
    if (rn < rows) ln_load(p, rows, rn, nv, lane, a);
    ln_row(p, rows, r, nv, lane, b);
    if (rn >= rows) break;
    r = rn;
  }
}

// per-head LayerNorm of q and k (head_dim 64) and / or 2-D RoPE, in place inside the fused qkv buffer; one warp per
// (row, head, q|k), two elements per lane.
// RoPE2D (UC/models/libs/croco/pos_embed.py:108-155; curope/kernels.cu computes the same): the first 32 dims rotate by
// the token's grid row y, the last 32 by its column x; inside a 32-dim half, element i < 16 pairs with i + 16:
//   out[i] = t[i] cos(a_i) - t[i+16] sin(a_i),  out[i+16] = t[i+16] cos(a_i) + t[i] sin(a_i),  a_i = pos * 100^(-i/16).
// Token (y, x) = ((row mod rope_n) / rope_w, (row mod rope_n) mod rope_w): every view carries the same patch grid
// (global_attention_transformer.py:499-503).
__global__ void __launch_bounds__(256) qk_norm_kernel(__nv_bfloat16* qkv, long long rows, int heads, const float* qw,
                                                     const float* qb, const float* kw, const float* kb, float eps,
                                                     int do_norm, int rope_n, int rope_w) {
  pdl_wait();
  const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long total = rows * heads * 2;
  if (wid >= total) return;
  const int lane = threadIdx.x & 31;
  const int which = (int)(wid % 2);
  const int head = (int)((wid / 2) % heads);
  const long long row = wid / (2 * heads);
  __nv_bfloat162* ptr =
      reinterpret_cast<__nv_bfloat162*>(qkv + row * (3LL * heads * 64) + (long long)which * heads * 64 + head * 64) + lane;
  float2 v = __bfloat1622float2(*ptr);
  if (do_norm) {
    const float mean = warp_sum(v.x + v.y) * (1.f / 64.f);
    const float a = v.x - mean, b = v.y - mean;
    const float rstd = rsqrtf(warp_sum(a * a + b * b) * (1.f / 64.f) + eps);
    const float* w = which ? kw : qw;
    const float* bb = which ? kb : qb;
    v = make_float2(a * rstd * w[2 * lane] + bb[2 * lane], b * rstd * w[2 * lane + 1] + bb[2 * lane + 1]);
  }
  if (rope_n > 0) {
    const int t = (int)(row % rope_n);
    const float pos = (float)((lane & 16) ? t % rope_w : t / rope_w);
    const float ox = __shfl_xor_sync(0xffffffffu, v.x, 8), oy = __shfl_xor_sync(0xffffffffu, v.y, 8);
    const int j = 2 * (lane & 7);
    float s0, c0, s1, c1;
    sincosf(pos * exp2f(-(float)j * (6.643856189774724f / 16.f)), &s0, &c0);        // 100^(-j/16)
    sincosf(pos * exp2f(-(float)(j + 1) * (6.643856189774724f / 16.f)), &s1, &c1);
    const float sg = (lane & 8) ? 1.f : -1.f;  // first half of the pair: - partner * sin, second half: + partner * sin
    v = make_float2(v.x * c0 + sg * ox * s0, v.y * c1 + sg * oy * s1);
  }
  *ptr = __floats2bfloat162_rn(v.x, v.y);
}

// ------------------------------------------------------------------------------------------------ patch im2col
__global__ void __launch_bounds__(256) patch_im2col_kernel(const PatchParams p) {
  pdl_wait();
  const int hp = p.H / p.P, wp = p.W / p.P;
  const int kgroups = p.Kpad >> 3;
  const long long total = 2LL * p.B * hp * wp * kgroups;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int kg = (int)(idx % kgroups);
  const long long row = idx / kgroups;
  const int px = (int)(row % wp);
  const int py = (int)((row / wp) % hp);
  const int img = (int)(row / ((long long)wp * hp));
  const int view = img / p.B, b = img - view * p.B;
  const void* base = view ? p.tgt : p.src;
  const int PP = p.P * p.P;
  const int K = 3 * PP;
  float f[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = kg * 8 + i;
    float val = 0.f;
    if (k < K) {
      const int c = k / PP;
      const int rem = k - c * PP;
      const int ky = rem / p.P, kx = rem - ky * p.P;
      const int y = py * p.P + ky, x = px * p.P + kx;
      const long long off = p.is_nhwc ? (((long long)b * p.H + y) * p.W + x) * 3 + c
                                      : (((long long)b * 3 + c) * p.H + y) * p.W + x;
      if (p.is_u8) {
        const float u = (float)reinterpret_cast<const uint8_t*>(base)[off];
        val = (u / 255.0f - p.mean[c]) * p.inv_std[c];
      } else {
        val = reinterpret_cast<const float*>(base)[off];
      }
    }
    f[i] = val;
  }
  *reinterpret_cast<uint4*>(p.out + row * p.Kpad + kg * 8) =
      make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// uint8 images, 14 x 14 patches (every UFM configuration): one warp per patch row of the operand, 32-bit index arithmetic
// with compile-time divisors, the normalisation through a 3 x 256-entry table built per block with the arithmetic of the
// generic kernel above (same bits), blocks walking PATCH_PER_WARP patches per warp so that the table is amortised.  The
// generic kernel spent its time in 64-bit divisions and an fp32 division per element (154 us at B = 16 for 22.6 MB in /
// 49 MB out).
constexpr int PATCH_PER_WARP = 8;
template <int P, bool NHWC>
__global__ void __launch_bounds__(256) patch_im2col_u8_kernel(const PatchParams p) {
  __shared__ float lut[3][256];
  for (int i = threadIdx.x; i < 768; i += 256) {
    const int c = i >> 8;
    lut[c][i & 255] = ((float)(i & 255) / 255.0f - p.mean[c]) * p.inv_std[c];
  }
  __syncthreads();
  pdl_wait();
  constexpr int PP = P * P, K = 3 * PP;
  const int hp = p.H / P, wp = p.W / P;
  const int n_tok = hp * wp;
  const int rows = 2 * p.B * n_tok;
  const int lane = threadIdx.x & 31;
  const int kgroups = p.Kpad >> 3;
  int row = (blockIdx.x * 8 + (threadIdx.x >> 5)) * PATCH_PER_WARP;
  for (int it = 0; it < PATCH_PER_WARP && row < rows; ++it, ++row) {
    const int img = row / n_tok, t = row - img * n_tok;
    const int py = t / wp, px = t - py * wp;
    const int view = img / p.B, b = img - view * p.B;
    const uint8_t* base = reinterpret_cast<const uint8_t*>(view ? p.tgt : p.src);
    // first pixel of the patch
    const uint8_t* b0 = NHWC ? base + (((size_t)b * p.H + py * P) * p.W + px * P) * 3
                             : base + ((size_t)b * 3 * p.H + py * P) * p.W + px * P;
    const size_t plane = (size_t)p.H * p.W;
    for (int kg = lane; kg < kgroups; kg += 32) {
      float f[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = kg * 8 + i;
        float val = 0.f;
        if (k < K) {
          const int c = k / PP, rem = k - c * PP;
          const int ky = rem / P, kx = rem - ky * P;
          const uint8_t u = NHWC ? b0[((size_t)ky * p.W + kx) * 3 + c] : b0[c * plane + (size_t)ky * p.W + kx];
          val = lut[c][u];
        }
        f[i] = val;
      }
      *reinterpret_cast<uint4*>(p.out + (size_t)row * p.Kpad + kg * 8) =
          make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
    }
  }
}

// rows in front of the patch tokens of every image: `prefix` [n_prefix, D] = (cls + its positional row), then the register
// tokens of the hub *_reg models (no positional embedding), precomputed at pack time
__global__ void cls_rows_kernel(float* x, const float* prefix, int n_prefix, int n_img, int seq_len, int D) {
  pdl_wait();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = n_prefix * D;
  if (idx >= n_img * per) return;
  const int img = idx / per, r = idx - img * per;
  x[(long long)img * seq_len * D + r] = prefix[r];
}

// ------------------------------------------------------------------------------------------------ pixel shuffle
// I = index type: unsigned 32-bit when every element index fits (the usual case; 64-bit divisions are software routines
// and made this copy issue-bound at a third of the HBM rate), long long otherwise
template <typename T, int VEC, typename I = long long>
__global__ void __launch_bounds__(256) shuffle_kernel(const T* in, T* out, int n, int h, int w, int k, int C) {
  pdl_wait();
  const int cg = C / VEC;
  const I total = (I)n * h * w * k * k * cg;
  const I idx = (I)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  // idx enumerates the INPUT contiguously: (row=(n,y,x), i, j, cgroup)
  const int g = (int)(idx % cg);
  const int j = (int)((idx / cg) % k);
  const int i = (int)((idx / ((I)cg * k)) % k);
  const I row = idx / ((I)cg * k * k);
  const int x = (int)(row % w);
  const int y = (int)((row / w) % h);
  const int nn = (int)(row / ((I)w * h));
  typedef typename std::conditional<sizeof(T) * VEC == 16, uint4, uint2>::type V;
  const V val = *reinterpret_cast<const V*>(in + idx * VEC);
  const I o = ((((I)nn * h * k + (y * k + i)) * (w * k)) + (x * k + j)) * C + g * VEC;
  *reinterpret_cast<V*>(out + o) = val;
}

// ------------------------------------------------------------------------------------------------ 3x3 s2 im2col
template <typename I>
__global__ void __launch_bounds__(256) im2col_s2_kernel(const __nv_bfloat16* in, __nv_bfloat16* out, int n, int h, int w,
                                                       int C, int oh, int ow) {
  pdl_wait();
  const int cg = C >> 3;
  const I total = (I)n * oh * ow * 9 * cg;
  const I idx = (I)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int g = (int)(idx % cg);
  const int tap = (int)((idx / cg) % 9);
  const I row = idx / ((I)cg * 9);
  const int ox = (int)(row % ow);
  const int oy = (int)((row / ow) % oh);
  const int nn = (int)(row / ((I)ow * oh));
  const int y = 2 * oy + tap / 3 - 1, x = 2 * ox + tap % 3 - 1;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (y >= 0 && y < h && x >= 0 && x < w)
    v = *reinterpret_cast<const uint4*>(in + (((I)nn * h + y) * w + x) * C + g * 8);
  *reinterpret_cast<uint4*>(out + idx * 8) = v;
}

// ------------------------------------------------------------------------------------------------ bilinear (ac=True)
// PyTorch upsample_bilinear2d, align_corners=True (dpt_block.py:251-253, dpt.py:304): src = dst*(in-1)/(out-1).
// Grid (x-chunks, oy, n): no 64-bit index arithmetic in the kernel (a flat 64-bit index costs four software divisions
// per thread and made the first version issue-bound at 20 % of HBM).  Each thread produces 16 bytes' worth of input
// channels (4 fp32 / 8 bf16) of one output pixel; a warp covers consecutive channels, then consecutive pixels.
template <int N>
struct Vec { float v[N]; };
// wa * a + wb * b with the contraction spelled out (one rounding of wa * a, then one fused multiply-add), so that every
// resize kernel produces the same bits whatever the compiler would have chosen to fuse
__device__ __forceinline__ float blend2(float wa, float a, float wb, float b) { return __fmaf_rn(wb, b, __fmul_rn(wa, a)); }
__device__ __forceinline__ Vec<4> ldv(const float* p) {
  const float4 t = __ldg(reinterpret_cast<const float4*>(p));
  return Vec<4>{{t.x, t.y, t.z, t.w}};
}
__device__ __forceinline__ Vec<8> ldv(const __nv_bfloat16* p) {
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
  const float2 c = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.z));
  const float2 d = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.w));
  return Vec<8>{{a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y}};
}
// One thread = one 16-byte channel group of a STRIP of RS_STRIP consecutive output pixels of one output row: the
// vertically interpolated source columns are kept in registers and shared by consecutive outputs (x2 upsampling
// touches a new source column only every other output pixel), and the index arithmetic is paid once per strip.  The
// first version (one output pixel per thread, four loads each) was issue-bound at 30 % of HBM with 75-80 % issue
// utilisation (ncu, profiles/r01_v4_simt_refine16.csv).
constexpr int RS_STRIP = 8;
template <typename T>
__global__ void __launch_bounds__(256) resize_ac_kernel(const T* __restrict__ in, int h, int w, int C, float sy, float sx,
                                                       int oh, int ow, float* __restrict__ out_f32,
                                                       __nv_bfloat16* __restrict__ out_bf16) {
  pdl_wait();
  constexpr int V = 16 / sizeof(T);
  const unsigned cg = (unsigned)C / V;
  const unsigned t = blockIdx.x * 256u + threadIdx.x;
  const unsigned strip = t / cg;
  const unsigned ox0 = strip * RS_STRIP;
  if (ox0 >= (unsigned)ow) return;
  const unsigned g = t - strip * cg;
  const int oy = blockIdx.y, nn = blockIdx.z;
  const float fy = sy * oy;
  const int y0 = (int)fy;
  const int y1 = y0 + (y0 < h - 1 ? 1 : 0);
  const float ly = fy - y0, hy = 1.f - ly;
  const T* r0 = in + ((size_t)nn * h + y0) * w * C + g * V;
  const T* r1 = in + ((size_t)nn * h + y1) * w * C + g * V;
  auto col = [&](int x) {
    const Vec<V> a = ldv(r0 + (size_t)x * C), b = ldv(r1 + (size_t)x * C);
    Vec<V> v;
#pragma unroll
    for (int i = 0; i < V; ++i) v.v[i] = blend2(hy, a.v[i], ly, b.v[i]);
    return v;
  };
  int c0 = -1, c1 = -1;
  Vec<V> v0, v1;
  size_t o = (((size_t)nn * oh + oy) * ow + ox0) * C + g * V;
#pragma unroll
  for (int k = 0; k < RS_STRIP; ++k, o += C) {
    const int ox = ox0 + k;
    if (ox >= ow) break;
    const float fx = sx * ox;
    const int x0 = (int)fx;
    const int x1 = x0 + (x0 < w - 1 ? 1 : 0);
    const float lx = fx - x0, hx = 1.f - lx;
    if (x0 != c0) {
      if (x0 == c1) v0 = v1;
      else v0 = col(x0);
      c0 = x0;
    }
    if (x1 != c1) {
      if (x1 == c0) v1 = v0;
      else v1 = col(x1);
      c1 = x1;
    }
    float r[V];
#pragma unroll
    for (int i = 0; i < V; ++i) r[i] = blend2(hx, v0.v[i], lx, v1.v[i]);
    if (out_f32) {
#pragma unroll
      for (int i = 0; i < V; i += 4) *reinterpret_cast<float4*>(out_f32 + o + i) = make_float4(r[i], r[i + 1], r[i + 2], r[i + 3]);
    }
    if (out_bf16) {
      if constexpr (V == 8)
        *reinterpret_cast<uint4*>(out_bf16 + o) =
            make_uint4(pack_bf16(r[0], r[1]), pack_bf16(r[2], r[3]), pack_bf16(r[4], r[5]), pack_bf16(r[6], r[7]));
      else
        *reinterpret_cast<uint2*>(out_bf16 + o) = make_uint2(pack_bf16(r[0], r[1]), pack_bf16(r[2], r[3]));
    }
  }
}

// Upsampling fast path (7 * sx < 4: the eight outputs of a strip touch at most five source columns, six with the +1
// neighbour).  All twelve source loads of the strip are issued up front (the generic kernel above fetches a column only
// when an output first needs it: five dependent round trips per thread, long-scoreboard-bound at 50-60 % of HBM with
// 146 instructions per 16-byte output, ncu simt_v6), the columns are blended vertically once, and the outputs are then
// emitted source column by source column, so every register index is static and nothing is moved or re-tested.
// Arithmetic (fy, fx, the two blends and their order) is the generic kernel's: results are bit-identical.
constexpr int RS_NC = 6;
template <typename T>
__global__ void __launch_bounds__(256, 3) resize_ac_up_kernel(const T* __restrict__ in, int h, int w, int C, float sy, float sx,
                                                             int oh, int ow, float* __restrict__ out_f32,
                                                             __nv_bfloat16* __restrict__ out_bf16) {
  pdl_wait();
  constexpr int V = 16 / sizeof(T);
  const unsigned cg = (unsigned)C / V;
  const unsigned t = blockIdx.x * 256u + threadIdx.x;
  const unsigned strip = t / cg;
  const int ox0 = (int)(strip * RS_STRIP);
  if (ox0 >= ow) return;
  const unsigned g = t - strip * cg;
  const int oy = blockIdx.y, nn = blockIdx.z;
  const float fy = sy * oy;
  const int y0 = (int)fy;
  const int y1 = y0 + (y0 < h - 1 ? 1 : 0);
  const float ly = fy - y0, hy = 1.f - ly;
  const T* r0 = in + ((size_t)nn * h + y0) * w * C + g * V;
  const T* r1 = in + ((size_t)nn * h + y1) * w * C + g * V;
  const int xb = (int)(sx * ox0);
  Vec<V> a[RS_NC], b[RS_NC];
#pragma unroll
  for (int j = 0; j < RS_NC; ++j) {
    const int x = min(xb + j, w - 1);
    a[j] = ldv(r0 + (size_t)x * C);
    b[j] = ldv(r1 + (size_t)x * C);
  }
#pragma unroll
  for (int j = 0; j < RS_NC; ++j)
#pragma unroll
    for (int i = 0; i < V; ++i) a[j].v[i] = blend2(hy, a[j].v[i], ly, b[j].v[i]);
  const int kend = min(RS_STRIP, ow - ox0);
  size_t o = (((size_t)nn * oh + oy) * ow + ox0) * C + g * V;
  int k = 0;
  float fx = sx * ox0;
  int x0 = xb;
#pragma unroll
  for (int j = 0; j < RS_NC - 1; ++j) {
    while (k < kend && x0 == xb + j) {
      const float lx = fx - x0, hx = 1.f - lx;
      float r[V];
#pragma unroll
      for (int i = 0; i < V; ++i) r[i] = blend2(hx, a[j].v[i], lx, a[j + 1].v[i]);
      if (out_f32) {
#pragma unroll
        for (int i = 0; i < V; i += 4) *reinterpret_cast<float4*>(out_f32 + o + i) = make_float4(r[i], r[i + 1], r[i + 2], r[i + 3]);
      }
      if (out_bf16) {
        if constexpr (V == 8)
          *reinterpret_cast<uint4*>(out_bf16 + o) =
              make_uint4(pack_bf16(r[0], r[1]), pack_bf16(r[2], r[3]), pack_bf16(r[4], r[5]), pack_bf16(r[6], r[7]));
        else
          *reinterpret_cast<uint2*>(out_bf16 + o) = make_uint2(pack_bf16(r[0], r[1]), pack_bf16(r[2], r[3]));
      }
      ++k;
      o += C;
      fx = sx * (ox0 + k);
      x0 = (int)fx;
    }
  }
}

// ------------------------------------------------------------------------------------------------ local refinement
// ufm.py:1012-1178 fused: one thread per source pixel.  Because the 25 bicubic samples of a pixel share one fractional
// offset and the score is linear in the sampled key, score[a][b] = sum_ty sum_tx wy[ty] wx[tx] D[a+ty][b+tx] with
// D[r][s] = <q, F2[y0-3+r][x0-3+s]> (zero outside the image = grid_sample padding_mode="zeros" per tap):
// 64 dot products of 16 channels, then a separable 8x8 -> 5x5 cubic filter, softmax over 25 and the expectation of
// the (dj, di) offsets.  Nothing is materialised (the reference writes 376 MB/pair here).
__device__ __forceinline__ void cubic_w(float t, float (&w)[4]) {
  const float A = -0.75f;
  const float x0 = t + 1.f, x1 = t, x2 = 1.f - t, x3 = 2.f - t;
  w[0] = ((A * x0 - 5.f * A) * x0 + 8.f * A) * x0 - 4.f * A;
  w[1] = ((A + 2.f) * x1 - (A + 3.f)) * x1 * x1 + 1.f;
  w[2] = ((A + 2.f) * x2 - (A + 3.f)) * x2 * x2 + 1.f;
  w[3] = ((A * x3 - 5.f * A) * x3 + 8.f * A) * x3 - 4.f * A;
}

// Descriptor maps are stored channel-group planar, [image][C/8][H][W][8] bf16: the 32 pixels of a warp then read 32
// consecutive 16-byte chunks per load instruction (4 L1 wavefronts) instead of 32 chunks 64 bytes apart (16 wavefronts)
// -- the gather of 64 taps x 16 channels per pixel is bound by L1 wavefronts, not by HBM.
// PRECISION NOTE: bf16 descriptors are NARROWER than the reference.  The classification head, the UNet combine and the
// refinement sit inside `with torch.autocast("cuda", torch.float32)` (ufm.py:921; :635 for the Base heads), which
// DISABLES autocast: the reference computes them in fp32 (TF32 matmuls / cuDNN convs on a GPU).  The measured cost
// (tests/test_gpu_vs_reference.py): on ufm_refine_assumed the engine is 2.0e-2 rel-L2 from the
// reference's fp32 run where the reference's own GPU recipe is 1.4e-2 (both have ~4 px max-abs outliers where the
// 25-way softmax flips); the refinement arithmetic itself is within 5e-4 px of the oracle on the same descriptors.
#ifndef REFINE_MIN_BLOCKS
#define REFINE_MIN_BLOCKS 6
#endif
// c + a.lo*b.lo + a.hi*b.hi for two packed bf16 pairs, accumulated in fp32 by the sm_100 mixed-precision FMA
// (fma.rn.f32.bf16 -> FHFMA.BF16 with .H0/.H1 operand selectors): the bf16 x bf16 products are exact in fp32, so this
// equals an fp32 dot product of the converted values without spending ALU slots on unpacking.
__device__ __forceinline__ float dot2_bf16(uint32_t a, uint32_t b, float c) {
  float d;
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1;\n\t"
      "mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.bf16 %0, al, bl, %3;\n\t"
      "fma.rn.f32.bf16 %0, ah, bh, %0;\n\t}"
      : "=&f"(d)
      : "r"(a), "r"(b), "f"(c));
  return d;
}

template <bool INTERIOR>
__device__ __forceinline__ void refine_rows(const uint4* __restrict__ f2, long long plane, int W, int H, int x0, int y0,
                                            const uint4 qa, const uint4 qb, const float (&wx)[4], const float (&wy)[4],
                                            float (&sc)[25]) {
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const int yy = y0 + r;
    if (yy < 0 || yy >= H) continue;  // the whole window row is padding: contributes zeros
    const uint4* row = f2 + ((long long)yy * W + x0);
    float D[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      float dd = 0.f;
      if (INTERIOR || (x0 + s >= 0 && x0 + s < W)) {
        const uint4 k0 = __ldg(row + s), k1 = __ldg(row + plane + s);
        float d0 = dot2_bf16(k0.x, qa.x, 0.f), d1 = dot2_bf16(k1.x, qb.x, 0.f);
        d0 = dot2_bf16(k0.y, qa.y, d0);
        d1 = dot2_bf16(k1.y, qb.y, d1);
        d0 = dot2_bf16(k0.z, qa.z, d0);
        d1 = dot2_bf16(k1.z, qb.z, d1);
        d0 = dot2_bf16(k0.w, qa.w, d0);
        d1 = dot2_bf16(k1.w, qb.w, d1);
        dd = d0 + d1;
      }
      D[s] = dd;
    }
    // E[bb] = sum_tx wx[tx] D[bb+tx]; window row r feeds score rows a = r-3 .. r with weight wy[r-a]
#pragma unroll
    for (int bb = 0; bb < 5; ++bb) {
      const float E = wx[0] * D[bb] + wx[1] * D[bb + 1] + wx[2] * D[bb + 2] + wx[3] * D[bb + 3];
#pragma unroll
      for (int a = 0; a < 5; ++a)
        if (r - a >= 0 && r - a < 4) sc[a * 5 + bb] = fmaf(wy[r - a], E, sc[a * 5 + bb]);
    }
  }
}

__global__ void __launch_bounds__(128, REFINE_MIN_BLOCKS) refine_kernel(const RefineParams p) {
  pdl_wait();
  const int x = blockIdx.x * 32 + (threadIdx.x & 31);
  const int y = blockIdx.y * 4 + (threadIdx.x >> 5);
  const int b = blockIdx.z;
  if (x >= p.W || y >= p.H) return;
  const long long plane = (long long)p.H * p.W;
  const uint4* d = reinterpret_cast<const uint4*>(p.desc);
  const uint4* f1 = d + (long long)(2 * b) * 2 * plane + (long long)y * p.W + x;  // group 0 of view 1; group 1 at +plane
  const uint4* f2 = d + (long long)(2 * b + 1) * 2 * plane;
  const uint4 qa = __ldg(f1), qb = __ldg(f1 + plane);  // the query stays packed: 8 registers
  const float fx = p.flow_in[(2LL * b) * plane + (long long)y * p.W + x];
  const float fy = p.flow_in[(2LL * b + 1) * plane + (long long)y * p.W + x];
  // same float sequence as the reference: normalise with align_corners=False then grid_sample's unnormalise
  const float gx = (((float)x + fx) + 0.5f) / (float)p.W * 2.f - 1.f;
  const float gy = (((float)y + fy) + 0.5f) / (float)p.H * 2.f - 1.f;
  const float ix = ((gx + 1.f) * (float)p.W - 1.f) * 0.5f;
  const float iy = ((gy + 1.f) * (float)p.H - 1.f) * 0.5f;
  const float flx = floorf(ix), fly = floorf(iy);
  float wx[4], wy[4];
  cubic_w(ix - flx, wx);
  cubic_w(iy - fly, wy);
  // clamp far-away targets so the integer window arithmetic cannot overflow; such windows are all-zero anyway
  const int x0 = (int)fminf(fmaxf(flx, -16.f), (float)p.W + 16.f) - 3;
  const int y0 = (int)fminf(fmaxf(fly, -16.f), (float)p.H + 16.f) - 3;
  float sc[25];
#pragma unroll
  for (int k = 0; k < 25; ++k) sc[k] = 0.f;
  // interior windows (all 8 columns inside the image) take a path without per-tap bounds predicates: one row pointer,
  // 16 loads at immediate offsets; border windows keep the predicated path (zero padding per tap, grid_sample "zeros")
  if (x0 >= 0 && x0 + 7 < p.W) refine_rows<true>(f2, plane, p.W, p.H, x0, y0, qa, qb, wx, wy, sc);
  else refine_rows<false>(f2, plane, p.W, p.H, x0, y0, qa, qb, wx, wy, sc);
  float mx = -INFINITY;
#pragma unroll
  for (int k = 0; k < 25; ++k) {
    sc[k] = sc[k] * p.inv_temperature + __ldg(p.bias25 + k);
    mx = fmaxf(mx, sc[k]);
  }
  float sum = 0.f, rx = 0.f, ry = 0.f;
#pragma unroll
  for (int k = 0; k < 25; ++k) {
    const float e = __expf(sc[k] - mx);
    sum += e;
    rx += e * (float)(k % 5 - 2);
    ry += e * (float)(k / 5 - 2);
  }
  const float inv = 1.f / sum;
  rx *= inv;
  ry *= inv;
  const long long o0 = (2LL * b) * plane + (long long)y * p.W + x, o1 = o0 + plane;
  p.flow_out[o0] = fx + rx;
  p.flow_out[o1] = fy + ry;
  if (p.residual) {
    p.residual[o0] = rx;
    p.residual[o1] = ry;
  }
  if (p.log_softmax) {
    const float lse = mx + __logf(sum);
    float* ls = p.log_softmax + ((long long)b * plane + (long long)y * p.W + x) * 25;
#pragma unroll
    for (int k = 0; k < 25; ++k) ls[k] = sc[k] - lse;
  }
}

// pixel_shuffle(k) of the classification head output (mlp_feature.py:99) into the channel-group planar descriptor
// layout: z[(img, ty, tx)][(i, j, c)] bf16 -> desc[img][c/8][ty*k+i][tx*k+j][c%8]; one 16-byte chunk per thread,
// input read contiguously.
template <int K, int CG>
__global__ void __launch_bounds__(256) desc_shuffle_kernel(const uint4* __restrict__ z, uint4* __restrict__ desc, int n, int h,
                                                          int w, int k_rt, int cg_rt) {
  pdl_wait();
  // one block row (blockIdx.y) = one token (img, ty, tx); threads enumerate its k*k*cg chunks contiguously.  K, CG > 0:
  // compile-time constants (divisions by constants); K == 0: runtime values.
  const int k = K ? K : k_rt, cg = CG ? CG : cg_rt;
  const int per_row = k * k * cg;
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= per_row) return;
  const unsigned row = blockIdx.y + blockIdx.z * gridDim.y;  // token index
  if (row >= (unsigned)(n * h * w)) return;
  const int g = c % cg;
  const int j = (c / cg) % k;
  const int i = c / (cg * k);
  const unsigned tx = row % (unsigned)w;
  const unsigned ty = (row / (unsigned)w) % (unsigned)h;
  const unsigned img = row / (unsigned)(w * h);
  const long long H = (long long)h * k, W = (long long)w * k;
  desc[(((long long)img * cg + g) * H + (ty * k + i)) * W + (tx * k + j)] = z[(long long)row * per_row + c];
}

__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const float* in, float* out, int n, int H, int W, int C) {
  pdl_wait();
  const long long total = (long long)n * H * W * C;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // output index (n,c,y,x)
  if (idx >= total) return;
  const int x = (int)(idx % W);
  const int y = (int)((idx / W) % H);
  const int c = (int)((idx / ((long long)W * H)) % C);
  const int nn = (int)(idx / ((long long)W * H * C));
  out[idx] = in[(((long long)nn * H + y) * W + x) * C + c];
}

inline unsigned blocks_for(long long total, int per) { return (unsigned)((total + per - 1) / per); }
inline int sm_count() {  // of the current device (persistent SIMT grids); cached per device
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int g_pdl = -1;  // -1: not decided yet (UFM_NO_PDL in the environment turns programmatic launches off)

}  // namespace

int pdl_mode() {
  if (g_pdl < 0) g_pdl = getenv("UFM_NO_PDL") ? 0 : 1;
  return g_pdl;
}
void set_pdl_mode(int v) { g_pdl = v ? 1 : 0; }

cudaError_t launch_layernorm(const LNParams& p, cudaStream_t stream) {
  if ((p.D & 127) || p.D > 128 * LN_MAXV) return cudaErrorInvalidValue;
  const long long rows = (long long)p.n_seq * p.tokens;
  if (rows == 0) return cudaSuccess;
  static const int per_sm = getenv("UFM_LN_BLOCKS") ? atoi(getenv("UFM_LN_BLOCKS")) : LN_BLOCKS_PER_SM;  // tuning knob
  const unsigned need = blocks_for(rows, 8), cap = per_sm > 0 ? (unsigned)(sm_count() * per_sm) : need;
  launch_k(layernorm_kernel, need < cap ? need : cap, 256, 0, stream, p);
  return cudaGetLastError();
}

cudaError_t launch_qk_norm(__nv_bfloat16* qkv, long long rows, int heads, const float* qw, const float* qb,
                           const float* kw, const float* kb, float eps, cudaStream_t stream, int do_norm, int rope_n,
                           int rope_w) {
  if (!do_norm && rope_n <= 0) return cudaSuccess;
  if (rope_n > 0 && rope_w <= 0) return cudaErrorInvalidValue;
  const long long warps = rows * heads * 2;
  launch_k(qk_norm_kernel, blocks_for(warps, 8), 256, 0, stream, qkv, rows, heads, qw, qb, kw, kb, eps, do_norm, rope_n, rope_w);
  return cudaGetLastError();
}

cudaError_t launch_patch_im2col(const PatchParams& p, cudaStream_t stream) {
  const long long rows = 2LL * p.B * (p.H / p.P) * (p.W / p.P);
  if (p.is_u8 && p.P == 14 && rows * p.Kpad < (1LL << 31) && !getenv("UFM_PATCH_GENERIC")) {
    const unsigned grid = blocks_for(rows, 8 * PATCH_PER_WARP);
    if (p.is_nhwc) launch_k(patch_im2col_u8_kernel<14, true>, grid, 256, 0, stream, p);
    else launch_k(patch_im2col_u8_kernel<14, false>, grid, 256, 0, stream, p);
    return cudaGetLastError();
  }
  const long long total = 2LL * p.B * (p.H / p.P) * (p.W / p.P) * (p.Kpad / 8);
  launch_k(patch_im2col_kernel, blocks_for(total, 256), 256, 0, stream, p);
  return cudaGetLastError();
}

cudaError_t launch_cls_rows(float* x, const float* prefix, int n_prefix, int n_img, int seq_len, int D, cudaStream_t s) {
  launch_k(cls_rows_kernel, blocks_for((long long)n_img * n_prefix * D, 256), 256, 0, s, x, prefix, n_prefix, n_img, seq_len, D);
  return cudaGetLastError();
}

cudaError_t launch_shuffle_bf16(const __nv_bfloat16* in, __nv_bfloat16* out, int n, int h, int w, int k, int C,
                                cudaStream_t s) {
  if (C % 8) return cudaErrorInvalidValue;
  const long long total = (long long)n * h * w * k * k * (C / 8);
  if (total * 8 < (1LL << 31)) launch_k(shuffle_kernel<__nv_bfloat16, 8, unsigned>, blocks_for(total, 256), 256, 0, s, in, out, n, h, w, k, C);
  else launch_k(shuffle_kernel<__nv_bfloat16, 8>, blocks_for(total, 256), 256, 0, s, in, out, n, h, w, k, C);
  return cudaGetLastError();
}
cudaError_t launch_shuffle_f32(const float* in, float* out, int n, int h, int w, int k, int C, cudaStream_t s) {
  if (C % 4) return cudaErrorInvalidValue;
  const long long total = (long long)n * h * w * k * k * (C / 4);
  if (total * 4 < (1LL << 31)) launch_k(shuffle_kernel<float, 4, unsigned>, blocks_for(total, 256), 256, 0, s, in, out, n, h, w, k, C);
  else launch_k(shuffle_kernel<float, 4>, blocks_for(total, 256), 256, 0, s, in, out, n, h, w, k, C);
  return cudaGetLastError();
}

cudaError_t launch_im2col_s2(const __nv_bfloat16* in, __nv_bfloat16* out, int n, int h, int w, int C, int oh, int ow,
                             cudaStream_t s) {
  if (C % 8) return cudaErrorInvalidValue;
  const long long total = (long long)n * oh * ow * 9 * (C / 8);
  if (total * 8 < (1LL << 31) && (long long)n * h * w * C < (1LL << 31))
    launch_k(im2col_s2_kernel<unsigned>, blocks_for(total, 256), 256, 0, s, in, out, n, h, w, C, oh, ow);
  else launch_k(im2col_s2_kernel<long long>, blocks_for(total, 256), 256, 0, s, in, out, n, h, w, C, oh, ow);
  return cudaGetLastError();
}

cudaError_t launch_resize_ac(const float* in_f32, const __nv_bfloat16* in_bf16, int n, int h, int w, int C, int rh, int rw,
                             int oh, int ow, float* out_f32, __nv_bfloat16* out_bf16, cudaStream_t s) {
  if (C % (in_f32 ? 4 : 8) || oh > 65535 || n > 65535) return cudaErrorInvalidValue;
  const float sy = rh > 1 ? (float)(h - 1) / (float)(rh - 1) : 0.f;
  const float sx = rw > 1 ? (float)(w - 1) / (float)(rw - 1) : 0.f;
  if (n == 0) return cudaSuccess;
  // a strip's last source column is at most floor(7 sx) + 1 past its first: the up-sampling kernel holds RS_NC = 6
  const bool up = 7.f * sx < 3.998f && !getenv("UFM_RESIZE_GENERIC");
  if (in_f32) {
    dim3 grid(blocks_for((long long)((ow + RS_STRIP - 1) / RS_STRIP) * (C / 4), 256), oh, n);
    if (up) launch_k(resize_ac_up_kernel<float>, grid, 256, 0, s, in_f32, h, w, C, sy, sx, oh, ow, out_f32, out_bf16);
    else launch_k(resize_ac_kernel<float>, grid, 256, 0, s, in_f32, h, w, C, sy, sx, oh, ow, out_f32, out_bf16);
  } else {
    dim3 grid(blocks_for((long long)((ow + RS_STRIP - 1) / RS_STRIP) * (C / 8), 256), oh, n);
    if (up) launch_k(resize_ac_up_kernel<__nv_bfloat16>, grid, 256, 0, s, in_bf16, h, w, C, sy, sx, oh, ow, out_f32, out_bf16);
    else launch_k(resize_ac_kernel<__nv_bfloat16>, grid, 256, 0, s, in_bf16, h, w, C, sy, sx, oh, ow, out_f32, out_bf16);
  }
  return cudaGetLastError();
}

cudaError_t launch_desc_shuffle(const __nv_bfloat16* z, __nv_bfloat16* desc, int n, int h, int w, int k, int C, cudaStream_t s) {
  if (C % 8) return cudaErrorInvalidValue;
  const long long rows = (long long)n * h * w;
  if (rows == 0) return cudaSuccess;
  const int cg = C / 8, per_row = k * k * cg;
  const unsigned gy = rows > 32768 ? 32768u : (unsigned)rows;
  dim3 grid((per_row + 255) / 256, gy, (unsigned)((rows + gy - 1) / gy));
  const uint4* zi = reinterpret_cast<const uint4*>(z);
  uint4* di = reinterpret_cast<uint4*>(desc);
  if (k == 14 && cg == 2) launch_k(desc_shuffle_kernel<14, 2>, grid, 256, 0, s, zi, di, n, h, w, k, cg);
  else launch_k(desc_shuffle_kernel<0, 0>, grid, 256, 0, s, zi, di, n, h, w, k, cg);
  return cudaGetLastError();
}

cudaError_t launch_refine(const RefineParams& p, cudaStream_t s) {
  dim3 grid((p.W + 31) / 32, (p.H + 3) / 4, p.B);
  launch_k(refine_kernel, grid, 128, 0, s, p);
  return cudaGetLastError();
}

cudaError_t launch_nhwc_to_nchw(const float* in, float* out, int n, int H, int W, int C, cudaStream_t s) {
  const long long total = (long long)n * H * W * C;
  launch_k(nhwc_to_nchw_kernel, blocks_for(total, 256), 256, 0, s, in, out, n, H, W, C);
  return cudaGetLastError();
}

UFM_DEFINE_PDL_EARLY_SETTER(simt)

}  // namespace ufm
