// SIMT kernels of the UNet fine-feature branch of UFM-Refine (uniflowmatch/models/unet_encoder.py:26-72,
// uniflowmatch/models/ufm.py:817-825, 916-918, 967-982).  The 3x3 convolutions with >= 64 input channels and the
// ConvTranspose GEMMs run on the tcgen05 implicit-GEMM kernel (gemm.cu); what is here is the memory-bound glue:
//   unet_stem_kernel     image normalisation + the 3 -> 64 channel first conv + ReLU (K = 27 is no tensor-core shape)
//   unet_pool_kernel     2x2/s2 max pool out of the skip half of a concat buffer
//   unet_up_kernel       ConvTranspose(k = s = 2) scatter + nearest resize to the skip size, into the concat buffer
//   unet_combine_kernel  final 1x1 conv (64 -> 16) + the per-pixel combine with the classification features, written
//                        straight into the channel-group planar descriptor layout the refinement kernel reads
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"
#include "launch.cuh"

namespace ufm {
namespace {

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack2(uint32_t u) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
}
__device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint4 max8(uint4 a, uint4 b) {
  return make_uint4(max2(a.x, b.x), max2(a.y, b.y), max2(a.z, b.z), max2(a.w, b.w));
}
// ATen upsample_nearest (mode "nearest"): src = min(floor(dst * (in / out)), in - 1), fp32 arithmetic
__device__ __forceinline__ int nearest_src(int dst, int in, int out) {
  const float scale = (float)in / (float)out;
  const int s = (int)floorf((float)dst * scale);
  return s < in - 1 ? s : in - 1;
}

// ------------------------------------------------------------------------------------------------ stem
// Generic version (any width): 64 pixels x 4 channel quarters per block; one thread = 16 output channels of one pixel.
// One LDS.128 feeds only 4 FMAs here, so it runs at the shared-memory rate; used when W % 4 != 0.
__global__ void __launch_bounds__(256) unet_stem_generic_kernel(const UnetStemParams p) {
  pdl_wait();
  __shared__ __align__(16) float ws[27][64];
  __shared__ float bs[64];
  for (int i = threadIdx.x; i < 27 * 64; i += 256) {
    const int co = i / 27, k = i - co * 27;  // weight is [co][ci][ky][kx]
    ws[k][co] = p.w[i];
  }
  if (threadIdx.x < 64) bs[threadIdx.x] = p.b[threadIdx.x];
  __syncthreads();
  const int q = threadIdx.x & 3;
  const int pix = blockIdx.x * 64 + (threadIdx.x >> 2);
  if (pix >= p.H * p.W) return;
  const int img = blockIdx.y;  // (b, view)
  const int b = img >> 1, view = img & 1;
  const void* base = view ? p.tgt : p.src;
  const int y = pix / p.W, x = pix - y * p.W;
  float in[27];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int yy = y + ky - 1, xx = x + kx - 1;
        float v = 0.f;  // zero padding of the NORMALISED image
        if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) {
          const long long off = p.is_nhwc ? (((long long)b * p.H + yy) * p.W + xx) * 3 + c
                                          : (((long long)b * 3 + c) * p.H + yy) * p.W + xx;
          if (p.is_u8)
            v = ((float)reinterpret_cast<const uint8_t*>(base)[off] / 255.0f - p.mean[c]) * p.inv_std[c];
          else
            v = reinterpret_cast<const float*>(base)[off];
        }
        in[c * 9 + ky * 3 + kx] = v;
      }
  float acc[16];
#pragma unroll
  for (int o = 0; o < 16; ++o) acc[o] = bs[q * 16 + o];
#pragma unroll
  for (int k = 0; k < 27; ++k) {
    const float4* wr = reinterpret_cast<const float4*>(&ws[k][q * 16]);
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const float4 w4 = wr[v];
      acc[v * 4 + 0] = fmaf(in[k], w4.x, acc[v * 4 + 0]);
      acc[v * 4 + 1] = fmaf(in[k], w4.y, acc[v * 4 + 1]);
      acc[v * 4 + 2] = fmaf(in[k], w4.z, acc[v * 4 + 2]);
      acc[v * 4 + 3] = fmaf(in[k], w4.w, acc[v * 4 + 3]);
    }
  }
  uint32_t u[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) u[i] = pack2(fmaxf(acc[2 * i], 0.f), fmaxf(acc[2 * i + 1], 0.f));
  uint4* dst = reinterpret_cast<uint4*>(p.out + ((long long)img * p.H * p.W + pix) * 64 + q * 16);
  dst[0] = make_uint4(u[0], u[1], u[2], u[3]);
  dst[1] = make_uint4(u[4], u[5], u[6], u[7]);
}

// Fast version (W % 4 == 0): a block covers 256 consecutive pixels of one image; warp w computes channel quarter q = w & 3
// (so every weight load is a warp-wide broadcast) for 4 consecutive pixels per thread -> one LDS.128 feeds 16 FMAs.
// The 256 x 64 bf16 tile is staged in shared memory (16-byte chunks XOR-swizzled against the 512-byte lane stride)
// and leaves as one contiguous 32 KB run.
// Input taps are fetched branch-free (clamped address + select) so that the 18 loads of one input channel are in
// flight together; uint8 pixels are normalised through a 3 x 256 table built per block (no per-tap division).
template <bool U8, bool NHWC>
__global__ void __launch_bounds__(256, 2) unet_stem_kernel(const UnetStemParams p) {
  pdl_wait();
  __shared__ __align__(16) float ws[27][64];
  __shared__ float bs[64];
  __shared__ float lut[U8 ? 3 * 256 : 1];
  __shared__ __align__(16) uint4 tile[256 * 8];
  for (int i = threadIdx.x; i < 27 * 64; i += 256) {
    const int co = i / 27, k = i - co * 27;  // weight is [co][ci][ky][kx]
    ws[k][co] = p.w[i];
  }
  if (threadIdx.x < 64) bs[threadIdx.x] = p.b[threadIdx.x];
  if (U8) {
    for (int i = threadIdx.x; i < 3 * 256; i += 256) {
      const int c = i >> 8;
      lut[i] = ((float)(i & 255) / 255.0f - p.mean[c]) * p.inv_std[c];  // base.py:215-229, as patch_im2col_kernel
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3;
  const int lp = ((warp >> 2) * 32 + lane) * 4;  // first of this thread's 4 pixels inside the tile
  const int plane = p.H * p.W;
  const int pix0 = blockIdx.x * 256 + lp;
  const int img = blockIdx.y;  // (b, view)
  const int b = img >> 1, view = img & 1;
  const uint8_t* base8 = reinterpret_cast<const uint8_t*>(view ? p.tgt : p.src);
  const float* base32 = reinterpret_cast<const float*>(view ? p.tgt : p.src);
  if (pix0 < plane) {
    const int y = pix0 / p.W, x0 = pix0 - y * p.W;  // W % 4 == 0: the 4 pixels share the row
    float acc[4][16];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int o = 0; o < 16; ++o) acc[j][o] = bs[q * 16 + o];
    int xc[6], yc[3];
    bool xok[6], yok[3];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const int xx = x0 + i - 1;
      xok[i] = (unsigned)xx < (unsigned)p.W;
      xc[i] = min(max(xx, 0), p.W - 1);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int yy = y + k - 1;
      yok[k] = (unsigned)yy < (unsigned)p.H;
      yc[k] = min(max(yy, 0), p.H - 1);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float in[3][6];
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          const int off = NHWC ? ((b * p.H + yc[ky]) * p.W + xc[i]) * 3 + c : ((b * 3 + c) * p.H + yc[ky]) * p.W + xc[i];
          const float val = U8 ? lut[c * 256 + base8[off]] : base32[off];
          in[ky][i] = (xok[i] && yok[ky]) ? val : 0.f;  // zero padding of the NORMALISED image
        }
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float4* wr = reinterpret_cast<const float4*>(&ws[c * 9 + ky * 3 + kx][q * 16]);
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            const float4 w4 = wr[v];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              acc[j][v * 4 + 0] = fmaf(in[ky][j + kx], w4.x, acc[j][v * 4 + 0]);
              acc[j][v * 4 + 1] = fmaf(in[ky][j + kx], w4.y, acc[j][v * 4 + 1]);
              acc[j][v * 4 + 2] = fmaf(in[ky][j + kx], w4.z, acc[j][v * 4 + 2]);
              acc[j][v * 4 + 3] = fmaf(in[ky][j + kx], w4.w, acc[j][v * 4 + 3]);
            }
          }
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t u[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) u[i] = pack2(fmaxf(acc[j][2 * i], 0.f), fmaxf(acc[j][2 * i + 1], 0.f));
      const int px = lp + j, sw = (px >> 2) & 7;
      tile[px * 8 + ((q * 2) ^ sw)] = make_uint4(u[0], u[1], u[2], u[3]);
      tile[px * 8 + ((q * 2 + 1) ^ sw)] = make_uint4(u[4], u[5], u[6], u[7]);
    }
  }
  __syncthreads();
  uint4* dst = reinterpret_cast<uint4*>(p.out + ((long long)img * plane + (long long)blockIdx.x * 256) * 64);
  const int valid = min(256, plane - blockIdx.x * 256) * 8;  // 16-byte chunks of this tile that exist
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int i = it * 256 + threadIdx.x;
    if (i < valid) {
      const int px = i >> 3;
      dst[i] = tile[px * 8 + ((i & 7) ^ ((px >> 2) & 7))];
    }
  }
}

// ------------------------------------------------------------------------------------------------ max pool
__global__ void __launch_bounds__(256) unet_pool_kernel(const bf16* __restrict__ in, long long in_pitch, int n, int H, int W,
                                                       int C, bf16* __restrict__ out, int oh, int ow) {
  pdl_wait();
  const int cg = C >> 3;
  const long long total = (long long)n * oh * ow * cg;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int g = (int)(idx % cg);
  const long long opix = idx / cg;
  const int ox = (int)(opix % ow);
  const int oy = (int)((opix / ow) % oh);
  const int img = (int)(opix / ((long long)ow * oh));
  const bf16* s = in + (((long long)img * H + 2 * oy) * W + 2 * ox) * in_pitch + g * 8;
  const uint4 a = *reinterpret_cast<const uint4*>(s);
  const uint4 b = *reinterpret_cast<const uint4*>(s + in_pitch);
  const uint4 c = *reinterpret_cast<const uint4*>(s + (long long)W * in_pitch);
  const uint4 d = *reinterpret_cast<const uint4*>(s + (long long)W * in_pitch + in_pitch);
  *reinterpret_cast<uint4*>(out + opix * C + g * 8) = max8(max8(a, b), max8(c, d));
}

// ------------------------------------------------------------------------------------------------ ConvT scatter
// g[(img, y, x)][(ky, kx, c)] (the ConvTranspose as a GEMM, bias already added) -> out[img, oy, ox, off + c] where
// (oy, ox) walks the SKIP's size and reads the nearest pixel of the (2h, 2w) transposed-conv result.
__global__ void __launch_bounds__(256) unet_up_kernel(const bf16* __restrict__ g, int n, int h, int w, int C,
                                                     bf16* __restrict__ out, long long out_pitch, int out_off, int oh,
                                                     int ow) {
  pdl_wait();
  const int cg = C >> 3;
  const long long total = (long long)n * oh * ow * cg;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int gi = (int)(idx % cg);
  const long long opix = idx / cg;
  const int ox = (int)(opix % ow);
  const int oy = (int)((opix / ow) % oh);
  const int img = (int)(opix / ((long long)ow * oh));
  const int Y = nearest_src(oy, 2 * h, oh), X = nearest_src(ox, 2 * w, ow);
  const long long row = ((long long)img * h + (Y >> 1)) * w + (X >> 1);
  const int col = (((Y & 1) << 1) | (X & 1)) * C + gi * 8;
  *reinterpret_cast<uint4*>(out + opix * out_pitch + out_off + gi * 8) =
      *reinterpret_cast<const uint4*>(g + row * (4LL * C) + col);
}

// ------------------------------------------------------------------------------------------------ combine
// One thread = CP pixels (128 apart, so every planar load / store stays coalesced): each broadcast weight load from
// shared memory feeds 4 * CP FMAs instead of 4 (the one-pixel version ran at the LDS rate).
constexpr int CP = 2;
__global__ void __launch_bounds__(128) unet_combine_kernel(const UnetCombineParams p) {
  pdl_wait();
  __shared__ __align__(16) float wf[64][16];
  __shared__ __align__(16) float w1[32][32];
  __shared__ __align__(16) float w2[32][16];
  __shared__ float bfin[16], b1[32], b2[16];
  const int c2in = p.method == 1 ? 32 : 16;
  for (int i = threadIdx.x; i < 64 * 16; i += 128) (&wf[0][0])[i] = p.wf[i];
  for (int i = threadIdx.x; i < 32 * 32; i += 128) (&w1[0][0])[i] = p.method == 1 ? p.w1[i] : 0.f;
  for (int i = threadIdx.x; i < c2in * 16; i += 128) (&w2[0][0])[i] = p.w2[i];
  if (threadIdx.x < 16) {
    bfin[threadIdx.x] = p.bf[threadIdx.x];
    b2[threadIdx.x] = p.b2[threadIdx.x];
  }
  if (threadIdx.x < 32) b1[threadIdx.x] = p.method == 1 ? p.b1[threadIdx.x] : 0.f;
  __syncthreads();
  const long long plane = (long long)p.H * p.W;
  const long long pix0 = (long long)blockIdx.x * (128 * CP) + threadIdx.x;
  if (pix0 >= plane) return;
  const int img = blockIdx.y;
  long long pix[CP];
#pragma unroll
  for (int j = 0; j < CP; ++j) pix[j] = min(pix0 + 128 * j, plane - 1);  // clamped duplicates are not stored
  // final_conv: 64 -> 16
  float f[CP][16];
#pragma unroll
  for (int j = 0; j < CP; ++j)
#pragma unroll
    for (int o = 0; o < 16; ++o) f[j][o] = bfin[o];
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) {
    uint4 xv[CP];
#pragma unroll
    for (int j = 0; j < CP; ++j) xv[j] = reinterpret_cast<const uint4*>(p.x0 + ((long long)img * plane + pix[j]) * 64)[ch];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float xs[CP];
#pragma unroll
      for (int j = 0; j < CP; ++j) {
        const uint32_t w = e < 2 ? xv[j].x : e < 4 ? xv[j].y : e < 6 ? xv[j].z : xv[j].w;
        const float2 t = unpack2(w);
        xs[j] = (e & 1) ? t.y : t.x;
      }
#pragma unroll
      for (int o4 = 0; o4 < 4; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&wf[ch * 8 + e][o4 * 4]);
#pragma unroll
        for (int j = 0; j < CP; ++j) {
          f[j][o4 * 4 + 0] = fmaf(xs[j], a.x, f[j][o4 * 4 + 0]);
          f[j][o4 * 4 + 1] = fmaf(xs[j], a.y, f[j][o4 * 4 + 1]);
          f[j][o4 * 4 + 2] = fmaf(xs[j], a.z, f[j][o4 * 4 + 2]);
          f[j][o4 * 4 + 3] = fmaf(xs[j], a.w, f[j][o4 * 4 + 3]);
        }
      }
    }
  }
  if (p.unet_feat) {
#pragma unroll
    for (int j = 0; j < CP; ++j)
      if (pix0 + 128 * j < plane) {
#pragma unroll
        for (int o = 0; o < 16; ++o) p.unet_feat[((long long)img * 16 + o) * plane + pix[j]] = f[j][o];
      }
  }
  // classification features: two 16-byte chunks per pixel of the planar layout [img][2][H][W][8]
  float v[CP][16];
#pragma unroll
  for (int j = 0; j < CP; ++j)
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      const uint4 cv = *reinterpret_cast<const uint4*>(p.desc_cls + (((long long)img * 2 + g) * plane + pix[j]) * 8);
      const uint32_t uu[4] = {cv.x, cv.y, cv.z, cv.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 t = unpack2(uu[k]);
        v[j][g * 8 + k * 2] = t.x;
        v[j][g * 8 + k * 2 + 1] = t.y;
      }
    }
  float o[CP][16];
#pragma unroll
  for (int j = 0; j < CP; ++j)
#pragma unroll
    for (int k = 0; k < 16; ++k) o[j][k] = b2[k];
  if (p.method == 1) {
    // conv2(relu(conv1(cat(cls, unet))))
    float h[CP][32];
#pragma unroll
    for (int j = 0; j < CP; ++j)
#pragma unroll
      for (int k = 0; k < 32; ++k) h[j][k] = b1[k];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
#pragma unroll
      for (int o4 = 0; o4 < 8; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&w1[i][o4 * 4]);
#pragma unroll
        for (int j = 0; j < CP; ++j) {
          const float vi = i < 16 ? v[j][i & 15] : f[j][i & 15];
          h[j][o4 * 4 + 0] = fmaf(vi, a.x, h[j][o4 * 4 + 0]);
          h[j][o4 * 4 + 1] = fmaf(vi, a.y, h[j][o4 * 4 + 1]);
          h[j][o4 * 4 + 2] = fmaf(vi, a.z, h[j][o4 * 4 + 2]);
          h[j][o4 * 4 + 3] = fmaf(vi, a.w, h[j][o4 * 4 + 3]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) {
#pragma unroll
      for (int o4 = 0; o4 < 4; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&w2[i][o4 * 4]);
#pragma unroll
        for (int j = 0; j < CP; ++j) {
          const float hv = fmaxf(h[j][i], 0.f);
          o[j][o4 * 4 + 0] = fmaf(hv, a.x, o[j][o4 * 4 + 0]);
          o[j][o4 * 4 + 1] = fmaf(hv, a.y, o[j][o4 * 4 + 1]);
          o[j][o4 * 4 + 2] = fmaf(hv, a.z, o[j][o4 * 4 + 2]);
          o[j][o4 * 4 + 3] = fmaf(hv, a.w, o[j][o4 * 4 + 3]);
        }
      }
    }
  } else {
    // conv2(cls * tanh(unet))
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float m[CP];
#pragma unroll
      for (int j = 0; j < CP; ++j) m[j] = v[j][i] * tanhf(f[j][i]);
#pragma unroll
      for (int o4 = 0; o4 < 4; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&w2[i][o4 * 4]);
#pragma unroll
        for (int j = 0; j < CP; ++j) {
          o[j][o4 * 4 + 0] = fmaf(m[j], a.x, o[j][o4 * 4 + 0]);
          o[j][o4 * 4 + 1] = fmaf(m[j], a.y, o[j][o4 * 4 + 1]);
          o[j][o4 * 4 + 2] = fmaf(m[j], a.z, o[j][o4 * 4 + 2]);
          o[j][o4 * 4 + 3] = fmaf(m[j], a.w, o[j][o4 * 4 + 3]);
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < CP; ++j)
    if (pix0 + 128 * j < plane) {
#pragma unroll
      for (int g = 0; g < 2; ++g)
        *reinterpret_cast<uint4*>(p.desc + (((long long)img * 2 + g) * plane + pix[j]) * 8) =
            make_uint4(pack2(o[j][g * 8 + 0], o[j][g * 8 + 1]), pack2(o[j][g * 8 + 2], o[j][g * 8 + 3]),
                       pack2(o[j][g * 8 + 4], o[j][g * 8 + 5]), pack2(o[j][g * 8 + 6], o[j][g * 8 + 7]));
    }
}

inline unsigned blocks_for(long long total, int per) { return (unsigned)((total + per - 1) / per); }

}  // namespace

cudaError_t launch_unet_stem(const UnetStemParams& p, cudaStream_t s) {
  if (p.W % 4 == 0) {
    // 32-bit input offsets in the fast kernel
    if ((long long)p.B * p.H * p.W * 3 >= (1LL << 31)) return cudaErrorInvalidValue;
    dim3 grid(blocks_for((long long)p.H * p.W, 256), 2 * p.B);
    if (p.is_u8 && p.is_nhwc) launch_k(unet_stem_kernel<true, true>, grid, 256, 0, s, p);
    else if (p.is_u8) launch_k(unet_stem_kernel<true, false>, grid, 256, 0, s, p);
    else if (p.is_nhwc) launch_k(unet_stem_kernel<false, true>, grid, 256, 0, s, p);
    else launch_k(unet_stem_kernel<false, false>, grid, 256, 0, s, p);
  } else {
    dim3 grid(blocks_for((long long)p.H * p.W, 64), 2 * p.B);
    launch_k(unet_stem_generic_kernel, grid, 256, 0, s, p);
  }
  return cudaGetLastError();
}

cudaError_t launch_unet_pool(const __nv_bfloat16* in, long long in_pitch, int n, int H, int W, int C, __nv_bfloat16* out,
                             cudaStream_t s) {
  if (C % 8 || in_pitch % 8) return cudaErrorInvalidValue;
  const int oh = H / 2, ow = W / 2;
  launch_k(unet_pool_kernel, blocks_for((long long)n * oh * ow * (C / 8), 256), 256, 0, s, in, in_pitch, n, H, W, C, out, oh, ow);
  return cudaGetLastError();
}

cudaError_t launch_unet_up(const __nv_bfloat16* g, int n, int h, int w, int C, __nv_bfloat16* out, long long out_pitch,
                           int out_off, int oh, int ow, cudaStream_t s) {
  if (C % 8 || out_pitch % 8 || out_off % 8) return cudaErrorInvalidValue;
  launch_k(unet_up_kernel, blocks_for((long long)n * oh * ow * (C / 8), 256), 256, 0, s, g, n, h, w, C, out, out_pitch, out_off,
                                                                                 oh, ow);
  return cudaGetLastError();
}

cudaError_t launch_unet_combine(const UnetCombineParams& p, cudaStream_t s) {
  if (p.method != 1 && p.method != 2) return cudaErrorInvalidValue;
  dim3 grid(blocks_for((long long)p.H * p.W, 128 * CP), p.n);
  launch_k(unet_combine_kernel, grid, 128, 0, s, p);
  return cudaGetLastError();
}

UFM_DEFINE_PDL_EARLY_SETTER(unet)

}  // namespace ufm
