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
// 64 pixels x 4 channel quarters per block; one thread = 16 output channels of one pixel.
__global__ void __launch_bounds__(256) unet_stem_kernel(const UnetStemParams p) {
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

// ------------------------------------------------------------------------------------------------ max pool
__global__ void __launch_bounds__(256) unet_pool_kernel(const bf16* __restrict__ in, long long in_pitch, int n, int H, int W,
                                                       int C, bf16* __restrict__ out, int oh, int ow) {
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
__global__ void __launch_bounds__(128) unet_combine_kernel(const UnetCombineParams p) {
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
  const long long pix = (long long)blockIdx.x * 128 + threadIdx.x;
  if (pix >= plane) return;
  const int img = blockIdx.y;
  // final_conv: 64 -> 16
  float f[16];
#pragma unroll
  for (int o = 0; o < 16; ++o) f[o] = bfin[o];
  const uint4* xs = reinterpret_cast<const uint4*>(p.x0 + ((long long)img * plane + pix) * 64);
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) {
    const uint4 v = xs[ch];
    const uint32_t uu[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 xv = unpack2(uu[j]);
      const int c = ch * 8 + j * 2;
#pragma unroll
      for (int o4 = 0; o4 < 4; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&wf[c][o4 * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&wf[c + 1][o4 * 4]);
        f[o4 * 4 + 0] = fmaf(xv.y, b.x, fmaf(xv.x, a.x, f[o4 * 4 + 0]));
        f[o4 * 4 + 1] = fmaf(xv.y, b.y, fmaf(xv.x, a.y, f[o4 * 4 + 1]));
        f[o4 * 4 + 2] = fmaf(xv.y, b.z, fmaf(xv.x, a.z, f[o4 * 4 + 2]));
        f[o4 * 4 + 3] = fmaf(xv.y, b.w, fmaf(xv.x, a.w, f[o4 * 4 + 3]));
      }
    }
  }
  if (p.unet_feat) {
#pragma unroll
    for (int o = 0; o < 16; ++o) p.unet_feat[((long long)img * 16 + o) * plane + pix] = f[o];
  }
  // classification features of this pixel: two 16-byte chunks of the planar layout [img][2][H][W][8]
  float v[32];
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    const uint4 cv = *reinterpret_cast<const uint4*>(p.desc_cls + (((long long)img * 2 + g) * plane + pix) * 8);
    const uint32_t uu[4] = {cv.x, cv.y, cv.z, cv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 t = unpack2(uu[j]);
      v[g * 8 + j * 2] = t.x;
      v[g * 8 + j * 2 + 1] = t.y;
    }
  }
  float o[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) o[k] = b2[k];
  if (p.method == 1) {
    // conv2(relu(conv1(cat(cls, unet))))
#pragma unroll
    for (int k = 0; k < 16; ++k) v[16 + k] = f[k];
    float h[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) h[k] = b1[k];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
#pragma unroll
      for (int o4 = 0; o4 < 8; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&w1[i][o4 * 4]);
        h[o4 * 4 + 0] = fmaf(v[i], a.x, h[o4 * 4 + 0]);
        h[o4 * 4 + 1] = fmaf(v[i], a.y, h[o4 * 4 + 1]);
        h[o4 * 4 + 2] = fmaf(v[i], a.z, h[o4 * 4 + 2]);
        h[o4 * 4 + 3] = fmaf(v[i], a.w, h[o4 * 4 + 3]);
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float hv = fmaxf(h[i], 0.f);
#pragma unroll
      for (int o4 = 0; o4 < 4; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&w2[i][o4 * 4]);
        o[o4 * 4 + 0] = fmaf(hv, a.x, o[o4 * 4 + 0]);
        o[o4 * 4 + 1] = fmaf(hv, a.y, o[o4 * 4 + 1]);
        o[o4 * 4 + 2] = fmaf(hv, a.z, o[o4 * 4 + 2]);
        o[o4 * 4 + 3] = fmaf(hv, a.w, o[o4 * 4 + 3]);
      }
    }
  } else {
    // conv2(cls * tanh(unet))
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float m = v[i] * tanhf(f[i]);
#pragma unroll
      for (int o4 = 0; o4 < 4; ++o4) {
        const float4 a = *reinterpret_cast<const float4*>(&w2[i][o4 * 4]);
        o[o4 * 4 + 0] = fmaf(m, a.x, o[o4 * 4 + 0]);
        o[o4 * 4 + 1] = fmaf(m, a.y, o[o4 * 4 + 1]);
        o[o4 * 4 + 2] = fmaf(m, a.z, o[o4 * 4 + 2]);
        o[o4 * 4 + 3] = fmaf(m, a.w, o[o4 * 4 + 3]);
      }
    }
  }
#pragma unroll
  for (int g = 0; g < 2; ++g)
    *reinterpret_cast<uint4*>(p.desc + (((long long)img * 2 + g) * plane + pix) * 8) =
        make_uint4(pack2(o[g * 8 + 0], o[g * 8 + 1]), pack2(o[g * 8 + 2], o[g * 8 + 3]), pack2(o[g * 8 + 4], o[g * 8 + 5]),
                   pack2(o[g * 8 + 6], o[g * 8 + 7]));
}

inline unsigned blocks_for(long long total, int per) { return (unsigned)((total + per - 1) / per); }

}  // namespace

cudaError_t launch_unet_stem(const UnetStemParams& p, cudaStream_t s) {
  dim3 grid(blocks_for((long long)p.H * p.W, 64), 2 * p.B);
  unet_stem_kernel<<<grid, 256, 0, s>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_unet_pool(const __nv_bfloat16* in, long long in_pitch, int n, int H, int W, int C, __nv_bfloat16* out,
                             cudaStream_t s) {
  if (C % 8 || in_pitch % 8) return cudaErrorInvalidValue;
  const int oh = H / 2, ow = W / 2;
  unet_pool_kernel<<<blocks_for((long long)n * oh * ow * (C / 8), 256), 256, 0, s>>>(in, in_pitch, n, H, W, C, out, oh, ow);
  return cudaGetLastError();
}

cudaError_t launch_unet_up(const __nv_bfloat16* g, int n, int h, int w, int C, __nv_bfloat16* out, long long out_pitch,
                           int out_off, int oh, int ow, cudaStream_t s) {
  if (C % 8 || out_pitch % 8 || out_off % 8) return cudaErrorInvalidValue;
  unet_up_kernel<<<blocks_for((long long)n * oh * ow * (C / 8), 256), 256, 0, s>>>(g, n, h, w, C, out, out_pitch, out_off,
                                                                                 oh, ow);
  return cudaGetLastError();
}

cudaError_t launch_unet_combine(const UnetCombineParams& p, cudaStream_t s) {
  if (p.method != 1 && p.method != 2) return cudaErrorInvalidValue;
  dim3 grid(blocks_for((long long)p.H * p.W, 128), p.n);
  unet_combine_kernel<<<grid, 128, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace ufm
