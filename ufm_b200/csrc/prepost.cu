// Pre/post-processing kernels for image pairs that are NOT at the inference resolution (reference:
// uniflowmatch/utils/flow_resizing.py).  At the native resolution none of this runs (the patch gather reads the raw
// images and the engine writes the API tensors directly).
//
//   prep_resize_kernel   ResizeToFixedManipulation.__call__ (flow_resizing.py:276-354): F.interpolate(bilinear,
//                        align_corners=False, antialias=True) of the normalised image, fused with the normalisation of
//                        predict_correspondences_batched (base.py:215-229): u8 (v/255 - mean)/std or an fp32 affine
//                        re-normalisation, read from BHWC or BCHW, written as normalised fp32 BCHW at (H, W).
//                        Weights follow ATen's anti-alias triangle filter (UpSampleKernel.cpp
//                        _compute_indices_min_size_weights_aa / UpSample.cuh _compute_weights): support = max(scale, 1),
//                        taps [int(c - s + .5), int(c + s + .5)), w = max(0, 1 - |(j + .5 - c) / max(scale, 1)|) / sum.
//   unmap_kernel         unmap_predicted_flow (flow_resizing.py:749-877) and unmap_predicted_channels (:955-1010) in
//                        one pass: crop the represented region, NEAREST-resample flow / covisibility onto the source
//                        grid, bilinear-resample the (x + .5, y + .5) coordinate grid (align_corners=False), scale the
//                        end points by the source / target size ratios, subtract, paste into zeros.
#include "kernels.h"
#include "launch.cuh"

namespace ufm {

namespace {

constexpr int kMaxTaps = 64;  // per axis; a down-scale factor up to ~31x

struct AaAxis {
  int lo, n;
  float center, invscale, inv_total;
};

// ATen: scale = in / out (float); support = scale >= 1 ? scale : 1; center = scale * (i + 0.5)
__device__ __forceinline__ AaAxis aa_axis(int i, int in, float scale) {
  AaAxis a;
  const float support = scale >= 1.f ? scale : 1.f;
  a.invscale = scale >= 1.f ? 1.f / scale : 1.f;
  a.center = scale * ((float)i + 0.5f);
  a.lo = max((int)(a.center - support + 0.5f), 0);
  a.n = min((int)(a.center + support + 0.5f), in) - a.lo;
  a.n = min(a.n, kMaxTaps);
  float tot = 0.f;
  for (int j = 0; j < a.n; ++j) tot += fmaxf(0.f, 1.f - fabsf(((float)(j + a.lo) - a.center + 0.5f) * a.invscale));
  a.inv_total = tot != 0.f ? 1.f / tot : 0.f;
  return a;
}
__device__ __forceinline__ float aa_w(const AaAxis& a, int j) {
  return fmaxf(0.f, 1.f - fabsf(((float)(j + a.lo) - a.center + 0.5f) * a.invscale)) * a.inv_total;
}

template <typename T, bool NHWC>
__global__ void __launch_bounds__(128) prep_resize_kernel(const PrepParams p) {
  pdl_wait();
  // uint8 input: the normalisation (v / 255 - mean) / std of base.py:215-229 costs two fp32 divisions per tap and channel
  // (~100 per output pixel: the kernel ran at 5 % of HBM bandwidth).  The 3 x 256 possible results are computed once per
  // block with the same arithmetic -- bit-identical -- and looked up from shared memory.
  __shared__ float lut[sizeof(T) == 1 ? 768 : 1];
  if (sizeof(T) == 1) {
    for (int i = threadIdx.x; i < 768; i += 128) {
      const int c = i >> 8;
      lut[i] = ((float)(i & 255) / 255.0f - p.mean[c]) / p.std[c];
    }
    __syncthreads();
  }
  const int ox = blockIdx.x * 128 + threadIdx.x;
  const int oy = blockIdx.y, b = blockIdx.z;
  if (ox >= p.W) return;
  const AaAxis ax = aa_axis(ox, p.W0, p.scale_x), ay = aa_axis(oy, p.H0, p.scale_y);
  const T* img = reinterpret_cast<const T*>(p.in) + (size_t)b * 3 * p.H0 * p.W0;
  const size_t plane0 = (size_t)p.H0 * p.W0;
  float acc[3] = {0.f, 0.f, 0.f};
  // separable like ATen's CPU kernel: horizontal pass per source row (fp32), then the vertical weights
  for (int jy = 0; jy < ay.n; ++jy) {
    const int y = ay.lo + jy;
    float row[3] = {0.f, 0.f, 0.f};
    for (int jx = 0; jx < ax.n; ++jx) {
      const int x = ax.lo + jx;
      const float w = aa_w(ax, jx);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const size_t idx = NHWC ? ((size_t)y * p.W0 + x) * 3 + c : c * plane0 + (size_t)y * p.W0 + x;
        // base.py:215-229: (u8 / 255 - mean) / std (table above), or an affine re-normalisation of an fp32 image
        float v;
        if (sizeof(T) == 1) v = lut[(c << 8) + (int)img[idx]];
        else v = (float)img[idx] * p.mul[c] + p.add[c];
        row[c] = fmaf(w, v, row[c]);
      }
    }
    const float wy = aa_w(ay, jy);
#pragma unroll
    for (int c = 0; c < 3; ++c) acc[c] = fmaf(wy, row[c], acc[c]);
  }
  const size_t plane = (size_t)p.H * p.W;
  float* o = p.out + (size_t)b * 3 * plane + (size_t)oy * p.W + ox;
#pragma unroll
  for (int c = 0; c < 3; ++c) o[c * plane] = acc[c];
}

// F.interpolate(mode="nearest"): src = min(floor(dst * (in / out)), in - 1)
__device__ __forceinline__ int nearest_src(int dst, float scale, int in) { return min((int)floorf((float)dst * scale), in - 1); }

__global__ void __launch_bounds__(128) unmap_kernel(const UnmapParams p) {
  pdl_wait();
  const int X = blockIdx.x * 128 + threadIdx.x;
  const int Y = blockIdx.y, b = blockIdx.z;
  if (X >= p.W0) return;
  const size_t plane0 = (size_t)p.H0 * p.W0, plane = (size_t)p.H * p.W;
  const size_t o = (size_t)Y * p.W0 + X;
  float fx = 0.f, fy = 0.f, cv = 0.f;
  const int y = Y - p.src0[0], x = X - p.src0[2];
  const int sh = p.src0[1] - p.src0[0], sw = p.src0[3] - p.src0[2];
  if (y >= 0 && y < sh && x >= 0 && x < sw) {
    const int rh = p.rep[1] - p.rep[0], rw = p.rep[3] - p.rep[2];
    const float ys = (float)rh / (float)sh, xs = (float)rw / (float)sw;  // ATen: scale = in / out
    const int iy = p.rep[0] + nearest_src(y, ys, rh), ix = p.rep[2] + nearest_src(x, xs, rw);
    const size_t in = (size_t)iy * p.W + ix;
    // bilinear (align_corners=False) resample of the coordinate grid sc[.., yy, xx] = (xx + 0.5, yy + 0.5)
    float sy = ys * ((float)y + 0.5f) - 0.5f, sx = xs * ((float)x + 0.5f) - 0.5f;
    sy = sy < 0.f ? 0.f : sy;
    sx = sx < 0.f ? 0.f : sx;
    const int y0 = (int)sy, x0 = (int)sx;
    const int y1 = y0 + (y0 < rh - 1 ? 1 : 0), x1 = x0 + (x0 < rw - 1 ? 1 : 0);
    const float ly = sy - (float)y0, lx = sx - (float)x0;
    const float hy = 1.f - ly, hx = 1.f - lx;
    const float gx0 = (float)x0 + 0.5f, gx1 = (float)x1 + 0.5f, gy0 = (float)y0 + 0.5f, gy1 = (float)y1 + 0.5f;
    const float scx = hy * (hx * gx0 + lx * gx1) + ly * (hx * gx0 + lx * gx1);
    const float scy = hy * (hx * gy0 + lx * gy0) + ly * (hx * gy1 + lx * gy1);
    if (p.flow) {
      const float* fl = p.flow + (size_t)b * 2 * plane;
      const float tx = fl[in] + scx, ty = fl[plane + in] + scy;
      fx = (tx * p.tgt_mul[0] + (float)p.src1[2]) - (scx * p.src_mul[0] + (float)p.src0[2]);
      fy = (ty * p.tgt_mul[1] + (float)p.src1[0]) - (scy * p.src_mul[1] + (float)p.src0[0]);
    }
    if (p.cov) cv = p.cov[(size_t)b * plane + in];
  }
  if (p.flow_out) {
    float* fo = p.flow_out + (size_t)b * 2 * plane0;
    fo[o] = fx;
    fo[plane0 + o] = fy;
  }
  if (p.cov_out) p.cov_out[(size_t)b * plane0 + o] = cv * p.cov_mul[b % p.cov_mul_n];
}

// ---------------------------------------------------------------------------------------------------------------
// Uncertainty-head adaptors (UC/models/prediction_heads/adaptors.py): AdaptorMap splits the head's raw
// channels in adaptor order (base.py:186-210); per adaptor:
//   MaskAdaptor (:980-985)              mask = sigmoid(x)
//   Covariance2DAdaptor exp_tanh (:944-965)  cov = (e^c1, e^c2, tanh(s) e^((c1+c2)/2)), log_det = c1 + c2 + log(1 - tanh^2 + 1e-8),
//                                       inv = (e^-c1, e^-c2, -tanh(s) e^(-(c1+c2)/2)) / (1 - tanh^2 + 1e-8)
//   ConfidenceAdaptor (:899-927)        exp: vmin + min(e^x, vmax - vmin); sigmoid: sigmoid(x) (vmax - vmin) + vmin;
//                                       softmax: softmax over the H*W pixels of an image, times H*W
__global__ void __launch_bounds__(1024) conf_softmax_stats_kernel(const float* raw, int C, int ch, long long plane, float* stats) {
  pdl_wait();
  const float* x = raw + ((long long)blockIdx.x * C + ch) * plane;
  float m = -INFINITY, s = 0.f;
  for (long long i = threadIdx.x; i < plane; i += 1024) {
    const float v = x[i];
    const float nm = fmaxf(m, v);
    s = s * __expf(m - nm) + __expf(v - nm);
    m = nm;
  }
  __shared__ float sm[32], ss[32];
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    const float nm = fmaxf(m, om);
    s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
    m = nm;
  }
  if ((threadIdx.x & 31) == 0) {
    sm[threadIdx.x >> 5] = m;
    ss[threadIdx.x >> 5] = s;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    m = sm[threadIdx.x];
    s = ss[threadIdx.x];
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
      const float nm = fmaxf(m, om);
      s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
      m = nm;
    }
    if (threadIdx.x == 0) {
      stats[2 * blockIdx.x] = m;
      stats[2 * blockIdx.x + 1] = s;
    }
  }
}

__global__ void __launch_bounds__(256) unc_adaptor_kernel(const UncAdaptParams p) {
  pdl_wait();
  const long long plane = (long long)p.H * p.W;
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= plane) return;
  const int b = blockIdx.y;
  const float* raw = p.raw + (long long)b * p.C * plane + i;
  for (int a = 0; a < p.n; ++a) {
    const float* x = raw + (long long)p.ch[a] * plane;
    switch (p.kind[a]) {
      case ADAPT_MASK:
        p.mask[(long long)b * plane + i] = 1.0f / (1.0f + expf(-x[0]));
        break;
      case ADAPT_COV_EXP_TANH: {
        const float c1 = x[0], c2 = x[plane], t = tanhf(x[2 * plane]);
        const float de = (c1 + c2) / 2, om = 1.f - t * t + 1e-8f;
        float* cov = p.cov + (long long)b * 3 * plane + i;
        cov[0] = expf(c1);
        cov[plane] = expf(c2);
        cov[2 * plane] = t * expf(de);
        p.cov_log_det[(long long)b * plane + i] = c1 + c2 + logf(om);
        const float ic = 1.f / om;
        float* inv = p.cov_inv + (long long)b * 3 * plane + i;
        inv[0] = ic * expf(-c1);
        inv[plane] = ic * expf(-c2);
        inv[2 * plane] = ic * (-t * expf(-de));
        break;
      }
      case ADAPT_CONF_EXP:
        p.conf[(long long)b * plane + i] = p.vmin[a] + fminf(expf(x[0]), p.vmax[a] - p.vmin[a]);
        break;
      case ADAPT_CONF_SIGMOID:
        p.conf[(long long)b * plane + i] = (1.0f / (1.0f + expf(-x[0]))) * (p.vmax[a] - p.vmin[a]) + p.vmin[a];
        break;
      case ADAPT_CONF_SOFTMAX:
        p.conf[(long long)b * plane + i] = expf(x[0] - p.conf_stats[2 * b]) / p.conf_stats[2 * b + 1] * (float)plane;
        break;
      default:
        break;
    }
  }
}

}  // namespace

cudaError_t launch_prep_resize(const PrepParams& p, cudaStream_t s) {
  if (p.B <= 0) return cudaSuccess;
  if (p.H > 65535 || p.B > 65535) return cudaErrorInvalidValue;
  // the tap loops are bounded by kMaxTaps per axis: refuse a down-scale they cannot represent instead of truncating it
  if (2.f * p.scale_x + 2.f > (float)kMaxTaps || 2.f * p.scale_y + 2.f > (float)kMaxTaps) return cudaErrorInvalidValue;
  dim3 grid((p.W + 127) / 128, p.H, p.B);
  if (p.is_u8) {
    if (p.nhwc) launch_k(prep_resize_kernel<uint8_t, true>, grid, 128, 0, s, p);
    else launch_k(prep_resize_kernel<uint8_t, false>, grid, 128, 0, s, p);
  } else {
    if (p.nhwc) launch_k(prep_resize_kernel<float, true>, grid, 128, 0, s, p);
    else launch_k(prep_resize_kernel<float, false>, grid, 128, 0, s, p);
  }
  return cudaGetLastError();
}

cudaError_t launch_unmap(const UnmapParams& p, cudaStream_t s) {
  if (p.B <= 0) return cudaSuccess;
  if (p.H0 > 65535 || p.B > 65535) return cudaErrorInvalidValue;
  dim3 grid((p.W0 + 127) / 128, p.H0, p.B);
  launch_k(unmap_kernel, grid, 128, 0, s, p);
  return cudaGetLastError();
}

cudaError_t launch_unc_adaptors(const UncAdaptParams& p, cudaStream_t s) {
  if (p.B <= 0 || p.n <= 0) return cudaSuccess;
  if (p.B > 65535) return cudaErrorInvalidValue;
  const long long plane = (long long)p.H * p.W;
  for (int a = 0; a < p.n; ++a)
    if (p.kind[a] == ADAPT_CONF_SOFTMAX) {
      if (!p.conf_stats) return cudaErrorInvalidValue;
      launch_k(conf_softmax_stats_kernel, p.B, 1024, 0, s, p.raw, p.C, p.ch[a], plane, p.conf_stats);
    }
  dim3 grid((unsigned)((plane + 255) / 256), p.B);
  launch_k(unc_adaptor_kernel, grid, 256, 0, s, p);
  return cudaGetLastError();
}

UFM_DEFINE_PDL_EARLY_SETTER(prepost)

}  // namespace ufm
