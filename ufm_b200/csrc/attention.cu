// Flash-style attention forward for sm_100a (head_dim 64, non-causal, no mask), tcgen05 + TMEM + TMA.
//
// Replaces every F.scaled_dot_product_attention site of the reference: the DINOv2 SDPA wrapper
// (UC/models/encoders/dinov2.py:128-141, 1 + h*w tokens per image) and Attention.forward
// (UC/models/utils/transformer_blocks.py:174-176; 2*h*w tokens per pair for global attention, h*w per view on the
// odd layers of alternating attention).
//
// One CTA handles 256 query rows of one (sequence, head) as two 128-row tiles that ping-pong on the tensor core
// (S_t = Q_t K_j^T and O_t += P_t V_j issued by one thread; two softmax warpgroups, one row per thread):
//   warp 0      TMA: Q0,Q1 once; K_j,V_j through a 3-stage ring, straight out of the fused qkv buffer
//               ([tokens, 3, heads, 64]) with a 4-D tensor map, so no head-major copy of q/k/v is ever made
//   warp 1      MMA issuer.  S: A=Q (K-major SW128), B=K (K-major).  PV: A=P (K-major SW128, written by the softmax
//               threads), B=V as MN-major SW128 (the row-major [kv,64] tile as loaded), fp32 accumulators in TMEM
//   warp 2      TMEM allocator: S0,S1 (128 cols each), O0,O1 (64 cols each)
//   warps 4-7 / 8-11  softmax warpgroup for tile 0 / 1: online softmax in fp32 with exp2 and lazy rescale
//               (O is rescaled in TMEM only when the row max grows by more than 2^8), P -> bf16 -> swizzled smem
#include <stdio.h>

#include "kernels.h"
#include "ptx.cuh"

namespace ufm {

namespace {

constexpr int kThreads = 384;
constexpr int TILE_BYTES = 128 * 64 * 2;  // 16 KB: one [128 x 64] bf16 SW128 tile
constexpr int KV_STAGES = 3;
constexpr int OFF_Q = 0;                                   // 2 tiles
constexpr int OFF_K = OFF_Q + 2 * TILE_BYTES;              // KV_STAGES tiles
constexpr int OFF_V = OFF_K + KV_STAGES * TILE_BYTES;      // KV_STAGES tiles
constexpr int OFF_P = OFF_V + KV_STAGES * TILE_BYTES;      // 2 x (2 tiles)
constexpr int OFF_BAR = OFF_P + 4 * TILE_BYTES;
constexpr int SMEM_BYTES = 1024 + OFF_BAR + 256;
constexpr int TMEM_COLS = 512;
constexpr uint32_t TM_S = 0, TM_O = 256;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kThreads, 1) attn_kernel(const __grid_constant__ AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars;                   // 1
  uint64_t* kv_full = bars + 1;              // KV_STAGES
  uint64_t* kv_empty = kv_full + KV_STAGES;  // KV_STAGES
  uint64_t* s_full = kv_empty + KV_STAGES;   // 2
  uint64_t* p_full = s_full + 2;             // 2
  uint64_t* o_done = p_full + 2;             // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int qblk = blockIdx.x, head = blockIdx.y, seq = blockIdx.z;
  const int q0 = qblk * 256;
  const int n_kv = (p.seq_len + 127) / 128;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&p.tmQKV);
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 128);
      mbar_init(&o_done[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, 2 * TILE_BYTES);
      tma_load_4d(smem + OFF_Q, &p.tmQKV, q_full, 0, head, q0, seq);
      tma_load_4d(smem + OFF_Q + TILE_BYTES, &p.tmQKV, q_full, 0, head, q0 + 128, seq);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < n_kv; ++j) {
        mbar_wait(&kv_empty[stage], phase ^ 1);
        mbar_expect_tx(&kv_full[stage], 2 * TILE_BYTES);
        tma_load_4d(smem + OFF_K + stage * TILE_BYTES, &p.tmQKV, &kv_full[stage], 0, p.heads + head, j * 128, seq);
        tma_load_4d(smem + OFF_V + stage * TILE_BYTES, &p.tmQKV, &kv_full[stage], 0, 2 * p.heads + head, j * 128, seq);
        if (++stage == KV_STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);  // B = V, MN-major
      const uint32_t q_addr = smem_u32(smem + OFF_Q);
      const uint32_t k_addr = smem_u32(smem + OFF_K);
      const uint32_t v_addr = smem_u32(smem + OFF_V);
      const uint32_t p_addr = smem_u32(smem + OFF_P);
      auto issue_s = [&](int t, int stage) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_ss(tmem_base + TM_S + t * 128, umma_desc_sw128(q_addr + t * TILE_BYTES + kk * 32),
                  umma_desc_sw128(k_addr + stage * TILE_BYTES + kk * 32), idesc_s, kk != 0);
        umma_commit(&s_full[t]);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      issue_s(0, 0);
      issue_s(1, 0);
      for (int j = 0; j < n_kv; ++j) {
        const int stage = j % KV_STAGES;
        const int nstage = (j + 1) % KV_STAGES;
        const uint32_t nparity = ((j + 1) / KV_STAGES) & 1;
        for (int t = 0; t < 2; ++t) {
          mbar_wait(&p_full[t], j & 1);
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_ss(tmem_base + TM_O + t * 64,
                    umma_desc_sw128(p_addr + t * 2 * TILE_BYTES + (ks >> 2) * TILE_BYTES + (ks & 3) * 32),
                    umma_desc_sw128(v_addr + stage * TILE_BYTES + ks * 2048), idesc_pv, (j | ks) != 0);
          umma_commit(&o_done[t]);
          if (t == 1) umma_commit(&kv_empty[stage]);
          if (j + 1 < n_kv) {
            if (t == 0) {
              mbar_wait(&kv_full[nstage], nparity);
              tc_fence_after();
            }
            issue_s(t, nstage);
          }
        }
      }
    }
  } else if (warp >= 4) {
    const int t = (warp - 4) >> 2;  // q tile
    const int q = warp & 3;         // TMEM lane quadrant
    const int r = q * 32 + lane;    // row within the tile
    const uint32_t lane_addr = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t s_addr = lane_addr + TM_S + t * 128;
    const uint32_t o_addr = lane_addr + TM_O + t * 64;
    uint8_t* p_row = smem + OFF_P + t * 2 * TILE_BYTES + (r >> 3) * 1024 + (r & 7) * 128;
    const int sw = r & 7;
    float m_ref = -INFINITY, l = 0.f;
    const float sl2 = p.scale_log2;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(&s_full[t], j & 1);
      tc_fence_after();
      const int kv_left = p.seq_len - j * 128;  // valid keys in this tile (>= 1)
      // pass 1: row max over the valid keys
      float tmax = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(s_addr + c * 32, v);
        tmem_ld_wait();
        if (kv_left >= (c + 1) * 32) {
#pragma unroll
          for (int i = 0; i < 32; ++i) tmax = fmaxf(tmax, __uint_as_float(v[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i < kv_left) tmax = fmaxf(tmax, __uint_as_float(v[i]));
        }
      }
      if (j == 0) {
        m_ref = tmax;
      } else {
        const bool need = (tmax - m_ref) * sl2 > 8.0f;
        const bool any = __any_sync(0xffffffffu, need);
        mbar_wait(&o_done[t], (j - 1) & 1);  // PV_{j-1} finished: P buffer free, O stable
        tc_fence_after();
        if (any) {
          float alpha = 1.0f;
          if (need) {
            alpha = ex2((m_ref - tmax) * sl2);
            m_ref = tmax;
          }
          l *= alpha;
#pragma unroll 1
          for (int c = 0; c < 2; ++c) {
            uint32_t v[32];
            tmem_ld32(o_addr + c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st32(o_addr + c * 32, v);
          }
          tmem_st_wait();
        }
      }
      // pass 2: p = exp2((s - m_ref) * scale*log2e), row sum, bf16 P tile into swizzled smem
      const float mb = m_ref * sl2;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(s_addr + c * 32, v);
        tmem_ld_wait();
        float pv[32];
        if (kv_left >= (c + 1) * 32) {
#pragma unroll
          for (int i = 0; i < 32; ++i) pv[i] = ex2(fmaf(__uint_as_float(v[i]), sl2, -mb));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            pv[i] = (c * 32 + i < kv_left) ? ex2(fmaf(__uint_as_float(v[i]), sl2, -mb)) : 0.f;
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) s += pv[i];
        l += s;
        uint8_t* atom = p_row + (c >> 1) * TILE_BYTES;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = (c & 1) * 4 + g;
          *reinterpret_cast<uint4*>(atom + ((chunk ^ sw) << 4)) =
              make_uint4(pack_bf16(pv[8 * g], pv[8 * g + 1]), pack_bf16(pv[8 * g + 2], pv[8 * g + 3]),
                         pack_bf16(pv[8 * g + 4], pv[8 * g + 5]), pack_bf16(pv[8 * g + 6], pv[8 * g + 7]));
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(&p_full[t]);
    }
    // epilogue: O / l -> bf16
    mbar_wait(&o_done[t], (n_kv - 1) & 1);
    tc_fence_after();
    const int row = q0 + t * 128 + r;
    const float inv_l = 1.0f / l;
    __nv_bfloat16* op = p.out + ((long long)seq * p.seq_len + row) * p.out_pitch + head * 64;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t v[32];
      __syncwarp();  // tcgen05.ld is .sync.aligned
      tmem_ld32(o_addr + c * 32, v);
      tmem_ld_wait();
      if (row < p.seq_len) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[8 * g + i]) * inv_l;
          *reinterpret_cast<uint4*>(op + c * 32 + g * 8) =
              make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

}  // namespace

cudaError_t launch_attention(const AttnParams& p, cudaStream_t stream) {
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    attr = true;
  }
  dim3 grid((p.seq_len + 255) / 256, p.heads, p.n_seq);
  attn_kernel<<<grid, kThreads, SMEM_BYTES, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace ufm
