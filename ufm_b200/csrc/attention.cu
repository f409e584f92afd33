// Flash-style attention forward for sm_100a (head_dim 64, non-causal, no mask), tcgen05 + TMEM + TMA.
//
// Replaces every F.scaled_dot_product_attention site of the reference: the DINOv2 SDPA wrapper
// (UC/models/encoders/dinov2.py:128-141, 1 + h*w tokens per image) and Attention.forward
// (UC/models/utils/transformer_blocks.py:174-176; 2*h*w tokens per pair for global attention, h*w per view on the
// odd layers of alternating attention).
//
// One CTA = one 128-row query tile of one (sequence, head); TWO CTAs are resident per SM (256 TMEM columns and
// ~112 KB of shared memory each), so the tensor pipe, the MUFU (exp2) pipe and the TMA engine of an SM are shared by
// two independent tiles: while one tile is in its softmax the other one's MMAs run, and the prologue / epilogue of a
// CTA overlaps the main loop of its neighbour.  With head_dim 64 the kernel is bound by the 16 exp2/clk/SM of the
// MUFU pipe (1024 clk per 128x128 tile against 512 clk of MMA), not by the tensor pipe.
//   warp 0      TMA: Q once; K_j,V_j through a 3-stage ring, straight out of the fused qkv buffer
//               ([tokens, 3, heads, 64]) with a 4-D tensor map, so no head-major copy of q/k/v is ever made
//   warp 1      MMA issuer (one thread).  S = Q K_j^T: A=Q, B=K_j (both K-major SW128 smem).  O += P V_j: A=P read
//               from TENSOR MEMORY (bf16 pairs written there by the softmax threads, no smem round trip), B=V_j as
//               MN-major SW128 (the row-major [kv,64] tile as loaded).  S_{j+1} is issued as soon as the softmax
//               threads have pulled S_j into registers, so it runs underneath the exponentials of step j.
//   warp 2      TMEM allocator: S (128 cols), P (64 cols, bf16x2), O (64 cols)
//   warps 4-7   softmax warpgroup, one query row per thread, the whole 128-wide S row in registers (setmaxnreg):
//               online softmax in fp32 with exp2 and lazy rescale (O is rescaled in TMEM only when the row max grows by
//               more than 2^8), P -> bf16 -> TMEM
#include <stdio.h>

#include "kernels.h"
#include "ptx.cuh"

namespace ufm {

namespace {

constexpr int kThreads = 256;
constexpr int TILE_BYTES = 128 * 64 * 2;  // 16 KB: one [128 x 64] bf16 SW128 tile
constexpr int KV_STAGES = 3;
constexpr int OFF_Q = 0;
constexpr int OFF_K = OFF_Q + TILE_BYTES;
constexpr int OFF_V = OFF_K + KV_STAGES * TILE_BYTES;
constexpr int OFF_BAR = OFF_V + KV_STAGES * TILE_BYTES;
constexpr int SMEM_BYTES = OFF_BAR + 128;  // 114816 B: two CTAs per SM
constexpr int TMEM_COLS = 256;
constexpr uint32_t TM_S = 0, TM_P = 128, TM_O = 192;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ void tmem_ld32_at(uint32_t taddr, uint32_t* v) {
  tmem_ld32(taddr, *reinterpret_cast<uint32_t(*)[32]>(v));
}
__device__ __forceinline__ void tmem_st32_at(uint32_t taddr, const uint32_t* v) {
  tmem_st32(taddr, *reinterpret_cast<const uint32_t(*)[32]>(v));
}

__global__ void __launch_bounds__(kThreads, 2) attn_kernel(const __grid_constant__ AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars;                   // 1
  uint64_t* kv_full = bars + 1;              // KV_STAGES
  uint64_t* kv_empty = kv_full + KV_STAGES;  // KV_STAGES
  uint64_t* s_full = kv_empty + KV_STAGES;   // S_j in TMEM            (MMA -> softmax)
  uint64_t* s_free = s_full + 1;             // S_j pulled to registers (softmax -> MMA), 4 warp arrivals
  uint64_t* p_full = s_free + 1;             // P_j in TMEM             (softmax -> MMA), 4 warp arrivals
  uint64_t* o_done = p_full + 1;             // O += P_j V_j finished   (MMA -> softmax)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int qblk = blockIdx.x, head = blockIdx.y, seq = blockIdx.z;
  const int q0 = qblk * 128;
  const int n_kv = (p.seq_len + 127) / 128;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) {
      printf("ufm: attention shared memory window is not 1024-byte aligned\n");
      __trap();
    }
    tma_prefetch_desc(&p.tmQKV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    setmaxnreg_dec<48>();
    if (warp == 0 && lane == 0) {
      mbar_expect_tx(q_full, TILE_BYTES);
      tma_load_4d(smem + OFF_Q, &p.tmQKV, q_full, 0, head, q0, seq);
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
    } else if (warp == 1 && lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);  // B = V, MN-major
      const uint32_t q_addr = smem_u32(smem + OFF_Q);
      const uint32_t k_addr = smem_u32(smem + OFF_K);
      const uint32_t v_addr = smem_u32(smem + OFF_V);
      auto issue_s = [&](int stage) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_ss(tmem_base + TM_S, umma_desc_sw128(q_addr + kk * 32),
                  umma_desc_sw128(k_addr + stage * TILE_BYTES + kk * 32), idesc_s, kk != 0);
        umma_commit(s_full);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < n_kv; ++j) {
        const int stage = j % KV_STAGES;
        if (j + 1 < n_kv) {
          const int nstage = (j + 1) % KV_STAGES;
          mbar_wait(&kv_full[nstage], ((j + 1) / KV_STAGES) & 1);
          mbar_wait(s_free, j & 1);  // S_j is in the softmax threads' registers: its TMEM columns may be overwritten
          tc_fence_after();
          issue_s(nstage);
        }
        mbar_wait(p_full, j & 1);
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          umma_ts(tmem_base + TM_O, tmem_base + TM_P + ks * 8, umma_desc_sw128(v_addr + stage * TILE_BYTES + ks * 2048),
                  idesc_pv, (j | ks) != 0);
        umma_commit(o_done);
        umma_commit(&kv_empty[stage]);
      }
    }
  } else {
    setmaxnreg_inc<200>();
    const int q = warp & 3;         // TMEM lane quadrant
    const int r = q * 32 + lane;    // row within the tile
    const uint32_t lane_addr = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t s_addr = lane_addr + TM_S;
    const uint32_t p_addr = lane_addr + TM_P;
    const uint32_t o_addr = lane_addr + TM_O;
    float m_ref = -INFINITY, l = 0.f;
    const float sl2 = p.scale_log2;
    for (int j = 0; j < n_kv; ++j) {
      uint32_t s[128];
      mbar_wait(s_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32_at(s_addr + c * 32, s + c * 32);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_free);
      const int kv_left = p.seq_len - j * 128;  // valid keys in this tile (>= 1)
      if (kv_left < 128) {
#pragma unroll
        for (int i = 0; i < 128; ++i)
          if (i >= kv_left) s[i] = 0xff800000u;  // -inf
      }
      // 4 independent max chains (one 128-deep dependent chain keeps the warp off the MUFU pipe for ~500 clk per tile)
      float mx[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) mx[i] = __uint_as_float(s[i]);
#pragma unroll
      for (int i = 4; i < 128; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(s[i]));
      const float tmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      if (j == 0) {
        m_ref = tmax;
      } else {
        const bool need = (tmax - m_ref) * sl2 > 8.0f;
        if (__any_sync(0xffffffffu, need)) {
          mbar_wait(o_done, (j - 1) & 1);  // PV_{j-1} finished: O stable
          tc_fence_after();
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
      // p = exp2((s - m_ref) * scale*log2e), row sum, bf16 pairs
      const float mb = m_ref * sl2;
      float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        const float p0 = ex2(fmaf(__uint_as_float(s[2 * i]), sl2, -mb));
        const float p1 = ex2(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -mb));
        sum0 += p0;
        sum1 += p1;
        s[i] = pack_bf16(p0, p1);
      }
      l += sum0 + sum1;
      if (j > 0) {
        mbar_wait(o_done, (j - 1) & 1);  // PV_{j-1} has consumed the previous P
        tc_fence_after();
      }
      tmem_st32_at(p_addr, s);
      tmem_st32_at(p_addr + 32, s + 32);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    // epilogue: O / l -> bf16
    mbar_wait(o_done, (n_kv - 1) & 1);
    tc_fence_after();
    const int row = q0 + r;
    const float inv_l = 1.0f / l;
    __nv_bfloat16* op = p.out + ((long long)seq * p.seq_len + row) * p.out_pitch + head * 64;
    uint32_t v[64];
    tmem_ld32_at(o_addr, v);
    tmem_ld32_at(o_addr + 32, v + 32);
    tmem_ld_wait();
    if (row < p.seq_len) {
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[8 * g + i]) * inv_l;
        *reinterpret_cast<uint4*>(op + g * 8) =
            make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
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
  dim3 grid((p.seq_len + 127) / 128, p.heads, p.n_seq);
  attn_kernel<<<grid, kThreads, SMEM_BYTES, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace ufm
