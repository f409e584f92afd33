// Flash-style attention forward for sm_100a (head_dim 64, non-causal, no mask), tcgen05 + TMEM + TMA.
//
// Replaces every F.scaled_dot_product_attention site of the reference: the DINOv2 SDPA wrapper
// (UC/models/encoders/dinov2.py:128-141, 1 + h*w tokens per image) and Attention.forward
// (UC/models/utils/transformer_blocks.py:174-176; 2*h*w tokens per pair for global attention, h*w per view on the
// odd layers of alternating attention).
//
// PERSISTENT kernel, two CTAs resident per SM (256 TMEM columns and ~97 KB of shared memory each): every CTA walks
// a strided list of work items (one item = one 128-row query tile of one (sequence, head)), so barrier setup, the
// TMEM allocation and the first TMA round trip are paid once per CTA instead of once per tile, and the K/V/Q
// prefetch of the next tile runs underneath the last iterations of the current one.  The tensor pipe, the MUFU
// (exp2) pipe and the TMA engine of an SM are shared by two independent tile streams.  With head_dim 64 the kernel
// is bound by the 16 exp2/clk/SM of the MUFU pipe (1024 clk per 128x128 tile against 512 clk of MMA), not by the
// tensor pipe.
//   warp 0      TMA: Q through a 2-deep ring, K_j and V_j through separate 2-deep rings, straight out of the fused
//               qkv buffer ([tokens, 3, heads, 64]) with a 4-D tensor map: no head-major copy of q/k/v is ever made
//   warp 1      MMA issuer (one thread).  S = Q K_j^T: A=Q, B=K_j (both K-major SW128 smem).  O += P V_j: A=P read
//               from TENSOR MEMORY (bf16 pairs written there by the softmax threads, no smem round trip), B=V_j as
//               MN-major SW128 (the row-major [kv,64] tile as loaded).  S_{g+1} is issued as soon as the softmax
//               threads have pulled S_g into registers, so it runs underneath the exponentials of step g -- also
//               across tile boundaries.  A ragged last K/V tile with <= 64 keys is run as N = 64 / K = 64 MMAs.
//   warp 2      TMEM allocator: S (128 cols), P (64 cols, bf16x2), O (64 cols)
//   warps 4-7   softmax warpgroup, one query row per thread, the whole S row in registers: 3-input max tree,
//               packed f32x2 FMA / ADD around the exp2, online softmax with lazy rescale (O is rescaled in TMEM only
//               when the row max grows by more than 2^8), P -> bf16 -> TMEM; then the tile epilogue O / l -> bf16.
#include <stdio.h>
#include <stdlib.h>

#include "kernels.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace ufm {

namespace {

constexpr int kThreads = 256;
constexpr int TILE_BYTES = 128 * 64 * 2;  // 16 KB: one [128 x 64] bf16 SW128 tile
constexpr int Q_STAGES = 2, K_STAGES = 3, V_STAGES = 2;
constexpr int OFF_Q = 0;
constexpr int OFF_K = OFF_Q + Q_STAGES * TILE_BYTES;
constexpr int OFF_V = OFF_K + K_STAGES * TILE_BYTES;
constexpr int OFF_BAR = OFF_V + V_STAGES * TILE_BYTES;
constexpr int SMEM_BYTES = OFF_BAR + 256;  // 114944 B: two CTAs per SM
constexpr int TMEM_COLS = 256;
constexpr uint32_t TM_S = 0, TM_P = 128, TM_O = 192;
int g_attn_poly = 132;  // low 4 bits: every n-th pair of exponentials of a full tile runs on the FMA pipe (0 = never);
                         // bit 7: streamed softmax steps with a speculative reference maximum; ufm_debug_flag 2

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// packed f32x2 arithmetic (one issue slot for two lanes of work)
__device__ __forceinline__ uint64_t pack2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

__device__ __forceinline__ uint64_t add2_rm(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rm.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// exp2 of two values on the FMA / ALU pipes (Cody-Waite split + degree-3 minimax polynomial of 2^f on [0,1), relative
// error 8.8e-5: far below the bf16 rounding of P), so that a fixed share of the exponentials does not occupy the
// MUFU pipe, which bounds this kernel.  x + 1.5*2^23 rounded DOWN leaves floor(x) in the low mantissa bits; the
// integer add of (bits << 23) puts it into the exponent of the polynomial value.
__device__ __forceinline__ void ex2_poly2(float x0, float x1, float& p0, float& p1) {
  constexpr float kMagic = 12582912.0f;
  x0 = fmaxf(x0, -126.0f);
  x1 = fmaxf(x1, -126.0f);
  const uint64_t x = pack2(x0, x1);
  const uint64_t t = add2_rm(x, pack2(kMagic, kMagic));
  const uint64_t fl = add2(t, pack2(-kMagic, -kMagic));
  float fl0, fl1;
  unpack2(fl, fl0, fl1);
  const uint64_t f = add2(x, pack2(-fl0, -fl1));
  uint64_t r = fma2(f, pack2(0.077119089663028717041015625f, 0.077119089663028717041015625f),
                    pack2(0.227564394474029541015625f, 0.227564394474029541015625f));
  r = fma2(r, f, pack2(0.695146143436431884765625f, 0.695146143436431884765625f));
  r = fma2(r, f, pack2(1.0f, 1.0f));
  float t0, t1, r0, r1;
  unpack2(t, t0, t1);
  unpack2(r, r0, r1);
  p0 = __uint_as_float(__float_as_uint(r0) + (__float_as_uint(t0) << 23));
  p1 = __uint_as_float(__float_as_uint(r1) + (__float_as_uint(t1) << 23));
}

__device__ __forceinline__ void tmem_ld32_at(uint32_t taddr, uint32_t* v) {
  tmem_ld32(taddr, *reinterpret_cast<uint32_t(*)[32]>(v));
}
__device__ __forceinline__ void tmem_st32_at(uint32_t taddr, const uint32_t* v) {
  tmem_st32(taddr, *reinterpret_cast<const uint32_t(*)[32]>(v));
}

// max over N values as a tree of 3-input maxima (FMNMX3): depth ~log3(N) instead of a few long dependent chains
template <int N>
__device__ __forceinline__ float tree_max(const float (&v)[N]) {
  if constexpr (N == 1) {
    return v[0];
  } else if constexpr (N == 2) {
    return fmaxf(v[0], v[1]);
  } else {
    constexpr int M = (N + 2) / 3;
    float w[M];
#pragma unroll
    for (int i = 0; i < N / 3; ++i) w[i] = fmaxf(fmaxf(v[3 * i], v[3 * i + 1]), v[3 * i + 2]);
    if constexpr (N % 3 == 1) w[M - 1] = v[N - 1];
    if constexpr (N % 3 == 2) w[M - 1] = fmaxf(v[N - 2], v[N - 1]);
    return tree_max<M>(w);
  }
}

struct Bars {
  uint64_t q_full[Q_STAGES], q_free[Q_STAGES];
  uint64_t k_full[K_STAGES], k_empty[K_STAGES];
  uint64_t v_full[V_STAGES], v_empty[V_STAGES];
  uint64_t s_full;  // S_g in TMEM             (MMA -> softmax)
  uint64_t s_free;  // S_g pulled to registers  (softmax -> MMA), 4 warp arrivals
  uint64_t p_full;  // P_g in TMEM              (softmax -> MMA), 4 warp arrivals
  uint64_t o_done;  // O += P_g V_g finished    (MMA -> softmax)
  uint32_t tmem_slot;
};
static_assert(sizeof(Bars) <= 256, "barrier block");

// One softmax step on NC (= 128, or 64 for a short ragged tail) key columns.  g = global step index of this CTA
// (runs on across tiles: it is the phase counter of the s/p/o barriers), j = step index within the tile.
template <int NC, int VAR>
__device__ __forceinline__ void softmax_step(Bars* b, uint32_t s_addr, uint32_t p_addr, uint32_t o_addr, uint32_t g, int j,
                                             int kv_left, float sl2, float& m_ref, float& l, int lane) {
  constexpr int POLY = VAR & 15;
  constexpr bool SCALAR = (VAR & 16) != 0, SPLIT_LD = (VAR & 32) != 0 && NC == 128, EARLY_ST = (VAR & 64) != 0 && NC == 128;
  uint32_t s[NC];
  mbar_wait(&b->s_full, g & 1);
  tc_fence_after();
  float tmax;
  if constexpr (SPLIT_LD) {
    // second half of the row is still in flight while the maximum of the first half is taken
    tmem_ld32_at(s_addr, s);
    tmem_ld32_at(s_addr + 32, s + 32);
    tmem_ld_wait();
    tmem_ld32_at(s_addr + 64, s + 64);
    tmem_ld32_at(s_addr + 96, s + 96);
    if (kv_left < 64) {
#pragma unroll
      for (int i = 0; i < 64; ++i)
        if (i >= kv_left) s[i] = 0xff800000u;
    }
    const float ma = tree_max<64>(*reinterpret_cast<const float(*)[64]>(s));
    tmem_ld_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&b->s_free);
    if (kv_left < NC) {
#pragma unroll
      for (int i = 64; i < NC; ++i)
        if (i >= kv_left) s[i] = 0xff800000u;
    }
    tmax = fmaxf(ma, tree_max<64>(*reinterpret_cast<const float(*)[64]>(s + 64)));
  } else {
#pragma unroll
    for (int c = 0; c < NC / 32; ++c) tmem_ld32_at(s_addr + c * 32, s + c * 32);
    tmem_ld_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&b->s_free);
    if (kv_left < NC) {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (i >= kv_left) s[i] = 0xff800000u;  // -inf
    }
    tmax = tree_max<NC>(*reinterpret_cast<const float(*)[NC]>(s));
  }
  if (j == 0) {
    m_ref = tmax;
  } else {
    const bool need = (tmax - m_ref) * sl2 > 8.0f;
    if (__any_sync(0xffffffffu, need)) {
      mbar_wait(&b->o_done, (g - 1) & 1);  // PV_{g-1} finished: O stable
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
  // p = exp2(s * scale*log2e - m_ref * scale*log2e), row sum, bf16 pairs
  const float mb = -m_ref * sl2;
  const uint64_t sl2_2 = pack2(sl2, sl2), mb_2 = pack2(mb, mb);
  uint64_t sum_a = pack2(0.f, 0.f), sum_b = sum_a;
  float fs0 = 0.f, fs1 = 0.f;
#pragma unroll
  for (int i = 0; i < NC / 2; ++i) {
    float x0, x1, p0, p1;
    const bool poly = POLY > 0 && (i % (POLY > 0 ? POLY : 1)) == POLY - 1;
    if (SCALAR && !poly) {
      x0 = fmaf(__uint_as_float(s[2 * i]), sl2, mb);
      x1 = fmaf(__uint_as_float(s[2 * i + 1]), sl2, mb);
    } else {
      unpack2(fma2(pack2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), sl2_2, mb_2), x0, x1);
    }
    if (poly) {
      ex2_poly2(x0, x1, p0, p1);
    } else {
      p0 = ex2(x0);
      p1 = ex2(x1);
    }
    if (SCALAR) {
      fs0 += p0;
      fs1 += p1;
    } else {
      if (i & 1) sum_b = add2(sum_b, pack2(p0, p1));
      else sum_a = add2(sum_a, pack2(p0, p1));
    }
    s[i] = pack_bf16(p0, p1);
    if (EARLY_ST && i == NC / 4 - 1) {
      // first half of P leaves for tensor memory while the second half is still being exponentiated
      if (j > 0) {
        mbar_wait(&b->o_done, (g - 1) & 1);  // PV_{g-1} has consumed the previous P
        tc_fence_after();
      }
      tmem_st32_at(p_addr, s);
    }
  }
  {
    float a0, a1;
    unpack2(add2(sum_a, sum_b), a0, a1);
    l += (a0 + a1) + (fs0 + fs1);
  }
  if constexpr (EARLY_ST) {
    tmem_st32_at(p_addr + 32, s + 32);
  } else {
    if (j > 0) {
      mbar_wait(&b->o_done, (g - 1) & 1);  // PV_{g-1} has consumed the previous P
      tc_fence_after();
    }
#pragma unroll
    for (int c = 0; c < NC / 64; ++c) tmem_st32_at(p_addr + c * 32, s + c * 32);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(&b->p_full);
}

// Exponentials of one 32-column chunk of a score row (in place in s[0..32)): p = exp2(s * sl2 + mb), every POLY-th pair on
// the FMA pipe; returns the chunk's row sum in (sum_a, sum_b) and the bf16 pairs in pk[0..16).
template <int POLY>
__device__ __forceinline__ void exp_chunk32(uint32_t* s, uint32_t* pk, uint64_t sl2_2, uint64_t mb_2, uint64_t& sum_a,
                                            uint64_t& sum_b, int phase) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float x0, x1, p0, p1;
    const bool poly = POLY > 0 && ((i + phase) % (POLY > 0 ? POLY : 1)) == POLY - 1;
    unpack2(fma2(pack2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), sl2_2, mb_2), x0, x1);
    if (poly) {
      ex2_poly2(x0, x1, p0, p1);
    } else {
      p0 = ex2(x0);
      p1 = ex2(x1);
    }
    if (i & 1) sum_b = add2(sum_b, pack2(p0, p1));
    else sum_a = add2(sum_a, pack2(p0, p1));
    pk[i] = pack_bf16(p0, p1);
  }
}

// STREAMED softmax step for a full 128-key tile that is not the first of its query tile (j > 0): the row maximum is not
// waited for.  The running reference maximum m_ref of the previous steps is used speculatively -- the online softmax
// only needs ANY reference within 2^8 of the true maximum (lazy rescale), and it changes on a few percent of the steps --
// so the exponentials of chunk 0 start as soon as its 32 columns have arrived from tensor memory, while chunks 1..3 are
// still in flight and the max tree (FMNMX3, ALU pipe) runs beside the FFMA2 / MUFU stream instead of in front of it.
// S leaves tensor memory in four 32-column chunks and P goes back in four 16-column chunks; these loads / stores are
// ordered (volatile), which pins the chunk pipeline in the instruction stream: the FMA-pipe work of chunk k+1 is issued
// between the MUFU instructions of chunk k instead of all FMA work first and one long MUFU burst at the end (what ptxas
// makes of the one-shot version: ncu source view of r01, 25 % of a softmax warp's time in an FMA-only phase and 38 % in
// a MUFU-only phase).  s_free is signalled once chunk 3 is in registers and the speculation is confirmed; if the row
// maximum did grow by more than 2^8 (any lane of the warp), nothing has been published yet and the step is redone by
// the one-shot path below (S is still in tensor memory, P is simply overwritten).
// Shared-memory addresses of the barriers a softmax warp touches, computed once per kernel (see opaque_u32)
struct SoftmaxBars {
  uint32_t s_full, s_free, p_full, o_done;
};

template <int VAR>
__device__ __forceinline__ bool softmax_step_stream(const SoftmaxBars& sb, uint32_t s_addr, uint32_t p_addr, uint32_t g, float sl2,
                                                    float m_ref, float& l, int lane) {
  constexpr int POLY = VAR & 15;
  uint32_t c0[32], c1[32], c2[32], c3[32];
  uint32_t pk0[16], pk1[16];
  mbar_wait_a(sb.s_full, g & 1);
  tc_fence_after();
  tmem_ld32(s_addr, c0);
  tmem_ld32(s_addr + 32, c1);
  tmem_ld_wait();
  tmem_ld32(s_addr + 64, c2);
  const float mb = -m_ref * sl2;
  const uint64_t sl2_2 = pack2(sl2, sl2), mb_2 = pack2(mb, mb);
  uint64_t sum_a = pack2(0.f, 0.f), sum_b = sum_a, sum_c = sum_a, sum_d = sum_a;
  float tmax = fmaxf(tree_max<32>(*reinterpret_cast<const float(*)[32]>(c0)),
                     tree_max<32>(*reinterpret_cast<const float(*)[32]>(c1)));
  exp_chunk32<POLY>(c0, pk0, sl2_2, mb_2, sum_a, sum_b, 0);
  tmem_ld_wait();  // chunk 2 has arrived
  tmem_ld32(s_addr + 96, c3);
  tmax = fmaxf(tmax, tree_max<32>(*reinterpret_cast<const float(*)[32]>(c2)));
  exp_chunk32<POLY>(c1, pk1, sl2_2, mb_2, sum_c, sum_d, 1);
  tmem_ld_wait();  // chunk 3 has arrived: the whole row is in registers
  tmax = fmaxf(tmax, tree_max<32>(*reinterpret_cast<const float(*)[32]>(c3)));
  if (__any_sync(0xffffffffu, (tmax - m_ref) * sl2 > 8.0f)) return false;  // rare: redo with a rescale (one-shot path)
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive_a(sb.s_free);
  // The first P chunks are held back until here: PV_{g-1} (issued when the LAST softmax warp finished step g-1, 256 clk of
  // MMA behind whatever the co-resident CTA queued) is usually still running when chunk 0 is done -- storing P right
  // away put 8.5 % of a softmax warp's samples on this wait (ncu source view, round 2).
  mbar_wait_a(sb.o_done, (g - 1) & 1);  // PV_{g-1} has consumed the previous P (j > 0 on this path)
  tc_fence_after();
  tmem_st16(p_addr, pk0);
  tmem_st16(p_addr + 16, pk1);
  exp_chunk32<POLY>(c2, pk0, sl2_2, mb_2, sum_a, sum_b, 2);
  tmem_st16(p_addr + 32, pk0);
  exp_chunk32<POLY>(c3, pk1, sl2_2, mb_2, sum_c, sum_d, 3);
  tmem_st16(p_addr + 48, pk1);
  {
    float a0, a1;
    unpack2(add2(add2(sum_a, sum_b), add2(sum_c, sum_d)), a0, a1);
    l += a0 + a1;
  }
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive_a(sb.p_full);
  return true;
}

template <int VAR>
__global__ void __launch_bounds__(kThreads, 2) attn_kernel(const __grid_constant__ AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Bars* b = reinterpret_cast<Bars*>(smem + OFF_BAR);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_kv = (p.seq_len + 127) / 128;
  const int kv_tail = p.seq_len - (n_kv - 1) * 128;  // valid keys in the last K/V tile (1..128)
  const bool tail64 = kv_tail <= 64;
  const int n_my = (p.n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles this CTA walks (>= 1)

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) {
      printf("ufm: attention shared memory window is not 1024-byte aligned\n");
      __trap();
    }
    tma_prefetch_desc(&p.tmQKV);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < Q_STAGES; ++i) {
      mbar_init(&b->q_full[i], 1);
      mbar_init(&b->q_free[i], 1);
    }
    for (int i = 0; i < K_STAGES; ++i) {
      mbar_init(&b->k_full[i], 1);
      mbar_init(&b->k_empty[i], 1);
    }
    for (int i = 0; i < V_STAGES; ++i) {
      mbar_init(&b->v_full[i], 1);
      mbar_init(&b->v_empty[i], 1);
    }
    mbar_init(&b->s_full, 1);
    mbar_init(&b->s_free, 4);
    mbar_init(&b->p_full, 4);
    mbar_init(&b->o_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(&b->tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = b->tmem_slot;
  pdl_wait();  // barrier init + TMEM allocation above overlap the previous kernel's tail (launch.cuh)

  if (warp < 4) {
    setmaxnreg_dec<48>();
    if (warp == 0 && lane == 0) {
      // ---------------------------------------------------------------- TMA producer
      // K runs one step ahead of V: K_{g+1} is needed when softmax step g starts (S_{g+1} is issued then), V_g only
      // when it ends, and a V stage frees one step later than a K stage -- loading in the order K_{g+1}, V_g keeps
      // two softmax steps of lead on both rings instead of gating K behind the V wait.
      struct Cur { int t, j, head, seq, qblk; uint32_t g; };
      auto decode = [&](Cur& c) {
        const int w = blockIdx.x + c.t * gridDim.x;
        c.qblk = w % p.n_qblk;
        const int hs = w / p.n_qblk;
        c.head = hs % p.heads;
        c.seq = hs / p.heads;
      };
      auto advance = [&](Cur& c) {
        ++c.g;
        if (++c.j == n_kv) {
          c.j = 0;
          ++c.t;
          if (c.t < n_my) decode(c);
        }
      };
      auto load_k = [&](Cur& c) {
        if (c.j == 0) {
          const int qs = c.t % Q_STAGES;
          if (c.t >= Q_STAGES) mbar_wait(&b->q_free[qs], ((c.t / Q_STAGES) - 1) & 1);
          mbar_expect_tx(&b->q_full[qs], TILE_BYTES);
          tma_load_4d(smem + OFF_Q + qs * TILE_BYTES, &p.tmQKV, &b->q_full[qs], 0, c.head, c.qblk * 128, c.seq);
        }
        const int ks = c.g % K_STAGES;
        mbar_wait(&b->k_empty[ks], ((c.g / K_STAGES) & 1) ^ 1);
        mbar_expect_tx(&b->k_full[ks], TILE_BYTES);
        tma_load_4d(smem + OFF_K + ks * TILE_BYTES, &p.tmQKV, &b->k_full[ks], 0, p.heads + c.head, c.j * 128, c.seq);
        advance(c);
      };
      auto load_v = [&](Cur& c) {
        const int vs = c.g % V_STAGES;
        mbar_wait(&b->v_empty[vs], ((c.g / V_STAGES) & 1) ^ 1);
        mbar_expect_tx(&b->v_full[vs], TILE_BYTES);
        tma_load_4d(smem + OFF_V + vs * TILE_BYTES, &p.tmQKV, &b->v_full[vs], 0, 2 * p.heads + c.head, c.j * 128, c.seq);
        advance(c);
      };
      Cur ck{0, 0, 0, 0, 0, 0}, cv{0, 0, 0, 0, 0, 0};
      decode(ck);
      decode(cv);
      load_k(ck);
      while (cv.t < n_my) {
        if (ck.t < n_my) load_k(ck);
        load_v(cv);
      }
    } else if (warp == 1) {
      // ---------------------------------------------------------------- MMA issuer
      // The whole warp walks the loop (uniform control flow keeps descriptors in uniform registers, see gemm.cu); one
      // elected lane issues the MMAs and the commits.
      constexpr uint32_t idesc_s128 = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_s64 = umma_idesc_bf16(128, 64, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);  // B = V, MN-major
      constexpr uint32_t hi = umma_desc_hi_sw128(1024);
      const uint32_t q_lo0 = umma_desc_lo(smem_u32(smem + OFF_Q));
      const uint32_t k_lo0 = umma_desc_lo(smem_u32(smem + OFF_K));
      const uint32_t v_lo0 = umma_desc_lo(smem_u32(smem + OFF_V));
      const bool leader = elect_one();
      const uint32_t G = (uint32_t)n_my * (uint32_t)n_kv;
      uint32_t gs = 0;       // next S step to issue
      int js = 0, ts = 0;    // its step-in-tile / tile index
      auto issue_s = [&]() {
        const int ks = gs % K_STAGES, qs = ts % Q_STAGES;
        mbar_wait(&b->k_full[ks], (gs / K_STAGES) & 1);
        if (js == 0) mbar_wait(&b->q_full[qs], (ts / Q_STAGES) & 1);
        if (gs > 0) mbar_wait(&b->s_free, (gs - 1) & 1);  // S_{g-1} is in registers: its TMEM columns may be overwritten
        tc_fence_after();
        const bool last = js == n_kv - 1;
        if (leader) {
          const uint32_t idesc = (last && tail64) ? idesc_s64 : idesc_s128;
          const uint32_t q_lo = q_lo0 + qs * (TILE_BYTES >> 4), k_lo = k_lo0 + ks * (TILE_BYTES >> 4);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_ss(tmem_base + TM_S, umma_desc_words(q_lo + kk * 2, hi), umma_desc_words(k_lo + kk * 2, hi), idesc, kk != 0);
          umma_commit(&b->s_full);
          umma_commit(&b->k_empty[ks]);
          if (last) umma_commit(&b->q_free[qs]);
        }
        ++gs;
        if (++js == n_kv) {
          js = 0;
          ++ts;
        }
      };
      issue_s();
      int jp = 0;
      for (uint32_t g = 0; g < G; ++g) {
        if (gs < G) issue_s();
        const int vs = g % V_STAGES;
        mbar_wait(&b->v_full[vs], (g / V_STAGES) & 1);
        mbar_wait(&b->p_full, g & 1);
        tc_fence_after();
        if (leader) {
          const uint32_t v_lo = v_lo0 + vs * (TILE_BYTES >> 4);
          if (jp == n_kv - 1 && tail64) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              umma_ts(tmem_base + TM_O, tmem_base + TM_P + ks * 8, umma_desc_words(v_lo + ks * (2048 >> 4), hi), idesc_pv,
                      (jp | ks) != 0);
          } else {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              umma_ts(tmem_base + TM_O, tmem_base + TM_P + ks * 8, umma_desc_words(v_lo + ks * (2048 >> 4), hi), idesc_pv,
                      (jp | ks) != 0);
          }
          umma_commit(&b->o_done);
          umma_commit(&b->v_empty[vs]);
        }
        if (++jp == n_kv) jp = 0;
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax warpgroup + tile epilogue
    setmaxnreg_inc<200>();
    const int q = warp & 3;       // TMEM lane quadrant
    const int r = q * 32 + lane;  // row within the tile
    const uint32_t lane_addr = opaque_u32(tmem_base + (uint32_t(q * 32) << 16));
    const uint32_t s_addr = lane_addr + TM_S;
    const uint32_t p_addr = lane_addr + TM_P;
    const uint32_t o_addr = lane_addr + TM_O;
    const SoftmaxBars sb{opaque_u32(smem_u32(&b->s_full)), opaque_u32(smem_u32(&b->s_free)), opaque_u32(smem_u32(&b->p_full)),
                         opaque_u32(smem_u32(&b->o_done))};
    const float sl2 = p.scale_log2;
    uint32_t g = 0;
    for (int t = 0; t < n_my; ++t) {
      const int w = blockIdx.x + t * gridDim.x;
      const int qblk = w % p.n_qblk, hs = w / p.n_qblk;
      const int head = hs % p.heads, seq = hs / p.heads;
      float m_ref = -INFINITY, l = 0.f;
      for (int j = 0; j < n_kv - 1; ++j, ++g) {
        if constexpr ((VAR & 128) != 0) {
          if (j > 0 && softmax_step_stream<VAR>(sb, s_addr, p_addr, g, sl2, m_ref, l, lane)) continue;
        }
        softmax_step<128, VAR>(b, s_addr, p_addr, o_addr, g, j, 128, sl2, m_ref, l, lane);
      }
      if (tail64) softmax_step<64, 0>(b, s_addr, p_addr, o_addr, g, n_kv - 1, kv_tail, sl2, m_ref, l, lane);
      else softmax_step<128, 0>(b, s_addr, p_addr, o_addr, g, n_kv - 1, kv_tail, sl2, m_ref, l, lane);
      // epilogue: O / l -> bf16 (PV_0 of the next tile cannot start before all four warps passed this point: it needs
      // their p_full arrivals)
      mbar_wait(&b->o_done, g & 1);
      ++g;
      tc_fence_after();
      const int row = qblk * 128 + r;
      const float inv_l = 1.0f / l;
      __nv_bfloat16* op = p.out + ((long long)seq * p.seq_len + row) * p.out_pitch + head * 64;
      uint32_t v[64];
      tmem_ld32_at(o_addr, v);
      tmem_ld32_at(o_addr + 32, v + 32);
      tmem_ld_wait();
      tc_fence_before();
      if (row < p.seq_len) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[8 * c + i]) * inv_l;
          *reinterpret_cast<uint4*>(op + c * 8) =
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

void set_attention_poly(int every) { g_attn_poly = every; }

cudaError_t launch_attention(const AttnParams& p_in, cudaStream_t stream) {
  static int n_sm = 0;
  if (!n_sm) {
    cudaError_t e = cudaSuccess;
    if (getenv("UFM_ATTN_POLY")) g_attn_poly = atoi(getenv("UFM_ATTN_POLY"));  // tuning
#define UFM_ATTN_VARIANTS(X) X(0) X(4) X(128) X(132) X(131) X(130) X(133) X(134) X(136)
#define X(v) if ((e = cudaFuncSetAttribute(attn_kernel<v>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES)) != cudaSuccess) return e;
    UFM_ATTN_VARIANTS(X)
#undef X
    int dev = 0, v = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    n_sm = v;
  }
  AttnParams p = p_in;
  p.n_qblk = (p.seq_len + 127) / 128;
  p.n_items = p.n_seq * p.heads * p.n_qblk;
  if (p.n_items <= 0) return cudaSuccess;
  const int grid = p.n_items < 2 * n_sm ? p.n_items : 2 * n_sm;
  bool launched = false;
#define X(v) if (!launched && g_attn_poly == v) { launch_k(attn_kernel<v>, dim3(grid), dim3(kThreads), SMEM_BYTES, stream, p); launched = true; }
  UFM_ATTN_VARIANTS(X)
#undef X
  if (!launched) return cudaErrorInvalidValue;
  return cudaGetLastError();
}

UFM_DEFINE_PDL_EARLY_SETTER(attention)

}  // namespace ufm
