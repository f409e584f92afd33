// Persistent warp-specialised tcgen05 GEMM / implicit-GEMM 3x3 convolution for sm_100a.
//
// Replaces every nn.Linear of the reference's transformers (UC/models/utils/transformer_blocks.py:75-78,137-145;
// hub DINOv2 qkv/proj/fc1/fc2), the patch-embed conv (as im2col GEMM), the DPT 1x1 / 3x3 convs
// (UC/models/prediction_heads/dpt.py:107-172,271-277; UC/models/libs/croco/dpt_block.py:34-69,128-146,210-218)
// and the classification-head MLP (UC/models/prediction_heads/mlp_feature.py:55-70).
//
// Roles (384 threads, 1 CTA / SM, grid = #SMs, static round-robin tile schedule, n fastest so the CTAs
// running concurrently share A tiles through L2 and the whole weight panel stays L2 resident):
//   warp 0     TMA producer: A box (64ch x tw x th x 1) and B box (64 x BN), 128B-swizzled, mbarrier ring
//   warp 1     MMA issuer: one thread, tcgen05.mma cta_group::1 kind::f16, M=128 N=BN K=16, fp32 accum in TMEM
//   warp 2     TMEM allocator (2 accumulator stages x BN columns)
//   warps 4-11 epilogue: tcgen05.ld 32x32b.x32 -> registers -> fused bias/GELU/LayerScale/residual/ReLU -> global
// Pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue), so the epilogue of tile i
// overlaps the mainloop of tile i+1.
#include <stdio.h>
#include <stdlib.h>

#include "kernels.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace ufm {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 128 + kEpiWarps * 32;

// CTAS = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) computes a 256 x BN tile; each CTA stages its own 128 rows
// of A and only HALF of the B tile, which cuts the L2 -> SM operand traffic per FLOP by a third (the single-CTA
// 128 x 256 tile needs 96 B/clk/SM at the MMA floor; the chip-wide L2 cap is about 43 B/clk/SM).
//
// HALO (3x3 convs): the A operand of all 9 taps comes from ONE resident (8+2) x (16+2)-pixel x 64-channel halo tile per
// channel chunk instead of 9 shifted 8 x 16 boxes: the tile's 128 rows are 16 image rows of 8 pixels, one 128B-swizzle
// atom (8 rows x 128 B) per image row, so the operand of tap (dy, dx) is the same smem with the descriptor start moved
// by ((dy+1)*10 + dx+1) * 128 bytes and an atom stride (SBO) of one halo row, 10 * 128 bytes.  The swizzle XOR is a
// function of the shared-memory address bits, which TMA (writing) and the MMA unit (reading) share, so a start that is
// only 128-byte aligned reads back consistently.  L2 -> SM operand traffic per (tile, chunk): 23 KB + 9 B tiles instead
// of 9 x (16 KB + B tile); the A ring and the B ring are then independent pipelines.
constexpr int HALO_TW = 8, HALO_TH = 16;
constexpr int HALO_ROW_BYTES = (HALO_TW + 2) * 128;
constexpr int HALO_TX_BYTES = (HALO_TH + 2) * HALO_ROW_BYTES;  // 23040

// R2 (CTAS = 2, BN = 256, plain GEMM): every CTA owns TWO 128-row blocks, i.e. a pair computes a 512 x 256 tile with ONE
// accumulator set (row block r -> TMEM columns r * 256).  One B half-tile (128 x 64) per stage feeds both row blocks, so a
// pair moves 96 KB of operands per 512 x 256 x 64 MACs instead of 2 x 64 KB: -25 % L2 -> SM bytes per FLOP, which is what
// bounds the 256 x 256 pair tile (it is also the tile cuBLAS picks for these shapes: nvjet 256x256 / 256x224
// per CTA, 2-CTA, profiles/r02_gemm_probe_ncu.csv).  The price is that the epilogue of an item no longer overlaps the
// next item's main loop through a second accumulator stage; what is recovered (rows2 bit 1): the next item's MMAs of row
// block 0 start as soon as accumulator 0 is drained and run ahead by the depth of the smem ring while accumulator 1 is
// still being drained.
template <int BN, int CTAS, bool HALO, bool R2 = false>
struct Cfg {
  static constexpr int A_BYTES = HALO ? 24576 : (R2 ? 2 : 1) * BM * BK * 2;    // halo tile padded to the 1024-byte swizzle period
  static constexpr int A_TX = HALO ? HALO_TX_BYTES : (R2 ? 2 : 1) * BM * BK * 2;
  static constexpr int B_ROWS = BN / CTAS;
  static constexpr int B_BYTES = B_ROWS * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = HALO ? (B_BYTES == 32768 ? 2 : B_BYTES == 8192 ? 9 : 5) : R2 ? 4 : (STAGE_BYTES == 49152) ? 3 : 5;  // B (and A) ring
  static constexpr int A_STAGES = HALO ? 3 : STAGES;  // a halo lasts 36 MMAs: three in flight cover the TMA latency
  static constexpr int BAR_BYTES = 512;
  // epilogue slabs: 8 warps x 2 x (32 rows x 128 B) (R2: one slab per warp, the smem goes to the operand ring); TAIL mode
  // reuses them as [acc][half][row][kTailMax] floats
  static constexpr int SLABS = R2 ? 1 : 2;
  static constexpr int SCRATCH_BYTES = kEpiWarps * SLABS * 4096;
  // R2: [bias | scale] of the item's BN columns, staged once per item by the epilogue warps (the per-column epilogue
  // constants are then LDS broadcasts instead of L2 round trips: with ~226 KB of smem carved out there is no L1 left)
  static constexpr int COLC_BYTES = R2 ? 2 * BN * 4 : 0;
  static constexpr int ALIGN_SLACK = R2 ? 0 : 1024;  // R2 relies on the 1024-byte alignment of the dynamic smem window (checked)
  static constexpr int SMEM_BYTES = ALIGN_SLACK + A_STAGES * A_BYTES + STAGES * B_BYTES + BAR_BYTES + SCRATCH_BYTES + COLC_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
};

// GELU(erf) = x * Phi(x) with Phi from Abramowitz-Stegun 26.2.17 (|error| < 7.5e-8, far below the bf16 rounding of the
// stored result): gelu(x) = relu(x) - |x| * phi(x) * poly(t), t = 1 / (1 + 0.2316419 |x|).  13 instructions, 2 MUFU
// (the libdevice erff path costs about twice that and made the fc1 epilogue pace the tensor pipe).
__device__ __forceinline__ float gelu_erf(float x) {
  const float ax = fabsf(x);
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.2316419f, ax, 1.0f)));
  float pl = fmaf(t, 1.330274429f, -1.821255978f);
  pl = fmaf(pl, t, 1.781477937f);
  pl = fmaf(pl, t, -0.356563782f);
  pl = fmaf(pl, t, 0.319381530f);
  pl *= t;
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * -0.72134752044448170f));  // exp(-x^2/2)
  return fmaf(-ax * 0.3989422804014327f, pl * e, fmaxf(x, 0.f));
}

// The same function for two values with packed f32x2 arithmetic (fma.rn.f32x2 / mul.rn.f32x2: one issue slot for two lanes;
// every operation is the IEEE round-to-nearest counterpart of the scalar one in the same order, so the bits are the same):
// 19 instead of 26 issue slots per pair in the fc1 epilogue, which paces that GEMM (GEMM_GELU_PACKED).
__device__ __forceinline__ uint64_t f2pack(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void f2unpack(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t f2fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t f2mul(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ void gelu_erf2(float& x0, float& x1) {
  const float ax0 = fabsf(x0), ax1 = fabsf(x1);
  const uint64_t ax = f2pack(ax0, ax1), x = f2pack(x0, x1);
  float d0, d1, t0, t1;
  f2unpack(f2fma(f2pack(0.2316419f, 0.2316419f), ax, f2pack(1.0f, 1.0f)), d0, d1);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(d0));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(d1));
  const uint64_t t = f2pack(t0, t1);
  uint64_t pl = f2fma(t, f2pack(1.330274429f, 1.330274429f), f2pack(-1.821255978f, -1.821255978f));
  pl = f2fma(pl, t, f2pack(1.781477937f, 1.781477937f));
  pl = f2fma(pl, t, f2pack(-0.356563782f, -0.356563782f));
  pl = f2fma(pl, t, f2pack(0.319381530f, 0.319381530f));
  pl = f2mul(pl, t);
  float a0, a1, e0, e1;
  f2unpack(f2mul(f2mul(x, x), f2pack(-0.72134752044448170f, -0.72134752044448170f)), a0, a1);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(a0));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(a1));
  const uint64_t r = f2fma(f2mul(f2pack(-ax0, -ax1), f2pack(0.3989422804014327f, 0.3989422804014327f)), f2mul(pl, f2pack(e0, e1)),
                           f2pack(fmaxf(x0, 0.f), fmaxf(x1, 0.f)));
  f2unpack(r, x0, x1);
}

// Work item -> (row-tile index, first column, width).  Items below split_from are whole BN-wide tiles; the tiles of the
// last, partly filled wave (item >= split_from) are cut into 2^split_log2 column slices so that they spread over the idle
// scheduling units: with 604 tiles on 74 CTA pairs the ninth wave holds 12 tiles, as 48 quarter tiles it costs about a
// third of a wave instead of a whole one (host side: engine.cu build_gemm).
struct TileRef {
  int mt, n0, ncols;
};
__device__ __forceinline__ TileRef tile_ref(const GemmParams& p, int item, int split_from, int BN) {
  int tile = item, sub = 0, ncols = BN;
  if (item >= split_from) {
    const int r = item - split_from;
    tile = split_from + (r >> p.split_log2);
    sub = r & ((1 << p.split_log2) - 1);
    ncols = BN >> p.split_log2;
  }
  const int tn = tile % p.tiles_n;
  return TileRef{tile / p.tiles_n, tn * BN + sub * ncols, ncols};
}

// R2 launches: items >= r2_items are HALF items -- ordinary 256 x BN pair tiles (row block 0 / accumulator 0 only) over the
// rows behind the last whole wave of 512-row items, each cut into 2^split_log2 column slices like the last wave of the
// 256-row kernel, so that the remainder rides in the same launch instead of a second, serialised one.
struct HalfRef {
  int mt, n0, ncols;  // pair-tile (256-row) index, first column, width
};
__device__ __forceinline__ HalfRef half_ref(const GemmParams& p, int item, int BN) {
  const int h = item - p.r2_items;
  const int tile = h >> p.split_log2, sub = h & ((1 << p.split_log2) - 1);
  const int ncols = BN >> p.split_log2;
  return HalfRef{p.r2_half_mt0 + tile / p.tiles_n, (tile % p.tiles_n) * BN + sub * ncols, ncols};
}

template <int BN, int CTAS, bool HALO, bool R2 = false>
__global__ void __launch_bounds__(kThreads, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  static_assert(!R2 || (CTAS == 2 && BN == 256 && !HALO), "R2 tiles are CTA-pair, BN = 256, plain GEMM");
  using C = Cfg<BN, CTAS, HALO, R2>;
  constexpr int RB = R2 ? 2 : 1;  // 128-row blocks per CTA and work item
  const uint32_t cta_rank = (CTAS == 2) ? cluster_ctarank() : 0u;
  const int unit = blockIdx.x / CTAS;          // scheduling unit: a CTA (CTAS 1) or a CTA pair (CTAS 2)
  const int num_units = gridDim.x / CTAS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + C::A_STAGES * C::A_BYTES;
  uint8_t* epi_smem = sB + C::STAGES * C::B_BYTES;  // 8 warps x 2 x 4 KB TMA slabs (1024-byte aligned)
  uint64_t* full = reinterpret_cast<uint64_t*>(epi_smem + C::SCRATCH_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* resbar = tempty + 2;  // [8 warps][2 slabs] residual-slab "full" barriers
  uint64_t* afull = resbar + 2 * kEpiWarps;  // HALO: the A ring's own barriers (full/empty then belong to the B ring)
  uint64_t* aempty = afull + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aempty + 3);
  float* colc = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(full) + C::BAR_BYTES);  // R2: [bias BN | scale BN]
  float* scratch = reinterpret_cast<float*>(epi_smem);  // TAIL mode: [acc][half][row][4] floats

  // logical warp (role) index: physical warp + warp_shift mod 12; a shift of 4 keeps the TMEM lane quadrant (warp & 3)
  int warp = (threadIdx.x >> 5) + p.warp_shift;
  warp = warp >= kThreads / 32 ? warp - kThreads / 32 : warp;
  const int lane = threadIdx.x & 31;

  if (R2 && threadIdx.x == 0 && (smem_u32(smem_raw) & 1023u)) {
    printf("ufm: gemm shared memory window is not 1024-byte aligned\n");
    __trap();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    if (p.epi_mode != EPI_DIRECT && !(p.epi_mode == EPI_TMA_MIX && !p.out_f32)) tma_prefetch_desc(&p.tmOut);
    if ((p.epi_mode == EPI_TMA_F32 || p.epi_mode == EPI_TMA_MIX) && p.res1 && !(p.flags & GEMM_RED_INPLACE))
      tma_prefetch_desc(&p.tmRes);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], kEpiWarps * CTAS);  // pair: the leader's barrier also collects the peer's epilogue warps
    }
    for (int i = 0; i < 2 * kEpiWarps; ++i) mbar_init(&resbar[i], 1);
    for (int i = 0; i < 3; ++i) {
      mbar_init(&afull[i], 1);
      mbar_init(&aempty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (CTAS == 2) tmem_alloc2<C::TMEM_COLS>(tmem_slot);
    else tmem_alloc<C::TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  if constexpr (CTAS == 2) cluster_sync_all();  // peer barriers initialised before any remote arrive / multicast commit
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped the tail of the previous kernel
  pdl_wait();

  // a scheduling unit owns CTAS consecutive 128-row tiles (one per CTA) x one BN-wide column tile; a CTA whose row tile
  // lies past the end loads zero-filled boxes (image index out of range) and stores nothing
  const int tiles_m1 = p.Nimg * p.tiles_y * p.tiles_x;
  const int tiles_m = (tiles_m1 + CTAS * RB - 1) / (CTAS * RB);  // R2: an item = two consecutive pair tiles of the same columns
  const int total_tiles = tiles_m * p.tiles_n;
  const int split_from = (p.split_log2 && !R2) ? p.split_from : total_tiles;
  const int total_items = R2 ? p.r2_items + (p.r2_half_tiles << p.split_log2) : split_from + ((total_tiles - split_from) << p.split_log2);
  const int k_iters = p.taps * p.cchunks;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0, astage = 0;
      uint32_t phase = 0, aphase = 0;
      for (int item = unit; item < total_items; item += num_units) {
        const TileRef tr = tile_ref(p, item, split_from, BN);
        const int tm = tr.mt * CTAS + (int)cta_rank;
        const int tx = tm % p.tiles_x;
        const int ty = (tm / p.tiles_x) % p.tiles_y;
        const int img = tm / (p.tiles_x * p.tiles_y);
        const int x0 = tx * p.tw, y0 = ty * p.th, n0 = tr.n0 + (int)cta_rank * (tr.ncols / CTAS);
        // a column slice loads only its own rows of B (tmBs: the same weight matrix with a (BN / CTAS) >> split box)
        const bool sliced = tr.ncols != BN;
        const CUtensorMap* tmB = sliced ? &p.tmBs : &p.tmB;
        const uint32_t b_tx = sliced ? (uint32_t)(C::B_BYTES >> p.split_log2) : (uint32_t)C::B_BYTES;
        if constexpr (HALO) {
          for (int cc = 0; cc < p.cchunks; ++cc) {
            mbar_wait(&aempty[astage], aphase ^ 1);
            if constexpr (CTAS == 2) {
              if (cta_rank == 0) mbar_expect_tx(&afull[astage], 2 * C::A_TX);
              tma2_load_4d(sA + astage * C::A_BYTES, &p.tmA, smem_u32(&afull[astage]) & kPeerBitMask, cc * BK, x0 - 1,
                           y0 - 1, img);
            } else {
              mbar_expect_tx(&afull[astage], C::A_TX);
              tma_load_4d(sA + astage * C::A_BYTES, &p.tmA, &afull[astage], cc * BK, x0 - 1, y0 - 1, img);
            }
            if (++astage == C::A_STAGES) {
              astage = 0;
              aphase ^= 1;
            }
            for (int tap = 0; tap < 9; ++tap) {
              mbar_wait(&empty[stage], phase ^ 1);
              if constexpr (CTAS == 2) {
                if (cta_rank == 0) mbar_expect_tx(&full[stage], 2 * C::B_BYTES);
                tma2_load_2d(sB + stage * C::B_BYTES, &p.tmB, smem_u32(&full[stage]) & kPeerBitMask,
                             tap * p.cin_pad + cc * BK, n0);
              } else {
                mbar_expect_tx(&full[stage], C::B_BYTES);
                tma_load_2d(sB + stage * C::B_BYTES, &p.tmB, &full[stage], tap * p.cin_pad + cc * BK, n0);
              }
              if (++stage == C::STAGES) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
          continue;
        }
        if constexpr (R2) {
          if (item >= p.r2_items) {
            // half item: one 128-row block per CTA, the B half-tile (or its column slice) of the pair
            const HalfRef hr = half_ref(p, item, BN);
            const int hx = (hr.mt * CTAS + (int)cta_rank) * BM;
            const int hn0 = hr.n0 + (int)cta_rank * (hr.ncols / CTAS);
            const CUtensorMap* tmBh = p.split_log2 ? &p.tmBs : &p.tmB;
            const uint32_t hb_tx = (uint32_t)(C::B_BYTES >> p.split_log2);
            for (int k = 0; k < k_iters; ++k) {
              mbar_wait(&empty[stage], phase ^ 1);
              if (cta_rank == 0) mbar_expect_tx(&full[stage], 2 * (BM * BK * 2 + hb_tx));
              const uint32_t lbar = smem_u32(&full[stage]) & kPeerBitMask;
              tma2_load_4d(sA + stage * C::A_BYTES, &p.tmA, lbar, k * BK, hx, 0, 0);
              tma2_load_2d(sB + stage * C::B_BYTES, tmBh, lbar, k * BK, hn0);
              if (++stage == C::STAGES) {
                stage = 0;
                phase ^= 1;
              }
            }
            continue;
          }
          // row block r of this CTA = 128-row tile (2 * mt + r) * 2 + rank; both blocks and the B half-tile of the pair land
          // on the leader's barrier
          // (plain row matrix: X = rows, Y = Nimg = 1, tw = 128; boxes past the last row are zero-filled by TMA)
          const int bx0 = ((tr.mt * 2) * CTAS + (int)cta_rank) * BM, bx1 = bx0 + CTAS * BM;
          for (int k = 0; k < k_iters; ++k) {
            mbar_wait(&empty[stage], phase ^ 1);
            if (cta_rank == 0) mbar_expect_tx(&full[stage], 2 * (C::A_TX + C::B_BYTES));
            const uint32_t lbar = smem_u32(&full[stage]) & kPeerBitMask;
            tma2_load_4d(sA + stage * C::A_BYTES, &p.tmA, lbar, k * BK, bx0, 0, 0);
            tma2_load_4d(sA + stage * C::A_BYTES + BM * BK * 2, &p.tmA, lbar, k * BK, bx1, 0, 0);
            tma2_load_2d(sB + stage * C::B_BYTES, &p.tmB, lbar, k * BK, n0);
            if (++stage == C::STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
          continue;
        }
        for (int k = 0; k < k_iters; ++k) {
          const int tap = k / p.cchunks;
          const int cc = k - tap * p.cchunks;
          int dy = 0, dx = 0;
          if (p.taps == 9) {
            dy = tap / 3 - 1;
            dx = tap % 3 - 1;
          }
          mbar_wait(&empty[stage], phase ^ 1);
          if constexpr (CTAS == 2) {
            // both CTAs' boxes complete on the LEADER's full barrier, which expects the bytes of the whole pair
            if (cta_rank == 0) mbar_expect_tx(&full[stage], 2 * (C::A_TX + b_tx));
            const uint32_t lbar = smem_u32(&full[stage]) & kPeerBitMask;
            tma2_load_4d(sA + stage * C::A_BYTES, &p.tmA, lbar, cc * BK, x0 + dx, y0 + dy, img);
            tma2_load_2d(sB + stage * C::B_BYTES, tmB, lbar, tap * p.cin_pad + cc * BK, n0);
          } else {
            mbar_expect_tx(&full[stage], C::A_TX + b_tx);
            tma_load_4d(sA + stage * C::A_BYTES, &p.tmA, &full[stage], cc * BK, x0 + dx, y0 + dy, img);
            tma_load_2d(sB + stage * C::B_BYTES, tmB, &full[stage], tap * p.cin_pad + cc * BK, n0);
          }
          if (++stage == C::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // The whole warp walks the loops (uniform control flow keeps the operand descriptors in uniform registers); one
    // elected lane issues the MMAs and commits.  With `if (lane == 0)` around the loop instead, every tcgen05.mma cost
    // an ELECT / R2UR.BROADCAST / branch sequence and ~25 dependent uniform-datapath instructions: ~600 clk per group of
    // four MMAs, more than the 256-512 clk the tensor pipe needs for them.
    if (cta_rank == 0) {
      constexpr uint32_t idesc_full = umma_idesc_bf16(BM * CTAS, BN, 0, 0);
      const uint32_t idesc_slice = umma_idesc_bf16(BM * CTAS, BN >> p.split_log2, 0, 0);
      constexpr uint32_t hi_b = umma_desc_hi_sw128(1024);
      constexpr uint32_t hi_a = umma_desc_hi_sw128(HALO ? HALO_ROW_BYTES : 1024);
      const uint32_t a_lo0 = umma_desc_lo(smem_u32(sA));
      const uint32_t b_lo0 = umma_desc_lo(smem_u32(sB));
      const bool leader = elect_one();
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0, astage = 0;
      uint32_t accphase = 0, aphase = 0;
      for (int item = unit; item < total_items; item += num_units) {
        if constexpr (R2) {
          // both accumulators belong to this item: row block r -> TMEM columns [r * BN, (r + 1) * BN).  With run-ahead the
          // first `ahead` k-steps of row block 0 are issued as soon as accumulator 0 is drained, without releasing their
          // smem stages; row block 1 catches up (and releases them) once accumulator 1 is drained as well.
          auto mma_block = [&](int st_, int r, uint32_t accum0, uint32_t idesc_) {
            const uint32_t a_lo = a_lo0 + st_ * (C::A_BYTES >> 4) + r * ((BM * BK * 2) >> 4);
            const uint32_t b_lo = b_lo0 + st_ * (C::B_BYTES >> 4);
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk)
              umma2_ss(tmem_base + r * BN, umma_desc_words(a_lo + kk * 2, hi_a), umma_desc_words(b_lo + kk * 2, hi_b), idesc_,
                       (accum0 | (uint32_t)kk) != 0 ? 1u : 0u);
          };
          if (item >= p.r2_items) {
            // half item: row block 0 into accumulator 0 (a column slice fills its first columns); accumulator 1 is not
            // touched any more -- half items come after every 512-row item of this unit
            mbar_wait(&tempty[0], accphase ^ 1);
            tc_fence_after();
            for (int k = 0; k < k_iters; ++k) {
              mbar_wait(&full[stage], phase);
              tc_fence_after();
              if (leader) {
                mma_block(stage, 0, (uint32_t)k, idesc_slice);
                umma2_commit_mc(&empty[stage]);
                if (k == k_iters - 1) umma2_commit_mc(&tfull[0]);
              }
              if (++stage == C::STAGES) {
                stage = 0;
                phase ^= 1;
              }
            }
            accphase ^= 1;
            continue;
          }
          const int ahead = (p.rows2 & 2) ? (k_iters < C::STAGES ? k_iters : C::STAGES) : 0;
          mbar_wait(&tempty[0], accphase ^ 1);
          tc_fence_after();
          {
            int s2 = stage;
            uint32_t ph2 = phase;
            for (int k = 0; k < ahead; ++k) {
              mbar_wait(&full[s2], ph2);
              tc_fence_after();
              if (leader) mma_block(s2, 0, (uint32_t)k, idesc_full);
              if (++s2 == C::STAGES) {
                s2 = 0;
                ph2 ^= 1;
              }
            }
          }
          mbar_wait(&tempty[1], accphase ^ 1);
          tc_fence_after();
          for (int k = 0; k < k_iters; ++k) {
            if (k >= ahead) {
              mbar_wait(&full[stage], phase);
              tc_fence_after();
            }
            if (leader) {
              if (k >= ahead) mma_block(stage, 0, (uint32_t)k, idesc_full);
              mma_block(stage, 1, (uint32_t)k, idesc_full);
              umma2_commit_mc(&empty[stage]);
              if (k == k_iters - 1) {
                umma2_commit_mc(&tfull[0]);
                umma2_commit_mc(&tfull[1]);
              }
            }
            if (++stage == C::STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
          accphase ^= 1;
          continue;
        }
        const uint32_t idesc = (HALO || item < split_from) ? idesc_full : idesc_slice;  // a slice fills the first columns
        mbar_wait(&tempty[acc], accphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        if constexpr (HALO) {
          for (int cc = 0; cc < p.cchunks; ++cc) {
            mbar_wait(&afull[astage], aphase);
            const uint32_t halo_lo = a_lo0 + astage * (C::A_BYTES >> 4);
#pragma unroll 1
            for (int ty3 = 0; ty3 < 3; ++ty3) {
#pragma unroll
              for (int tx3 = 0; tx3 < 3; ++tx3) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                if (leader) {
                  const uint32_t a_lo = halo_lo + (ty3 * (HALO_TW + 2) + tx3) * (128 >> 4);
                  const uint32_t b_lo = b_lo0 + stage * (C::B_BYTES >> 4);
#pragma unroll
                  for (int kk = 0; kk < BK / 16; ++kk) {
                    const uint32_t accum = (cc | ty3 | tx3 | kk) != 0 ? 1u : 0u;
                    if constexpr (CTAS == 2)
                      umma2_ss(d_tmem, umma_desc_words(a_lo + kk * 2, hi_a), umma_desc_words(b_lo + kk * 2, hi_b), idesc, accum);
                    else
                      umma_ss(d_tmem, umma_desc_words(a_lo + kk * 2, hi_a), umma_desc_words(b_lo + kk * 2, hi_b), idesc, accum);
                  }
                  if constexpr (CTAS == 2) umma2_commit_mc(&empty[stage]);
                  else umma_commit(&empty[stage]);
                }
                if (++stage == C::STAGES) {
                  stage = 0;
                  phase ^= 1;
                }
              }
            }
            if (leader) {
              if constexpr (CTAS == 2) {
                umma2_commit_mc(&aempty[astage]);
                if (cc == p.cchunks - 1) umma2_commit_mc(&tfull[acc]);
              } else {
                umma_commit(&aempty[astage]);
                if (cc == p.cchunks - 1) umma_commit(&tfull[acc]);
              }
            }
            if (++astage == C::A_STAGES) {
              astage = 0;
              aphase ^= 1;
            }
          }
        } else {
          for (int k = 0; k < k_iters; ++k) {
            mbar_wait(&full[stage], phase);
            tc_fence_after();
            if (leader) {
              const uint32_t a_lo = a_lo0 + stage * (C::A_BYTES >> 4);
              const uint32_t b_lo = b_lo0 + stage * (C::B_BYTES >> 4);
#pragma unroll
              for (int kk = 0; kk < BK / 16; ++kk) {
                const uint32_t accum = (k | kk) != 0 ? 1u : 0u;
                if constexpr (CTAS == 2)
                  umma2_ss(d_tmem, umma_desc_words(a_lo + kk * 2, hi_a), umma_desc_words(b_lo + kk * 2, hi_b), idesc, accum);
                else
                  umma_ss(d_tmem, umma_desc_words(a_lo + kk * 2, hi_a), umma_desc_words(b_lo + kk * 2, hi_b), idesc, accum);
              }
              if constexpr (CTAS == 2) {
                umma2_commit_mc(&empty[stage]);
                if (k == k_iters - 1) umma2_commit_mc(&tfull[acc]);
              } else {
                umma_commit(&empty[stage]);
                if (k == k_iters - 1) umma_commit(&tfull[acc]);
              }
            }
            if (++stage == C::STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        acc ^= 1;
        if (acc == 0) accphase ^= 1;
      }
    }
  } else if (warp >= 4) {
    const int e = warp - 4;
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    const int half = e >> 2;
    // BN = 64 (thin convs): the four half-0 warps take all 64 columns, the half-1 warps only keep the barrier protocol
    constexpr int HALF_COLS = BN == 64 ? 64 : BN / 2;
    const int Nlim = (BN == 64 && half == 1) ? 0 : p.N;  // column bound of this warp
    const bool tail = (p.flags & GEMM_TAIL) != 0;
    const int emode = p.epi_mode;
    // TMA epilogues: two 4 KB slabs per warp = 32 rows x 128 bytes in the 128B-swizzle pattern (16-byte chunk j of row r
    // lives at chunk j ^ (r & 7)), so a thread touches only its own row, conflict free, and the slab is the smem image of
    // a (32 fp32 | 64 bf16) x 32-row TMA box.
    uint8_t* wbuf = epi_smem + e * (C::SLABS * 4096);
    uint64_t* rbar = resbar + e * 2;
    uint32_t rph0 = 0, rph1 = 0;
    const uint32_t sw = lane & 7;
    uint32_t tempty_remote[2] = {0u, 0u};
    if constexpr (CTAS == 2) {
      tempty_remote[0] = mapa_cluster(smem_u32(&tempty[0]), 0);
      tempty_remote[1] = mapa_cluster(smem_u32(&tempty[1]), 0);
    }
    if constexpr (R2) {
      // ---- 512 x 256 pair tiles: nothing overlaps the epilogue through a second accumulator set, so the accumulator is
      // handed back as early as possible: all 128 columns of this warp leave tensor memory in one batch of loads, tempty is
      // signalled, and only then the bias / GELU / LayerScale math, the slab and the TMA store (or reduce-add) run -- under
      // the next item's main loop.  Column constants come from smem (staged per item by the eight epilogue warps).
      const bool f32 = emode == EPI_TMA_F32;
      const bool red = (p.flags & GEMM_RED_INPLACE) != 0;
      const int tid8 = e * 32 + lane;  // 0..255 over the eight epilogue warps
      int it2 = 0;
      int item = unit;
      for (; item < p.r2_items; item += num_units, ++it2) {
        const TileRef tr = tile_ref(p, item, split_from, BN);
        const int n0 = tr.n0;
        // stage bias / scale of columns [n0, n0 + BN): every warp has finished reading the previous item's values
        asm volatile("bar.sync 2, 256;" ::: "memory");
        colc[tid8] = (p.bias && n0 + tid8 < p.N) ? __ldg(p.bias + n0 + tid8) : 0.f;
        colc[BN + tid8] = (p.scale && n0 + tid8 < p.N) ? __ldg(p.scale + n0 + tid8) : 1.f;
        asm volatile("bar.sync 2, 256;" ::: "memory");
        const int col_off = half * (BN / 2);
        const int c0 = n0 + col_off;
        const float* bcol = colc + col_off;
        const float* scol = colc + BN + col_off;
        for (int rb = 0; rb < 2; ++rb) {
          // plain row matrix (X = rows, Y = Nimg = 1, tw = 128; checked by the host): rows past the end are clipped by TMA
          const int wx = ((tr.mt * 2 + rb) * CTAS + (int)cta_rank) * BM + 32 * q;
          constexpr int wy = 0, img = 0;
          const uint32_t acc_addr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(rb * BN + col_off);
          if (f32 && red && rb == 0 && lane == 0 && (p.flags & GEMM_RED_PREFETCH)) {
            // x slabs of both row blocks into L2 under the main loop (see the 256-row kernel)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              if (c0 + 32 * c >= p.N) break;
              tma_prefetch_4d(&p.tmOut, c0 + 32 * c, wx, 0, 0);
              tma_prefetch_4d(&p.tmOut, c0 + 32 * c, wx + CTAS * BM, 0, 0);
            }
          }
          mbar_wait(&tfull[rb], (uint32_t)(it2 & 1));
          tc_fence_after();
          uint32_t vr[128];
#pragma unroll
          for (int c = 0; c < 4; ++c) tmem_ld32(acc_addr + c * 32, *reinterpret_cast<uint32_t(*)[32]>(vr + c * 32));
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster_relaxed(tempty_remote[rb]);
          uint8_t* rowp = wbuf + lane * 128;
          if (f32) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              if (c0 + 32 * c >= p.N) break;
              if (lane == 0) bulk_wait_read<0>();  // single slab: the previous store has read it
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 b4 = *reinterpret_cast<const float4*>(bcol + 32 * c + 4 * j);
                const float4 s4 = *reinterpret_cast<const float4*>(scol + 32 * c + 4 * j);
                float4 v = make_float4(__uint_as_float(vr[32 * c + 4 * j]) + b4.x, __uint_as_float(vr[32 * c + 4 * j + 1]) + b4.y,
                                       __uint_as_float(vr[32 * c + 4 * j + 2]) + b4.z, __uint_as_float(vr[32 * c + 4 * j + 3]) + b4.w);
                if (p.flags & GEMM_GELU) {
                  v.x = gelu_erf(v.x); v.y = gelu_erf(v.y); v.z = gelu_erf(v.z); v.w = gelu_erf(v.w);
                }
                if (p.scale) {
                  v.x *= s4.x; v.y *= s4.y; v.z *= s4.z; v.w *= s4.w;
                }
                *reinterpret_cast<float4*>(rowp + ((uint32_t(j) ^ sw) << 4)) = v;
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                if (red) tma_reduce_add_4d(&p.tmOut, wbuf, c0 + 32 * c, wx, wy, img);
                else tma_store_4d(&p.tmOut, wbuf, c0 + 32 * c, wx, wy, img);
                bulk_commit_group();
              }
            }
          } else {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (c0 + 64 * c >= p.N) break;
              if (lane == 0) bulk_wait_read<0>();
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float v[8];
                const float4 b0 = *reinterpret_cast<const float4*>(bcol + 64 * c + 8 * j);
                const float4 b1 = *reinterpret_cast<const float4*>(bcol + 64 * c + 8 * j + 4);
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(vr[64 * c + 8 * j + i]) + bb[i];
                if (p.flags & GEMM_GELU_PACKED) {
#pragma unroll
                  for (int i = 0; i < 8; i += 2) gelu_erf2(v[i], v[i + 1]);
                } else if (p.flags & GEMM_GELU) {
#pragma unroll
                  for (int i = 0; i < 8; ++i) v[i] = gelu_erf(v[i]);
                }
                if (p.scale) {
                  const float4 s0 = *reinterpret_cast<const float4*>(scol + 64 * c + 8 * j);
                  const float4 s1 = *reinterpret_cast<const float4*>(scol + 64 * c + 8 * j + 4);
                  v[0] *= s0.x; v[1] *= s0.y; v[2] *= s0.z; v[3] *= s0.w;
                  v[4] *= s1.x; v[5] *= s1.y; v[6] *= s1.z; v[7] *= s1.w;
                }
                if (p.flags & GEMM_RELU_BF16) {
#pragma unroll
                  for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
                }
                *reinterpret_cast<uint4*>(rowp + ((uint32_t(j) ^ sw) << 4)) =
                    make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_4d(&p.tmOut, wbuf, c0 + 64 * c, wx, wy, img);
                bulk_commit_group();
              }
            }
          }
        }
      }
      // ---- half items (the rows behind the last whole wave of 512-row items): row block 0 / accumulator 0 of a 256-row pair
      // tile, possibly a column slice; chunk by chunk, the accumulator is released once all chunks are out ----
      for (; item < total_items; item += num_units, ++it2) {
        const HalfRef hr = half_ref(p, item, BN);
        const int n0 = hr.n0;
        asm volatile("bar.sync 2, 256;" ::: "memory");
        colc[tid8] = (p.bias && n0 + tid8 < p.N) ? __ldg(p.bias + n0 + tid8) : 0.f;
        colc[BN + tid8] = (p.scale && n0 + tid8 < p.N) ? __ldg(p.scale + n0 + tid8) : 1.f;
        asm volatile("bar.sync 2, 256;" ::: "memory");
        // columns of this warp: 128+ columns are shared by the two warp halves, a 64-column slice is taken by half 0 alone
        // (half 1 only keeps the barrier protocol)
        const int per_half = hr.ncols >= 128 ? (hr.ncols >> 1) : hr.ncols;
        const int col_off = hr.ncols >= 128 ? half * per_half : 0;
        const int c0 = n0 + col_off;
        const int cend = (hr.ncols < 128 && half == 1) ? c0 : (c0 + per_half < p.N ? c0 + per_half : p.N);
        const float* bcol = colc + col_off;
        const float* scol = colc + BN + col_off;
        const int wx = (hr.mt * CTAS + (int)cta_rank) * BM + 32 * q;
        const uint32_t acc_addr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(col_off);
        mbar_wait(&tfull[0], (uint32_t)(it2 & 1));
        tc_fence_after();
        uint8_t* rowp = wbuf + lane * 128;
#pragma unroll 1
        for (int cc = c0; cc < cend; cc += 32) {
          // 32 accumulator columns: one fp32 slab, or half of a 64-column bf16 slab
          const int o = cc - c0;
          uint32_t vr[32];
          tmem_ld32(acc_addr + o, vr);
          tmem_ld_wait();
          const bool first = f32 || (o & 32) == 0, last = f32 || (o & 32) != 0;
          if (first) {
            if (lane == 0) bulk_wait_read<0>();  // single slab: the previous store has read it
            __syncwarp();
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b4 = *reinterpret_cast<const float4*>(bcol + o + 4 * j);
            float4 v = make_float4(__uint_as_float(vr[4 * j]) + b4.x, __uint_as_float(vr[4 * j + 1]) + b4.y,
                                   __uint_as_float(vr[4 * j + 2]) + b4.z, __uint_as_float(vr[4 * j + 3]) + b4.w);
            if (p.flags & GEMM_GELU_PACKED) {
              gelu_erf2(v.x, v.y);
              gelu_erf2(v.z, v.w);
            } else if (p.flags & GEMM_GELU) {
              v.x = gelu_erf(v.x); v.y = gelu_erf(v.y); v.z = gelu_erf(v.z); v.w = gelu_erf(v.w);
            }
            if (p.scale) {
              const float4 s4 = *reinterpret_cast<const float4*>(scol + o + 4 * j);
              v.x *= s4.x; v.y *= s4.y; v.z *= s4.z; v.w *= s4.w;
            }
            if (f32) {
              *reinterpret_cast<float4*>(rowp + ((uint32_t(j) ^ sw) << 4)) = v;
            } else {
              if (p.flags & GEMM_RELU_BF16) {
                v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
              }
              // 8-byte half of 16-byte chunk (o & 32) / 8 + j / 2 of the 64-column bf16 slab row
              const uint32_t chunk = uint32_t(((o & 32) >> 3) + (j >> 1));
              *reinterpret_cast<uint2*>(rowp + ((chunk ^ sw) << 4) + ((j & 1) << 3)) =
                  make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
            }
          }
          if (last) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (f32 && red) tma_reduce_add_4d(&p.tmOut, wbuf, cc, wx, 0, 0);
              else tma_store_4d(&p.tmOut, wbuf, f32 ? cc : cc - 32, wx, 0, 0);
              bulk_commit_group();
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster_relaxed(tempty_remote[0]);
      }
      if (lane == 0) bulk_wait<0>();
    } else {
    int iter = 0;
    for (int item = unit; item < total_items; item += num_units, ++iter)
    for (int rb = 0; rb < RB; ++rb) {
      // accumulator slot and its barrier phase: alternating stages per item, or (R2) one stage per row block of the item
      const int acc = R2 ? rb : (iter & 1);
      const uint32_t accphase = R2 ? (uint32_t)(iter & 1) : (uint32_t)((iter >> 1) & 1);
      const TileRef tr = tile_ref(p, item, split_from, BN);
      const int tm = (tr.mt * RB + rb) * CTAS + (int)cta_rank;
      const int tx = tm % p.tiles_x;
      const int ty = (tm / p.tiles_x) % p.tiles_y;
      const int img = tm / (p.tiles_x * p.tiles_y);
      const int n0 = tr.n0;
      // columns of this warp: a tile (or slice) of 128+ columns is shared by the two warp halves, a 64-column one is
      // taken by half 0 alone (half 1 only keeps the barrier protocol)
      const int per_half = tr.ncols >= 128 ? (tr.ncols >> 1) : tr.ncols;
      const int col_off = tr.ncols >= 128 ? half * per_half : 0;
      const int c0 = n0 + col_off;  // first column of this warp
      const int Nl = (tr.ncols < 128 && half == 1) ? 0 : (Nlim < c0 + per_half ? Nlim : c0 + per_half);  // its column bound
      const uint32_t acc_addr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(acc * BN + col_off);

      if (emode == EPI_TMA_F32 || emode == EPI_TMA_MIX) {
        // ---- fp32 output (+ optional fp32 residual) through TMA: 32-column slabs -------------------------------------
        // EPI_TMA_MIX adds a second, coalesced pass over the finished slab (second residual with its own strides, bf16
        // copy of the result) so that layers with two residuals / two outputs leave the row-per-thread direct path too.
        const bool mix = emode == EPI_TMA_MIX;
        const int wx = tx * p.tw + ((32 * q) & (p.tw - 1));
        const int wy = ty * p.th + ((32 * q) >> p.tw_log2);
        int nch = (Nl - c0 + 31) >> 5;
        nch = nch < 0 ? 0 : (nch > HALF_COLS / 32 ? HALF_COLS / 32 : nch);
        const bool red = (p.flags & GEMM_RED_INPLACE) != 0;  // x += v by the L2 (TMA reduce-add): no residual slab
        const bool has_res = p.res1 != nullptr && !red;
        if (lane == 0) {
          bulk_wait_read<0>();  // the previous tile's stores have left both slabs
          if (red && (p.flags & GEMM_RED_PREFETCH)) {
            // in-place reduce-add: pull the x slabs of this tile into L2 while the MMAs run, so that the L2's read-modify-write
            // does not wait for DRAM when the stores arrive
            for (int c = 0; c < nch; ++c) tma_prefetch_4d(&p.tmOut, c0 + 32 * c, wx, wy, img);
          }
          if (has_res) {
            for (int c = 0; c < nch; ++c) {
              if (c < 2) {
                mbar_expect_tx(&rbar[c], 4096);
                tma_load_4d(wbuf + c * 4096, &p.tmRes, &rbar[c], c0 + 32 * c, wx, wy, img);
              } else {
                tma_prefetch_4d(&p.tmRes, c0 + 32 * c, wx, wy, img);  // pull the rest into L2 under the mainloop
              }
            }
          }
        }
        if (mix && p.res2) {
          // res2 rows of this tile do not depend on the accumulator: pull them into L2 while the MMAs still run
          const int li = 32 * q + lane;
          const int ly = li >> p.tw_log2, lx = li & (p.tw - 1);
          const int x = tx * p.tw + lx, y = ty * p.th + ly;
          if (x < p.X && y < p.Y && img < p.Nimg) {
            const float* rp2 = p.res2 + (long long)img * p.r2N + (long long)y * p.r2Y + (long long)x * p.r2X + c0;
            for (int c = 0; c < nch; ++c) prefetch_l2(rp2 + 32 * c);
          }
        }
        __syncwarp();
        mbar_wait(&tfull[acc], accphase);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < nch; ++c) {
          uint8_t* buf = wbuf + (c & 1) * 4096;
          uint8_t* rowp = buf + lane * 128;
          // EPI_TMA_MIX: second-pass addressing (8 lanes per slab row, 4 rows per warp instruction); the res2 loads are
          // issued here so that their latency runs under the residual wait, the TMEM read and the first pass
          float4 r2v[8];
          const int colg = c0 + 32 * c + 4 * (lane & 7);
          if (mix && p.res2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int li = 32 * q + 4 * i + (lane >> 3);
              const int x = tx * p.tw + (li & (p.tw - 1)), y = ty * p.th + (li >> p.tw_log2);
              r2v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (colg < Nl && x < p.X && y < p.Y && img < p.Nimg)
                r2v[i] = *reinterpret_cast<const float4*>(p.res2 + (long long)img * p.r2N + (long long)y * p.r2Y +
                                                          (long long)x * p.r2X + colg);
            }
          }
          if (has_res) {
            if (c >= 1 && c + 1 < nch && lane == 0) {
              bulk_wait_read<0>();  // slab (c+1)&1 was stored by chunk c-1
              mbar_expect_tx(&rbar[(c + 1) & 1], 4096);
              tma_load_4d(wbuf + ((c + 1) & 1) * 4096, &p.tmRes, &rbar[(c + 1) & 1], c0 + 32 * (c + 1), wx, wy, img);
            }
            if (c & 1) {
              mbar_wait(&rbar[1], rph1);
              rph1 ^= 1;
            } else {
              mbar_wait(&rbar[0], rph0);
              rph0 ^= 1;
            }
          } else if (c >= 2) {
            if (lane == 0) bulk_wait_read<1>();
            __syncwarp();
          }
          const int gcol = c0 + c * 32;
          uint32_t vr[32];
          tmem_ld32(acc_addr + c * 32, vr);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 v = make_float4(__uint_as_float(vr[4 * j]), __uint_as_float(vr[4 * j + 1]),
                                   __uint_as_float(vr[4 * j + 2]), __uint_as_float(vr[4 * j + 3]));
            if (p.bias) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + gcol + 4 * j));
              v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
            }
            if (p.flags & GEMM_GELU) {
              v.x = gelu_erf(v.x); v.y = gelu_erf(v.y); v.z = gelu_erf(v.z); v.w = gelu_erf(v.w);
            }
            if (p.scale) {
              const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + gcol + 4 * j));
              v.x *= s4.x; v.y *= s4.y; v.z *= s4.z; v.w *= s4.w;
            }
            float4* sp = reinterpret_cast<float4*>(rowp + ((j ^ sw) << 4));
            if (has_res) {
              const float4 r4 = *sp;
              v.x += r4.x; v.y += r4.y; v.z += r4.z; v.w += r4.w;
            }
            *sp = v;
          }
          if (mix) {
            // transposed pass: 8 lanes cover the 128 bytes of one slab row, a warp instruction 4 rows -> full-line global
            // accesses for res2 (fp32, own strides) and the bf16 output; the fp32 sum goes back into the slab.  The adds
            // keep the direct epilogue's order ((acc + bias) * scale + res1) + res2, so both paths give the same bits.
            __syncwarp();
            const int ch = lane & 7;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int row = 4 * i + (lane >> 3);
              const int li = 32 * q + row;
              const int x = tx * p.tw + (li & (p.tw - 1)), y = ty * p.th + (li >> p.tw_log2);
              if (!(colg < Nl && x < p.X && y < p.Y && img < p.Nimg)) continue;
              float4* sp = reinterpret_cast<float4*>(buf + row * 128 + ((uint32_t(ch) ^ uint32_t(row & 7)) << 4));
              float4 v = *sp;
              if (p.res2) {
                v.x += r2v[i].x; v.y += r2v[i].y; v.z += r2v[i].z; v.w += r2v[i].w;
                if (p.out_f32) *sp = v;
              }
              if (p.out_bf16) {
                if (p.flags & GEMM_RELU_BF16) {
                  v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
                }
                *reinterpret_cast<uint2*>(p.out_bf16 + (long long)img * p.sN + (long long)y * p.sY + (long long)x * p.sX +
                                          p.out_off + colg) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
              }
            }
          }
          if (!mix || p.out_f32) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (red) tma_reduce_add_4d(&p.tmOut, buf, gcol, wx, wy, img);
              else tma_store_4d(&p.tmOut, buf, gcol, wx, wy, img);
              bulk_commit_group();
            }
          } else {
            __syncwarp();  // the slab may be refilled (TMA residual load) once every lane has read it
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CTAS == 2) mbar_arrive_cluster_relaxed(tempty_remote[acc]);
          else mbar_arrive(&tempty[acc]);
        }
      } else if (emode == EPI_TMA_BF16) {
        // ---- bf16 output through TMA: 64-column slabs ----------------------------------------------------------------
        const int wx = tx * p.tw + ((32 * q) & (p.tw - 1));
        const int wy = ty * p.th + ((32 * q) >> p.tw_log2);
        int nch = (Nl - c0 + 63) >> 6;
        nch = nch < 0 ? 0 : (nch > HALF_COLS / 64 ? HALF_COLS / 64 : nch);
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
        mbar_wait(&tfull[acc], accphase);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < nch; ++c) {
          uint8_t* buf = wbuf + (c & 1) * 4096;
          uint8_t* rowp = buf + lane * 128;
          const int gcol = c0 + c * 64;
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            // bias of these 32 columns before the tensor-memory load (see the fp32 path)
            float4 bq[8];
            const bool colok = gcol + h2 * 32 < Nl;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              bq[j] = (p.bias && colok) ? __ldg(reinterpret_cast<const float4*>(p.bias + gcol + h2 * 32 + 4 * j))
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
            uint32_t vr[32];
            tmem_ld32(acc_addr + c * 64 + h2 * 32, vr);
            tmem_ld_wait();
            if (colok) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(vr[8 * j + i]);
                const int col = gcol + h2 * 32 + 8 * j;
                if (p.bias) {
                  const float4 b0 = bq[2 * j], b1 = bq[2 * j + 1];
                  v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
                  v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
                }
                if (p.flags & GEMM_GELU_PACKED) {
#pragma unroll
                  for (int i = 0; i < 8; i += 2) gelu_erf2(v[i], v[i + 1]);
                } else if (p.flags & GEMM_GELU) {
#pragma unroll
                  for (int i = 0; i < 8; ++i) v[i] = gelu_erf(v[i]);
                }
                if (p.scale) {
                  const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.scale + col));
                  const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.scale + col + 4));
                  v[0] *= s0.x; v[1] *= s0.y; v[2] *= s0.z; v[3] *= s0.w;
                  v[4] *= s1.x; v[5] *= s1.y; v[6] *= s1.z; v[7] *= s1.w;
                }
                if (p.flags & GEMM_RELU_BF16) {
#pragma unroll
                  for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
                }
                *reinterpret_cast<uint4*>(rowp + ((uint32_t(h2 * 4 + j) ^ sw) << 4)) =
                    make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
              }
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_4d(&p.tmOut, buf, gcol, wx, wy, img);
            bulk_commit_group();
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CTAS == 2) mbar_arrive_cluster_relaxed(tempty_remote[acc]);
          else mbar_arrive(&tempty[acc]);
        }
      } else {
        // ---- direct epilogue: one output row per thread (strided / broadcast residuals, dual outputs, TAIL mode) ------
        const int r = q * 32 + lane;
        const int ly = r / p.tw;
        const int lx = r - ly * p.tw;
        const int x = tx * p.tw + lx, y = ty * p.th + ly;
        const bool valid = (x < p.X) && (y < p.Y) && (img < p.Nimg);
        const long long off = (long long)img * p.sN + (long long)y * p.sY + (long long)x * p.sX + p.out_off;
        const long long off2 = (long long)img * p.r2N + (long long)y * p.r2Y + (long long)x * p.r2X;

        // the residual rows of this tile do not depend on the accumulator: pull them into L2 while the MMAs still run
        if (valid && !tail) {
          const int pc0 = c0;
          if (p.res1) {
#pragma unroll
            for (int c = 0; c < HALF_COLS / 32; ++c)
              if (pc0 + c * 32 < Nl) prefetch_l2(p.res1 + off + pc0 + c * 32);
          }
          if (p.res2) {
#pragma unroll
            for (int c = 0; c < HALF_COLS / 32; ++c)
              if (pc0 + c * 32 < Nl) prefetch_l2(p.res2 + off2 + pc0 + c * 32);
          }
        }
        mbar_wait(&tfull[acc], accphase);
        tc_fence_after();
        float part[kTailMax] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
        for (int c = 0; c < HALF_COLS / 32; ++c) {
          const int col0 = col_off + c * 32;
          const int gcol = n0 + col0;
          if (gcol >= Nl) break;
          uint32_t vr[32];
          __syncwarp();  // tcgen05.ld is .sync.aligned: reconverge after the per-row predicated body
          // issue the residual loads first so that their latency overlaps the TMEM read
          float4 rr[8];
          const bool has_r1 = valid && !tail && p.res1 != nullptr;
          if (has_r1) {
            const float4* rp = reinterpret_cast<const float4*>(p.res1 + off + gcol);
#pragma unroll
            for (int j = 0; j < 8; ++j) rr[j] = rp[j];
          }
          tmem_ld32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t(acc * BN + col0), vr);
          tmem_ld_wait();
          if (!valid) continue;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(vr[j]);
          if (p.bias) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + gcol + j));
              v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
            }
          }
          if (tail) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
#pragma unroll
            for (int o = 0; o < kTailMax; ++o) {
              if (o >= p.tail_out) break;
              const float* w = p.w2 + (long long)o * p.N + gcol;
              float s = 0.f;
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 w4 = __ldg(reinterpret_cast<const float4*>(w + j));
                s += v[j] * w4.x + v[j + 1] * w4.y + v[j + 2] * w4.z + v[j + 3] * w4.w;
              }
              part[o] += s;
            }
            continue;
          }
          if (p.flags & GEMM_GELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
          }
          if (p.scale) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + gcol + j));
              v[j] *= s4.x; v[j + 1] *= s4.y; v[j + 2] *= s4.z; v[j + 3] *= s4.w;
            }
          }
          if (has_r1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[4 * j] += rr[j].x; v[4 * j + 1] += rr[j].y; v[4 * j + 2] += rr[j].z; v[4 * j + 3] += rr[j].w;
            }
          }
          if (p.res2) {
            const float4* rp = reinterpret_cast<const float4*>(p.res2 + off2 + gcol);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 r4 = rp[j];
              v[4 * j] += r4.x; v[4 * j + 1] += r4.y; v[4 * j + 2] += r4.z; v[4 * j + 3] += r4.w;
            }
          }
          if (p.out_f32) {
            float4* op = reinterpret_cast<float4*>(p.out_f32 + off + gcol);
#pragma unroll
            for (int j = 0; j < 8; ++j) op[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
          if (p.out_bf16) {
            if (p.flags & GEMM_RELU_BF16) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
            }
            uint4* op = reinterpret_cast<uint4*>(p.out_bf16 + off + gcol);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              op[j] = make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                                 pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7]));
            }
          }
        }
        // accumulator stage drained -> hand it back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CTAS == 2) mbar_arrive_cluster_relaxed(tempty_remote[acc]);
          else mbar_arrive(&tempty[acc]);
        }

        if (tail) {
          // combine the two column halves of each row through smem (double-buffered by accumulator stage)
          float* sc = scratch + acc * (2 * 128 * kTailMax);
          float4* mine = reinterpret_cast<float4*>(sc + (half * 128 + r) * kTailMax);
          mine[0] = make_float4(part[0], part[1], part[2], part[3]);
          mine[1] = make_float4(part[4], part[5], part[6], part[7]);
          asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
          if (half == 0 && valid) {
            const float4* oth = reinterpret_cast<const float4*>(sc + (128 + r) * kTailMax);
            const float4 o4 = oth[0], o5 = oth[1];
            float yv[kTailMax] = {part[0] + o4.x, part[1] + o4.y, part[2] + o4.z, part[3] + o4.w,
                                  part[4] + o5.x, part[5] + o5.y, part[6] + o5.z, part[7] + o5.w};
            const long long plane = (long long)p.Y * p.X;
#pragma unroll
            for (int o = 0; o < kTailMax; ++o) {
              if (o >= p.tail_out) break;
              float t = yv[o] + p.tail_b2[o];
              const long long idx = ((long long)img * p.tail_out + o) * plane + (long long)y * p.X + x;
              if (p.out2) p.out2[idx] = t;
              if (p.flags & GEMM_TAIL_SIGMOID) t = 1.0f / (1.0f + __expf(-t));
              else t = t * p.tail_scale[o] + p.tail_shift[o];
              // FlowAdaptor.output_normalized_coordinate (adaptors.py:110-118): the head predicts the landing position
              if ((p.flags & GEMM_TAIL_COORD) && o < 2) t -= (float)(o == 0 ? x : y);
              p.out_f32[idx] = t;
            }
          }
        }
      }
    }
    if (emode != EPI_DIRECT && lane == 0) bulk_wait<0>();
    }  // !R2
  }

  tc_fence_before();
  if constexpr (CTAS == 2) {
    cluster_sync_all();  // neither CTA may exit (or free TMEM) while the pair's MMAs / remote arrives are in flight
    if (warp == 2) tmem_dealloc2<C::TMEM_COLS>(tmem_base);
  } else {
    __syncthreads();
    if (warp == 2) tmem_dealloc<C::TMEM_COLS>(tmem_base);
  }
}

// ufm_debug_flag 0: 0 = always the direct epilogue; 1 = TMA epilogues where eligible, EPI_TMA_MIX and the reduce-add in-place
// residual included (default); 2 = TMA epilogues without EPI_TMA_MIX; | 4 = in-place residuals through a residual slab
// (TMA load + add in the SM) instead of the TMA reduce-add store
int g_gemm_epi = 1;
// ufm_debug_flag 1 / UFM_PAIR: bit 0 = CTA-pair (cta_group::2) tiles for BN = 256, bit 1 = also for BN = 128 (the N = 128 processor
// convs of the DPT heads: each CTA then loads half of the weight tile, 39 instead of 69 B/clk/SM of operand traffic against the
// ~43 B/clk/SM the L2 delivers; neutral in round 1 (808 vs 795 TFLOP/s, the issue loop was the limit then), +12 - 14 % now)
int g_gemm_pair = -1;

template <int BN, int CTAS, bool HALO, bool R2 = false>
static cudaError_t launch_one(const GemmParams& p, int grid, cudaStream_t stream) {
  using C = Cfg<BN, CTAS, HALO, R2>;
  static bool attr = false;
  if (!attr) {
    cudaError_t err = cudaFuncSetAttribute(gemm_kernel<BN, CTAS, HALO, R2>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
    if (err != cudaSuccess) return err;
    attr = true;
  }
  return launch_kc(gemm_kernel<BN, CTAS, HALO, R2>, CTAS, dim3(grid), dim3(kThreads), C::SMEM_BYTES, stream, p);
}

}  // namespace

void set_gemm_epilogue_variant(int v) { g_gemm_epi = v; }
int gemm_epilogue_variant() { return g_gemm_epi; }
void set_gemm_pair_mode(int v) { g_gemm_pair = v; }
// ufm_debug_flag 3: 1 = halo-resident A operand for the 3x3 convs.  (Bring-up note: the operand descriptors must NOT set
// the base-offset field for the 128-byte-aligned tap starts; with it set every tap but the centre one reads garbage.)
int g_gemm_halo = 1;
void set_gemm_halo_mode(int v) { g_gemm_halo = v; }
int gemm_halo_mode() { return g_gemm_halo; }
int gemm_pair_mode() {
  if (g_gemm_pair < 0) g_gemm_pair = getenv("UFM_PAIR") ? atoi(getenv("UFM_PAIR")) : 3;
  return g_gemm_pair;
}
int g_gemm_split = 1;  // ufm_debug_flag 5
void set_gemm_split_mode(int v) { g_gemm_split = v; }
int gemm_split_mode() { return g_gemm_split; }
// ufm_debug_flag 6 / UFM_ROWS2: bit 0 = 512-row pair tiles, bit 1 = run-ahead, bit 2 = also with GELU, bit 3 = only K >= 2048 or
// (K >= 1024 and N >= 2048), bit 4 = only K >= 2048.  Default 27: the long-K GEMMs (fc2), +10 - 16 % on every box measured;
// the K = 1024 qkv GEMM gains 5 % on some boxes and loses 2 % on others, short-K shapes lose (profiles/r02_gemm_bench_fold.txt)
int g_gemm_rows2 = -1;
void set_gemm_rows2_mode(int v) { g_gemm_rows2 = v; }
int gemm_rows2_mode() {
  if (g_gemm_rows2 < 0) g_gemm_rows2 = getenv("UFM_ROWS2") ? atoi(getenv("UFM_ROWS2")) : 27;
  return g_gemm_rows2;
}
int g_gemm_gelu_packed = -1;  // ufm_debug_flag 9 / UFM_GELU_PACKED (tuning); -1 = default
void set_gemm_gelu_packed(int v) { g_gemm_gelu_packed = v; }
int gemm_gelu_packed() {
  if (g_gemm_gelu_packed < 0) g_gemm_gelu_packed = getenv("UFM_GELU_PACKED") ? atoi(getenv("UFM_GELU_PACKED")) : 1;
  return g_gemm_gelu_packed;
}
int g_gemm_red_prefetch = -1;  // ufm_debug_flag 10 / UFM_RED_PREFETCH (tuning)
void set_gemm_red_prefetch(int v) { g_gemm_red_prefetch = v; }
int gemm_red_prefetch() {
  if (g_gemm_red_prefetch < 0) g_gemm_red_prefetch = getenv("UFM_RED_PREFETCH") ? atoi(getenv("UFM_RED_PREFETCH")) : 0;
  return g_gemm_red_prefetch;
}
int g_gemm_warp_shift = 0;  // ufm_debug_flag 7
void set_gemm_warp_shift(int v) { g_gemm_warp_shift = v; }
int gemm_warp_shift() { return g_gemm_warp_shift; }

cudaError_t launch_gemm(const GemmParams& p, int BN, int num_sms, cudaStream_t stream) {
  const int ctas = p.ctas == 2 ? 2 : 1;
  const int tiles_m1 = p.Nimg * p.tiles_y * p.tiles_x;
  if (p.rows2) {
    // r2_items whole 512-row items, then the remaining rows as (column-sliced) 256-row half items in the same launch
    if (ctas != 2 || BN != 256 || p.halo || p.taps != 1 || (p.flags & GEMM_TAIL) || p.r2_items < 0 || p.r2_half_tiles < 0 ||
        p.r2_items % p.tiles_n || p.r2_half_tiles % p.tiles_n || p.split_log2 > 2 ||
        (p.r2_items / p.tiles_n) * 4 + (p.r2_half_tiles / p.tiles_n) * 2 < tiles_m1 || p.r2_half_mt0 != p.r2_items / p.tiles_n * 2 ||
        (p.epi_mode != EPI_TMA_BF16 && !(p.epi_mode == EPI_TMA_F32 && (!p.res1 || (p.flags & GEMM_RED_INPLACE)))))
      return cudaErrorInvalidValue;
    const int items = p.r2_items + (p.r2_half_tiles << p.split_log2);
    if (items <= 0) return cudaSuccess;
    const int g2 = items * 2 < num_sms ? items * 2 : num_sms / 2 * 2;
    return launch_one<256, 2, false, true>(p, g2, stream);
  }
  const int total = ((tiles_m1 + ctas - 1) / ctas) * p.tiles_n;
  if (total <= 0) return cudaSuccess;
  int grid = total * ctas < num_sms ? total * ctas : num_sms / ctas * ctas;
  if (p.halo) {
    if (p.taps != 9 || p.tw != HALO_TW || p.th != HALO_TH) return cudaErrorInvalidValue;
    if (ctas == 2 && BN == 256) return launch_one<256, 2, true>(p, grid, stream);
    if (ctas == 2 && BN == 128) return launch_one<128, 2, true>(p, grid, stream);
    if (BN == 256) return launch_one<256, 1, true>(p, grid, stream);
    if (BN == 128) return launch_one<128, 1, true>(p, grid, stream);
    if (BN == 64) return launch_one<64, 1, true>(p, grid, stream);
    return cudaErrorInvalidValue;
  }
  if (ctas == 2) {
    if (BN == 256) return launch_one<256, 2, false>(p, grid, stream);
    if (BN == 128) return launch_one<128, 2, false>(p, grid, stream);
    return cudaErrorInvalidValue;
  }
  if (BN == 256) return launch_one<256, 1, false>(p, grid, stream);
  if (BN == 128) return launch_one<128, 1, false>(p, grid, stream);
  return cudaErrorInvalidValue;
}

UFM_DEFINE_PDL_EARLY_SETTER(gemm)

}  // namespace ufm
