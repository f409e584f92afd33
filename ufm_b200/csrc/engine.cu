// UFM engine: owns packed weights and workspace on one B200 and sequences the hand-written kernels for
// UniFlowMatchConfidence.forward (uniflowmatch/models/ufm.py:562-662) and
// UniFlowMatchClassificationRefinement.forward (ufm.py:843-1009) behind the C ABI of include/ufm_b200.h.
// There is no CPU path: creation fails without an sm_100 device.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/ufm_b200.h"
#include "kernels.h"
#include "launch.cuh"

using namespace ufm;
typedef __nv_bfloat16 bf16;

namespace {
constexpr int kGraphMaxPairs = 4;  // ufm_forward / ufm_forward_host replay a CUDA graph up to this batch size

// ufm_debug_flag 8 / UFM_BRANCHES: independent branches of a launch-bound pass (<= kGraphMaxPairs pairs) on side streams
int g_branch_mode = -1;
void set_branch_mode(int v) { g_branch_mode = v; }
int branch_mode() {
  if (g_branch_mode < 0) g_branch_mode = getenv("UFM_BRANCHES") ? atoi(getenv("UFM_BRANCHES")) : 1;
  return g_branch_mode;
}

struct Tensor {
  void* ptr = nullptr;
  int dtype = UFM_F32;
  std::vector<int64_t> shape;
  std::vector<float> host;  // host copy of small fp32 tensors (epilogue constants)
  int64_t numel() const {
    int64_t n = 1;
    for (auto s : shape) n *= s;
    return n;
  }
  int64_t bytes() const { return numel() * (dtype == UFM_F32 ? 4 : dtype == UFM_BF16 ? 2 : 1); }
};

struct AView {
  const void* p;
  int C, X, Y, N;
  long long sx, sy, sn;  // element strides
  int tw, th;
};

void pick_tile(int W, int H, int* tw, int* th) {
  static const int cand[7][2] = {{16, 8}, {8, 16}, {32, 4}, {4, 32}, {64, 2}, {128, 1}, {2, 64}};
  long long best = -1;
  for (auto& c : cand) {
    const long long tiles = (long long)((W + c[0] - 1) / c[0]) * ((H + c[1] - 1) / c[1]);
    if (best < 0 || tiles < best) {
      best = tiles;
      *tw = c[0];
      *th = c[1];
    }
  }
}

AView rows_view(const void* p, long long M, int C) { return AView{p, C, (int)M, 1, 1, C, M * C, M * C, 128, 1}; }
AView image_view(const void* p, int N, int H, int W, int C) {
  AView a{p, C, W, H, N, C, (long long)W * C, (long long)H * W * C, 16, 8};
  pick_tile(W, H, &a.tw, &a.th);
  return a;
}

struct Epi {
  const float* bias = nullptr;
  const float* scale = nullptr;
  const float* res1 = nullptr;
  const float* res2 = nullptr;
  long long r2n = 0, r2y = 0, r2x = 0;
  float* out_f32 = nullptr;
  bf16* out_bf16 = nullptr;
  long long sn = 0, sy = 0, sx = 0, out_off = 0;
  uint32_t flags = 0;
  const float* w2 = nullptr;
  float b2[kTailMax] = {0, 0, 0, 0, 0, 0, 0, 0}, tscale[kTailMax] = {1, 1, 1, 1, 1, 1, 1, 1}, tshift[kTailMax] = {0, 0, 0, 0, 0, 0, 0, 0};
  int tail_out = 0;
  float* out2 = nullptr;
};

bool build_gemm(GemmParams* gp, int* BN, const AView& a_in, const void* wt, int n_out, int taps, int cin_pad, const Epi& e,
                char* err, int errlen, int num_sms = 0, int rows2 = 0, long long r2_rows = 0) {
  memset(gp, 0, sizeof(*gp));
  gp->warp_shift = gemm_warp_shift();
  AView a = a_in;
  const bool halo = taps == 9 && (gemm_halo_mode() & 1);
  if (halo) {  // gemm.cu HALO: 8 x 16-pixel tiles, one (8+2) x (16+2) halo box per channel chunk
    a.tw = 8;
    a.th = 16;
  }
  gp->halo = halo ? 1 : 0;
  if (a.tw * a.th != 128 || (a.tw & (a.tw - 1))) {
    snprintf(err, errlen, "tile %dx%d must cover 128 rows with a power-of-two width", a.tw, a.th);
    return false;
  }
  const uint64_t adims[4] = {(uint64_t)a.C, (uint64_t)a.X, (uint64_t)a.Y, (uint64_t)a.N};
  const uint64_t astr[4] = {2, (uint64_t)a.sx * 2, (uint64_t)a.sy * 2, (uint64_t)a.sn * 2};
  const uint32_t abox[4] = {64, (uint32_t)a.tw + (halo ? 2u : 0u), (uint32_t)a.th + (halo ? 2u : 0u), 1};
  if (!make_tmap_bf16(&gp->tmA, a.p, 4, adims, astr, abox, err, errlen)) return false;
  const bool tail = (e.flags & GEMM_TAIL) != 0;
  // TAIL launches (the 420 x 560 processor conv with ReLU + 1x1 + adaptor in the epilogue) are epilogue-heavy: with the
  // epilogue on the LOW physical warps the TMA / MMA warps win the issue arbitration (highest warp id first) and are not
  // delayed behind eight busy epilogue warps: proc.c2 0.301 -> 0.265 ms (profiles/r02_gemm_bench_rows2.txt, w4 column);
  // neutral to slightly negative elsewhere, so only here
  if (tail && gp->warp_shift == 0) gp->warp_shift = 4;
  *BN = (!tail && n_out % 256 == 0) ? 256 : (halo && !tail && n_out <= 64) ? 64 : 128;
  const uint64_t bdims[2] = {(uint64_t)taps * cin_pad, (uint64_t)n_out};
  const uint64_t bstr[2] = {2, (uint64_t)taps * cin_pad * 2};
  gp->ctas = ((*BN == 256 && (gemm_pair_mode() & 1)) || (*BN == 128 && (gemm_pair_mode() & 2))) ? 2 : 1;
  const uint32_t bbox[2] = {64, (uint32_t)(*BN / gp->ctas)};
  if (!make_tmap_bf16(&gp->tmB, wt, 2, bdims, bstr, bbox, err, errlen)) return false;
  if (tail) {
    if (n_out > 128 || e.tail_out < 1 || e.tail_out > kTailMax || n_out % 32) {
      snprintf(err, errlen, "tail epilogue needs N<=128, N%%32==0, 1..4 outputs (N=%d out=%d)", n_out, e.tail_out);
      return false;
    }
  } else if (n_out % 32) {
    snprintf(err, errlen, "GEMM N=%d must be a multiple of 32", n_out);
    return false;
  }
  gp->tw = a.tw; gp->th = a.th;
  gp->tw_log2 = 0;
  while ((1 << gp->tw_log2) < a.tw) ++gp->tw_log2;
  gp->X = a.X; gp->Y = a.Y; gp->Nimg = a.N;
  gp->tiles_x = (a.X + a.tw - 1) / a.tw;
  gp->tiles_y = (a.Y + a.th - 1) / a.th;
  gp->N = n_out;
  gp->tiles_n = (n_out + *BN - 1) / *BN;
  gp->taps = taps;
  gp->cchunks = (a.C + 63) / 64;
  gp->cin_pad = cin_pad;
  // ---- last-wave column slices (gemm.cu tile_ref): the static schedule runs ceil(T / U) waves of T tiles on U units; when
  // the last wave holds R <= U/4 (U/2) tiles, each is cut into 4 (2) column slices of >= 64 columns so that the wave
  // costs a fraction of a tile time (UFM-Base encoder proj / fc2: 604 tiles on 74 pairs, R = 12 -> 48 quarter tiles) ----
  if (!halo && !tail && taps == 1 && num_sms > 0 && gemm_split_mode() && !rows2) {
    const int tiles_m1 = a.N * gp->tiles_y * gp->tiles_x;
    const int T = ((tiles_m1 + gp->ctas - 1) / gp->ctas) * gp->tiles_n;
    const int U = num_sms / gp->ctas;
    const int R = U > 0 ? T % U : 0;
    int sl = 0;
    if (T > U && R > 0) {
      if (*BN / 4 >= 64 && R * 4 <= U) sl = 2;
      else if (*BN / 2 >= 64 && R * 2 <= U) sl = 1;
    }
    if (sl) {
      gp->split_from = T - R;
      gp->split_log2 = sl;
      const uint32_t sbox[2] = {64, (uint32_t)((*BN / gp->ctas) >> sl)};
      if (!make_tmap_bf16(&gp->tmBs, wt, 2, bdims, bstr, sbox, err, errlen)) return false;
    }
  }
  if (rows2 && num_sms > 0) {
    // rows2 launch (gemm.cu R2): r2_rows rows as 512-row items, the rest as half items (256-row pair tiles) behind them in
    // the same launch; one partly filled wave of half items is column-sliced like the last wave above
    const int U = num_sms / 2;
    gp->r2_items = (int)(r2_rows / 512) * gp->tiles_n;
    gp->r2_half_mt0 = (int)(r2_rows / 256);
    const int R = (int)((a.X - r2_rows + 255) / 256) * gp->tiles_n;
    gp->r2_half_tiles = R;
    int sl = 0;
    if (gemm_split_mode() && R > 0) {
      if (R * 4 <= U) sl = 2;
      else if (R * 2 <= U) sl = 1;
    }
    if (sl) {
      gp->split_log2 = sl;
      const uint32_t sbox[2] = {64, (uint32_t)((*BN / gp->ctas) >> sl)};
      if (!make_tmap_bf16(&gp->tmBs, wt, 2, bdims, bstr, sbox, err, errlen)) return false;
    }
  }
  if (gp->cchunks * 64 > cin_pad && taps > 1) {
    snprintf(err, errlen, "conv weight pitch %d too small for C=%d", cin_pad, a.C);
    return false;
  }
  // ---- epilogue path: TMA slabs when the output (and residual) is one dense-row tensor the box geometry can address ----
  gp->epi_mode = EPI_DIRECT;
  bool red_inplace = false;
  if (!tail && gemm_epilogue_variant() && !e.res2 && (e.out_f32 != nullptr) != (e.out_bf16 != nullptr) &&
      !(e.out_bf16 && e.res1)) {
    const bool f32 = e.out_f32 != nullptr;
    const long long es = f32 ? 4 : 2;
    const int bw = a.tw < 32 ? a.tw : 32;
    // strides of extent-1 dimensions are never used for addressing: give them any legal value
    const long long sx = e.sx, sy = a.Y > 1 ? e.sy : e.sx * a.X, sn = a.N > 1 ? e.sn : sy * a.Y;
    const char* base = f32 ? reinterpret_cast<const char*>(e.out_f32) : reinterpret_cast<const char*>(e.out_bf16);
    base += e.out_off * es;
    const bool ok = sx > 0 && sy > 0 && sn > 0 && (sx * es) % 16 == 0 && (sy * es) % 16 == 0 && (sn * es) % 16 == 0 &&
                    (reinterpret_cast<uintptr_t>(base) & 15) == 0 && sx >= n_out;
    if (ok) {
      const uint64_t od[4] = {(uint64_t)n_out, (uint64_t)a.X, (uint64_t)a.Y, (uint64_t)a.N};
      const uint64_t os[4] = {(uint64_t)es, (uint64_t)(sx * es), (uint64_t)(sy * es), (uint64_t)(sn * es)};
      const uint32_t ob[4] = {f32 ? 32u : 64u, (uint32_t)bw, (uint32_t)(32 / bw), 1u};
      if (!make_tmap(&gp->tmOut, f32, base, 4, od, os, ob, err, errlen)) return false;
      if (f32 && e.res1) {
        const char* rb = reinterpret_cast<const char*>(e.res1) + e.out_off * es;
        if ((reinterpret_cast<uintptr_t>(rb) & 15) != 0) {
          snprintf(err, errlen, "residual base not 16-byte aligned");
          return false;
        }
        if (!make_tmap(&gp->tmRes, true, rb, 4, od, os, ob, err, errlen)) return false;
      }
      gp->epi_mode = f32 ? EPI_TMA_F32 : EPI_TMA_BF16;
      red_inplace = f32 && e.res1 && e.res1 == e.out_f32 && !(gemm_epilogue_variant() & 4);
    }
  }
  // EPI_TMA_MIX: what the two plain TMA modes cannot take (a second residual, a bf16 copy next to / instead of the fp32
  // output of a residual add) as long as res1 / out_f32 have the dense-row geometry of a TMA box
  // (a bf16-only output with a residual stays on the direct path: measured 0.283 ms direct against 0.306 ms here for
  // the 120x160 RCU conv - without an fp32 store the direct path already hides behind the MMAs)
  if (gp->epi_mode == EPI_DIRECT && !tail && (gemm_epilogue_variant() & 1) && e.out_f32 && (e.res2 || e.out_bf16) &&
      n_out % 32 == 0) {
    const int bw = a.tw < 32 ? a.tw : 32;
    const long long sx = e.sx, sy = a.Y > 1 ? e.sy : e.sx * a.X, sn = a.N > 1 ? e.sn : sy * a.Y;
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    bool ok = sx > 0 && sy > 0 && sn > 0 && sx % 4 == 0 && sy % 4 == 0 && sn % 4 == 0 && e.out_off % 4 == 0 && sx >= n_out;
    if (e.out_f32) ok = ok && al16(e.out_f32);
    if (e.res1) ok = ok && al16(e.res1);
    if (e.out_bf16) ok = ok && al16(e.out_bf16);
    if (e.res2) ok = ok && al16(e.res2) && e.r2x % 4 == 0 && e.r2y % 4 == 0 && e.r2n % 4 == 0;
    if (ok) {
      const uint64_t od[4] = {(uint64_t)n_out, (uint64_t)a.X, (uint64_t)a.Y, (uint64_t)a.N};
      const uint64_t os[4] = {4u, (uint64_t)(sx * 4), (uint64_t)(sy * 4), (uint64_t)(sn * 4)};
      const uint32_t ob[4] = {32u, (uint32_t)bw, (uint32_t)(32 / bw), 1u};
      if (e.out_f32 && !make_tmap(&gp->tmOut, true, reinterpret_cast<const char*>(e.out_f32) + e.out_off * 4, 4, od, os, ob,
                                  err, errlen))
        return false;
      if (e.res1 && !make_tmap(&gp->tmRes, true, reinterpret_cast<const char*>(e.res1) + e.out_off * 4, 4, od, os, ob, err,
                               errlen))
        return false;
      gp->epi_mode = EPI_TMA_MIX;
    }
  }
  if (rows2) {
    // 512 x 256 pair tiles (gemm.cu R2): TMA epilogue without a residual slab load
    const bool epi_ok = gp->epi_mode == EPI_TMA_BF16 || (gp->epi_mode == EPI_TMA_F32 && (!e.res1 || red_inplace));
    if (halo || tail || taps != 1 || *BN != 256 || gp->ctas != 2 || !epi_ok || r2_rows <= 0 || r2_rows % 512 || r2_rows > a.X ||
        a.Y != 1 || a.N != 1 || a.tw != 128) {
      snprintf(err, errlen, "rows2 tiles need a plain CTA-pair BN=256 row-matrix GEMM with a TMA epilogue");
      return false;
    }
    gp->rows2 = rows2;
  }
  gp->flags = e.flags | (red_inplace ? GEMM_RED_INPLACE : 0u);
  if ((e.flags & GEMM_GELU) && gp->epi_mode == EPI_TMA_BF16 && gemm_gelu_packed()) gp->flags |= GEMM_GELU_PACKED;
  if (red_inplace && gemm_red_prefetch()) gp->flags |= GEMM_RED_PREFETCH;
  gp->bias = e.bias; gp->scale = e.scale; gp->res1 = e.res1; gp->res2 = e.res2;
  gp->out_f32 = e.out_f32; gp->out_bf16 = e.out_bf16;
  gp->sN = e.sn; gp->sY = e.sy; gp->sX = e.sx; gp->out_off = e.out_off;
  gp->r2N = e.r2n; gp->r2Y = e.r2y; gp->r2X = e.r2x;
  gp->w2 = e.w2; gp->tail_out = e.tail_out; gp->out2 = e.out2;
  for (int i = 0; i < kTailMax; ++i) {
    gp->tail_b2[i] = e.b2[i];
    gp->tail_scale[i] = e.tscale[i];
    gp->tail_shift[i] = e.tshift[i];
  }
  return true;
}

// Rows of a plain row-matrix GEMM that run as 512 x 256 pair tiles (gemm.cu R2): the largest count of whole 512-row items
// whose (items x column tiles) fill the pair units an integral number of times; the remaining rows follow as 256-row half
// items in the same launch.  0 = not eligible.
long long rows2_rows(const AView& a, int n_out, int taps, const Epi& e, int num_sms) {
  if (!(gemm_rows2_mode() & 1) || !(gemm_pair_mode() & 1) || !gemm_epilogue_variant()) return 0;
  if (taps != 1 || a.Y != 1 || a.N != 1 || a.tw != 128 || a.th != 1 || n_out % 256 || a.C % 64) return 0;
  if ((e.flags & GEMM_TAIL) || e.res2 || (e.out_f32 != nullptr) == (e.out_bf16 != nullptr)) return 0;
  if (e.out_bf16 && e.res1) return 0;
  if (e.out_f32 && e.res1 && (e.res1 != e.out_f32 || (gemm_epilogue_variant() & 4))) return 0;  // reduce-add in place only
  // heavy per-element epilogues (GELU) need the overlap of the double-buffered kernel
  if ((e.flags & GEMM_GELU) && !(gemm_rows2_mode() & 4)) return 0;
  // bits 3 / 4: shape rules (profiles/r02_gemm_bench_fold.txt).  Short-K items cannot hide the single accumulator set's drain
  // (the second row block's accumulator is released only after the first one's epilogue math): K = 768 loses up to 25 % on
  // some boxes, K = 1024 is +5 % .. -2 % depending on the box, K >= 3072 gains 10 - 16 % everywhere
  if ((gemm_rows2_mode() & 8) && !(a.C >= 2048 || (a.C >= 1024 && n_out >= 2048))) return 0;
  if ((gemm_rows2_mode() & 16) && a.C < 2048) return 0;
  const int units = num_sms / 2, tiles_n = n_out / 256;
  const long long items_m = a.X / 512;
  int g = units, t = tiles_n;  // step = units / gcd(units, tiles_n)
  while (t) { const int r = g % t; g = t; t = r; }
  const long long step = units / g;
  const long long r2 = items_m / step * step * 512;
  // the rows behind the whole waves run as half items without a second accumulator stage: worth it for one (sliced) wave of
  // them, not when a large share of the matrix would take that path (e.g. 12 pairs: 34 % of the rows)
  if (r2 > 0 && (a.X - r2 + 255) / 256 * tiles_n > units) return 0;
  return r2;
}

bool build_attn(AttnParams* ap, const void* qkv, int n_seq, int seq_len, int heads, float scale, void* out, char* err,
                int errlen) {
  memset(ap, 0, sizeof(*ap));
  const uint64_t pitch = (uint64_t)3 * heads * 64;
  const uint64_t dims[4] = {64, (uint64_t)3 * heads, (uint64_t)seq_len, (uint64_t)n_seq};
  const uint64_t str[4] = {2, 128, pitch * 2, pitch * 2 * (uint64_t)seq_len};
  const uint32_t box[4] = {64, 1, 128, 1};
  if (!make_tmap_bf16(&ap->tmQKV, qkv, 4, dims, str, box, err, errlen)) return false;
  ap->n_seq = n_seq; ap->seq_len = seq_len; ap->heads = heads;
  ap->scale_log2 = scale * 1.4426950408889634f;
  ap->out = reinterpret_cast<bf16*>(out);
  ap->out_pitch = (long long)heads * 64;
  return true;
}

}  // namespace

// C-ABI entry points run on the engine's device and restore the caller's current device on return (a multi-GPU
// process must not find its current device switched by a predict on another GPU).
struct DevGuard {
  int prev = -1, dev;
  explicit DevGuard(int d) : dev(d) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DevGuard() {
    if (prev >= 0 && prev != dev) cudaSetDevice(prev);
  }
};

struct ufm_engine {
  ufm_config cfg;
  int device = 0;
  int num_sms = 148;
  int max_pairs = 0;
  bool prepared = false;
  bool keep_aux = false;
  int view_index = 1;  // row of info.viewpos the non-reference view adds (ufm_set_view_index)
  int launches = 0;
  int last_pairs = 0;
  std::string err;
  std::map<std::string, Tensor> w;
  std::map<std::string, Tensor> buf;    // workspace
  std::map<std::string, Tensor> stage;  // views of the last forward (not owned)
  cudaStream_t st = nullptr;
  // Every pass shares one workspace: a pass on a different stream than the previous one first waits for the event the
  // previous pass recorded at its end (ufm_forward on the caller's stream, ufm_forward_host on the legacy stream or
  // the private graph stream may follow each other in any order).
  cudaEvent_t ev_done = nullptr;
  cudaStream_t done_stream = nullptr;
  bool have_done = false, capturing = false;
  bool order_after_previous(cudaStream_t s) {
    if (capturing) return true;
    if (have_done && s != done_stream) return ck(cudaStreamWaitEvent(s, ev_done, 0), "order after the previous pass");
    return true;
  }
  bool mark_done(cudaStream_t s) {
    if (capturing) return true;
    if (!ev_done && !ck(cudaEventCreateWithFlags(&ev_done, cudaEventDisableTiming), "event")) return false;
    if (!ck(cudaEventRecord(ev_done, s), "record end of pass")) return false;
    done_stream = s;
    have_done = true;
    return true;
  }
  int forward_impl(const void* src, const void* tgt, int pairs, int in_dtype, int in_layout, float* flow_out,
                   float* covis_out, cudaStream_t stream);
  // early programmatic-launch trigger for launch-latency-bound passes only (<= 2 pairs; launch.cuh).  The flag is a
  // constant per translation unit and per device; it is read when a kernel executes, so it is (re)set in stream order in
  // front of every pass or graph replay whose batch size asks for the other value.
  bool set_pdl_early(int pairs, cudaStream_t s) {
    static int cur[64];
    static bool init = false;
    if (!init) {
      for (int& v : cur) v = -1;
      init = true;
    }
    if (capturing || device >= 64) return true;
    const int want = (pairs <= 2 && !getenv("UFM_NO_PDL_EARLY")) ? 1 : 0;
    if (cur[device] == want) return true;
    if (!ck(ufm_set_pdl_early_gemm(want, s), "pdl flag") || !ck(ufm_set_pdl_early_attention(want, s), "pdl flag") ||
        !ck(ufm_set_pdl_early_simt(want, s), "pdl flag") || !ck(ufm_set_pdl_early_prepost(want, s), "pdl flag") ||
        !ck(ufm_set_pdl_early_unet(want, s), "pdl flag"))
      return false;
    cur[device] = want;
    return true;
  }
  // NVTX ranges per stage of the pass (UFM_NVTX=1): "ufm_forward" > encoder / info_sharing / flow_head / uncertainty_head /
  // classification_head / unet_branch / local_refinement.  The tracing aid the reference has for this path is the
  // CUDA-event decorator of UC/models/encoders/utils.py:11-86; ufm_set_profile gives its per-kernel-family equivalent and
  // these ranges name the stages in an ncu / Nsight timeline.
  bool nvtx = getenv("UFM_NVTX") != nullptr;
  int nvtx_depth = 0;
  void nvtx_stage(const char* name) {  // closes the previous stage range and opens `name` (nullptr: only close)
    if (!nvtx) return;
    if (nvtx_depth > 1) {
      nvtxRangePop();
      --nvtx_depth;
    }
    if (name) {
      nvtxRangePushA(name);
      ++nvtx_depth;
    }
  }
  // ufm_forward_host: result copies run on their own stream as soon as each output is final (flow after the flow head
  // or the refinement, covisibility after the uncertainty head), under the rest of the pass
  cudaStream_t copy_st = nullptr;
  cudaEvent_t ev_flow = nullptr, ev_cov = nullptr;
  bool mark_outputs = false;
  int ln_reverse = getenv("UFM_LN_FWD") ? 0 : 1;
  // (mean, std) of the encoder's data_norm_type; an all-zero config entry means the "dinov2" constants
  float img_mean(int i) const {
    static const float m[3] = {0.485f, 0.456f, 0.406f};
    return cfg.img_std[0] == 0.f ? m[i] : cfg.img_mean[i];
  }
  float img_std(int i) const {
    static const float s[3] = {0.229f, 0.224f, 0.225f};
    return cfg.img_std[0] == 0.f ? s[i] : cfg.img_std[i];
  }
  // ufm_forward_host, small batches (launch-bound: ~330 launches and as many tensor-map encodes per pass): the pass is
  // captured once per (pairs, dtype, layout, keep_aux) into a CUDA graph over the engine-owned staging buffers and replayed
  struct GraphEntry { cudaGraphExec_t exec = nullptr; std::map<std::string, Tensor> stage; int launches = 0; };
  std::map<std::vector<int>, GraphEntry> graphs;
  cudaStream_t graph_st = nullptr;

  // ---- concurrent branches for launch-bound passes (<= kGraphMaxPairs pairs; ufm_debug_flag 8) ----
  // With one or two pairs most launches of the DPT heads fill 4 - 40 of the 148 SMs and are pure latency chains (two heads
  // x ~36 launches x ~28 us at one pair = a third of the pass).  The pass then forks: the uncertainty head, the
  // classification head (+ UNet branch) of UFM-Refine and the four input_process chains of each head run on side streams
  // ("lanes") and join where their results are consumed.  Fork / join are event record + wait pairs, which stream capture
  // turns into graph edges, so the replayed graph carries the same parallelism.  The side-lane head works on its own copy
  // of the head workspace (buffer names with the suffix "@u", sized for kGraphMaxPairs).  Same kernels on the same
  // operands: results are bit-identical to the sequential pass (tests/test_gpu_parity.py::test_branches_bit_identical).
  static constexpr int kLanes = 8;  // 0: classification head; 1-3: flow-head chains; 4: uncertainty head; 5-7: its chains
  cudaStream_t lane_st[kLanes] = {};
  std::vector<cudaEvent_t> sync_ev;
  size_t sync_used = 0;
  bool branches = false;  // set per pass
  std::string bsfx;       // workspace suffix of the branch being issued ("" or "@u")
  cudaEvent_t next_sync() {
    if (sync_used == sync_ev.size()) {
      cudaEvent_t e = nullptr;
      if (!ck(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "sync event")) return nullptr;
      sync_ev.push_back(e);
    }
    return sync_ev[sync_used++];
  }
  // an event on the current stream that lanes started later wait for (nullptr when not branching)
  cudaEvent_t fork_point() {
    if (!branches) return nullptr;
    cudaEvent_t ev = next_sync();
    if (ev && !ck(cudaEventRecord(ev, st), "fork record")) return nullptr;
    return ev;
  }
  // the launches that follow go to lane i, ordered after `fork`
  bool lane_begin(int i, cudaEvent_t fork, cudaStream_t* saved) {
    *saved = st;
    if (!branches) return true;
    if (!fork) return false;  // fork_point failed (error already recorded)
    if (!lane_st[i] && !ck(cudaStreamCreateWithFlags(&lane_st[i], cudaStreamNonBlocking), "lane stream")) return false;
    if (!ck(cudaStreamWaitEvent(lane_st[i], fork, 0), "fork wait")) return false;
    st = lane_st[i];
    return true;
  }
  // back to the stream saved by lane_begin; *done marks the end of the lane's work (nullptr when not branching)
  bool lane_end(cudaStream_t saved, cudaEvent_t* done) {
    *done = nullptr;
    if (!branches) return true;
    cudaEvent_t ev = next_sync();
    if (!ev || !ck(cudaEventRecord(ev, st), "join record")) return false;
    *done = ev;
    st = saved;
    return true;
  }
  bool lane_join(cudaEvent_t done) { return !done || ck(cudaStreamWaitEvent(st, done, 0), "join wait"); }

  // Launch-bound regime (small batches: ~330 launches and as many tensor-map encodes per pass): the pass over the
  // engine-owned staging buffers (in_src / in_tgt -> out_flow / covis) as a CUDA graph, captured once per (pairs, dtype,
  // layout, keep_aux) on the private graph stream and replayed on `s` (ufm_forward: the caller's stream, around two
  // device-to-device copies; ufm_forward_host: the graph stream, around the H2D / D2H copies).  The caller has ordered `s`
  // after the previous pass and filled the staging inputs on `s`.
  bool graph_pass(int pairs, int in_dtype, int in_layout, cudaStream_t s) {
    void* dsrc = buf["in_src"].ptr;
    void* dtgt = buf["in_tgt"].ptr;
    float* dflow = reinterpret_cast<float*>(buf["out_flow"].ptr);
    float* dcov = cfg.unc_out_ch > 0 ? reinterpret_cast<float*>(buf["covis"].ptr) : nullptr;
    const std::vector<int> key = {pairs, in_dtype, in_layout, keep_aux ? 1 : 0, view_index, branch_mode()};
    auto it = graphs.find(key);
    if (it == graphs.end()) {
      // first call with this shape: a plain pass (it also sets the lazily initialised kernel attributes and produces this
      // call's results), then the capture on the graph stream
      if (forward(dsrc, dtgt, pairs, in_dtype, in_layout, dflow, dcov, s)) return false;
      if (!ck(cudaStreamSynchronize(s), "sync before capture")) return false;
      if (!graph_st && !ck(cudaStreamCreateWithFlags(&graph_st, cudaStreamNonBlocking), "graph stream")) return false;
      GraphEntry ge;
      cudaGraph_t graph = nullptr;
      const std::map<std::string, Tensor> stage_keep = stage;
      const int launches_keep = launches;
      if (!ck(cudaStreamBeginCapture(graph_st, cudaStreamCaptureModeThreadLocal), "begin capture")) return false;
      capturing = true;
      const int rc = forward(dsrc, dtgt, pairs, in_dtype, in_layout, dflow, dcov, graph_st);
      capturing = false;
      const cudaError_t ce = cudaStreamEndCapture(graph_st, &graph);
      st = s;
      if (rc || ce != cudaSuccess) {
        if (graph) cudaGraphDestroy(graph);
        if (rc) return false;
      }
      if (!ck(ce, "end capture")) return false;
      const cudaError_t ie = cudaGraphInstantiate(&ge.exec, graph, 0);
      cudaGraphDestroy(graph);
      if (!ck(ie, "graph instantiate")) return false;
      ge.stage = stage_keep;
      ge.launches = launches_keep;
      graphs[key] = ge;
      return true;
    }
    if (!set_pdl_early(pairs, s)) return false;
    if (!ck(cudaGraphLaunch(it->second.exec, s), "graph launch")) return false;
    stage = it->second.stage;
    launches = it->second.launches;
    last_pairs = pairs;
    st = s;
    return mark_done(s);
  }

  // ---- per-kernel-family timing (CUDA events on the launching stream; read by bench.py for the roofline) ----
  bool profile = false;
  struct ProfRec { int cat; double flops, bytes; cudaEvent_t a, b; std::string what; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_used = 0;
  double flops_override = -1.0;  // algorithmic FLOPs of the next GEMM launch when they differ from the executed ones
  cudaEvent_t next_event() {
    if (ev_used == ev_pool.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      ev_pool.push_back(e);
    }
    return ev_pool[ev_used++];
  }
  template <class F>
  bool run(int cat, double flops, double bytes, const char* what, F&& f) {
    ++launches;
    ProfRec r{cat, flops, bytes, nullptr, nullptr, what};
    if (profile) {
      r.a = next_event();
      cudaEventRecord(r.a, st);
    }
    const cudaError_t e = f();
    if (profile) {
      r.b = next_event();
      cudaEventRecord(r.b, st);
      prof.push_back(r);
    }
    return ck(e, what);
  }

  int fail(const char* fmt, ...) {
    char tmp[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(tmp, sizeof(tmp), fmt, ap);
    va_end(ap);
    err = tmp;
    return 1;
  }
  bool ck(cudaError_t e, const char* what) {
    if (e != cudaSuccess) {
      fail("%s: %s", what, cudaGetErrorString(e));
      return false;
    }
    return true;
  }
  const Tensor* W(const std::string& n) {
    auto it = w.find(n);
    if (it == w.end()) {
      fail("missing weight '%s'", n.c_str());
      return nullptr;
    }
    return &it->second;
  }
  const float* Wf(const std::string& n) {
    const Tensor* t = W(n);
    return t ? reinterpret_cast<const float*>(t->ptr) : nullptr;
  }
  bool has(const std::string& n) { return w.count(n) != 0; }
  bool alloc(const std::string& n, int dtype, int64_t numel) {
    Tensor t;
    t.dtype = dtype;
    t.shape = {numel};
    if (numel <= 0) numel = 1;
    if (!ck(cudaMalloc(&t.ptr, (size_t)numel * (dtype == UFM_F32 ? 4 : 2)), ("cudaMalloc " + n).c_str())) return false;
    buf[n] = t;
    return true;
  }
  template <typename T>
  T* B(const char* n) {
    if (!bsfx.empty()) {
      auto it = buf.find(std::string(n) + bsfx);
      if (it != buf.end()) return reinterpret_cast<T*>(it->second.ptr);
    }
    return reinterpret_cast<T*>(buf[n].ptr);
  }
  void set_stage(const char* name, void* p, int dtype, std::vector<int64_t> shape) {
    Tensor t;
    t.ptr = p;
    t.dtype = dtype;
    t.shape = shape;
    stage[name] = t;
  }

  // ---- launch helpers -------------------------------------------------------------------------------------------
  bool gemm(const AView& a, const std::string& wname, int taps, const Epi& e) {
    const Tensor* wt = W(wname);
    if (!wt) return false;
    const int n_out = (int)wt->shape[0];
    const int ktot = (int)wt->shape[1];
    char eb[512];
    double fl = 2.0 * a.N * a.X * a.Y * (double)n_out * taps * a.C;
    if (flops_override >= 0) fl = flops_override;
    flops_override = -1.0;
    const long long r2 = rows2_rows(a, n_out, taps, e, num_sms);
    if (r2 > 0) {
      GemmParams g2;
      int BN2;
      if (!build_gemm(&g2, &BN2, a, wt->ptr, n_out, taps, ktot / taps, e, eb, sizeof(eb), num_sms, gemm_rows2_mode(), r2)) {
        fail("gemm %s (rows2): %s", wname.c_str(), eb);
        return false;
      }
      return run(UFM_CAT_GEMM, fl, 0.0, ("launch gemm " + wname + " [512-row tiles]").c_str(),
                 [&] { return launch_gemm(g2, BN2, num_sms, st); });
    }
    GemmParams gp;
    int BN;
    if (!build_gemm(&gp, &BN, a, wt->ptr, n_out, taps, ktot / taps, e, eb, sizeof(eb), num_sms)) {
      fail("gemm %s: %s", wname.c_str(), eb);
      return false;
    }
    return run(taps == 9 ? UFM_CAT_CONV : UFM_CAT_GEMM, fl, 0.0, ("launch gemm " + wname).c_str(),
               [&] { return launch_gemm(gp, BN, num_sms, st); });
  }
  bool attention(const void* qkv, int n_seq, int seq_len, int heads, void* out) {
    AttnParams ap;
    char eb[512];
    if (!build_attn(&ap, qkv, n_seq, seq_len, heads, 0.125f, out, eb, sizeof(eb))) {
      fail("attention: %s", eb);
      return false;
    }
    return run(UFM_CAT_ATTN, 4.0 * n_seq * heads * (double)seq_len * seq_len * 64, 0.0, "launch attention",
               [&] { return launch_attention(ap, st); });
  }
  bool layernorm(const float* x, int n_seq, int src_seq_len, int skip, int tokens, int D, const std::string& pre,
                 int remap_B, bf16* ob, long long ob_pitch, long long ob_col, float* of = nullptr, long long of_pitch = 0,
                 long long of_col = 0, bool no_norm = false) {
    LNParams p;
    p.x = x; p.D = D; p.eps = 1e-6f;
    // an absent norm (CrossAttentionBlock.norm_y = nn.Identity() when norm_cross_tokens=False, transformer_blocks.py:458;
    // norm_intermediate=False, dinov2.py:259-261 / global_attention_transformer.py:549) runs the same kernel as a plain
    // fp32 -> bf16 conversion
    const bool identity = no_norm || (pre.size() > 4 && pre.compare(pre.size() - 4, 4, ".lny") == 0 && !has(pre + ".w"));
    p.w = identity ? nullptr : Wf(pre + ".w"); p.b = identity ? nullptr : Wf(pre + ".b");
    if (!identity && (!p.w || !p.b)) return false;
    p.n_seq = n_seq; p.src_seq_len = src_seq_len; p.src_skip = skip; p.tokens = tokens; p.remap_B = remap_B;
    p.reverse = ln_reverse;
    p.out_bf16 = ob; p.ob_pitch = ob_pitch; p.ob_col = ob_col;
    p.out_f32 = of; p.of_pitch = of_pitch; p.of_col = of_col;
    const double rows = (double)n_seq * tokens;
    return run(UFM_CAT_LN, 0.0, rows * D * (4 + (ob ? 2 : 0) + (of ? 4 : 0)), ("layernorm " + pre).c_str(),
               [&] { return launch_layernorm(p, st); });
  }

  // One pre-LN transformer block over x [n_rows, D] (fp32 residual stream), attention over n_seq sequences of seq_len.
  bool block(const std::string& pre, float* x, int n_seq, int seq_len, int D, int heads, int hidden, bool qk_norm,
             int rope_n = 0, int rope_w = 0) {
    const long long M = (long long)n_seq * seq_len;
    bf16* xn = B<bf16>("xn");
    bf16* qkv = B<bf16>("qkv");
    bf16* att = B<bf16>("att");
    bf16* h = B<bf16>("h");
    if (!layernorm(x, 1, (int)M, 0, (int)M, D, pre + ".ln1", 0, xn, D, 0)) return false;
    {
      Epi e;
      e.bias = has(pre + ".qkv.b") ? Wf(pre + ".qkv.b") : nullptr;
      e.out_bf16 = qkv; e.sx = 3LL * D; e.sy = e.sn = 0;
      if (!gemm(rows_view(xn, M, D), pre + ".qkv.w", 1, e)) return false;
    }
    if (qk_norm || rope_n > 0) {
      // q/k LayerNorm (transformer_blocks.py:164) and RoPE2D (:166-171) in one in-place pass over the q and k columns
      const float *qw = nullptr, *qb = nullptr, *kw = nullptr, *kb = nullptr;
      if (qk_norm) {
        qw = Wf(pre + ".qn.w"); qb = Wf(pre + ".qn.b"); kw = Wf(pre + ".kn.w"); kb = Wf(pre + ".kn.b");
        if (!qw || !qb || !kw || !kb) return false;
      }
      if (!run(UFM_CAT_LN, 0.0, (double)M * heads * 64 * 2 * 4, "qk_norm/rope", [&] {
            return launch_qk_norm(qkv, M, heads, qw, qb, kw, kb, 1e-6f, st, qk_norm ? 1 : 0, rope_n, rope_w);
          }))
        return false;
    }
    if (!attention(qkv, n_seq, seq_len, heads, att)) return false;
    {
      Epi e;
      e.bias = Wf(pre + ".proj.b");
      e.scale = has(pre + ".ls1") ? Wf(pre + ".ls1") : nullptr;
      e.res1 = x; e.out_f32 = x; e.sx = D;
      if (!gemm(rows_view(att, M, D), pre + ".proj.w", 1, e)) return false;
    }
    if (!layernorm(x, 1, (int)M, 0, (int)M, D, pre + ".ln2", 0, xn, D, 0)) return false;
    {
      Epi e;
      e.bias = Wf(pre + ".fc1.b");
      e.flags = GEMM_GELU;
      e.out_bf16 = h; e.sx = hidden;
      if (!gemm(rows_view(xn, M, D), pre + ".fc1.w", 1, e)) return false;
    }
    {
      Epi e;
      e.bias = Wf(pre + ".fc2.b");
      e.scale = has(pre + ".ls2") ? Wf(pre + ".ls2") : nullptr;
      e.res1 = x; e.out_f32 = x; e.sx = D;
      if (!gemm(rows_view(h, M, hidden), pre + ".fc2.w", 1, e)) return false;
    }
    return true;
  }

  // One depth of MultiViewCrossAttentionTransformer (UC/models/info_sharing/cross_attention_transformer.py:418-437) for two
  // views on the VIEW-MAJOR residual stream y [2][B][n][d]: both views are updated from the previous depth's features,
  // so the K/V projections of both branches (from norm_y of the OTHER view) are taken first; then per view
  // CrossAttentionBlock.forward (UC/models/utils/transformer_blocks.py:489-519): self-attention, cross-attention, MLP.
  // The cross q / k / v land in one fused [rows, 3, heads, 64] buffer, so the attention kernel is the same one.
  bool cross_depth(int i, float* y, int Bp, int n, int D, int heads, int hidden, bool qk_norm) {
    const long long Mv = (long long)Bp * n;
    bf16* xn = B<bf16>("xn");
    bf16* qkv = B<bf16>("qkv");
    bf16* cqkv = B<bf16>("cqkv");
    bf16* att = B<bf16>("att");
    bf16* h = B<bf16>("h");
    for (int v = 0; v < 2; ++v) {
      const std::string pre = "info.v" + std::to_string(v) + "." + std::to_string(i);
      if (!layernorm(y + (long long)(1 - v) * Mv * D, 1, (int)Mv, 0, (int)Mv, D, pre + ".lny", 0, xn, D, 0)) return false;
      Epi e;
      e.bias = has(pre + ".ckv.b") ? Wf(pre + ".ckv.b") : nullptr;
      e.out_bf16 = cqkv + (long long)v * Mv * 3 * D; e.sx = 3LL * D; e.out_off = D;
      if (!gemm(rows_view(xn, Mv, D), pre + ".ckv.w", 1, e)) return false;
    }
    for (int v = 0; v < 2; ++v) {
      const std::string pre = "info.v" + std::to_string(v) + "." + std::to_string(i);
      float* x = y + (long long)v * Mv * D;
      bf16* cq = cqkv + (long long)v * Mv * 3 * D;
      // x += ls1(attn(norm1 x))
      if (!layernorm(x, 1, (int)Mv, 0, (int)Mv, D, pre + ".ln1", 0, xn, D, 0)) return false;
      {
        Epi e;
        e.bias = has(pre + ".qkv.b") ? Wf(pre + ".qkv.b") : nullptr;
        e.out_bf16 = qkv; e.sx = 3LL * D;
        if (!gemm(rows_view(xn, Mv, D), pre + ".qkv.w", 1, e)) return false;
      }
      if (qk_norm && !run(UFM_CAT_LN, 0.0, (double)Mv * heads * 64 * 2 * 4, "qk_norm", [&] {
            return launch_qk_norm(qkv, Mv, heads, Wf(pre + ".qn.w"), Wf(pre + ".qn.b"), Wf(pre + ".kn.w"), Wf(pre + ".kn.b"),
                                  1e-6f, st);
          }))
        return false;
      if (!attention(qkv, Bp, n, heads, att)) return false;
      {
        Epi e;
        e.bias = Wf(pre + ".proj.b");
        e.scale = has(pre + ".ls1") ? Wf(pre + ".ls1") : nullptr;
        e.res1 = x; e.out_f32 = x; e.sx = D;
        if (!gemm(rows_view(att, Mv, D), pre + ".proj.w", 1, e)) return false;
      }
      // x += ls2(cross_attn(norm2 x, norm_y y, norm_y y))
      if (!layernorm(x, 1, (int)Mv, 0, (int)Mv, D, pre + ".ln2", 0, xn, D, 0)) return false;
      {
        Epi e;
        e.bias = has(pre + ".cq.b") ? Wf(pre + ".cq.b") : nullptr;
        e.out_bf16 = cq; e.sx = 3LL * D;
        if (!gemm(rows_view(xn, Mv, D), pre + ".cq.w", 1, e)) return false;
      }
      if (qk_norm && !run(UFM_CAT_LN, 0.0, (double)Mv * heads * 64 * 2 * 4, "qk_norm", [&] {
            return launch_qk_norm(cq, Mv, heads, Wf(pre + ".cqn.w"), Wf(pre + ".cqn.b"), Wf(pre + ".ckn.w"),
                                  Wf(pre + ".ckn.b"), 1e-6f, st);
          }))
        return false;
      if (!attention(cq, Bp, n, heads, att)) return false;
      {
        Epi e;
        e.bias = Wf(pre + ".cproj.b");
        e.scale = has(pre + ".ls2") ? Wf(pre + ".ls2") : nullptr;
        e.res1 = x; e.out_f32 = x; e.sx = D;
        if (!gemm(rows_view(att, Mv, D), pre + ".cproj.w", 1, e)) return false;
      }
      // x += ls3(mlp(norm3 x))
      if (!layernorm(x, 1, (int)Mv, 0, (int)Mv, D, pre + ".ln3", 0, xn, D, 0)) return false;
      {
        Epi e;
        e.bias = Wf(pre + ".fc1.b");
        e.flags = GEMM_GELU;
        e.out_bf16 = h; e.sx = hidden;
        if (!gemm(rows_view(xn, Mv, D), pre + ".fc1.w", 1, e)) return false;
      }
      {
        Epi e;
        e.bias = Wf(pre + ".fc2.b");
        e.scale = has(pre + ".ls3") ? Wf(pre + ".ls3") : nullptr;
        e.res1 = x; e.out_f32 = x; e.sx = D;
        if (!gemm(rows_view(h, Mv, hidden), pre + ".fc2.w", 1, e)) return false;
      }
    }
    return true;
  }

  // 3x3 conv (pad 1, stride 1) on an NHWC bf16 image through the implicit-GEMM kernel
  bool conv3(const bf16* in, int N, int H, int Wd, int C, const std::string& wname, const float* bias, const float* res1,
             const float* res2, float* of, bf16* ob, bool relu_bf16) {
    const Tensor* wt = W(wname);
    if (!wt) return false;
    const int n_out = (int)wt->shape[0];
    Epi e;
    e.bias = bias; e.res1 = res1; e.res2 = res2;
    e.out_f32 = of; e.out_bf16 = ob;
    e.sx = n_out; e.sy = (long long)Wd * n_out; e.sn = (long long)H * Wd * n_out;
    e.r2x = e.sx; e.r2y = e.sy; e.r2n = e.sn;
    e.flags = relu_bf16 ? GEMM_RELU_BF16 : 0;
    return gemm(image_view(in, N, H, Wd, C), wname, 9, e);
  }

  bool dpt_head(const std::string& hp, int Bp, int out_ch, bool sigmoid, const float* tscale, const float* tshift,
                bf16* dpt_out, float* out_final, float* out_raw, int lane_base);
  bool unet_branch(const void* src, const void* tgt, int in_dtype, int in_layout, int Bp);
  int forward(const void* src, const void* tgt, int pairs, int in_dtype, int in_layout, float* flow_out, float* covis_out,
              cudaStream_t stream);
};

// UC/models/prediction_heads/dpt.py:180-232 (DPTFeature.forward) + :285-311 (DPTRegressionProcessor.forward) + adaptor.
bool ufm_engine::dpt_head(const std::string& hp, int Bp, int out_ch, bool sigmoid, const float* tscale,
                          const float* tshift, bf16* dpt_out, float* out_final, float* out_raw, int lane_base) {
  const ufm_config& c = cfg;
  const int hpn = c.img_h / c.patch, wpn = c.img_w / c.patch, n = hpn * wpn;
  const int D = c.enc_dim, d = c.info_dim, fd = c.dpt_feature_dim, hid = c.dpt_hidden;
  const int* ld = c.dpt_layer_dims;
  const int h3 = (hpn - 1) / 2 + 1, w3 = (wpn - 1) / 2 + 1;
  // DPTFeature.hooks (dpt.py:198-205): input stage k reads entry hooks[k] of [enc_last, inter0, inter1, final]
  const bf16* feats[4] = {B<bf16>("enc_last"), B<bf16>("inter0"), B<bf16>("inter1"), B<bf16>("final")};
  const int fdim[4] = {D, d, d, d};
  const int* hooks = hp == "hu" ? c.unc_hooks : c.dpt_hooks;
  int hk[4] = {0, 1, 2, 3};
  if (c.hooks_set)
    for (int k = 0; k < 4; ++k) hk[k] = hooks[k];
  const bf16 *F0 = feats[hk[0]], *F1 = feats[hk[1]], *F2 = feats[hk[2]], *F3 = feats[hk[3]];
  const int C0 = fdim[hk[0]], C1 = fdim[hk[1]], C2 = fdim[hk[2]], C3 = fdim[hk[3]];
  auto tok_view = [&](const bf16* p, int C) { return AView{p, C, n, 1, Bp, C, (long long)n * C, 2LL * n * C, 128, 1}; };
  bf16 *G = B<bf16>("G"), *G1 = B<bf16>("G1");
  bf16 *T0 = B<bf16>("T0"), *T1 = B<bf16>("T1"), *T2 = B<bf16>("T2"), *T3 = B<bf16>("T3"), *T3c = B<bf16>("T3col"),
       *T3b = B<bf16>("T3b");
  float* Lf[4] = {B<float>("L0f"), B<float>("L1f"), B<float>("L2f"), B<float>("L3f")};
  bf16* Lr[4] = {B<bf16>("L0r"), B<bf16>("L1r"), B<bf16>("L2r"), B<bf16>("L3r")};
  const int LH[4] = {4 * hpn, 2 * hpn, hpn, h3}, LW[4] = {4 * wpn, 2 * wpn, wpn, w3};

  // The four input_process chains are independent: in a branching pass (launch-bound batches) chains 1-3 run on the
  // lanes lane_base .. lane_base + 2 and join in front of the fusion blocks.
  const cudaEvent_t chains_fork = fork_point();
  cudaStream_t home = st;
  cudaEvent_t chain_done[3] = {nullptr, nullptr, nullptr};
  // input_process.0: 1x1 (D -> ld0) folded with ConvT 4x4/s4 into one GEMM, pixel shuffle, 3x3 -> fd (dpt.py:107-125)
  {
    Epi e;
    e.bias = Wf(hp + ".in0.b"); e.out_bf16 = G; e.sx = 16LL * ld[0]; e.sn = (long long)n * e.sx;
    flops_override = 2.0 * Bp * n * ((double)C0 * ld[0] + 16.0 * ld[0] * ld[0]);  // reference: 1x1 then ConvT 4x4/s4
    if (!gemm(tok_view(F0, C0), hp + ".in0.w", 1, e)) return false;
    if (!run(UFM_CAT_MOVE, 0.0, 4.0 * Bp * n * 16 * ld[0], "shuffle0",
             [&] { return launch_shuffle_bf16(G, T0, Bp, hpn, wpn, 4, ld[0], st); }))
      return false;
    if (!conv3(T0, Bp, LH[0], LW[0], ld[0], hp + ".rn0.w", nullptr, nullptr, nullptr, Lf[0], Lr[0], true)) return false;
  }
  // input_process.1: 1x1 + ConvT 2x2/s2 (dpt.py:127-145); its GEMM result has a buffer of its own (G1) so that it can
  // run beside chain 0
  {
    if (!lane_begin(lane_base, chains_fork, &home)) return false;
    Epi e;
    e.bias = Wf(hp + ".in1.b"); e.out_bf16 = G1; e.sx = 4LL * ld[1]; e.sn = (long long)n * e.sx;
    flops_override = 2.0 * Bp * n * ((double)C1 * ld[1] + 4.0 * ld[1] * ld[1]);  // reference: 1x1 then ConvT 2x2/s2
    if (!gemm(tok_view(F1, C1), hp + ".in1.w", 1, e)) return false;
    if (!run(UFM_CAT_MOVE, 0.0, 4.0 * Bp * n * 4 * ld[1], "shuffle1",
             [&] { return launch_shuffle_bf16(G1, T1, Bp, hpn, wpn, 2, ld[1], st); }))
      return false;
    if (!conv3(T1, Bp, LH[1], LW[1], ld[1], hp + ".rn1.w", nullptr, nullptr, nullptr, Lf[1], Lr[1], true)) return false;
    if (!lane_end(home, &chain_done[0])) return false;
  }
  // input_process.2: 1x1 (dpt.py:147-155)
  {
    if (!lane_begin(lane_base + 1, chains_fork, &home)) return false;
    Epi e;
    e.bias = Wf(hp + ".in2.b"); e.out_bf16 = T2; e.sx = ld[2]; e.sn = (long long)n * e.sx;
    if (!gemm(tok_view(F2, C2), hp + ".in2.w", 1, e)) return false;
    if (!conv3(T2, Bp, LH[2], LW[2], ld[2], hp + ".rn2.w", nullptr, nullptr, nullptr, Lf[2], Lr[2], true)) return false;
    if (!lane_end(home, &chain_done[1])) return false;
  }
  // input_process.3: 1x1, 3x3 stride 2 (as im2col + GEMM) (dpt.py:157-172)
  {
    if (!lane_begin(lane_base + 2, chains_fork, &home)) return false;
    Epi e;
    e.bias = Wf(hp + ".in3a.b"); e.out_bf16 = T3; e.sx = ld[3]; e.sn = (long long)n * e.sx;
    if (!gemm(tok_view(F3, C3), hp + ".in3a.w", 1, e)) return false;
    if (!run(UFM_CAT_MOVE, 0.0, 2.0 * Bp * ld[3] * (n + 9.0 * h3 * w3), "im2col_s2",
             [&] { return launch_im2col_s2(T3, T3c, Bp, hpn, wpn, ld[3], h3, w3, st); }))
      return false;
    Epi e2;
    e2.bias = Wf(hp + ".in3b.b"); e2.out_bf16 = T3b; e2.sx = ld[3];
    if (!gemm(rows_view(T3c, (long long)Bp * h3 * w3, 9 * ld[3]), hp + ".in3b.w", 1, e2)) return false;
    if (!conv3(T3b, Bp, LH[3], LW[3], ld[3], hp + ".rn3.w", nullptr, nullptr, nullptr, Lf[3], Lr[3], true)) return false;
    if (!lane_end(home, &chain_done[2])) return false;
  }
  for (cudaEvent_t ev : chain_done)
    if (!lane_join(ev)) return false;
  // fusion blocks (dpt_block.py:225-255): refinenet4 (no resConfUnit1) ... refinenet1
  bf16* bufA = B<bf16>("bufA");
  bf16* bufO = B<bf16>("bufO");
  float* sF = B<float>("sF");
  bf16* sR = B<bf16>("sR");
  float* low = B<float>("low");
  float* path = B<float>("path");
  for (int r = 4; r >= 1; --r) {
    const int k = r - 1;  // skip level
    const int H = LH[k], Wd = LW[k];
    const std::string rp = hp + ".ref" + std::to_string(r);
    const float* sumF;
    const bf16* sumR;
    if (r == 4) {
      sumF = Lf[3];
      sumR = Lr[3];
    } else {
      // output = path + resConfUnit1(skip):  skip + conv2(relu(conv1(relu(skip)))) + path
      if (!conv3(Lr[k], Bp, H, Wd, fd, rp + ".rcu1.c1.w", Wf(rp + ".rcu1.c1.b"), nullptr, nullptr, nullptr, bufA, true))
        return false;
      if (!conv3(bufA, Bp, H, Wd, fd, rp + ".rcu1.c2.w", Wf(rp + ".rcu1.c2.b"), Lf[k], path, sF, sR, true)) return false;
      sumF = sF;
      sumR = sR;
    }
    // resConfUnit2
    if (!conv3(sumR, Bp, H, Wd, fd, rp + ".rcu2.c1.w", Wf(rp + ".rcu2.c1.b"), nullptr, nullptr, nullptr, bufA, true))
      return false;
    if (!conv3(bufA, Bp, H, Wd, fd, rp + ".rcu2.c2.w", Wf(rp + ".rcu2.c2.b"), sumF, nullptr, nullptr, bufO, false))
      return false;
    // out_conv (1x1) commuted in front of the x2 bilinear upsample (both linear; interpolation weights sum to 1)
    {
      Epi e;
      e.bias = Wf(rp + ".out.b"); e.out_f32 = low; e.sx = fd;
      if (!gemm(rows_view(bufO, (long long)Bp * H * Wd, fd), rp + ".out.w", 1, e)) return false;
    }
    int oh = 2 * H, ow = 2 * Wd;
    if (r == 4) {  // crop to the next skip's shape (dpt.py:213)
      oh = LH[2] < oh ? LH[2] : oh;
      ow = LW[2] < ow ? LW[2] : ow;
    }
    const double up_bytes = (double)Bp * fd * (4.0 * H * Wd + (r > 1 ? 4.0 : 2.0) * oh * ow);
    if (!run(UFM_CAT_RESIZE, 0.0, up_bytes, "upsample", [&] {
          return launch_resize_ac(low, nullptr, Bp, H, Wd, fd, 2 * H, 2 * Wd, oh, ow, r > 1 ? path : nullptr,
                                  r > 1 ? nullptr : dpt_out, st);
        }))
      return false;
  }
  // processor (dpt.py:285-311): conv3x3 fd->hid @8x, bilinear to the image, conv3x3 + ReLU + 1x1 + adaptor (tail epilogue)
  const int H8 = 8 * hpn, W8 = 8 * wpn;
  bf16* c1 = B<bf16>("c1");
  bf16* rs = B<bf16>("rs");
  if (!conv3(dpt_out, Bp, H8, W8, fd, hp + ".proc.c1.w", Wf(hp + ".proc.c1.b"), nullptr, nullptr, nullptr, c1, false))
    return false;
  if (!run(UFM_CAT_RESIZE, 0.0, 2.0 * Bp * hid * ((double)H8 * W8 + (double)c.img_h * c.img_w), "resize", [&] {
        return launch_resize_ac(nullptr, c1, Bp, H8, W8, hid, c.img_h, c.img_w, c.img_h, c.img_w, nullptr, rs, st);
      }))
    return false;
  {
    Epi e;
    e.bias = Wf(hp + ".proc.c2.b");
    e.flags = GEMM_TAIL | (sigmoid ? GEMM_TAIL_SIGMOID : 0) | ((hp == "h1" && cfg.flow_coord) ? GEMM_TAIL_COORD : 0);
    e.w2 = Wf(hp + ".proc.out.w");
    const Tensor* b2 = W(hp + ".proc.out.b");
    if (!e.w2 || !b2) return false;
    if ((int)b2->host.size() < out_ch) {
      fail("%s.proc.out.b has %d entries, need %d", hp.c_str(), (int)b2->host.size(), out_ch);
      return false;
    }
    for (int i = 0; i < out_ch; ++i) {
      e.b2[i] = b2->host[i];
      e.tscale[i] = tscale ? tscale[i] : 1.f;
      e.tshift[i] = tshift ? tshift[i] : 0.f;
    }
    e.tail_out = out_ch;
    e.out_f32 = out_final;
    e.out2 = out_raw;
    if (!gemm(image_view(rs, Bp, c.img_h, c.img_w, hid), hp + ".proc.c2.w", 9, e)) return false;
  }
  return true;
}

// UNet fine-feature branch (uniflowmatch/models/unet_encoder.py:50-72 on both images, ufm.py:916-918) and the combine
// with the classification features (ufm.py:967-982).  Images run in (b, view) order, NHWC bf16; every cat(skip, x) is a
// concat buffer "ucat<l>" whose left half the encoder conv writes and whose right half the ConvTranspose scatter writes.
bool ufm_engine::unet_branch(const void* src, const void* tgt, int in_dtype, int in_layout, int Bp) {
  const ufm_config& c = cfg;
  const int N2 = 2 * Bp;
  static const int F[4] = {64, 128, 256, 512};
  int LH[5], LW[5];
  LH[0] = c.img_h; LW[0] = c.img_w;
  for (int l = 1; l < 5; ++l) { LH[l] = LH[l - 1] / 2; LW[l] = LW[l - 1] / 2; }
  bf16* cat[4] = {B<bf16>("ucat0"), B<bf16>("ucat1"), B<bf16>("ucat2"), B<bf16>("ucat3")};
  bf16 *T = B<bf16>("uT"), *P = B<bf16>("uP"), *X = B<bf16>("uX"), *G = B<bf16>("uG");
  // conv3x3 + bias + ReLU, dense NHWC input, output at a channel offset of a buffer with pixel pitch out_pitch
  auto conv = [&](const bf16* in, int l, int C, const std::string& wn, bf16* out, long long out_pitch, long long off) {
    const Tensor* wt = W(wn + ".w");
    if (!wt) return false;
    Epi e;
    e.bias = Wf(wn + ".b");
    if (!e.bias) return false;
    e.out_bf16 = out; e.sx = out_pitch; e.sy = LW[l] * out_pitch; e.sn = (long long)LH[l] * LW[l] * out_pitch;
    e.out_off = off;
    e.flags = GEMM_RELU_BF16;
    return gemm(image_view(in, N2, LH[l], LW[l], C), wn + ".w", 9, e);
  };
  {
    UnetStemParams sp;
    sp.src = src; sp.tgt = tgt; sp.B = Bp; sp.H = c.img_h; sp.W = c.img_w;
    sp.is_u8 = in_dtype == UFM_U8; sp.is_nhwc = in_layout == UFM_NHWC;
    for (int i = 0; i < 3; ++i) { sp.mean[i] = img_mean(i); sp.inv_std[i] = 1.0f / img_std(i); }
    sp.w = Wf("unet.stem.w"); sp.b = Wf("unet.stem.b"); sp.out = T;
    if (!sp.w || !sp.b) return false;
    const double px = (double)N2 * c.img_h * c.img_w;
    if (!run(UFM_CAT_MOVE, 2.0 * px * 64 * 27, px * (3.0 * (sp.is_u8 ? 1 : 4) + 128), "unet stem",
             [&] { return launch_unet_stem(sp, st); }))
      return false;
  }
  if (!conv(T, 0, 64, "unet.d0.c2", cat[0], 128, 0)) return false;
  for (int l = 1; l <= 4; ++l) {
    const int Cin = F[l - 1];
    const double pb = (double)N2 * LH[l] * LW[l] * Cin * 2;
    if (!run(UFM_CAT_MOVE, 0.0, 5.0 * pb, "unet pool",
             [&] { return launch_unet_pool(cat[l - 1], 2 * Cin, N2, LH[l - 1], LW[l - 1], Cin, P, st); }))
      return false;
    const std::string pre = l < 4 ? "unet.d" + std::to_string(l) : std::string("unet.bn");
    const int Cout = 2 * Cin;
    if (!conv(P, l, Cin, pre + ".c1", T, Cout, 0)) return false;
    if (l < 4) {
      if (!conv(T, l, Cout, pre + ".c2", cat[l], 2 * Cout, 0)) return false;
    } else {
      if (!conv(T, l, Cout, pre + ".c2", X, Cout, 0)) return false;
    }
  }
  for (int l = 3; l >= 0; --l) {
    const std::string pre = "unet.u" + std::to_string(l);
    const long long rows = (long long)N2 * LH[l + 1] * LW[l + 1];
    {
      Epi e;
      e.bias = Wf(pre + ".t.b");
      if (!e.bias) return false;
      e.out_bf16 = G; e.sx = 4LL * F[l];
      if (!gemm(rows_view(X, rows, 2 * F[l]), pre + ".t.w", 1, e)) return false;
    }
    const double ub = (double)N2 * LH[l] * LW[l] * F[l] * 2;
    if (!run(UFM_CAT_MOVE, 0.0, 2.0 * ub, "unet up", [&] {
          return launch_unet_up(G, N2, LH[l + 1], LW[l + 1], F[l], cat[l], 2 * F[l], F[l], LH[l], LW[l], st);
        }))
      return false;
    if (!conv(cat[l], l, 2 * F[l], pre + ".c1", T, F[l], 0)) return false;
    if (!conv(T, l, F[l], pre + ".c2", X, F[l], 0)) return false;
  }
  set_stage("unet_x0", X, UFM_BF16, {Bp, 2, c.img_h, c.img_w, 64});
  UnetCombineParams cp;
  cp.x0 = X; cp.desc_cls = B<bf16>("desc_cls"); cp.desc = B<bf16>("desc");
  cp.unet_feat = keep_aux ? B<float>("unet_feat") : nullptr;
  cp.n = N2; cp.H = c.img_h; cp.W = c.img_w; cp.method = c.unet;
  cp.wf = Wf("unet.final.w"); cp.bf = Wf("unet.final.b");
  cp.w1 = c.unet == 1 ? Wf("unet.conv1.w") : nullptr; cp.b1 = c.unet == 1 ? Wf("unet.conv1.b") : nullptr;
  cp.w2 = Wf("unet.conv2.w"); cp.b2 = Wf("unet.conv2.b");
  if (!cp.wf || !cp.bf || !cp.w2 || !cp.b2 || (c.unet == 1 && (!cp.w1 || !cp.b1))) return false;
  const double px = (double)N2 * c.img_h * c.img_w;
  const double fl = 2.0 * px * (64 * 16 + (c.unet == 1 ? 32 * 32 + 32 * 16 : 16 * 16));
  if (!run(UFM_CAT_MOVE, fl, px * (128 + 32 + 32 + (keep_aux ? 64 : 0)), "unet combine",
           [&] { return launch_unet_combine(cp, st); }))
    return false;
  if (keep_aux) set_stage("unet_feat", cp.unet_feat, UFM_F32, {Bp, 2, 16, c.img_h, c.img_w});
  return true;
}

int ufm_engine::forward(const void* src, const void* tgt, int pairs, int in_dtype, int in_layout, float* flow_out,
                        float* covis_out, cudaStream_t stream) {
  if (!order_after_previous(stream)) return 1;
  if (!set_pdl_early(pairs, stream)) return 1;
  if (nvtx) {
    nvtxRangePushA("ufm_forward");
    nvtx_depth = 1;
  }
  const int rc = forward_impl(src, tgt, pairs, in_dtype, in_layout, flow_out, covis_out, stream);
  for (; nvtx && nvtx_depth > 0; --nvtx_depth) nvtxRangePop();
  if (!mark_done(stream)) return 1;
  return rc;
}

int ufm_engine::forward_impl(const void* src, const void* tgt, int pairs, int in_dtype, int in_layout, float* flow_out,
                             float* covis_out, cudaStream_t stream) {
  if (!prepared) return fail("ufm_prepare was not called");
  if (pairs < 1 || pairs > max_pairs) return fail("pairs=%d outside [1, %d]", pairs, max_pairs);
  if (in_dtype != UFM_U8 && in_dtype != UFM_F32) return fail("input dtype must be u8 or f32");
  st = stream;
  launches = 0;
  branches = pairs <= kGraphMaxPairs && !profile && branch_mode();
  sync_used = 0;
  bsfx.clear();
  if (profile && prof.size() > (1u << 20)) return fail("profile records not drained: call ufm_profile_read");
  last_pairs = pairs;
  stage.clear();
  const ufm_config& c = cfg;
  const int Bp = pairs;
  const int R1 = 1 + c.enc_registers;  // rows in front of the patch tokens of every image: cls (+ register tokens)
  const int hpn = c.img_h / c.patch, wpn = c.img_w / c.patch, n = hpn * wpn, S = n + R1;
  const int D = c.enc_dim, d = c.info_dim;
  const int Kpad = (int)w["enc.patch.w"].shape[1];
  const long long Me = 2LL * Bp * S, Mi = 2LL * Bp * n;

  // ---- encoder (hub DINOv2 forward) ----
  nvtx_stage("encoder");
  {
    PatchParams pp;
    pp.src = src; pp.tgt = tgt; pp.B = Bp; pp.H = c.img_h; pp.W = c.img_w; pp.P = c.patch; pp.Kpad = Kpad;
    pp.is_u8 = in_dtype == UFM_U8; pp.is_nhwc = in_layout == UFM_NHWC;
    for (int i = 0; i < 3; ++i) {
      pp.mean[i] = img_mean(i);
      pp.inv_std[i] = 1.0f / img_std(i);
    }
    pp.out = B<bf16>("Apatch");
    const double px = 2.0 * Bp * c.img_h * c.img_w * 3;
    if (!run(UFM_CAT_MOVE, 0.0, px * (in_dtype == UFM_U8 ? 1 : 4) + 2.0 * 2 * Bp * n * Kpad, "patch_im2col",
             [&] { return launch_patch_im2col(pp, st); }))
      return 1;
  }
  float* x = B<float>("x");
  {
    Epi e;
    e.bias = Wf("enc.patch.b");
    const float* pos = Wf("enc.pos");
    if (!e.bias || !pos) return 1;
    e.res2 = pos + D; e.r2x = D; e.r2n = 0; e.r2y = 0;
    e.out_f32 = x; e.sx = D; e.sn = (long long)S * D; e.out_off = (long long)R1 * D;
    AView a{B<bf16>("Apatch"), Kpad, n, 1, 2 * Bp, Kpad, (long long)n * Kpad, (long long)n * Kpad, 128, 1};
    flops_override = 2.0 * 2 * Bp * n * (double)D * 3 * c.patch * c.patch;  // K = 588, not the padded 640
    if (!gemm(a, "enc.patch.w", 1, e)) return 1;
    if (!run(UFM_CAT_MOVE, 0.0, 2.0 * Bp * R1 * D * 4, "cls_rows",
             [&] { return launch_cls_rows(x, Wf("enc.cls"), R1, 2 * Bp, S, D, st); }))
      return 1;
  }
  bf16* enc_last = B<bf16>("enc_last");
  const bool enc_raw = c.enc_norm_intermediate == 0, info_raw = c.info_norm_intermediate == 0;
  for (int i = 0; i < c.enc_depth; ++i) {
    if (!block("enc." + std::to_string(i), x, 2 * Bp, S, D, c.enc_heads, 4 * D, false)) return 1;
    if (c.refine && i == c.enc_take_first) {
      // final norm, drop cls, (view,b) -> (b,view) order, into the left D columns of the classification-head input
      if (!layernorm(x, 2 * Bp, S, R1, n, D, "enc.norm", Bp, B<bf16>("Xcls"), c.cls_in, 0, nullptr, 0, 0, enc_raw)) return 1;
      set_stage("enc_first", B<bf16>("Xcls"), UFM_BF16, {Bp, 2, n, c.cls_in});
    }
    if (i == c.enc_take_last) {
      if (!layernorm(x, 2 * Bp, S, R1, n, D, "enc.norm", Bp, enc_last, D, 0, nullptr, 0, 0, enc_raw)) return 1;
      set_stage("enc_last", enc_last, UFM_BF16, {Bp, 2, n, D});
    }
  }
  // ---- info sharing ----
  nvtx_stage("info_sharing");
  float* y = B<float>("y");
  const bool cross = c.info_alternating == 2;
  {
    Epi e;
    e.bias = Wf("info.proj.b");
    if (!e.bias) return 1;
    if (!cross) {
      const Tensor* vt = W("info.viewpos");
      if (!vt) return 1;
      if (view_index < 0 || view_index >= (int)vt->shape[0]) return fail("view index %d outside the view_pos_table", view_index);
      e.res2 = reinterpret_cast<const float*>(vt->ptr);
      e.r2y = (long long)view_index * d; e.r2x = 0; e.r2n = 0;  // view 0 -> row 0, view 1 -> row view_index
    }
    // residual stream layout: pair-major [B][2][n][d] (one attention sequence = one pair's 2n tokens), or VIEW-major
    // [2][B][n][d] for the per-view branches of cross-attention info sharing
    e.out_f32 = y; e.sx = d;
    e.sy = cross ? (long long)Bp * n * d : (long long)n * d;
    e.sn = cross ? (long long)n * d : 2LL * n * d;
    AView a{enc_last, D, n, 2, Bp, D, (long long)n * D, 2LL * n * D, 128, 1};
    if (!gemm(a, "info.proj.w", 1, e)) return 1;
  }
  // final LayerNorm of the stream -> a pair-major [B,2,n,*] feature buffer (the layout the heads read)
  auto info_norm = [&](bf16* dst, long long pitch, long long col, bool raw = false) {
    return cross ? layernorm(y, 2 * Bp, n, 0, n, d, "info.norm", Bp, dst, pitch, col, nullptr, 0, 0, raw)
                 : layernorm(y, 1, (int)Mi, 0, (int)Mi, d, "info.norm", 0, dst, pitch, col, nullptr, 0, 0, raw);
  };
  for (int i = 0; i < c.info_depth; ++i) {
    if (cross) {
      if (!cross_depth(i, y, Bp, n, d, c.info_heads, c.info_hidden, c.info_qk_norm != 0)) return 1;
    } else {
      const bool per_view = c.info_alternating && (i & 1);
      if (!block("info." + std::to_string(i), y, per_view ? 2 * Bp : Bp, per_view ? n : 2 * n, d, c.info_heads,
                 c.info_hidden, c.info_qk_norm != 0, c.info_rope ? n : 0, wpn))
        return 1;
    }
    if (i == c.info_take0) {
      if (!info_norm(B<bf16>("inter0"), d, 0, info_raw)) return 1;
      set_stage("info_inter0", B<bf16>("inter0"), UFM_BF16, {Bp, 2, n, d});
    }
    if (i == c.info_take1) {
      if (!info_norm(B<bf16>("inter1"), d, 0, info_raw)) return 1;
      set_stage("info_inter1", B<bf16>("inter1"), UFM_BF16, {Bp, 2, n, d});
    }
  }
  if (!info_norm(B<bf16>("final"), d, 0)) return 1;
  set_stage("info_final", B<bf16>("final"), UFM_BF16, {Bp, 2, n, d});
  if (c.refine) {
    if (!info_norm(B<bf16>("Xcls"), c.cls_in, D)) return 1;
  }
  // ---- heads (view 1 only: "dual+single", ufm.py:180,700) ----
  // The flow head, the uncertainty head and (UFM-Refine) the classification head + UNet branch only read the encoder /
  // info-sharing features: a branching pass (launch-bound batches, see `branches`) issues the last two on side lanes and
  // joins them where their results are needed; otherwise they follow each other on the one stream, in this order.
  nvtx_stage("dpt_heads");
  const int H8 = 8 * hpn, W8 = 8 * wpn;
  float* flow_reg = c.refine ? B<float>("flow_reg") : flow_out;
  const cudaEvent_t heads_fork = fork_point();
  cudaStream_t main_st = st;
  cudaEvent_t unc_done = nullptr, cls_done = nullptr;

  auto flow_head = [&]() -> bool {
    if (!dpt_head("h1", Bp, c.flow_out_ch, false, c.flow_scale, c.flow_shift, B<bf16>("dpt1"), flow_reg,
                  B<float>("head1_raw"), 1))
      return false;
    set_stage("dpt1", B<bf16>("dpt1"), UFM_BF16, {Bp, H8, W8, c.dpt_feature_dim});
    set_stage("head1_raw", B<float>("head1_raw"), UFM_F32, {Bp, c.flow_out_ch, c.img_h, c.img_w});
    set_stage("flow_reg", flow_reg, UFM_F32, {Bp, 2, c.img_h, c.img_w});
    if (mark_outputs && !c.refine) cudaEventRecord(ev_flow, st);
    return true;
  };

  auto unc_head = [&]() -> bool {
    float* cov = covis_out ? covis_out : B<float>("covis");
    bf16* dptu = B<bf16>("dptu");
    float* headu_raw = B<float>("headu_raw");
    if (!lane_begin(4, heads_fork, &main_st)) return false;
    if (branches) bsfx = "@u";  // the side-lane head works on its own copy of the head workspace
    bool ok = true;
    if (c.unc_n_adaptors == 0) {
      // the single non_occluded_mask MaskAdaptor: sigmoid fused into the tail epilogue
      ok = dpt_head("hu", Bp, c.unc_out_ch, true, nullptr, nullptr, dptu, cov, headu_raw, 5);
      bsfx.clear();
      if (!ok) return false;
    } else {
      // several adaptors: the tail writes the raw channels, one elementwise kernel applies AdaptorMap (adaptors.py)
      ok = dpt_head("hu", Bp, c.unc_out_ch, false, nullptr, nullptr, dptu, headu_raw, nullptr, 5);
      bsfx.clear();
      if (!ok) return false;
      UncAdaptParams up;
      up.raw = headu_raw; up.B = Bp; up.C = c.unc_out_ch; up.H = c.img_h; up.W = c.img_w;
      up.n = c.unc_n_adaptors;
      int ch = 0;
      bool has_cov = false, has_conf = false;
      for (int a = 0; a < up.n; ++a) {
        up.kind[a] = c.unc_kind[a]; up.ch[a] = ch; up.vmin[a] = c.unc_vmin[a]; up.vmax[a] = c.unc_vmax[a];
        ch += c.unc_kind[a] == UFM_ADAPT_COV_EXP_TANH ? 3 : 1;
        has_cov = has_cov || c.unc_kind[a] == UFM_ADAPT_COV_EXP_TANH;
        has_conf = has_conf || c.unc_kind[a] >= UFM_ADAPT_CONF_EXP;
      }
      up.mask = cov;
      up.cov = has_cov ? B<float>("flow_cov") : nullptr;
      up.cov_inv = has_cov ? B<float>("flow_cov_inv") : nullptr;
      up.cov_log_det = has_cov ? B<float>("flow_cov_log_det") : nullptr;
      up.conf = has_conf ? B<float>("kp_conf") : nullptr;
      up.conf_stats = has_conf ? B<float>("conf_stats") : nullptr;
      const double plane = (double)c.img_h * c.img_w;
      if (!run(UFM_CAT_MOVE, 0.0, Bp * plane * 4.0 * (c.unc_out_ch + 1 + (has_cov ? 7 : 0) + (has_conf ? 1 : 0)),
               "unc adaptors", [&] { return launch_unc_adaptors(up, st); }))
        return false;
      if (has_cov) {
        set_stage("flow_cov", up.cov, UFM_F32, {Bp, 3, c.img_h, c.img_w});
        set_stage("flow_cov_inv", up.cov_inv, UFM_F32, {Bp, 3, c.img_h, c.img_w});
        set_stage("flow_cov_log_det", up.cov_log_det, UFM_F32, {Bp, 1, c.img_h, c.img_w});
      }
      if (has_conf) set_stage("keypoint_confidence", up.conf, UFM_F32, {Bp, c.img_h, c.img_w});
    }
    set_stage("dptu", dptu, UFM_BF16, {Bp, H8, W8, c.dpt_feature_dim});
    set_stage("headu_raw", headu_raw, UFM_F32, {Bp, c.unc_out_ch, c.img_h, c.img_w});
    if (mark_outputs) cudaEventRecord(ev_cov, st);
    return lane_end(main_st, &unc_done);
  };

  // Refine: classification head (ufm.py:954-966) and UNet branch (:916-918, :967-982) -> descriptors
  auto cls_head = [&]() -> bool {
    if (!lane_begin(0, heads_fork, &main_st)) return false;
    const long long Mc = Mi;
    {
      Epi e;
      e.bias = Wf("cls.fc1.b"); e.flags = GEMM_GELU; e.out_bf16 = B<bf16>("Hcls"); e.sx = c.cls_hidden;
      if (!gemm(rows_view(B<bf16>("Xcls"), Mc, c.cls_in), "cls.fc1.w", 1, e)) return false;
    }
    {
      Epi e;
      e.bias = Wf("cls.fc2.b"); e.out_bf16 = B<bf16>("Ycls"); e.sx = c.cls_in;
      if (!gemm(rows_view(B<bf16>("Hcls"), Mc, c.cls_hidden), "cls.fc2.w", 1, e)) return false;
    }
    const int zc = c.cls_patch * c.cls_patch * c.cls_out;
    {
      Epi e;
      e.bias = Wf("cls.lin.b"); e.out_bf16 = B<bf16>("Z"); e.sx = zc;
      if (!gemm(rows_view(B<bf16>("Ycls"), Mc, c.cls_in), "cls.lin.w", 1, e)) return false;
    }
    bf16* desc_dst = c.unet ? B<bf16>("desc_cls") : B<bf16>("desc");
    if (!run(UFM_CAT_MOVE, 0.0, 4.0 * Mi * zc, "shuffle desc", [&] {
          return launch_desc_shuffle(B<bf16>("Z"), desc_dst, 2 * Bp, hpn, wpn, c.cls_patch, c.cls_out, st);
        }))
      return false;
    if (c.unet) {
      if (!unet_branch(src, tgt, in_dtype, in_layout, Bp)) return false;
      set_stage("desc_cls", desc_dst, UFM_BF16, {Bp, 2, c.cls_out / 8, c.img_h, c.img_w, 8});
    }
    set_stage("desc", B<bf16>("desc"), UFM_BF16, {Bp, 2, c.cls_out / 8, c.img_h, c.img_w, 8});
    return lane_end(main_st, &cls_done);
  };

  if (branches) {  // side lanes first, so that their first launches are queued while the flow head is being issued
    if (c.unc_out_ch > 0 && !unc_head()) return 1;
    if (c.refine) {
      nvtx_stage("classification_head");
      if (!cls_head()) return 1;
    }
    if (!flow_head()) return 1;
  } else {
    if (!flow_head()) return 1;
    if (c.unc_out_ch > 0 && !unc_head()) return 1;
    if (c.refine) {
      nvtx_stage("classification_head_and_refinement");
      if (!cls_head()) return 1;
    }
  }
  // ---- Refine: local refinement (ufm.py:983-991, 1041-1178) ----
  if (c.refine) {
    if (!lane_join(cls_done)) return 1;
    RefineParams rp;
    rp.desc = B<bf16>("desc"); rp.B = Bp; rp.H = c.img_h; rp.W = c.img_w;
    rp.flow_in = flow_reg;
    rp.bias25 = Wf("cls.bias");
    if (!rp.bias25) return 1;
    rp.inv_temperature = 1.0f / c.temperature;
    rp.flow_out = flow_out;
    rp.residual = keep_aux ? B<float>("residual") : nullptr;
    rp.log_softmax = keep_aux ? B<float>("log_softmax") : nullptr;
    {
      // algorithmic bytes: F1, F2 (16 ch bf16) + flow in + flow out (+ residual, log-softmax when kept)
      const double plane = (double)c.img_h * c.img_w;
      const double rb = Bp * plane * (2 * 16 * 2 + 2 * 4 + 2 * 4 + (keep_aux ? 2 * 4 + 25 * 4 : 0));
      if (!run(UFM_CAT_REFINE, 0.0, rb, "refine", [&] { return launch_refine(rp, st); })) return 1;
    }
    if (keep_aux) {
      set_stage("residual", B<float>("residual"), UFM_F32, {Bp, 2, c.img_h, c.img_w});
      set_stage("log_softmax", B<float>("log_softmax"), UFM_F32, {Bp, c.img_h, c.img_w, 25});
    }
    if (mark_outputs) cudaEventRecord(ev_flow, st);
  }
  if (!lane_join(unc_done)) return 1;
  return 0;
}


// ================================================================================================ C ABI
extern "C" {

int ufm_create(const ufm_config* cfg, int device, ufm_engine** out) {
  if (!cfg || !out) return 1;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device >= ndev) {
    fprintf(stderr, "ufm_create: no CUDA device %d (this engine has no CPU fallback)\n", device);
    return 2;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10) {
    fprintf(stderr, "ufm_create: device %d is not sm_100 (compute capability %d.%d)\n", device, prop.major, prop.minor);
    return 3;
  }
  ufm_engine* e = new ufm_engine();
  e->cfg = *cfg;
  e->device = device;
  e->num_sms = prop.multiProcessorCount;
  DevGuard dev_guard(device);
  const ufm_config& c = e->cfg;
  if (c.img_h % c.patch || c.img_w % c.patch) {
    delete e;
    fprintf(stderr, "ufm_create: Input shape must be divisible by patch size\n");
    return 4;
  }
  if (c.enc_dim % 128 || c.info_dim % 128 || c.enc_dim / c.enc_heads != 64 || c.info_dim / c.info_heads != 64) {
    delete e;
    fprintf(stderr, "ufm_create: dims must be multiples of 128 with head_dim 64\n");
    return 5;
  }
  {
    int ch = 0, n_mask = 0, n_cov = 0, n_conf = 0;
    bool bad = c.unc_n_adaptors < 0 || c.unc_n_adaptors > 4 || c.unc_out_ch < 0 || c.unc_out_ch > kTailMax;
    for (int a = 0; !bad && a < c.unc_n_adaptors; ++a) {
      const int k = c.unc_kind[a];
      bad = k < UFM_ADAPT_MASK || k > UFM_ADAPT_CONF_SOFTMAX;
      ch += k == UFM_ADAPT_COV_EXP_TANH ? 3 : 1;
      n_mask += k == UFM_ADAPT_MASK;
      n_cov += k == UFM_ADAPT_COV_EXP_TANH;
      n_conf += k >= UFM_ADAPT_CONF_EXP;
    }
    bad = bad || n_mask > 1 || n_cov > 1 || n_conf > 1 || (c.unc_n_adaptors > 0 && ch != c.unc_out_ch) ||
          (c.unc_n_adaptors == 0 && c.unc_out_ch > 1);
    if (bad) {
      delete e;
      fprintf(stderr, "ufm_create: uncertainty adaptors must be at most one mask, one covariance and one confidence adaptor "
                      "whose channels add up to unc_out_ch\n");
      return 7;
    }
  }
  if (c.refine && (c.cls_out != 16 || c.refine_P != 5)) {
    delete e;
    fprintf(stderr, "ufm_create: refinement kernel is specialised for 16-channel descriptors and a 5x5 window\n");
    return 6;
  }
  if (c.info_rope && c.info_alternating != 0) {
    delete e;
    fprintf(stderr, "ufm_create: info_rope is defined for global-attention info sharing only\n");
    return 9;
  }
  if (c.unet && (!c.refine || c.unet < 0 || c.unet > 2 || c.img_h < 16 || c.img_w < 16)) {
    delete e;
    fprintf(stderr, "ufm_create: unet must be 0, 1 (conv) or 2 (modulate), on a Refine config with an image of at least 16x16\n");
    return 8;
  }
  *out = e;
  return 0;
}

void ufm_destroy(ufm_engine* e) {
  if (!e) return;
  DevGuard dev_guard(e->device);
  for (auto ev : e->ev_pool) cudaEventDestroy(ev);
  for (auto& kv : e->graphs) cudaGraphExecDestroy(kv.second.exec);
  if (e->graph_st) cudaStreamDestroy(e->graph_st);
  if (e->ev_done) cudaEventDestroy(e->ev_done);
  for (auto ev : e->sync_ev) cudaEventDestroy(ev);
  for (auto ls : e->lane_st)
    if (ls) cudaStreamDestroy(ls);
  if (e->copy_st) {
    cudaStreamDestroy(e->copy_st);
    cudaEventDestroy(e->ev_flow);
    cudaEventDestroy(e->ev_cov);
  }
  for (auto& kv : e->w) cudaFree(kv.second.ptr);
  for (auto& kv : e->buf) cudaFree(kv.second.ptr);
  delete e;
}

const char* ufm_last_error(ufm_engine* e) { return e ? e->err.c_str() : "null engine"; }

int ufm_set_weight(ufm_engine* e, const char* name, const void* ptr, int dtype, const int64_t* shape, int ndim) {
  if (!e) return 1;
  DevGuard dev_guard(e->device);
  Tensor t;
  t.dtype = dtype;
  t.shape.assign(shape, shape + ndim);
  auto it = e->w.find(name);
  if (it != e->w.end()) {
    cudaFree(it->second.ptr);
    e->w.erase(it);
  }
  if (!e->ck(cudaMalloc(&t.ptr, (size_t)t.bytes()), "cudaMalloc weight")) return 1;
  if (!e->ck(cudaMemcpy(t.ptr, ptr, (size_t)t.bytes(), cudaMemcpyDefault), "copy weight")) return 1;
  if (dtype == UFM_F32 && t.numel() <= 64) {
    t.host.resize((size_t)t.numel());
    if (!e->ck(cudaMemcpy(t.host.data(), t.ptr, (size_t)t.bytes(), cudaMemcpyDeviceToHost), "copy small weight")) return 1;
  }
  e->w[name] = t;
  return 0;
}

int ufm_prepare(ufm_engine* e, int max_pairs) {
  if (!e) return 1;
  DevGuard dev_guard(e->device);
  for (auto& kv : e->buf) cudaFree(kv.second.ptr);
  e->buf.clear();
  e->prepared = false;
  const ufm_config& c = e->cfg;
  const long long Bp = max_pairs;
  const long long hpn = c.img_h / c.patch, wpn = c.img_w / c.patch, n = hpn * wpn, S = n + 1 + c.enc_registers;
  const long long D = c.enc_dim, d = c.info_dim, fd = c.dpt_feature_dim, hid = c.dpt_hidden;
  const long long Me = 2 * Bp * S, Mi = 2 * Bp * n;
  if (!e->W("enc.patch.w")) return 1;
  const long long Kpad = e->w["enc.patch.w"].shape[1];
  auto mx = [](long long a, long long b) { return a > b ? a : b; };
  bool ok = true;
  ok = ok && e->alloc("Apatch", UFM_BF16, 2 * Bp * n * Kpad);
  ok = ok && e->alloc("x", UFM_F32, Me * D);
  ok = ok && e->alloc("y", UFM_F32, Mi * d);
  ok = ok && e->alloc("xn", UFM_BF16, mx(Me * D, Mi * d));
  ok = ok && e->alloc("qkv", UFM_BF16, mx(Me * 3 * D, Mi * 3 * d));
  if (c.info_alternating == 2) ok = ok && e->alloc("cqkv", UFM_BF16, Mi * 3 * d);
  ok = ok && e->alloc("att", UFM_BF16, mx(Me * D, Mi * d));
  ok = ok && e->alloc("h", UFM_BF16, mx(Me * 4 * D, Mi * c.info_hidden));
  ok = ok && e->alloc("enc_last", UFM_BF16, Mi * D);
  ok = ok && e->alloc("inter0", UFM_BF16, Mi * d);
  ok = ok && e->alloc("inter1", UFM_BF16, Mi * d);
  ok = ok && e->alloc("final", UFM_BF16, Mi * d);
  const int* ld = c.dpt_layer_dims;
  const long long h3 = (hpn - 1) / 2 + 1, w3 = (wpn - 1) / 2 + 1;
  ok = ok && e->alloc("G", UFM_BF16, Bp * n * mx(16 * ld[0], 4 * ld[1]));
  ok = ok && e->alloc("G1", UFM_BF16, Bp * n * 4 * ld[1]);
  ok = ok && e->alloc("T0", UFM_BF16, Bp * 16 * n * ld[0]);
  ok = ok && e->alloc("T1", UFM_BF16, Bp * 4 * n * ld[1]);
  ok = ok && e->alloc("T2", UFM_BF16, Bp * n * ld[2]);
  ok = ok && e->alloc("T3", UFM_BF16, Bp * n * ld[3]);
  ok = ok && e->alloc("T3col", UFM_BF16, Bp * h3 * w3 * 9 * ld[3]);
  ok = ok && e->alloc("T3b", UFM_BF16, Bp * h3 * w3 * ld[3]);
  const long long lv[4] = {16 * n, 4 * n, n, h3 * w3};
  const char* lfn[4] = {"L0f", "L1f", "L2f", "L3f"};
  const char* lrn[4] = {"L0r", "L1r", "L2r", "L3r"};
  for (int k = 0; k < 4; ++k) {
    ok = ok && e->alloc(lfn[k], UFM_F32, Bp * lv[k] * fd);
    ok = ok && e->alloc(lrn[k], UFM_BF16, Bp * lv[k] * fd);
  }
  ok = ok && e->alloc("bufA", UFM_BF16, Bp * 16 * n * fd);
  ok = ok && e->alloc("bufO", UFM_BF16, Bp * 16 * n * fd);
  ok = ok && e->alloc("sF", UFM_F32, Bp * 16 * n * fd);
  ok = ok && e->alloc("sR", UFM_BF16, Bp * 16 * n * fd);
  ok = ok && e->alloc("low", UFM_F32, Bp * 16 * n * fd);
  ok = ok && e->alloc("path", UFM_F32, Bp * 16 * n * fd);
  ok = ok && e->alloc("dpt1", UFM_BF16, Bp * 64 * n * fd);
  ok = ok && e->alloc("c1", UFM_BF16, Bp * 64 * n * hid);
  ok = ok && e->alloc("rs", UFM_BF16, Bp * c.img_h * c.img_w * hid);
  ok = ok && e->alloc("head1_raw", UFM_F32, Bp * c.flow_out_ch * c.img_h * c.img_w);
  if (c.unc_out_ch > 0) {
    // second copy of the head workspace for the uncertainty head of a branching pass (it runs beside the flow head)
    const long long Bs = Bp < kGraphMaxPairs ? Bp : kGraphMaxPairs;
    ok = ok && e->alloc("G@u", UFM_BF16, Bs * n * mx(16 * ld[0], 4 * ld[1]));
    ok = ok && e->alloc("G1@u", UFM_BF16, Bs * n * 4 * ld[1]);
    ok = ok && e->alloc("T0@u", UFM_BF16, Bs * 16 * n * ld[0]);
    ok = ok && e->alloc("T1@u", UFM_BF16, Bs * 4 * n * ld[1]);
    ok = ok && e->alloc("T2@u", UFM_BF16, Bs * n * ld[2]);
    ok = ok && e->alloc("T3@u", UFM_BF16, Bs * n * ld[3]);
    ok = ok && e->alloc("T3col@u", UFM_BF16, Bs * h3 * w3 * 9 * ld[3]);
    ok = ok && e->alloc("T3b@u", UFM_BF16, Bs * h3 * w3 * ld[3]);
    for (int k = 0; k < 4; ++k) {
      ok = ok && e->alloc(std::string(lfn[k]) + "@u", UFM_F32, Bs * lv[k] * fd);
      ok = ok && e->alloc(std::string(lrn[k]) + "@u", UFM_BF16, Bs * lv[k] * fd);
    }
    for (const char* nm : {"bufA@u", "bufO@u", "sR@u"}) ok = ok && e->alloc(nm, UFM_BF16, Bs * 16 * n * fd);
    for (const char* nm : {"sF@u", "low@u", "path@u"}) ok = ok && e->alloc(nm, UFM_F32, Bs * 16 * n * fd);
    ok = ok && e->alloc("c1@u", UFM_BF16, Bs * 64 * n * hid);
    ok = ok && e->alloc("rs@u", UFM_BF16, Bs * c.img_h * c.img_w * hid);
    ok = ok && e->alloc("dptu", UFM_BF16, Bp * 64 * n * fd);
    ok = ok && e->alloc("headu_raw", UFM_F32, Bp * c.unc_out_ch * c.img_h * c.img_w);
    ok = ok && e->alloc("covis", UFM_F32, Bp * c.img_h * c.img_w);
    for (int a = 0; a < c.unc_n_adaptors; ++a) {
      if (c.unc_kind[a] == UFM_ADAPT_COV_EXP_TANH) {
        ok = ok && e->alloc("flow_cov", UFM_F32, Bp * 3 * c.img_h * c.img_w);
        ok = ok && e->alloc("flow_cov_inv", UFM_F32, Bp * 3 * c.img_h * c.img_w);
        ok = ok && e->alloc("flow_cov_log_det", UFM_F32, Bp * c.img_h * c.img_w);
      } else if (c.unc_kind[a] >= UFM_ADAPT_CONF_EXP) {
        ok = ok && e->alloc("kp_conf", UFM_F32, Bp * c.img_h * c.img_w);
        ok = ok && e->alloc("conf_stats", UFM_F32, Bp * 2);
      }
    }
  }
  if (c.refine) {
    ok = ok && e->alloc("flow_reg", UFM_F32, Bp * 2 * c.img_h * c.img_w);
    ok = ok && e->alloc("Xcls", UFM_BF16, Mi * c.cls_in);
    ok = ok && e->alloc("Hcls", UFM_BF16, Mi * c.cls_hidden);
    ok = ok && e->alloc("Ycls", UFM_BF16, Mi * c.cls_in);
    ok = ok && e->alloc("Z", UFM_BF16, Mi * c.cls_patch * c.cls_patch * c.cls_out);
    ok = ok && e->alloc("desc", UFM_BF16, 2 * Bp * c.img_h * c.img_w * c.cls_out);
    ok = ok && e->alloc("residual", UFM_F32, Bp * 2 * c.img_h * c.img_w);
    ok = ok && e->alloc("log_softmax", UFM_F32, Bp * c.img_h * c.img_w * 25);
    if (c.unet) {
      const long long N2 = 2 * Bp, HW = (long long)c.img_h * c.img_w;
      int lh = c.img_h, lw = c.img_w;
      for (int l = 0; l < 4; ++l) {
        ok = ok && e->alloc("ucat" + std::to_string(l), UFM_BF16, N2 * lh * lw * (128LL << l));
        lh /= 2; lw /= 2;
      }
      for (const char* nm : {"uT", "uX", "uG"}) ok = ok && e->alloc(nm, UFM_BF16, N2 * HW * 64);
      ok = ok && e->alloc("uP", UFM_BF16, N2 * (c.img_h / 2) * (c.img_w / 2) * 64);
      ok = ok && e->alloc("desc_cls", UFM_BF16, N2 * HW * c.cls_out);
      ok = ok && e->alloc("unet_feat", UFM_F32, N2 * HW * 16);
    }
  }
  // staging buffers for ufm_forward_host
  ok = ok && e->alloc("in_src", UFM_F32, Bp * 3 * c.img_h * c.img_w);
  ok = ok && e->alloc("in_tgt", UFM_F32, Bp * 3 * c.img_h * c.img_w);
  ok = ok && e->alloc("out_flow", UFM_F32, Bp * 2 * c.img_h * c.img_w);
  if (!ok) return 1;
  e->max_pairs = max_pairs;
  e->prepared = true;
  return 0;
}

int ufm_forward(ufm_engine* e, const void* src, const void* tgt, int pairs, int in_dtype, int in_layout,
                float* flow_out, float* covis_out, void* stream) {
  if (!e) return 1;
  DevGuard dev_guard(e->device);
  if (!flow_out) return e->fail("flow_out is NULL");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (e->prepared && pairs >= 1 && pairs <= kGraphMaxPairs && !e->profile && !getenv("UFM_NO_GRAPH") &&
      (in_dtype == UFM_U8 || in_dtype == UFM_F32) && cudaStreamIsCapturing(s, &cap) == cudaSuccess &&
      cap == cudaStreamCaptureStatusNone) {
    // small batches are launch-bound: device-to-device copies into / out of the engine's staging buffers around a replay
    // of the captured pass (2 x 0.7 MB in, 2.8 MB out per pair: microseconds, against ~330 launches and tensor-map encodes)
    const ufm_config& c = e->cfg;
    const size_t in_bytes = (size_t)pairs * 3 * c.img_h * c.img_w * (in_dtype == UFM_U8 ? 1 : 4);
    const size_t plane = (size_t)c.img_h * c.img_w;
    float* dcov = c.unc_out_ch > 0 ? reinterpret_cast<float*>(e->buf["covis"].ptr) : nullptr;
    if (!e->order_after_previous(s)) return 1;
    if (!e->ck(cudaMemcpyAsync(e->buf["in_src"].ptr, src, in_bytes, cudaMemcpyDeviceToDevice, s), "D2D src")) return 1;
    if (!e->ck(cudaMemcpyAsync(e->buf["in_tgt"].ptr, tgt, in_bytes, cudaMemcpyDeviceToDevice, s), "D2D tgt")) return 1;
    if (!e->graph_pass(pairs, in_dtype, in_layout, s)) return 1;
    if (!e->ck(cudaMemcpyAsync(flow_out, e->buf["out_flow"].ptr, pairs * 2 * plane * 4, cudaMemcpyDeviceToDevice, s), "D2D flow"))
      return 1;
    if (covis_out && dcov &&
        !e->ck(cudaMemcpyAsync(covis_out, dcov, pairs * plane * 4, cudaMemcpyDeviceToDevice, s), "D2D covis"))
      return 1;
    return e->mark_done(s) ? 0 : 1;
  }
  return e->forward(src, tgt, pairs, in_dtype, in_layout, flow_out, covis_out, s);
}

int ufm_forward_host(ufm_engine* e, const void* src, const void* tgt, int pairs, int in_dtype, int in_layout,
                     float* flow_out, float* covis_out) {
  if (!e) return 1;
  DevGuard dev_guard(e->device);
  if (!e->prepared) return e->fail("ufm_prepare was not called");
  if (pairs < 1 || pairs > e->max_pairs) return e->fail("pairs=%d outside [1, %d]", pairs, e->max_pairs);
  const ufm_config& c = e->cfg;
  const size_t px = (size_t)pairs * 3 * c.img_h * c.img_w;
  const size_t in_bytes = px * (in_dtype == UFM_U8 ? 1 : 4);
  cudaStream_t st = nullptr;
  void* dsrc = e->buf["in_src"].ptr;
  void* dtgt = e->buf["in_tgt"].ptr;
  float* dflow = reinterpret_cast<float*>(e->buf["out_flow"].ptr);
  float* dcov = c.unc_out_ch > 0 ? reinterpret_cast<float*>(e->buf["covis"].ptr) : nullptr;
  const size_t plane0 = (size_t)c.img_h * c.img_w;
  if (pairs <= kGraphMaxPairs && !e->profile && !getenv("UFM_NO_GRAPH")) {
    // ---- launch-bound regime: replay a captured pass ----
    if (!e->graph_st && !e->ck(cudaStreamCreateWithFlags(&e->graph_st, cudaStreamNonBlocking), "graph stream")) return 1;
    cudaStream_t gs = e->graph_st;
    if (!e->order_after_previous(gs)) return 1;  // the staging buffers and the workspace may still be in use elsewhere
    if (!e->ck(cudaMemcpyAsync(dsrc, src, in_bytes, cudaMemcpyHostToDevice, gs), "H2D src")) return 1;
    if (!e->ck(cudaMemcpyAsync(dtgt, tgt, in_bytes, cudaMemcpyHostToDevice, gs), "H2D tgt")) return 1;
    if (!e->graph_pass(pairs, in_dtype, in_layout, gs)) return 1;
    if (!e->ck(cudaMemcpyAsync(flow_out, dflow, pairs * 2 * plane0 * 4, cudaMemcpyDeviceToHost, gs), "D2H flow")) return 1;
    if (covis_out && dcov)
      if (!e->ck(cudaMemcpyAsync(covis_out, dcov, pairs * plane0 * 4, cudaMemcpyDeviceToHost, gs), "D2H covis")) return 1;
    return e->ck(cudaStreamSynchronize(gs), "sync") ? 0 : 1;
  }
  if (!e->order_after_previous(st)) return 1;
  if (!e->ck(cudaMemcpyAsync(dsrc, src, in_bytes, cudaMemcpyHostToDevice, st), "H2D src")) return 1;
  if (!e->ck(cudaMemcpyAsync(dtgt, tgt, in_bytes, cudaMemcpyHostToDevice, st), "H2D tgt")) return 1;
  if (!e->copy_st) {
    if (!e->ck(cudaStreamCreateWithFlags(&e->copy_st, cudaStreamNonBlocking), "copy stream")) return 1;
    if (!e->ck(cudaEventCreateWithFlags(&e->ev_flow, cudaEventDisableTiming), "event")) return 1;
    if (!e->ck(cudaEventCreateWithFlags(&e->ev_cov, cudaEventDisableTiming), "event")) return 1;
  }
  e->mark_outputs = true;
  const int rc = e->forward(dsrc, dtgt, pairs, in_dtype, in_layout, dflow, dcov, st);
  e->mark_outputs = false;
  if (rc) return 1;
  const size_t plane = (size_t)c.img_h * c.img_w;
  const bool want_cov = covis_out && dcov;
  // queue the copies in the order the outputs become final: Base = flow, then covisibility; Refine = the reverse
  for (int k = 0; k < 2; ++k) {
    const bool do_flow = (k == 0) != (c.refine != 0);
    if (do_flow) {
      if (!e->ck(cudaStreamWaitEvent(e->copy_st, e->ev_flow, 0), "wait flow")) return 1;
      if (!e->ck(cudaMemcpyAsync(flow_out, dflow, pairs * 2 * plane * 4, cudaMemcpyDeviceToHost, e->copy_st), "D2H flow")) return 1;
    } else if (want_cov) {
      if (!e->ck(cudaStreamWaitEvent(e->copy_st, e->ev_cov, 0), "wait covis")) return 1;
      if (!e->ck(cudaMemcpyAsync(covis_out, dcov, pairs * plane * 4, cudaMemcpyDeviceToHost, e->copy_st), "D2H covis")) return 1;
    }
  }
  if (!e->ck(cudaStreamSynchronize(e->copy_st), "sync copies")) return 1;
  if (!e->ck(cudaStreamSynchronize(st), "sync")) return 1;
  return 0;
}

int ufm_set_keep_aux(ufm_engine* e, int keep) {
  if (!e) return 1;
  e->keep_aux = keep != 0;
  return 0;
}

int ufm_set_view_index(ufm_engine* e, int row) {
  if (!e) return 1;
  if (row < 0) return e->fail("view index %d < 0", row);
  e->view_index = row;
  return 0;
}

int ufm_stage_desc(ufm_engine* e, const char* name, int* dtype, int64_t* shape, int* ndim) {
  if (!e) return 1;
  auto it = e->stage.find(name);
  if (it == e->stage.end()) return e->fail("no stage '%s' in the last forward", name);
  *dtype = it->second.dtype;
  *ndim = (int)it->second.shape.size();
  for (int i = 0; i < *ndim; ++i) shape[i] = it->second.shape[i];
  return 0;
}

int ufm_stage_copy(ufm_engine* e, const char* name, void* dst, int64_t bytes) {
  if (!e) return 1;
  DevGuard dev_guard(e->device);
  auto it = e->stage.find(name);
  if (it == e->stage.end()) return e->fail("no stage '%s' in the last forward", name);
  if (bytes != it->second.bytes()) return e->fail("stage '%s' is %lld bytes, caller gave %lld", name,
                                                  (long long)it->second.bytes(), (long long)bytes);
  if (!e->ck(cudaStreamSynchronize(e->st), "sync")) return 1;
  if (!e->ck(cudaMemcpy(dst, it->second.ptr, (size_t)bytes, cudaMemcpyDefault), "stage copy")) return 1;
  return 0;
}

int ufm_last_launch_count(ufm_engine* e) { return e ? e->launches : -1; }

int ufm_set_profile(ufm_engine* e, int on) {
  if (!e) return 1;
  e->profile = on != 0;
  e->prof.clear();
  e->ev_used = 0;
  return 0;
}

int ufm_profile_read(ufm_engine* e, double* ms, double* flops, double* bytes, int64_t* launches) {
  if (!e) return 1;
  DevGuard dev_guard(e->device);
  if (!e->ck(cudaStreamSynchronize(e->st), "sync")) return 1;
  for (int i = 0; i < UFM_NUM_CATS; ++i) {
    ms[i] = flops[i] = bytes[i] = 0.0;
    launches[i] = 0;
  }
  // UFM_PROFILE_DUMP=<path>: also append one "label,family,ms,flops,bytes" line per launch (tuning aid)
  FILE* dump = nullptr;
  if (const char* path = getenv("UFM_PROFILE_DUMP")) dump = fopen(path, "a");
  for (auto& r : e->prof) {
    float t = 0.f;
    if (!e->ck(cudaEventElapsedTime(&t, r.a, r.b), "event elapsed")) {
      if (dump) fclose(dump);
      return 1;
    }
    if (dump) fprintf(dump, "%s,%d,%.4f,%.6g,%.6g\n", r.what.c_str(), r.cat, t, r.flops, r.bytes);
    ms[r.cat] += t;
    flops[r.cat] += r.flops;
    bytes[r.cat] += r.bytes;
    launches[r.cat] += 1;
  }
  if (dump) fclose(dump);
  e->prof.clear();
  e->ev_used = 0;
  return 0;
}

int ufm_debug_flag(int which, int value) {
  if (which == 0) {
    set_gemm_epilogue_variant(value);
    return 0;
  }
  if (which == 1) {
    set_gemm_pair_mode(value);
    return 0;
  }
  if (which == 2) {
    set_attention_poly(value);
    return 0;
  }
  if (which == 3) {
    set_gemm_halo_mode(value);
    return 0;
  }
  if (which == 4) {
    set_pdl_mode(value);
    return 0;
  }
  if (which == 5) {
    set_gemm_split_mode(value);
    return 0;
  }
  if (which == 6) {
    set_gemm_rows2_mode(value);
    return 0;
  }
  if (which == 7) {
    set_gemm_warp_shift(value);
    return 0;
  }
  if (which == 8) {
    set_branch_mode(value);
    return 0;
  }
  if (which == 9) {
    set_gemm_gelu_packed(value);
    return 0;
  }
  if (which == 10) {
    set_gemm_red_prefetch(value);
    return 0;
  }
  return 1;
}

// ---- kernel-level entry points ---------------------------------------------------------------------------------------
static int g_num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

int ufm_op_gemm(const void* a, int C, int X, int Y, int Nimg, int64_t a_sx, int64_t a_sy, int64_t a_sn, int tw, int th,
                const void* wt, int n_out, int taps, int cin_pad, const float* bias, const float* scale,
                const float* res1, const float* res2, int64_t r2n, int64_t r2y, int64_t r2x, float* out_f32,
                void* out_bf16, int64_t sn, int64_t sy, int64_t sx, int64_t out_off, uint32_t flags,
                const float* tail_w2, const float* tail_b2, const float* tail_scale, const float* tail_shift,
                int tail_out, float* out2, void* stream, char* err, int errlen) {
  AView av{a, C, X, Y, Nimg, a_sx, a_sy, a_sn, tw, th};
  if (tw <= 0) pick_tile(X, Y, &av.tw, &av.th);
  Epi e;
  e.bias = bias; e.scale = scale; e.res1 = res1; e.res2 = res2;
  e.r2n = r2n; e.r2y = r2y; e.r2x = r2x;
  e.out_f32 = out_f32; e.out_bf16 = reinterpret_cast<bf16*>(out_bf16);
  e.sn = sn; e.sy = sy; e.sx = sx; e.out_off = out_off;
  e.flags = flags;
  e.w2 = tail_w2; e.tail_out = tail_out; e.out2 = out2;
  for (int i = 0; i < tail_out && i < kTailMax; ++i) {
    e.b2[i] = tail_b2 ? tail_b2[i] : 0.f;
    e.tscale[i] = tail_scale ? tail_scale[i] : 1.f;
    e.tshift[i] = tail_shift ? tail_shift[i] : 0.f;
  }
  GemmParams gp;
  int BN;
  const long long r2 = rows2_rows(av, n_out, taps, e, g_num_sms());
  if (r2 > 0) {  // same as ufm_engine::gemm: 512-row pair tiles for the whole waves, the rest as half items of the same launch
    if (!build_gemm(&gp, &BN, av, wt, n_out, taps, cin_pad, e, err, errlen, g_num_sms(), gemm_rows2_mode(), r2)) return 1;
    cudaError_t c2 = launch_gemm(gp, BN, g_num_sms(), reinterpret_cast<cudaStream_t>(stream));
    if (c2 != cudaSuccess) {
      snprintf(err, errlen, "launch_gemm (rows2): %s", cudaGetErrorString(c2));
      return 1;
    }
    return 0;
  }
  if (!build_gemm(&gp, &BN, av, wt, n_out, taps, cin_pad, e, err, errlen, g_num_sms())) return 1;
  cudaError_t ce = launch_gemm(gp, BN, g_num_sms(), reinterpret_cast<cudaStream_t>(stream));
  if (ce != cudaSuccess) {
    snprintf(err, errlen, "launch_gemm: %s", cudaGetErrorString(ce));
    return 1;
  }
  return 0;
}

int ufm_op_attention(const void* qkv, int n_seq, int seq_len, int heads, float scale, void* out, void* stream, char* err,
                     int errlen) {
  AttnParams ap;
  if (!build_attn(&ap, qkv, n_seq, seq_len, heads, scale, out, err, errlen)) return 1;
  cudaError_t ce = launch_attention(ap, reinterpret_cast<cudaStream_t>(stream));
  if (ce != cudaSuccess) {
    snprintf(err, errlen, "launch_attention: %s", cudaGetErrorString(ce));
    return 1;
  }
  return 0;
}

int ufm_op_layernorm(const float* x, int n_seq, int src_seq_len, int src_skip, int tokens, int D, const float* w,
                     const float* b, float eps, int remap_B, void* out_bf16, int64_t ob_pitch, int64_t ob_col,
                     float* out_f32, int64_t of_pitch, int64_t of_col, void* stream) {
  LNParams p;
  p.x = x; p.D = D; p.eps = eps; p.w = w; p.b = b;
  p.n_seq = n_seq; p.src_seq_len = src_seq_len; p.src_skip = src_skip; p.tokens = tokens; p.remap_B = remap_B;
  p.reverse = 1;
  p.out_bf16 = reinterpret_cast<bf16*>(out_bf16); p.ob_pitch = ob_pitch; p.ob_col = ob_col;
  p.out_f32 = out_f32; p.of_pitch = of_pitch; p.of_col = of_col;
  return launch_layernorm(p, reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

int ufm_op_qk_norm(void* qkv, int64_t rows, int heads, const float* qw, const float* qb, const float* kw,
                   const float* kb, float eps, void* stream) {
  return launch_qk_norm(reinterpret_cast<bf16*>(qkv), rows, heads, qw, qb, kw, kb, eps,
                        reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

int ufm_op_rope(void* qkv, int64_t rows, int heads, int tokens_per_view, int grid_w, void* stream) {
  return launch_qk_norm(reinterpret_cast<bf16*>(qkv), rows, heads, nullptr, nullptr, nullptr, nullptr, 0.f,
                        reinterpret_cast<cudaStream_t>(stream), 0, tokens_per_view, grid_w) != cudaSuccess;
}

int ufm_op_resize_ac(const float* in_f32, const void* in_bf16, int n, int h, int w, int C, int rh, int rw, int oh, int ow,
                     float* out_f32, void* out_bf16, void* stream) {
  return launch_resize_ac(in_f32, reinterpret_cast<const bf16*>(in_bf16), n, h, w, C, rh, rw, oh, ow, out_f32,
                          reinterpret_cast<bf16*>(out_bf16), reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

int ufm_op_shuffle(const void* in, void* out, int is_f32, int n, int h, int w, int k, int C, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (is_f32)
    return launch_shuffle_f32(reinterpret_cast<const float*>(in), reinterpret_cast<float*>(out), n, h, w, k, C, s) != cudaSuccess;
  return launch_shuffle_bf16(reinterpret_cast<const bf16*>(in), reinterpret_cast<bf16*>(out), n, h, w, k, C, s) != cudaSuccess;
}

int ufm_op_im2col_s2(const void* in, void* out, int n, int h, int w, int C, int oh, int ow, void* stream) {
  return launch_im2col_s2(reinterpret_cast<const bf16*>(in), reinterpret_cast<bf16*>(out), n, h, w, C, oh, ow,
                          reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

int ufm_op_prep_resize(const void* in, int is_u8, int nhwc, int B, int H0, int W0, int H, int W, const float* mean,
                       const float* std, const float* mul, const float* add, float* out, void* stream) {
  PrepParams p;
  p.in = in; p.is_u8 = is_u8; p.nhwc = nhwc; p.B = B; p.H0 = H0; p.W0 = W0; p.H = H; p.W = W;
  p.scale_y = (float)H0 / (float)H;
  p.scale_x = (float)W0 / (float)W;
  for (int c = 0; c < 3; ++c) {
    p.mean[c] = mean ? mean[c] : 0.f;
    p.std[c] = std ? std[c] : 1.f;
    p.mul[c] = mul ? mul[c] : 1.f;
    p.add[c] = add ? add[c] : 0.f;
  }
  p.out = out;
  return launch_prep_resize(p, reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

int ufm_op_unmap(const float* flow, const float* cov, int B, int H, int W, int H0, int W0, const int* rep, const int* src0,
                 const int* src1, const float* src_mul, const float* tgt_mul, float* flow_out, float* cov_out,
                 const float* cov_mul, int cov_mul_n, void* stream) {
  UnmapParams p;
  p.flow = flow; p.cov = cov; p.B = B; p.H = H; p.W = W; p.H0 = H0; p.W0 = W0;
  for (int i = 0; i < 4; ++i) {
    p.rep[i] = rep[i];
    p.src0[i] = src0[i];
    p.src1[i] = src1[i];
  }
  if (p.rep[0] < 0 || p.rep[1] > H || p.rep[2] < 0 || p.rep[3] > W || p.rep[1] <= p.rep[0] || p.rep[3] <= p.rep[2] ||
      p.src0[0] < 0 || p.src0[1] > H0 || p.src0[2] < 0 || p.src0[3] > W0)
    return 1;
  for (int i = 0; i < 2; ++i) {
    p.src_mul[i] = src_mul[i];
    p.tgt_mul[i] = tgt_mul[i];
  }
  p.flow_out = flow_out; p.cov_out = cov_out;
  p.cov_mul_n = (cov_mul && cov_mul_n >= 1 && cov_mul_n <= 4) ? cov_mul_n : 1;
  for (int i = 0; i < 4; ++i) p.cov_mul[i] = (cov_mul && i < p.cov_mul_n) ? cov_mul[i] : 1.f;
  if ((flow == nullptr) != (flow_out == nullptr)) return 1;
  return launch_unmap(p, reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

int ufm_op_refine(const void* desc, int B, int H, int W, const float* flow_in, const float* bias25, float inv_temperature,
                  float* flow_out, float* residual, float* log_softmax, void* stream) {
  RefineParams p;
  p.desc = reinterpret_cast<const bf16*>(desc); p.B = B; p.H = H; p.W = W; p.flow_in = flow_in; p.bias25 = bias25; p.inv_temperature = inv_temperature;
  p.flow_out = flow_out; p.residual = residual; p.log_softmax = log_softmax;
  return launch_refine(p, reinterpret_cast<cudaStream_t>(stream)) != cudaSuccess;
}

}  // extern "C"

// ---- allocation-free 3x3 stride-2 convolution -------------------------------------------------------------------------
// The im2col operand of the one stride-2 convolution (im2col_s2_kernel -> gemm_kernel) normally sits in the engine's
// workspace or in a caller-allocated buffer.  This variant keeps it in a file-scope device array instead, so the entry
// point below allocates nothing.  Every GPU that loads the library holds the array until the process exits.  In host code
// `g_s2_scratch` is only the host-side stand-in of the symbol: the device address is taken once per device with
// cudaGetSymbolAddress (that device current) and that pointer is what the kernels and the tensor map receive.
static __device__ __align__(256) unsigned char g_s2_scratch[UFM_CONV3X3_S2_STATIC_MAX_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_s2_scratch_ptr[kMaxDevices] = {};
std::mutex g_s2_mutex;

// device address of g_s2_scratch on the CURRENT device (nullptr + message on failure)
void* s2_scratch_on_current_device(char* err, int errlen) {
  int dev = -1;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    snprintf(err, errlen, "ufm_op_conv3x3_s2_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(ce), dev);
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_s2_mutex);
  if (!g_s2_scratch_ptr[dev]) {
    void* p = nullptr;
    ce = cudaGetSymbolAddress(&p, g_s2_scratch);
    if (ce != cudaSuccess || !p) {
      snprintf(err, errlen, "ufm_op_conv3x3_s2_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(ce));
      return nullptr;
    }
    g_s2_scratch_ptr[dev] = p;
  }
  return g_s2_scratch_ptr[dev];
}
}  // namespace

extern "C" int ufm_op_conv3x3_s2_static(const void* in, int n, int h, int w, int C, const void* wt, int n_out,
                                        const float* bias, float* out_f32, void* stream, char* err, int errlen) {
  char local[8];
  if (!err || errlen <= 0) {
    err = local;
    errlen = (int)sizeof(local);
  }
  err[0] = 0;
  if (!in || !wt || !out_f32 || n <= 0 || h <= 0 || w <= 0 || n_out <= 0 || C <= 0 || (C & 63)) {
    snprintf(err, errlen, "ufm_op_conv3x3_s2_static: needs in / wt / out_f32, positive sizes and C a multiple of 64 (C = %d)", C);
    return 1;
  }
  const int oh = (h - 1) / 2 + 1, ow = (w - 1) / 2 + 1;
  const long long rows = (long long)n * oh * ow;
  const long long need = rows * 9 * C * (long long)sizeof(bf16);
  if (need > (long long)UFM_CONV3X3_S2_STATIC_MAX_BYTES) {
    snprintf(err, errlen,
             "ufm_op_conv3x3_s2_static: the im2col operand of n=%d h=%d w=%d C=%d needs %lld bytes, the static scratch holds "
             "%lld (UFM_CONV3X3_S2_STATIC_MAX_BYTES): use ufm_op_im2col_s2 + ufm_op_gemm with a buffer of your own",
             n, h, w, C, need, (long long)UFM_CONV3X3_S2_STATIC_MAX_BYTES);
    return 2;
  }
  bf16* col = reinterpret_cast<bf16*>(s2_scratch_on_current_device(err, errlen));
  if (!col) return 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // every element of col[rows][9 * C] is written (zeros for the padding taps) before the GEMM reads it: nothing of the
  // previous call's contents survives into this one, so nothing needs clearing
  cudaError_t ce = launch_im2col_s2(reinterpret_cast<const bf16*>(in), col, n, h, w, C, oh, ow, st);
  if (ce != cudaSuccess) {
    snprintf(err, errlen, "ufm_op_conv3x3_s2_static: im2col_s2: %s", cudaGetErrorString(ce));
    return 1;
  }
  const long long K = 9LL * C;
  return ufm_op_gemm(col, (int)K, (int)rows, 1, 1, K, rows * K, rows * K, 128, 1, wt, n_out, 1, (int)K, bias, nullptr, nullptr,
                     nullptr, 0, 0, 0, out_f32, nullptr, 0, 0, n_out, 0, 0u, nullptr, nullptr, nullptr, nullptr, 0, nullptr,
                     stream, err, errlen);
}
