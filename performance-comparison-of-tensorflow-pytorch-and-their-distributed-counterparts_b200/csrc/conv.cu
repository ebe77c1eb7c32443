// Convolution fprop / dgrad / wgrad as GEMMs over NHWC activations.
//   fp32  : exact FFMA GEMM (sgemm.cu) -- the parity mode
//   bf16  : tcgen05 GEMM (gemm_sm100.cuh), fp32 accumulation in TMEM
// 1x1 / stride-1 convolutions run directly on the activation matrix [N*H*W][C]; everything else goes through an
// explicit im2col / col2im staging buffer in the caller-provided workspace (round-1 path; the 4-D TMA implicit
// path validated by tools/umma_probe.cu replaces the staging for the bf16 3x3 convolutions).
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "conv_sm100.cuh"
#include "wgrad_sm100.cuh"
#include "conv_halo_sm100.cuh"
#include <vector>
#include <stdlib.h>
#include <utility>

namespace b200 {
int32_t launch_sgemm(const float*, const float*, float*, int, int, int, long long, long long, long long, bool, bool,
                     const float*, int, int, int, long long, cudaStream_t);
int sgemm_effective_splits(int K, int splits);
int32_t launch_im2col(const void*, void*, int, int, int, int, int, int, int, int, int, int, int, int, cudaStream_t);
int32_t launch_col2im(const float*, void*, int, int, int, int, int, int, int, int, int, int, int, int, int,
                      cudaStream_t);
int32_t launch_reduce_wgrad(const float*, int, long long, int, int, int, int, float*, cudaStream_t);
int32_t launch_reduce_plain(const float*, int, long long, int, int, long long, void*, long long, int, int,
                            cudaStream_t);

struct ConvGeom {
  int Ho, Wo, RSC, ldw;
  long long M;
  bool direct;
  size_t esz;
};
static inline ConvGeom geom(const b200_conv_desc* d) {
  ConvGeom g;
  g.Ho = (d->H + 2 * d->pad - d->R) / d->stride + 1;
  g.Wo = (d->W + 2 * d->pad - d->S) / d->stride + 1;
  g.RSC = d->R * d->S * d->C;
  g.ldw = (g.RSC + 7) / 8 * 8;
  g.M = (long long)d->N * g.Ho * g.Wo;
  g.direct = d->R == 1 && d->S == 1 && d->stride == 1 && d->pad == 0 && (d->C % 8 == 0);
  g.esz = d->dtype == B200_BF16 ? 2 : 4;
  return g;
}
static inline int32_t check_desc(const b200_conv_desc* d) {
  if (!d) return fail(B200_INVALID_ARG, "conv: null descriptor");
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->C <= 0 || d->K <= 0 || d->R <= 0 || d->S <= 0 || d->stride <= 0 ||
      d->pad < 0)
    return fail(B200_INVALID_ARG, "conv: non-positive dimension");
  if (d->dtype != B200_F32 && d->dtype != B200_BF16) return fail(B200_INVALID_ARG, "conv: unknown dtype");
  if (d->H + 2 * d->pad < d->R || d->W + 2 * d->pad < d->S) return fail(B200_INVALID_ARG, "conv: filter larger than input");
  if (d->dtype == B200_BF16 && d->K % 8 != 0)
    return fail(B200_UNSUPPORTED, "conv(bf16): output channels must be a multiple of 8 (got %d)", d->K);
  return B200_OK;
}

static inline int wgrad_splits(const b200_conv_desc* d, const ConvGeom& g) {
  const int bn = d->dtype == B200_BF16 ? (g.RSC > 128 ? 256 : (g.RSC > 64 ? 128 : 64)) : 128;
  const long long tiles = (long long)((d->K + 127) / 128) * ((g.RSC + bn - 1) / bn);
  long long s = (2LL * num_sms() + tiles - 1) / tiles;
  const long long kb = (g.M + 63) / 64;
  if (s > kb) s = kb;
  if (s > 512) s = 512;
  const long long slab = (long long)d->K * g.ldw * 4;
  while (s > 1 && s * slab > (512LL << 20)) --s;
  if (s < 1) s = 1;
  return (int)s;
}

// optional per-launch timing of the tcgen05 GEMM (bench.py roofline leg): CUDA events on the launching stream
struct GemmTiming {
  bool on = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;
  double flops = 0;
};
static GemmTiming& gemm_timing() {
  static GemmTiming t;
  return t;
}

static int32_t gemm_bf16_run(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                             long long ldc, bool a_mn, bool b_mn, int out_dtype, int accumulate, int splits,
                             long long split_stride, cudaStream_t st) {
  GemmProblem p;
  p.A = A; p.B = B; p.lda = lda; p.ldb = ldb; p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn;
  p.splits = splits;
  p.out.ptr = C; p.out.ld = ldc; p.out.split_stride = split_stride; p.out.out_fp32 = out_dtype == B200_F32;
  p.out.accumulate = accumulate;
  p.out.vec_ok = ((reinterpret_cast<uintptr_t>(C) & 15) == 0) && (ldc % (out_dtype == B200_F32 ? 4 : 8) == 0);
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) || (lda % 8) || (ldb % 8))
    return fail(B200_UNSUPPORTED, "gemm_bf16: operands must be 16-byte aligned with leading dimensions multiple of 8");
  GemmTiming& gt = gemm_timing();
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (gt.on) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
  }
  int rc = launch_gemm_sm100(p, st);
  if (gt.on) {
    cudaEventRecord(e1, st);
    gt.ev.emplace_back(e0, e1);
    gt.flops += 2.0 * M * N * K;
  }
  launch_counter() += 1;
  if (rc) return fail(B200_CUDA_ERROR, "gemm_bf16: launch failed (code %d: %s)", rc, cudaGetErrorString(cudaGetLastError()));
  return B200_OK;
}


// ------------------------------------------------------------------------------------------------ implicit path
// bf16, channel counts that are multiples of 64, filters up to 3x3, stride 1 or 2: everything in ResNet-50 except the
// 3-channel stem.  Other shapes keep the staged (im2col) path above.
static inline bool implicit_ok(const b200_conv_desc* d) {
  static const bool disabled = getenv("B200_NO_IMPLICIT_CONV") != nullptr;
  return !disabled && d->dtype == B200_BF16 && d->C % 64 == 0 && d->K % 64 == 0 && d->R * d->S <= CONV_MAX_TAPS &&
         (d->stride == 1 || d->stride == 2) && d->pad < d->R && d->pad < d->S;
}
static inline bool is_flat(const b200_conv_desc* d) { return d->R == 1 && d->S == 1 && d->stride == 1 && d->pad == 0; }

struct WgradPlan {
  PixelBox box;
  int Wq, Hq, Nq, bn_tile, nsplits, pb_total, pb_per;
};
static inline WgradPlan wgrad_plan(const b200_conv_desc* d, const ConvGeom& g) {
  WgradPlan w;
  if (is_flat(d)) { w.Wq = (int)g.M; w.Hq = 1; w.Nq = 1; }
  else { w.Wq = g.Wo; w.Hq = g.Ho; w.Nq = d->N; }
  // columns of the packed [K][R*S*C] gradient per tile: 256 (four 64-channel tap groups share one dY tile) when
  // C == 64, else 128 with twice as many pixels per k-block
  static const int force256 = getenv("B200_WGRAD_BN256") ? atoi(getenv("B200_WGRAD_BN256")) : 0;
  w.bn_tile = (d->C == 64 || (force256 && d->C % 256 == 0)) ? 256 : 128;
  w.box = choose_pixel_box(w.Wq, w.Hq, w.Nq, w.bn_tile == 256 ? 64 : 128, 16, d->stride);
  w.pb_total = w.box.tiles_w * w.box.tiles_h * w.box.tiles_n;
  const long long ctas = (long long)((g.RSC + w.bn_tile - 1) / w.bn_tile) * ((d->K + 127) / 128);
  long long s = (2LL * num_sms() + ctas - 1) / ctas;   // two waves of work items: measured best for the conv layers
  // Linear layers (a true [tokens][features] matrix: N = H = 1, thousands of rows; BERT) are tuned separately --
  // measured on the 768/2304/3072-wide layers at 8192 tokens.  The same rules applied to ResNet's 1x1 convolutions
  // cost 1.7 ms per step (A/B: 14.08 vs 15.83 ms), so they stay on the two-wave rule.
  static const bool old_splits = getenv("B200_WGRAD_OLD_SPLITS") != nullptr;
  const bool linear = is_flat(d) && d->N == 1 && d->H == 1 && g.M >= 4096 && !old_splits;
  if (linear && ctas * 100 >= 85LL * num_sms() && ctas <= num_sms()) {
    // the output tiles already fill one wave (768 x 3072: 144 tiles): run unsplit, the kernel then writes the
    // gradient in place and the slab reduction disappears
    s = 1;
  } else if (linear) {
    // 2304x768 / 768x768 gradients (108 / 36 tiles): the split count with the smallest rounds-per-split ratio
    // ceil(ctas*s / SMs) / s (108 tiles: 4 splits = 2.92 waves instead of 3 splits = 2.19 waves)
    long long best = 1;
    double best_cost = 1e30;
    for (long long c = 1; c <= 16; ++c) {
      const double cost = (double)((ctas * c + num_sms() - 1) / num_sms()) / (double)c + 0.01 * (double)c;
      if (cost < best_cost - 1e-9) { best_cost = cost; best = c; }
    }
    s = best;
  }
  if (s > w.pb_total) s = w.pb_total;
  if (s > 1024) s = 1024;
  const long long slab = (long long)d->K * g.ldw * 4;
  while (s > 1 && s * slab > (1024LL << 20)) --s;
  if (s < 1) s = 1;
  w.pb_per = (int)((w.pb_total + s - 1) / s);
  w.nsplits = (w.pb_total + w.pb_per - 1) / w.pb_per;
  return w;
}

// CTA pairs (tcgen05 cta_group::2) for 256-wide fprop / dgrad tiles: each CTA of a cluster loads half of the B tile and
// the leader issues one M = 256 MMA for both (conv_sm100_kernel<..., CG = 2>, validated by tools/umma2sm_probe.cu and
// tests/test_kernels_gpu.py::test_cta_pair_kernels).  OPT-IN (B200_2SM=1): measured on BERT's Linears it is not
// faster than the single-CTA kernel (fprop 31.0 -> 35.2 us, dgrad 34.6 -> 33.6 us at 8192 x 768 x 2304/3072) -- those
// GEMMs already run at 1.0-1.15 PFLOP/s against a measured sustained cuBLAS peak of 1.40, so B-operand traffic per SM
// was not the limiter it looked like.  Not combined with the fused BatchNorm statistics (their per-CTA partial rows
// assume the single-CTA tile walk).
// Two-pixel-tile work items for 256-wide fprop / dgrad tiles; both variants are built, parity-tested
// (tests/test_kernels_gpu.py::test_two_tile_kernel_variants) and OFF by default because neither beat the single-tile
// kernel when measured (tools/bert_kernel_bench.py, tools/conv_table.py vgg):
//   mode 3 (B200_DUAL_MIN_KB=<k-blocks>): one CTA computes both tiles -- a 256 x 256 output tile, B loaded once for
//           two MMAs.  Only the deepest reductions gain (512->512 3x3 dgrad 164 -> 158 us); 12-k-block Linears lose
//           20-40 % (QKV forward 30.7 -> 39.2 us) because the epilogue no longer overlaps the next main loop.
//   mode 2 (B200_2SM=1): CTA pairs, tcgen05 cta_group::2 -- see conv_sm100.cuh; fprop 31.0 -> 35.2 us.
// The deep layers already run at 0.85-0.88 of the nominal roofline (~1.3 PFLOP/s, the chip's sustained limit), so
// operand traffic per SM was not the limiter the first ncu reading suggested; what is left is per-tile latency on
// shallow reductions.
static inline int pair_mode_for(int kblocks) {
  static const bool two_sm = getenv("B200_2SM") != nullptr;
  static const int dual_min = getenv("B200_DUAL_MIN_KB") ? atoi(getenv("B200_DUAL_MIN_KB")) : 0;
  if (two_sm) return 2;
  if (dual_min > 0 && kblocks >= dual_min) return 3;
  return 0;
}
static inline int pair_items(int bn, long long pixel_tiles, int n_tiles, bool stats, int mode) {
  if (mode == 0 || bn != 256 || stats || pixel_tiles < 2) return 0;
  const long long items = ((pixel_tiles + 1) / 2) * n_tiles;
  return items > 0x3fffffffLL ? 0 : (int)items;
}

template <int MODE>
static int launch_conv_mode(int bn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                            const ConvParams& P, cudaStream_t st) {
  const int ctas = num_sms();   // persistent: one CTA per SM
  if constexpr (MODE == MODE_WGRAD) {
    if (bn == 256) return launch_conv_cfg<256, 4, MODE>(ta, tb, to, P, ctas, st);
    return launch_conv_cfg<128, 3, MODE>(ta, tb, to, P, ctas, st);
  } else {
    if (bn == 256 && P.total_pairs > 0 && P.pair_mode == 3) return launch_conv_cfg<256, 3, MODE, 3>(ta, tb, to, P, ctas, st);
    if (bn == 256 && P.total_pairs > 0) return launch_conv_cfg<256, 5, MODE, 2>(ta, tb, to, P, ctas, st);   // CTA pairs
    // Short reductions (up to four k-blocks per tile: the 1x1 convolutions with C <= 256) are bound by their epilogue
    // and take the variant with two staging tiles per epilogue group and one operand stage less; deep reductions keep
    // the stages.  Measured: 64->256 1x1 forward 59.5 -> 48.3 us (0.83 of its HBM roofline), 128->512 34 -> 28 us,
    // 256->1024 22.5 -> 18.4 us, while 512->256 / 512->1024 lose 2 us with the shallow variant.
    static const int shallow_kb = getenv("B200_SHALLOW_KB") ? atoi(getenv("B200_SHALLOW_KB")) : 4;
    const bool shallow = P.ntaps * P.cchunks <= shallow_kb || P.res_tma;   // the TMA residual needs two staging tiles
    if (bn == 256) return shallow ? launch_conv_cfg<256, 3, MODE, 1, 2>(ta, tb, to, P, ctas, st)
                                  : launch_conv_cfg<256, 4, MODE, 1, 1>(ta, tb, to, P, ctas, st);
    if (bn == 128) return shallow ? launch_conv_cfg<128, 5, MODE, 1, 2>(ta, tb, to, P, ctas, st)
                                  : launch_conv_cfg<128, 6, MODE, 1, 1>(ta, tb, to, P, ctas, st);
    return shallow ? launch_conv_cfg<64, 6, MODE, 1, 2>(ta, tb, to, P, ctas, st)
                   : launch_conv_cfg<64, 7, MODE, 1, 1>(ta, tb, to, P, ctas, st);
  }
}

static inline void timing_begin(cudaEvent_t& e0, cudaEvent_t& e1, cudaStream_t st) {
  if (gemm_timing().on) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
  }
}
static inline void timing_end(cudaEvent_t e0, cudaEvent_t e1, double flops, cudaStream_t st) {
  GemmTiming& gt = gemm_timing();
  if (gt.on) {
    cudaEventRecord(e1, st);
    gt.ev.emplace_back(e0, e1);
    gt.flops += flops;
  }
  launch_counter() += 1;
}

// Column-tile width of fprop / dgrad: as wide as the output allows.  (Measured on BERT's 768-column Linears, which
// fill only 1.3 waves with 256-wide tiles: 128-wide tiles = 2.6 waves of half the work are 25-30 % SLOWER, e.g. FFN2
// forward 39.5 -> 50.9 us -- the kernel is bound by operand traffic from L2 per MMA, which narrower tiles raise.)
// Exception, measured on the N = 256 / 160 px layer table after the MMA-issue fixes of round 2: layers with exactly 256
// output columns and a deep reduction (>= 16 k-blocks) run faster on 128-wide tiles -- 3x3 256 -> 256 at 10x10:
// 29.5 -> 26.8 us forward, 29.3 -> 26.4 us dgrad; 20x20 stride 2: 29.7 -> 26.9 / 48.0 -> 44.3 us; 1x1 1024 -> 256 forward
// 17.5 -> 15.8 us and the dgrad of 256 -> 1024 17.3 -> 15.8 us: a few hundred pixel tiles on 148 SMs are two uneven waves
// of long tiles.  Shallower 256-column layers (512 -> 256: 33.6 -> 36.0 us) and 512-column ones (3x3: +4 us) lose with
// them and keep 256, and so do problems with many pixel tiles (more than two per SM: VGG's 56x56 maps), where the wave
// shape does not matter and the narrower tile only adds activation traffic.  `red` = reduction channels (C for fprop,
// K for dgrad), `dgrad` selects which pixel grid counts.  B200_BN_DEEP256=256 restores the wide tile.
static inline int pick_bn(const b200_conv_desc* d, const ConvGeom& g, int cols, int red, bool dgrad = false) {
  static const int bn_deep256 = getenv("B200_BN_DEEP256") ? atoi(getenv("B200_BN_DEEP256")) : 128;
  const long long pixels = dgrad ? (long long)d->N * d->H * d->W : (long long)g.M;
  if (cols == 256 && d->R * d->S * (red / 64) >= 16 && pixels <= 256LL * num_sms() &&
      (bn_deep256 == 128 || bn_deep256 == 64))
    return bn_deep256;
  static const int cap = getenv("B200_BN_CAP") ? atoi(getenv("B200_BN_CAP")) : 256;      // experiment knob
  const int bn = cols > 128 ? 256 : (cols > 64 ? 128 : 64);
  return bn > cap ? cap : bn;
}

// Weights-stationary variant (conv_halo_sm100.cuh): the CTA's weight tile (all taps x channel chunks) stays in shared
// memory and only activations stream.  kind 1 = 1x1 stride-1 layers whose reduction fits the weight slots (4 k-blocks
// at 256 output columns, 8 at 128 / 64); kind 2 = "same" 3x3 stride-1 layers with 64 -> 64 channels, whose activations
// come as one halo tile per tile of bh full image rows and whose taps are descriptor shifts.  `cin` = reduction
// channels, `cout` = output columns (swapped for dgrad).  B200_WS = bit mask of the kinds to use.  Default 2: measured
// on the N = 256 / 160 px layer table, kind 2 takes 64 -> 64 3x3 at 40x40 from 53.6 / 52.6 us (fprop / dgrad) to
// 38.8 / 37.7 us, while kind 1 LOSES on every 1x1 class (64 -> 256: 48.5 -> 58.9 us, 256 -> 1024 at 10x10: 18.7 ->
// 21.3 us): those tiles have 1-4 k-blocks and are bound by the epilogue (fused statistics), not by operand traffic,
// and this kernel has one staging tile per epilogue group where the shallow variant of conv_sm100_kernel has two.
// kind 4 (mask bit 3, OFF) = "same" 3x3 layers whose weight tile does not fit: halo activations, weights streamed
// through a ring -- measured slower than the plain kernel (128 -> 128 at 20x20: 32.9 vs 30.2 us, 256 -> 256 at 10x10:
// 30.6 vs 29.6 us): with a weight barrier per tap the issue loop is no tighter and the activation re-reads it saves
// were L2 hits anyway.
// kind 3 (mask bit 2, fprop only) = "valid" 4 x 1 stride-1 layers with 64 -> 64 channels -- the row-pair
// staged stem: a tile is the usual bw x bh pixel box, its activations ONE box of bh + R - 1 rows, a tap a shift by bw rows.
// kind 3's pixel box: one image per box (the halo lives inside an image), the fewest tiles, then the tallest box (the
// halo rows are re-read once per box); bn = 0 when no box satisfies the stage limits
static inline PixelBox ws3_box(const b200_conv_desc* d, const ConvGeom& g) {
  PixelBox best{0, 0, 0, 0, 0, 0};
  long long best_tiles = -1;
  for (int bw = 1; bw <= g.Wo && bw <= GEMM_BM; ++bw)
    for (int bh = 1; bh <= g.Ho && bw * bh <= GEMM_BM; ++bh) {
      if ((bh + d->R - 1) * bw > HALO_ROWS || (d->R - 1) * bw + GEMM_BM > HALO_ROWS || bh + d->R - 1 > 256) continue;
      const long long tw = (g.Wo + bw - 1) / bw, th = (g.Ho + bh - 1) / bh, tiles = tw * th;
      if (best_tiles < 0 || tiles < best_tiles || (tiles == best_tiles && bh > best.bh)) {
        best_tiles = tiles;
        best = PixelBox{bw, bh, 1, (int)tw, (int)th, d->N};
      }
    }
  return best;
}
static inline int ws_kind(const b200_conv_desc* d, const ConvGeom& g, int cin, int cout, bool dgrad = false) {
  static const int mask = getenv("B200_WS") ? atoi(getenv("B200_WS")) : 6;
  if (mask == 0 || d->stride != 1 || cin % 64 != 0 || cout % 8 != 0) return 0;
  const int bn = pick_bn(d, g, cout, cin, dgrad);
  const int chunks = cin / 64;
  if ((mask & 4) && !dgrad && d->S == 1 && d->R == 4 && d->pad == 0 && bn == 64 && chunks == 1) {
    const PixelBox b = ws3_box(d, g);
    return (b.bn == 1 && b.bw * b.bh * 10 >= GEMM_BM * 7) ? 3 : 0;            // at least 70 % of the MMA rows are pixels
  }
  if (d->R == 1 && d->S == 1 && d->pad == 0) {
    if (!(mask & 1)) return 0;
    return chunks <= (bn == 256 ? 4 : 8) ? 1 : 0;
  }
  if (d->R != 3 || d->S != 3 || d->pad != 1) return 0;
  const bool resident = bn == 64 && chunks == 1;          // kind 2; otherwise kind 4: halo activations, streaming weights
  if (resident ? !(mask & 2) : !(mask & 8)) return 0;
  if (g.Wo != d->W || g.Ho != d->H) return 0;
  const int wh = d->W + d->S - 1;
  if ((d->R - 1) * wh + d->S - 1 + GEMM_BM > HALO_ROWS) return 0;           // the last tap's view must stay inside the stage
  const int bh = GEMM_BM / wh;
  if (bh < 1 || (bh + d->R - 1) * wh > HALO_ROWS || bh + d->R - 1 > 256) return 0;
  const int bh_eff = bh < d->H ? bh : d->H;
  if (bh_eff * d->W * 10 < GEMM_BM * 7) return 0;                            // at least 70 % of the MMA rows are pixels
  return resident ? 2 : 4;
}
static inline PixelBox halo_box(const b200_conv_desc* d) {
  const int wh = d->W + d->S - 1;
  int bh = GEMM_BM / wh;
  if (bh > d->H) bh = d->H;
  return PixelBox{wh, bh, 1, 1, (d->H + bh - 1) / bh, d->N};
}

// fprop (mode 0) and dgrad (mode 1) of one pixel grid / tap set
static inline PixelBox fprop_box(const b200_conv_desc* d, const ConvGeom& g) {
  const bool flat = is_flat(d);
  const int Wq = flat ? (int)g.M : g.Wo, Hq = flat ? 1 : g.Ho, Nq = flat ? 1 : d->N;
  const int ws = ws_kind(d, g, d->C, d->K);
  if (ws == 2 || ws == 4) return halo_box(d);
  if (ws == 3) return ws3_box(d, g);
  return choose_pixel_box(Wq, Hq, Nq, GEMM_BM, 1, d->stride);
}

template <int MODE>
static int launch_conv_ws(int kind, int bn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                          const ConvParams& P, const HaloParams& HP, cudaStream_t st) {
  const int ctas = num_sms();
  if (kind == 2) return launch_conv_ws_cfg<64, 9, MODE, 3, HALO_ROWS, 9>(ta, tb, to, P, HP, ctas, st);
  if (kind == 4) {
    if (bn == 256) return launch_conv_ws_cfg<256, 4, MODE, 2, HALO_ROWS, 9, false>(ta, tb, to, P, HP, ctas, st);
    if (bn == 128) return launch_conv_ws_cfg<128, 6, MODE, 3, HALO_ROWS, 9, false>(ta, tb, to, P, HP, ctas, st);
    return launch_conv_ws_cfg<64, 8, MODE, 3, HALO_ROWS, 9, false>(ta, tb, to, P, HP, ctas, st);
  }
  if (kind == 3) {
    if constexpr (MODE == MODE_FPROP) {
      return launch_conv_ws_cfg<64, 4, MODE, 3, HALO_ROWS, 4>(ta, tb, to, P, HP, ctas, st);
    }
    return 250;
  }
  if (bn == 256) return launch_conv_ws_cfg<256, 4, MODE, 4, GEMM_BM, 1>(ta, tb, to, P, HP, ctas, st);
  if (bn == 128) return launch_conv_ws_cfg<128, 8, MODE, 4, GEMM_BM, 1>(ta, tb, to, P, HP, ctas, st);
  return launch_conv_ws_cfg<64, 8, MODE, 6, GEMM_BM, 1>(ta, tb, to, P, HP, ctas, st);
}

static int32_t implicit_fprop(const b200_conv_desc* d, const ConvGeom& g, const void* x, const void* w, void* y,
                              float* stats, cudaStream_t st, const float* ep_scale = nullptr,
                              const float* ep_shift = nullptr, int ep_relu = 0, const void* ep_res = nullptr) {
  ConvParams P{};
  const bool flat = is_flat(d);
  const int Wq = flat ? (int)g.M : g.Wo, Hq = flat ? 1 : g.Ho, Nq = flat ? 1 : d->N;
  const PixelBox b = fprop_box(d, g);
  P.bw = b.bw; P.bh = b.bh; P.bn = b.bn; P.tiles_w = b.tiles_w; P.tiles_h = b.tiles_h; P.tiles_n = b.tiles_n;
  P.a_mul = d->stride;
  P.ntaps = d->R * d->S;
  P.cchunks = d->C / 64;
  for (int r = 0; r < d->R; ++r)
    for (int s2 = 0; s2 < d->S; ++s2) P.taps[r * d->S + s2] = ConvTap{s2 - d->pad, r - d->pad, (r * d->S + s2) * d->C};
  P.out = y; P.ldo = d->K; P.Wout = Wq; P.Hout = Hq; P.Nout = Nq; P.o_mul = 1; P.o_off_w = 0; P.o_off_h = 0;
  P.ncols = d->K; P.out_fp32 = 0; P.accumulate = 0;
  P.vec_ok = (reinterpret_cast<uintptr_t>(y) & 15) == 0 && d->K % 8 == 0;
  CUtensorMap ta, tb;
  const int ws = ws_kind(d, g, d->C, d->K);
  const bool halo = ws >= 2;
  HaloParams HP{};
  HP.wh = b.bw; HP.rows = b.bw * b.bh * b.bn;                    // kind 1: the plain pixel box, no shifts
  if (halo) {
    HP.rows = (b.bh + d->R - 1) * b.bw; HP.org_w = -d->pad; HP.org_h = -d->pad;
    for (int r = 0; r < d->R; ++r)
      for (int s2 = 0; s2 < d->S; ++s2) HP.toff[r * d->S + s2] = r * b.bw + s2;
  }
  int rc = flat ? make_tmap_bf16_nhwc(&ta, x, 1, 1, (int)g.M, d->C, b.bw, b.bh, b.bn, 1)
                : make_tmap_bf16_nhwc(&ta, x, d->N, d->H, d->W, d->C, b.bw, halo ? b.bh + d->R - 1 : b.bh, b.bn, d->stride);
  if (rc) return fail(B200_CUDA_ERROR, "conv_fprop: activation tensor map failed (%d)", rc);
  const int bn = pick_bn(d, g, d->K, d->C);
  const long long tiles = (long long)b.tiles_w * b.tiles_h * b.tiles_n;
  P.n_tiles = (d->K + bn - 1) / bn;
  P.pair_mode = ws ? 0 : pair_mode_for(d->R * d->S * (d->C / 64));
  P.total_pairs = pair_items(bn, tiles, P.n_tiles, stats != nullptr, P.pair_mode);
  rc = make_tmap_bf16_2d(&tb, w, g.RSC, d->K, g.ldw, 64, (P.total_pairs > 0 && P.pair_mode == 2) ? bn / 2 : bn);
  if (rc) return fail(B200_CUDA_ERROR, "conv_fprop: weight tensor map failed (%d)", rc);
  P.m_tiles = 1;
  if (tiles * P.n_tiles > 0x7fffffffLL) return fail(B200_UNSUPPORTED, "conv_fprop: too many tiles");
  P.total_tiles = (int)(tiles * P.n_tiles);
  CUtensorMap to;
  static const bool no_tma_store = getenv("B200_NO_TMA_STORE") != nullptr;
  P.tma_store = !no_tma_store && P.vec_ok;
  P.stats = stats;
  P.ep_scale = ep_scale; P.ep_shift = ep_shift; P.ep_relu = ep_relu; P.ep_res = ep_res;
  if (ep_scale && !P.tma_store) return fail(B200_UNSUPPORTED, "conv_fprop_affine: needs the TMA-store epilogue");
  // residual of an eval-mode block output: fetched by TMA into the staging tile when every tile has all of its
  // column groups (B200_RES_TMA=0: per-thread loads, measured LSU-bound)
  static const bool res_tma_on = !(getenv("B200_RES_TMA") && atoi(getenv("B200_RES_TMA")) == 0);
  P.res_tma = 0;
  if (ep_res && res_tma_on && !ws && P.total_pairs == 0 && bn >= 128 && d->K % bn == 0 &&
      (reinterpret_cast<uintptr_t>(ep_res) & 15) == 0) {
    rc = make_tmap_bf16_nhwc(&P.tmR, ep_res, Nq, Hq, Wq, d->K, b.bw, b.bh, b.bn, 1);
    if (rc) return fail(B200_CUDA_ERROR, "conv_fprop_affine: residual tensor map failed (%d)", rc);
    P.res_tma = 1;
  }
  if (stats && !P.tma_store) return fail(B200_UNSUPPORTED, "conv_fprop_bnstats: needs the TMA-store epilogue");
  rc = make_tmap_bf16_nhwc(&to, y, Nq, Hq, Wq, d->K, b.bw, b.bh, b.bn, 1);
  if (rc) return fail(B200_CUDA_ERROR, "conv_fprop: output tensor map failed (%d)", rc);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  timing_begin(e0, e1, st);
  rc = ws ? launch_conv_ws<MODE_FPROP>(ws, bn, ta, tb, to, P, HP, st) : launch_conv_mode<MODE_FPROP>(bn, ta, tb, to, P, st);
  timing_end(e0, e1, 2.0 * g.M * d->K * g.RSC, st);
  if (rc) return fail(B200_CUDA_ERROR, "conv_fprop: launch failed (%d)", rc);
  return B200_OK;
}

static int32_t implicit_dgrad(const b200_conv_desc* d, const ConvGeom& g, const void* dy, const void* w, void* dx,
                              int accumulate, cudaStream_t st) {
  const bool flat = is_flat(d);
  CUtensorMap tb;
  int rc = make_tmap_bf16_2d(&tb, w, g.RSC, d->K, g.ldw, 64, 64);
  if (rc) return fail(B200_CUDA_ERROR, "conv_dgrad: weight tensor map failed (%d)", rc);
  const int bn = pick_bn(d, g, d->C, d->K, true);
  const int classes = d->stride == 1 ? 1 : 4;
  static const bool no_tma_store = getenv("B200_NO_TMA_STORE") != nullptr;
  static const bool no_tma_reduce = getenv("B200_NO_TMA_REDUCE") != nullptr;
  static const bool no_strided_tma = getenv("B200_NO_STRIDED_TMA_STORE") != nullptr;
  const bool vec_ok_dx = (reinterpret_cast<uintptr_t>(dx) & 15) == 0 && d->C % 8 == 0;
  const bool tma_out = !no_tma_store && vec_ok_dx && (d->stride == 1 || !no_strided_tma) && (!accumulate || !no_tma_reduce);
  if (classes > 1 && !accumulate && tma_out) {
    // Parity classes that no filter tap reaches (three of the four for a 1x1 stride-2 convolution) are plain zeros:
    // clear the gradient once instead of running a GEMM launch per empty class that only stores zeros.
    bool any_empty = false;
    for (int cls = 0; cls < classes; ++cls) {
      const int ph = cls / 2, pw = cls % 2;
      int nt = 0;
      for (int r = 0; r < d->R; ++r)
        for (int s2 = 0; s2 < d->S; ++s2)
          if ((ph + d->pad - r) % d->stride == 0 && (pw + d->pad - s2) % d->stride == 0) ++nt;
      if (nt == 0) any_empty = true;
    }
    if (any_empty) {
      cudaMemsetAsync(dx, 0, (size_t)d->N * d->H * d->W * d->C * 2, st);
      launch_counter() += 1;
    }
  }
  for (int cls = 0; cls < classes; ++cls) {
    const int ph = d->stride == 1 ? 0 : cls / 2, pw = d->stride == 1 ? 0 : cls % 2;
    ConvParams P{};
    int Wq, Hq, Nq;
    if (flat) { Wq = (int)g.M; Hq = 1; Nq = 1; }
    else {
      Wq = (d->W - pw + d->stride - 1) / d->stride;
      Hq = (d->H - ph + d->stride - 1) / d->stride;
      Nq = d->N;
    }
    if (Wq <= 0 || Hq <= 0) continue;
    P.ntaps = 0;
    for (int r = 0; r < d->R; ++r) {
      const int vh = ph + d->pad - r;
      if (vh % d->stride) continue;
      for (int s2 = 0; s2 < d->S; ++s2) {
        const int vw = pw + d->pad - s2;
        if (vw % d->stride) continue;
        P.taps[P.ntaps++] = ConvTap{vw / d->stride, vh / d->stride, (r * d->S + s2) * d->C};
      }
    }
    if (P.ntaps == 0 && (accumulate || (classes > 1 && tma_out))) continue;  // nothing to add / already cleared
    // weights-stationary variant: dgrad of a stride-1 convolution is again one (roles of C and K swapped)
    const int ws = (classes == 1 && tma_out) ? ws_kind(d, g, d->K, d->C, true) : 0;
    const bool halo = ws == 2 || ws == 4;
    HaloParams HP{};
    const PixelBox b = halo ? halo_box(d) : choose_pixel_box(Wq, Hq, Nq, GEMM_BM, 1, 1);
    HP.wh = b.bw; HP.rows = b.bw * b.bh * b.bn;
    if (ws == 1) { HP.org_w = P.ntaps ? P.taps[0].dw : 0; HP.org_h = P.ntaps ? P.taps[0].dh : 0; }
    if (halo) {
      int mw = 1 << 30, mh = 1 << 30;
      for (int t = 0; t < P.ntaps; ++t) { if (P.taps[t].dw < mw) mw = P.taps[t].dw; if (P.taps[t].dh < mh) mh = P.taps[t].dh; }
      HP.wh = b.bw; HP.rows = (b.bh + d->R - 1) * b.bw; HP.org_w = mw; HP.org_h = mh;
      for (int t = 0; t < P.ntaps; ++t) HP.toff[t] = (P.taps[t].dh - mh) * b.bw + (P.taps[t].dw - mw);
    }
    P.bw = b.bw; P.bh = b.bh; P.bn = b.bn; P.tiles_w = b.tiles_w; P.tiles_h = b.tiles_h; P.tiles_n = b.tiles_n;
    P.a_mul = 1;
    P.cchunks = d->K / 64;
    P.out = dx; P.ldo = d->C;
    P.Wout = flat ? (int)g.M : d->W; P.Hout = flat ? 1 : d->H; P.Nout = flat ? 1 : d->N;
    P.o_mul = d->stride; P.o_off_w = pw; P.o_off_h = ph;
    P.ncols = d->C; P.out_fp32 = 0; P.accumulate = accumulate;
    P.vec_ok = (reinterpret_cast<uintptr_t>(dx) & 15) == 0 && d->C % 8 == 0;
    CUtensorMap ta;
    rc = flat ? make_tmap_bf16_nhwc(&ta, dy, 1, 1, (int)g.M, d->K, b.bw, b.bh, b.bn, 1)
              : make_tmap_bf16_nhwc(&ta, dy, d->N, g.Ho, g.Wo, d->K, b.bw, halo ? b.bh + d->R - 1 : b.bh, b.bn, 1);
    if (rc) return fail(B200_CUDA_ERROR, "conv_dgrad: gradient tensor map failed (%d)", rc);
    const long long tiles = (long long)b.tiles_w * b.tiles_h * b.tiles_n;
    P.n_tiles = (d->C + bn - 1) / bn;
    P.m_tiles = 1;
    if (tiles * P.n_tiles > 0x7fffffffLL) return fail(B200_UNSUPPORTED, "conv_dgrad: too many tiles");
    P.total_tiles = (int)(tiles * P.n_tiles);
    P.pair_mode = ws ? 0 : pair_mode_for(P.ntaps * P.cchunks);
    P.total_pairs = pair_items(bn, tiles, P.n_tiles, false, P.pair_mode);
    CUtensorMap to;
    // bf16 outputs leave through TMA stores (TMA reduce-add when accumulating); for a parity class of a stride-2
    // layer the output map walks every second pixel (elementStrides) starting at the class's (ph, pw)
    P.tma_store = tma_out;
    rc = make_tmap_bf16_nhwc(&to, dx, P.Nout, P.Hout, P.Wout, d->C, b.bw, b.bh, b.bn, d->stride);
    if (rc) return fail(B200_CUDA_ERROR, "conv_dgrad: output tensor map failed (%d)", rc);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    timing_begin(e0, e1, st);
    rc = ws ? launch_conv_ws<MODE_DGRAD>(ws, bn, ta, tb, to, P, HP, st) : launch_conv_mode<MODE_DGRAD>(bn, ta, tb, to, P, st);
    timing_end(e0, e1, 2.0 * Wq * Hq * Nq * (double)d->C * d->K * P.ntaps, st);
    if (rc) return fail(B200_CUDA_ERROR, "conv_dgrad: launch failed (%d)", rc);
  }
  return B200_OK;
}

static int32_t implicit_wgrad(const b200_conv_desc* d, const ConvGeom& g, const void* x, const void* dy, float* dw,
                              void* ws, cudaStream_t st) {
  const WgradPlan wp = wgrad_plan(d, g);
  const bool flat = is_flat(d);
  ConvParams P{};
  P.bw = wp.box.bw; P.bh = wp.box.bh; P.bn = wp.box.bn;
  P.tiles_w = wp.box.tiles_w; P.tiles_h = wp.box.tiles_h; P.tiles_n = wp.box.tiles_n;
  P.ntaps = d->R * d->S;
  for (int r = 0; r < d->R; ++r)
    for (int s2 = 0; s2 < d->S; ++s2) P.taps[r * d->S + s2] = ConvTap{s2 - d->pad, r - d->pad, (r * d->S + s2) * d->C};
  // unsplit 1x1: the packed [K][C] slab IS the OIHW gradient, so the kernel writes dw directly
  const bool direct = wp.nsplits == 1 && d->R * d->S == 1 && g.ldw == d->C && (reinterpret_cast<uintptr_t>(dw) & 15) == 0;
  P.out = direct ? (void*)dw : ws; P.ldo = g.ldw; P.split_stride = (long long)d->K * g.ldw;
  P.ncols = g.RSC; P.cdim = d->C; P.out_fp32 = 1; P.accumulate = 0;
  P.vec_ok = (reinterpret_cast<uintptr_t>(P.out) & 15) == 0;   // ldw % 8 == 0 and column tiles start at multiples of 256
  P.nsplits = wp.nsplits; P.pb_total = wp.pb_total; P.pb_per_split = wp.pb_per; P.x_mul = d->stride;
  P.kr = wp.box.bw * wp.box.bh * wp.box.bn;
  P.mrows = d->K;
  CUtensorMap ta, tb;
  int rc = flat ? make_tmap_bf16_nhwc(&ta, dy, 1, 1, (int)g.M, d->K, P.bw, P.bh, P.bn, 1)
                : make_tmap_bf16_nhwc(&ta, dy, d->N, g.Ho, g.Wo, d->K, P.bw, P.bh, P.bn, 1);
  if (rc) return fail(B200_CUDA_ERROR, "conv_wgrad: gradient tensor map failed (%d)", rc);
  rc = flat ? make_tmap_bf16_nhwc(&tb, x, 1, 1, (int)g.M, d->C, P.bw, P.bh, P.bn, 1)
            : make_tmap_bf16_nhwc(&tb, x, d->N, d->H, d->W, d->C, P.bw, P.bh, P.bn, d->stride);
  if (rc) return fail(B200_CUDA_ERROR, "conv_wgrad: activation tensor map failed (%d)", rc);
  P.n_tiles = (g.RSC + wp.bn_tile - 1) / wp.bn_tile;
  P.n_inner = d->C / wp.bn_tile > 0 && P.n_tiles % (d->C / wp.bn_tile) == 0 ? d->C / wp.bn_tile : 1;
  P.m_tiles = (d->K + 127) / 128;
  P.total_tiles = P.n_tiles * P.m_tiles * wp.nsplits;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  timing_begin(e0, e1, st);
  P.tma_store = 0;
  rc = launch_conv_mode<MODE_WGRAD>(wp.bn_tile, ta, tb, ta, P, st);
  timing_end(e0, e1, 2.0 * g.M * d->K * g.RSC, st);
  if (rc) return fail(B200_CUDA_ERROR, "conv_wgrad: launch failed (%d)", rc);
  if (direct) return B200_OK;
  return launch_reduce_wgrad((const float*)ws, wp.nsplits, P.split_stride, d->K, d->C, d->R * d->S, g.ldw, dw, st);
}

// ------------------------------------------------------------------------------------------------ wgrad, 2nd generation
// (wgrad_sm100.cuh: 256 x 256 tiles, split-K over a thread-block cluster reduced through distributed shared memory)
struct WgradPlan2 {
  PixelBox box;
  int Wq, Hq, Nq, tm, n_tiles, m_tiles, pb_total, pb_per, cluster, groups;
  bool direct;
};
static inline bool wgrad_v2_ok(const b200_conv_desc* d, const ConvGeom& g) {
  static const int force = getenv("B200_WGRAD_GEN") ? atoi(getenv("B200_WGRAD_GEN")) : 0;   // 1 / 2: force a generation
  // true matrices (BERT's Linear layers: thousands of rows, N = H = 1) keep the first-generation kernel and its
  // measured split rules
  const bool linear = is_flat(d) && d->N == 1 && d->H == 1;
  if (linear) return false;
  if (force) return force == 2;
  // Measured per layer class at N = 256 (tools/conv_table.py, graph replay): the 256-wide tiles win on every 3x3 and
  // stride-2 layer (1.2-1.5x), on the big 1x1 layers (K*C >= 2^19) and on the narrow-output ones (K <= 128, where the
  // first generation re-reads X for nothing); the mid-size 1x1 layers (64->256, 128->512, 512->256, 256<->1024) are
  // 5-15 % faster on the 128-wide tiles, whose epilogue overlaps the next tile's main loop.
  return d->R * d->S > 1 || d->stride > 1 || (long long)d->K * d->C >= (1LL << 19) || d->K <= 128;
}
static inline WgradPlan2 wgrad_plan2(const b200_conv_desc* d, const ConvGeom& g, const float* dw) {
  WgradPlan2 w;
  if (is_flat(d)) { w.Wq = (int)g.M; w.Hq = 1; w.Nq = 1; }
  else { w.Wq = g.Wo; w.Hq = g.Ho; w.Nq = d->N; }
  w.tm = d->K > 128 ? 256 : 128;
  w.n_tiles = (g.RSC + WG_TN - 1) / WG_TN;
  w.m_tiles = (d->K + w.tm - 1) / w.tm;
  w.box = choose_pixel_box(w.Wq, w.Hq, w.Nq, WG_KR, 16, d->stride);
  w.pb_total = w.box.tiles_w * w.box.tiles_h * w.box.tiles_n;
  const int tiles = w.n_tiles * w.m_tiles;
  // Splits: one wave of CTAs (one per SM).  A cluster reduces its ranks' accumulators through distributed shared
  // memory; measured, that pays for PAIRS only (DSMEM moves ~20 B/clk per SM, so a rank that gathers a 1/8 slice from
  // eight peers takes longer than writing its slab to L2; clusters of 4 / 8 also strand SMs of a GPC), so the
  // default is pairs of CTAs and one fp32 slab per pair.  B200_WGRAD_CLUSTER=1/2/4/8 overrides.
  static const int force_cs = getenv("B200_WGRAD_CLUSTER") ? atoi(getenv("B200_WGRAD_CLUSTER")) : 0;
  int s_want = num_sms() / tiles;
  if (s_want > w.pb_total / 2) s_want = w.pb_total / 2;     // at least two k-blocks per CTA
  if (s_want < 1) s_want = 1;
  int cs = force_cs > 0 ? force_cs : 2;
  // ... except where four CTAs per tile are the whole split (no slab, the gradient is written directly) and each still
  // reduces >= 16 k-blocks, so the four-way DSMEM exchange is small next to the main loop: 1024 -> 2048 1x1 stride 2
  // (25 k-blocks per CTA) 46.1 -> 32.6 us.  1x1 layers only: 512 -> 512 3x3 loses with it (46 -> 75 us; its direct
  // epilogue scatters into OIHW)
  if (force_cs <= 0 && d->R * d->S == 1 && s_want > 2 && s_want <= 4 && w.pb_total / 4 >= 16) cs = 4;
  while (cs > 1 && cs > s_want) cs >>= 1;
  w.cluster = cs;
  w.groups = s_want / cs > 0 ? s_want / cs : 1;
  const int S = w.cluster * w.groups;
  w.pb_per = (w.pb_total + S - 1) / S;
  w.direct = w.groups == 1 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0 && (d->R * d->S > 1 || g.ldw == d->C);
  static const bool dbg = getenv("B200_WGRAD_DEBUG") != nullptr;
  if (dbg)
    fprintf(stderr, "wgrad plan: K %d C %d RS %d stride %d | tiles %d x %d, k-blocks %d, s_want %d -> cluster %d x groups %d, direct %d\n",
            d->K, d->C, d->R * d->S, d->stride, w.m_tiles, w.n_tiles, w.pb_total, s_want, w.cluster, w.groups, (int)w.direct);
  return w;
}

static int32_t implicit_wgrad_v2(const b200_conv_desc* d, const ConvGeom& g, const void* x, const void* dy, float* dw,
                                 void* ws, cudaStream_t st) {
  const WgradPlan2 wp = wgrad_plan2(d, g, dw);
  const bool flat = is_flat(d);
  WgradParams P{};
  P.bw = wp.box.bw; P.bh = wp.box.bh; P.bn = wp.box.bn; P.tiles_w = wp.box.tiles_w; P.tiles_h = wp.box.tiles_h;
  P.kr = wp.box.bw * wp.box.bh * wp.box.bn;
  P.x_mul = d->stride;
  for (int r = 0; r < d->R; ++r)
    for (int s2 = 0; s2 < d->S; ++s2) P.taps[r * d->S + s2] = ConvTap{s2 - d->pad, r - d->pad, (r * d->S + s2) * d->C};
  P.cdim = d->C; P.ncols = g.RSC; P.mrows = d->K; P.n_tiles = wp.n_tiles; P.tm = wp.tm;
  P.pb_total = wp.pb_total; P.pb_per_split = wp.pb_per; P.cluster = wp.cluster;
  P.direct = wp.direct ? 1 : 0;
  P.rs = d->R * d->S;
  P.out = wp.direct ? dw : reinterpret_cast<float*>(ws);
  P.ldo = g.ldw; P.group_stride = (long long)d->K * g.ldw;
  if (!wp.direct && ((reinterpret_cast<uintptr_t>(ws) & 15) != 0)) return fail(B200_INVALID_ARG, "conv_wgrad: misaligned workspace");
  CUtensorMap ta, tb;
  int rc = flat ? make_tmap_bf16_nhwc(&ta, dy, 1, 1, (int)g.M, d->K, P.bw, P.bh, P.bn, 1)
                : make_tmap_bf16_nhwc(&ta, dy, d->N, g.Ho, g.Wo, d->K, P.bw, P.bh, P.bn, 1);
  if (rc) return fail(B200_CUDA_ERROR, "conv_wgrad: gradient tensor map failed (%d)", rc);
  rc = flat ? make_tmap_bf16_nhwc(&tb, x, 1, 1, (int)g.M, d->C, P.bw, P.bh, P.bn, 1)
            : make_tmap_bf16_nhwc(&tb, x, d->N, d->H, d->W, d->C, P.bw, P.bh, P.bn, d->stride);
  if (rc) return fail(B200_CUDA_ERROR, "conv_wgrad: activation tensor map failed (%d)", rc);
  static int attr_state[B200_MAX_DEVICES] = {0};
  if (cudaError_t e = ensure_dyn_smem(wgrad_sm100_kernel, WG_SMEM_BYTES, attr_state))
    return fail(B200_CUDA_ERROR, "conv_wgrad: shared-memory attribute: %s", cudaGetErrorString(e));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(wp.cluster * wp.groups, wp.n_tiles * wp.m_tiles, 1);
  cfg.blockDim = dim3(WG_THREADS);
  cfg.dynamicSmemBytes = WG_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = wp.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  timing_begin(e0, e1, st);
  cudaError_t le = cudaLaunchKernelEx(&cfg, wgrad_sm100_kernel, ta, tb, P);
  timing_end(e0, e1, 2.0 * g.M * d->K * g.RSC, st);
  if (le != cudaSuccess) return fail(B200_CUDA_ERROR, "conv_wgrad: launch failed (%s)", cudaGetErrorString(le));
  if (wp.direct) return B200_OK;
  return launch_reduce_wgrad((const float*)ws, wp.groups, P.group_stride, d->K, d->C, d->R * d->S, g.ldw, dw, st);
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_conv_weight_ld(int32_t R, int32_t S, int32_t C) { return (R * S * C + 7) / 8 * 8; }

extern "C" int64_t b200_conv_workspace_bytes(const b200_conv_desc* d, int32_t which) {
  if (check_desc(d)) return -1;
  const ConvGeom g = geom(d);
  if (implicit_ok(d)) {
    if (which == B200_CONV_FPROP || which == B200_CONV_DGRAD) return 0;
    if (which == B200_CONV_WGRAD) {
      // upper bound over both generations of the kernel (the second one needs one slab per cluster group, and none
      // when a single cluster covers the reduction)
      size_t need = (size_t)wgrad_plan(d, g).nsplits * d->K * g.ldw * 4;
      if (wgrad_v2_ok(d, g)) need = (size_t)wgrad_plan2(d, g, nullptr).groups * d->K * g.ldw * 4;
      return (int64_t)align_up(need, 256);
    }
    return -1;
  }
  size_t col = g.direct ? 0 : align_up((size_t)g.M * g.ldw * g.esz, 256);
  if (which == B200_CONV_FPROP) return (int64_t)col;
  if (which == B200_CONV_DGRAD) return g.direct ? 0 : (int64_t)align_up((size_t)g.M * g.ldw * 4, 256);
  if (which == B200_CONV_WGRAD) {
    const int s = wgrad_splits(d, g);
    return (int64_t)(col + align_up((size_t)s * d->K * g.ldw * 4, 256));
  }
  return -1;
}

extern "C" int32_t b200_conv_fprop(const b200_conv_desc* d, const void* x, const void* w, void* y, void* ws,
                                   int64_t ws_bytes, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!x || !w || !y) return fail(B200_INVALID_ARG, "conv_fprop: null pointer");
  const ConvGeom g = geom(d);
  cudaStream_t st = as_stream(stream);
  if (implicit_ok(d)) return implicit_fprop(d, g, x, w, y, nullptr, st);
  const void* A = x;
  long long lda = d->C;
  int Kred = d->C;
  if (!g.direct) {
    if (ws_bytes < b200_conv_workspace_bytes(d, B200_CONV_FPROP) || !ws)
      return fail(B200_WORKSPACE_TOO_SMALL, "conv_fprop: workspace too small");
    if (int32_t rc = launch_im2col(x, ws, d->dtype, d->N, d->H, d->W, d->C, d->R, d->S, d->stride, d->pad, g.Ho, g.Wo,
                                   g.ldw, st))
      return rc;
    A = ws;
    lda = g.ldw;
    Kred = g.ldw;
  }
  if (d->dtype == B200_F32)
    return launch_sgemm((const float*)A, (const float*)w, (float*)y, (int)g.M, d->K, Kred, lda, g.ldw, d->K, false,
                        false, nullptr, 0, 0, 1, 0, st);
  return gemm_bf16_run(A, w, y, (int)g.M, d->K, Kred, lda, g.ldw, d->K, false, false, B200_BF16, 0, 1, 0, st);
}

extern "C" int32_t b200_conv_dgrad(const b200_conv_desc* d, const void* dy, const void* w, void* dx,
                                   int32_t accumulate, void* ws, int64_t ws_bytes, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!dy || !w || !dx) return fail(B200_INVALID_ARG, "conv_dgrad: null pointer");
  const ConvGeom g = geom(d);
  cudaStream_t st = as_stream(stream);
  if (implicit_ok(d)) return implicit_dgrad(d, g, dy, w, dx, accumulate, st);
  if (g.direct) {
    // dx[M][C] (+)= dy[M][K] * W[K][C]
    if (d->dtype == B200_F32)
      return launch_sgemm((const float*)dy, (const float*)w, (float*)dx, (int)g.M, d->C, d->K, d->K, g.ldw, d->C,
                          false, true, nullptr, 0, accumulate, 1, 0, st);
    return gemm_bf16_run(dy, w, dx, (int)g.M, d->C, d->K, d->K, g.ldw, d->C, false, true, B200_BF16, accumulate, 1, 0,
                         st);
  }
  if (ws_bytes < b200_conv_workspace_bytes(d, B200_CONV_DGRAD) || !ws)
    return fail(B200_WORKSPACE_TOO_SMALL, "conv_dgrad: workspace too small");
  float* dcol = (float*)ws;  // [M][ldw] fp32 = dy[M][K] * W[K][RSC]
  int32_t rc;
  if (d->dtype == B200_F32)
    rc = launch_sgemm((const float*)dy, (const float*)w, dcol, (int)g.M, g.RSC, d->K, d->K, g.ldw, g.ldw, false, true,
                      nullptr, 0, 0, 1, 0, st);
  else
    rc = gemm_bf16_run(dy, w, dcol, (int)g.M, g.RSC, d->K, d->K, g.ldw, g.ldw, false, true, B200_F32, 0, 1, 0, st);
  if (rc) return rc;
  return launch_col2im(dcol, dx, d->dtype, d->N, d->H, d->W, d->C, d->R, d->S, d->stride, d->pad, g.Ho, g.Wo, g.ldw,
                       accumulate, st);
}

extern "C" int32_t b200_conv_wgrad(const b200_conv_desc* d, const void* x, const void* dy, float* dw, void* ws,
                                   int64_t ws_bytes, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!x || !dy || !dw) return fail(B200_INVALID_ARG, "conv_wgrad: null pointer");
  const ConvGeom g = geom(d);
  cudaStream_t st = as_stream(stream);
  if (ws_bytes < b200_conv_workspace_bytes(d, B200_CONV_WGRAD) || !ws)
    return fail(B200_WORKSPACE_TOO_SMALL, "conv_wgrad: workspace too small");
  if (implicit_ok(d)) return wgrad_v2_ok(d, g) ? implicit_wgrad_v2(d, g, x, dy, dw, ws, st) : implicit_wgrad(d, g, x, dy, dw, ws, st);
  const size_t col_bytes = g.direct ? 0 : align_up((size_t)g.M * g.ldw * g.esz, 256);
  const void* Bm = x;
  long long ldb = d->C;
  int Ncols = d->C;
  if (!g.direct) {
    if (int32_t rc = launch_im2col(x, ws, d->dtype, d->N, d->H, d->W, d->C, d->R, d->S, d->stride, d->pad, g.Ho, g.Wo,
                                   g.ldw, st))
      return rc;
    Bm = ws;
    ldb = g.ldw;
    Ncols = g.RSC;
  }
  float* slabs = reinterpret_cast<float*>(static_cast<uint8_t*>(ws) + col_bytes);
  const int want = wgrad_splits(d, g);
  const long long slab_stride = (long long)d->K * g.ldw;
  int eff;
  int32_t rc;
  // dWp[K][RSC] = dy^T[K][M] * col[M][RSC] : both operands have the reduction (pixel) index outermost
  if (d->dtype == B200_F32) {
    eff = sgemm_effective_splits((int)g.M, want);
    rc = launch_sgemm((const float*)dy, (const float*)Bm, slabs, d->K, Ncols, (int)g.M, d->K, ldb, g.ldw, true, true,
                      nullptr, 0, 0, want, slab_stride, st);
  } else {
    eff = gemm_effective_splits((int)g.M, want);
    rc = gemm_bf16_run(dy, Bm, slabs, d->K, Ncols, (int)g.M, d->K, ldb, g.ldw, true, true, B200_F32, 0, want,
                       slab_stride, st);
  }
  if (rc) return rc;
  return launch_reduce_wgrad(slabs, eff, slab_stride, d->K, d->C, d->R * d->S, g.ldw, dw, st);
}

extern "C" int32_t b200_gemm_bf16(const void* A, const void* B, void* C, int32_t M, int32_t N, int32_t K, int64_t lda,
                                  int64_t ldb, int64_t ldc, int32_t a_mn, int32_t b_mn, int32_t out_dtype,
                                  int32_t accumulate, int32_t splits, void* ws, int64_t ws_bytes, void* stream) {
  if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0) return fail(B200_INVALID_ARG, "gemm_bf16: bad arguments");
  cudaStream_t st = as_stream(stream);
  const int eff = gemm_effective_splits(K, splits);
  if (eff <= 1) return gemm_bf16_run(A, B, C, M, N, K, lda, ldb, ldc, a_mn != 0, b_mn != 0, out_dtype, accumulate, 1, 0, st);
  const long long slab = (long long)M * ldc;
  if (!ws || ws_bytes < (int64_t)eff * slab * 4) return fail(B200_WORKSPACE_TOO_SMALL, "gemm_bf16: split-K workspace too small");
  if (int32_t rc = gemm_bf16_run(A, B, ws, M, N, K, lda, ldb, ldc, a_mn != 0, b_mn != 0, B200_F32, 0, splits, slab, st))
    return rc;
  return launch_reduce_plain((const float*)ws, eff, slab, M, N, ldc, C, ldc, out_dtype, accumulate, st);
}

extern "C" int32_t b200_gemm_timing_enable(int32_t on) {
  GemmTiming& gt = gemm_timing();
  for (auto& pr : gt.ev) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  gt.ev.clear();
  gt.flops = 0;
  gt.on = on != 0;
  return B200_OK;
}
extern "C" int32_t b200_gemm_timing_read(double* total_ms, int64_t* launches, double* flops) {
  GemmTiming& gt = gemm_timing();
  double ms = 0;
  for (auto& pr : gt.ev) {
    if (cudaEventSynchronize(pr.second) != cudaSuccess) return fail(B200_CUDA_ERROR, "gemm_timing_read: event sync failed");
    float t = 0;
    cudaEventElapsedTime(&t, pr.first, pr.second);
    ms += t;
  }
  if (total_ms) *total_ms = ms;
  if (launches) *launches = (int64_t)gt.ev.size();
  if (flops) *flops = gt.flops;
  return B200_OK;
}

// ---- convolution forward with the following BatchNorm's batch statistics fused into the epilogue
extern "C" int64_t b200_conv_fprop_bnstats_rows(const b200_conv_desc* d) {
  if (check_desc(d)) return -1;
  static const bool off = getenv("B200_NO_FUSED_BNSTATS") != nullptr || getenv("B200_NO_TMA_STORE") != nullptr;
  if (off || !implicit_ok(d)) return 0;
  const ConvGeom g = geom(d);
  const PixelBox b = fprop_box(d, g);
  const int bn = pick_bn(d, g, d->K, d->C);
  const int n_tiles = (d->K + bn - 1) / bn;
  const long long total = (long long)b.tiles_w * b.tiles_h * b.tiles_n * n_tiles;
  if (total > 0x7fffffffLL) return 0;
  return conv_stats_grid((int)total, n_tiles, num_sms()) / n_tiles;   // rows of partial sums the launch writes
}

extern "C" int32_t b200_conv_fprop_bnstats(const b200_conv_desc* d, const void* x, const void* w, void* y,
                                           float* partials, int64_t partial_rows, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!x || !w || !y || !partials) return fail(B200_INVALID_ARG, "conv_fprop_bnstats: null pointer");
  const int64_t need = b200_conv_fprop_bnstats_rows(d);
  if (need <= 0) return fail(B200_UNSUPPORTED, "conv_fprop_bnstats: shape not on the implicit tcgen05 path");
  if (partial_rows < need) return fail(B200_WORKSPACE_TOO_SMALL, "conv_fprop_bnstats: %lld partial rows needed", (long long)need);
  if ((reinterpret_cast<uintptr_t>(partials) & 7) || (reinterpret_cast<uintptr_t>(y) & 15))
    return fail(B200_INVALID_ARG, "conv_fprop_bnstats: misaligned buffers");
  return implicit_fprop(d, geom(d), x, w, y, partials, as_stream(stream));
}

// ---- convolution forward with an eval-mode BatchNorm (+ReLU) folded into the epilogue
extern "C" int32_t b200_conv_fprop_affine(const b200_conv_desc* d, const void* x, const void* w, void* y,
                                          const float* scale, const float* shift, const void* residual,
                                          int32_t relu, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!x || !w || !y || !scale || !shift) return fail(B200_INVALID_ARG, "conv_fprop_affine: null pointer");
  if (b200_conv_fprop_bnstats_rows(d) <= 0) return fail(B200_UNSUPPORTED, "conv_fprop_affine: shape not on the implicit tcgen05 path");
  if ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(residual)) & 15)
    return fail(B200_INVALID_ARG, "conv_fprop_affine: misaligned output / residual");
  return implicit_fprop(d, geom(d), x, w, y, nullptr, as_stream(stream), scale, shift, relu, residual);
}
