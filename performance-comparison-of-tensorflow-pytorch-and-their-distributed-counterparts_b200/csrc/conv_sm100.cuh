// Implicit-GEMM convolution on tcgen05 for sm_100a: fprop, dgrad and wgrad over NHWC bf16 activations with NO
// im2col / col2im staging.  The A operand of each k-block is fetched straight from the activation tensor by a 4-D
// tiled TMA box {64 channels, bw, bh, bn pixels}:
//   * the filter tap (r,s) is a coordinate shift of the box; out-of-bounds rows/columns (the padding halo, ragged
//     tile tails, images past N) are zero-filled by the TMA unit;
//   * the convolution stride is the tensor map's elementStrides (the box walks every stride-th pixel);
//   * a 1x1/stride-1 convolution is the same code with the image flattened to one row of N*H*W pixels.
// The M tile of the MMA is the pixel box (up to 128 pixels; rows past the box are stale smem and only ever reach
// accumulator rows that are never stored).  Verified on hardware by tools/umma_probe.cu (4-D TMA cases).
//
//   MODE_FPROP : D[pix][cout] = sum_{tap,c} X[pix shifted by tap][c] * W[cout][tap][c]      (B K-major)
//   MODE_DGRAD : D[pix][cin]  = sum_{tap,k} dY[pix shifted][k]       * W[k][tap][cin]       (B MN-major)
//                stride-2 convolutions are decomposed into the four output-parity classes, each a stride-1 problem
//                over the taps of matching parity (no wasted MACs), written to the interleaved output positions
//   MODE_WGRAD : D[cout][(tap,cin)] = sum_{pixel boxes} dY[pix][cout]^T * X[pix shifted by tap][cin]
//                both operands MN-major (pixel index outermost), split-K over pixel boxes into fp32 slabs.
//                The N tile is 256 columns of the packed [K][R*S*C] gradient: each 64-column group is its own TMA
//                box of X with its own tap shift, so one dY tile feeds up to four taps (C=64) per MMA.
#pragma once
#include "sm100_ptx.cuh"
#include "gemm_sm100.cuh"

namespace b200 {

enum { MODE_FPROP = 0, MODE_DGRAD = 1, MODE_WGRAD = 2 };
constexpr int CONV_MAX_TAPS = 9;
constexpr int CONV_KR_MAX = 128;  // wgrad: most pixel rows per k-block (multiple of 16); 64 with 256-column tiles
constexpr int CONV_STAGING_TILE = 128 * 128;        // epilogue: one [128 rows][64 bf16] swizzled tile
// staging tiles per epilogue group (template parameter NSTG): 2 lets the TMA store of one column group overlap the
// packing of the next (what the short reductions need: their epilogue is the critical path), 1 leaves 32 KB more for
// operand stages (what the deep reductions need)
constexpr int CONV_STATS_RED_BYTES = 2 * 128 * 4;  // epilogue: per-group broadcast scratch (64 scales + 64 shifts of the affine epilogue)
// Threads: warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..9 = EIGHT epilogue warps in two groups of four.  A
// warp may only read the TMEM lanes of its quarter (warp % 4), so the two groups cover the same 128 accumulator rows
// and split the 64-column groups of a tile between them (group e takes column groups e, e+2, ...): with four warps
// the epilogue of a 128x256 tile (TMEM -> bf16 -> swizzled smem -> TMA store, plus the BatchNorm column sums) took
// longer than the main loop of every layer with fewer than ~4 k-blocks (the 64->256, 128->512, 256->1024 1x1
// convolutions ran at half of their HBM roofline); two groups halve it and give every scheduler two warps to hide
// the tcgen05.ld / shared-memory latencies behind each other.
constexpr int CONV_THREADS = 320;
constexpr int CONV_EPI_WARPS = 8;
// 64-column tiles have a single 64-channel column group: instead of leaving one epilogue group idle, the two groups
// take ALTERNATE tiles (group g drains accumulator buffer g), so an accumulator is released by four warps, not eight.
template <int BN, int CG>
__host__ __device__ constexpr bool conv_epi_alternates() { return BN == 64 && CG == 1; }
template <int BN, int CG>
__host__ __device__ constexpr int conv_tmem_empty_arrivals() {
  return CG == 2 ? 2 * CONV_EPI_WARPS : (conv_epi_alternates<BN, CG>() ? CONV_EPI_WARPS / 2 : CONV_EPI_WARPS);
}

struct ConvTap {
  int dw, dh;  // shift of the activation box (input-pixel units) for this tap
  int bcol;    // column offset of this tap in the packed weight matrix [K][R*S*C]
};

struct ConvParams {
  int bw, bh, bn;                    // pixel box extents
  int tiles_w, tiles_h, tiles_n;     // number of boxes along each pixel dimension
  int a_mul;                         // A box origin = tile origin * a_mul + tap shift
  int ntaps, cchunks;                // reduction: taps x 64-channel chunks (fprop/dgrad)
  ConvTap taps[CONV_MAX_TAPS];
  void* out;                         // fprop/dgrad: NHWC tensor; wgrad: fp32 slabs [split][K][ld]
  long long ldo;                     // elements per output pixel (fprop/dgrad) / per cout row (wgrad)
  long long split_stride;            // wgrad: elements per slab
  int Wout, Hout, Nout;              // output tensor extents (fprop/dgrad)
  int o_mul, o_off_w, o_off_h;       // output pixel = tile pixel * o_mul + offset
  int ncols;                         // valid output columns (cout / cin)
  int out_fp32, accumulate, vec_ok;
  // wgrad only
  int nsplits, pb_total, pb_per_split, x_mul;  // x_mul = conv stride (X box origin = tile origin * x_mul + shift)
  int kr;                                       // pixel rows per box = bw*bh*bn
  int mrows;                                    // valid cout rows
  int cdim;                                     // input channels C: output column j <-> tap j / C, channel j % C
  // persistent tile walk: tile = n_tile + n_tiles * (pixel tile | m_tile + m_tiles * (tap*nsplits + split))
  int n_tiles, m_tiles, total_tiles;
  int n_inner;                       // wgrad: column tiles per tap (n_tiles = n_inner * tap groups)
  int total_pairs;                   // CG = 2 / 3 kernels: ceil(pixel tiles / 2) * n_tiles work items (two pixel tiles each)
  int pair_mode;                     // 2 = CTA pair (cta_group::2), 3 = both pixel tiles in one CTA
  int tma_store;                     // fprop/dgrad: bf16 output written by TMA from a swizzled smem staging tile
  const float* ep_scale;             // fprop + tma_store (eval-mode BatchNorm folded into the epilogue):
  const float* ep_shift;             //   out = act(acc * ep_scale[col] + ep_shift[col]), act = ReLU when ep_relu
  int ep_relu;
  const void* ep_res;                //   optional residual tensor (same NHWC shape as the output), added before act
  float* stats;                      // fprop + tma_store: per-pixel-tile column sums [pixel tile][2][ncols] of the
                                     // bf16-rounded output (sum, sum of squares) for the BatchNorm that follows
  int res_tma;                       // ep_res arrives by TMA into the staging tile (tmR; kernels with two staging
                                     // tiles per epilogue group, ncols a multiple of the tile width)
  alignas(64) CUtensorMap tmR;       // residual tensor, same geometry as the output map
};

// wgrad stage geometry: 128-column tiles use 128-pixel k-blocks, 256-column tiles 64-pixel k-blocks (same bytes)
#define CONV_WGRAD_KR(BN_) ((BN_) == 256 ? 64 : 128)

template <int BN, int STAGES, int MODE, int CG = 1, int NSTG = 1>
constexpr int conv_smem_bytes() {
  constexpr int a = MODE == MODE_WGRAD ? 2 * CONV_WGRAD_KR(BN) * 128 : GEMM_BM * 128 * (CG == 3 ? 2 : 1);
  constexpr int b = MODE == MODE_WGRAD ? (BN / 64) * CONV_WGRAD_KR(BN) * 128 : BN * 128 / (CG == 2 ? 2 : 1);
  return STAGES * (a + b) + (MODE == MODE_WGRAD ? 0 : 2 * NSTG * CONV_STAGING_TILE + CONV_STATS_RED_BYTES) + 1024 + 256;
}

// Persistent kernel: grid = min(#tiles, #SMs); every CTA walks tiles blockIdx.x, blockIdx.x + gridDim.x, ...
// TMEM holds TWO accumulators (2*BN columns) so the epilogue of tile i (TMEM -> registers -> global) overlaps the
// TMA/MMA mainloop of tile i+1; the smem ring and its mbarrier phases run continuously across tiles.
struct ConvTile {
  int n0;                   // output-column tile origin
  int q0, p0, i0;           // pixel-box origin (fprop/dgrad)
  int m0, tap, split;       // wgrad
  int kb_begin, kb_end;
};
template <int BN, int MODE>
__device__ __forceinline__ ConvTile conv_decode_tile(const ConvParams& P, int tile) {
  ConvTile t{};
  if (MODE != MODE_WGRAD) {
    const int nt = tile % P.n_tiles;
    int rest = tile / P.n_tiles;
    t.n0 = nt * BN;
    const int tw = rest % P.tiles_w;
    rest /= P.tiles_w;
    const int th = rest % P.tiles_h;
    const int tn = rest / P.tiles_h;
    t.q0 = tw * P.bw;
    t.p0 = th * P.bh;
    t.i0 = tn * P.bn;
    t.kb_begin = 0;
    t.kb_end = P.ntaps * P.cchunks;
  } else {
    // order: column tile within a tap (fastest), cout tile, split, tap group.  Concurrent CTAs then work on
    // different pixel ranges (splits) instead of all hitting the same dY / X lines in L2 at once.
    const int ni = tile % P.n_inner;
    int rest = tile / P.n_inner;
    const int mt = rest % P.m_tiles;
    rest /= P.m_tiles;
    t.split = rest % P.nsplits;
    const int no = rest / P.nsplits;
    t.n0 = (no * P.n_inner + ni) * BN;
    t.m0 = mt * GEMM_BM;
    t.tap = t.n0 / P.cdim;
    t.kb_begin = t.split * P.pb_per_split;
    t.kb_end = t.kb_begin + P.pb_per_split;
    if (t.kb_end > P.pb_total) t.kb_end = P.pb_total;
  }
  return t;
}

// The epilogue warps' whole job, shared by the convolution kernels: for every tile of this CTA's walk, drain the
// accumulator (TMEM -> registers), apply the affine / residual / ReLU epilogue if asked, round to bf16, stage the tile
// in 128B-swizzled shared memory and hand it to the TMA unit (plain store or reduce-add), accumulate the fused
// BatchNorm column sums, release the accumulator; at the end write this CTA's row of partial statistics.
template <int BN, int MODE, int CG, int NSTG>
__device__ __forceinline__ void conv_epilogue_warps(const CUtensorMap& tmO, const ConvParams& P, uint8_t* smStage,
                                                    float* smRedBase, uint64_t* tmem_full, uint64_t* tmem_empty,
                                                    uint32_t tmem_base, int warp, int lane, int rank, int walk_id,
                                                    int walk_n, int walk_total, uint64_t* res_bar = nullptr) {
  auto tile_of = [&](int w, int r) -> int {
    if (CG == 1) return w;
    return (w % P.n_tiles) + P.n_tiles * (2 * (w / P.n_tiles) + r);
  };
  // ------------------------------------------------------------ epilogue warps (two groups of four)
  const int q = warp & 3;
  const int l = q * 32 + lane;  // accumulator row
  const int eg = (warp - 2) >> 2;              // epilogue group 0 / 1
  const int bar_id = 1 + eg;                   // named barrier of the group (128 threads)
  float* smRed = smRedBase + eg * 128;
  uint32_t it = 0;
  uint32_t stage_ctr = 0;                      // which of the group's two staging tiles comes next
  // fused BatchNorm statistics: this thread's running column sums over ALL tiles of this CTA (the launcher makes
  // gridDim a multiple of n_tiles, so a CTA only ever sees one output-column tile); slot gi = column group eg + 2*gi
  constexpr bool ALT = conv_epi_alternates<BN, CG>();
  constexpr int GPG = (BN / 64 + 1) / 2;       // column groups per epilogue group
  float st_acc[GPG][4];
#pragma unroll
  for (int g = 0; g < GPG; ++g) { st_acc[g][0] = 0.f; st_acc[g][1] = 0.f; st_acc[g][2] = 0.f; st_acc[g][3] = 0.f; }
  // Residual by TMA (eval-mode block outputs: conv3 + BatchNorm + add + ReLU in one kernel).  The group's work is a
  // stream of items (tile, column group); item k uses staging tile k & 1.  The residual tile of item k + 1 is fetched
  // into the OTHER staging tile while item k is processed (its previous store has drained by then), lands on
  // res_bar[eg * 2 + (k & 1)], and the threads add it from the same swizzled chunks they then overwrite in place.
  const bool res_tma = NSTG == 2 && CG == 1 && MODE == MODE_FPROP && P.res_tma && res_bar != nullptr;
  auto res_fetch = [&](int w_item, int gi_item, uint32_t k) {      // elected thread only
    const ConvTile Tn = conv_decode_tile<BN, MODE>(P, w_item);
    uint8_t* dst = smStage + (eg * NSTG + (k & 1u)) * CONV_STAGING_TILE;
    uint64_t* bar = &res_bar[eg * 2 + (k & 1u)];
    mbar_expect_tx(bar, static_cast<uint32_t>(P.bw * P.bh * P.bn) * 128u);     // the box, not the staging tile
    tma_load_4d(dst, &P.tmR, bar, Tn.n0 + (eg + 2 * gi_item) * 64, Tn.q0, Tn.p0, Tn.i0);
  };
  if (res_tma && ((warp - 2) & 3) == 0 && lane == 0 && walk_id < walk_total) res_fetch(walk_id, 0, 0u);
  for (int w = walk_id; w < walk_total; w += walk_n, ++it)
  for (int half = 0; half < (CG == 3 ? 2 : 1); ++half) {      // CG == 3: the item's two pixel tiles, one after the other
    const int tile = tile_of(w, CG == 3 ? half : rank);
    const ConvTile T = conv_decode_tile<BN, MODE>(P, tile);
    const uint32_t buf = CG == 3 ? (uint32_t)half : (it & 1), use = CG == 3 ? it : (it >> 1);
    if (ALT && buf != static_cast<uint32_t>(eg)) continue;      // the other group's tile
    const bool have_work = T.kb_end > T.kb_begin;
    mbar_wait(&tmem_full[buf], use & 1);
    tc_fence_after();
    bool row_ok;
    long long row_off;
    if (MODE != MODE_WGRAD) {
      const int iw = l % P.bw;
      const int t = l / P.bw;
      const int ih = t % P.bh;
      const int in = t / P.bh;
      const int ow = (T.q0 + iw) * P.o_mul + P.o_off_w;
      const int oh = (T.p0 + ih) * P.o_mul + P.o_off_h;
      const int on = T.i0 + in;
      row_ok = in < P.bn && ow < P.Wout && oh < P.Hout && on < P.Nout;
      row_off = ((static_cast<long long>(on) * P.Hout + oh) * P.Wout + ow) * P.ldo;
    } else {
      row_ok = (T.m0 + l) < P.mrows;
      row_off = static_cast<long long>(T.split) * P.split_stride + static_cast<long long>(T.m0 + l) * P.ldo;
    }
    if (MODE != MODE_WGRAD && P.tma_store) {
      // ---- TMA-store epilogue: TMEM -> registers -> bf16 -> 128B-swizzled smem tile [128 px][64 ch] -> one bulk
      // tensor store per 64-channel group (full 128-byte lines, bounds clipped by the TMA unit).  Two staging
      // tiles alternate so the store of group g overlaps the TMEM reads / packing of group g+1.
      const bool elected = (((warp - 2) & 3) == 0 && lane == 0);
#pragma unroll
      for (int gi = 0; gi < GPG; ++gi) {
        const int g = ALT ? gi : eg + 2 * gi;
        if (g >= BN / 64 || T.n0 + g * 64 >= P.ncols) break;     // uniform across the group
        uint8_t* stg = smStage + (eg * NSTG + (NSTG == 2 ? (stage_ctr & 1u) : 0u)) * CONV_STAGING_TILE;
        if (elected && !res_tma) tma_store_wait_read<NSTG - 1>();   // the store that last used this tile has finished reading it
        const bool affine = MODE == MODE_FPROP && P.ep_scale != nullptr;
        if (affine) {   // this group's 64 per-channel scales and shifts -> smem (read back as broadcasts)
          const int col = T.n0 + g * 64 + (l & 63);
          smRed[l] = l < 64 ? P.ep_scale[col] : P.ep_shift[col];
        }
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        // residual (eval-mode block output): this thread's 64 channels of its pixel, issued before the TMEM reads
        uint4 resv[8];
        const bool has_res = affine && P.ep_res != nullptr && !res_tma;
        if (has_res && row_ok) {
          const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(P.ep_res) +
                                                           row_off + T.n0 + g * 64);
#pragma unroll
          for (int j = 0; j < 8; ++j) resv[j] = rp[j];
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t r[32];
          if (have_work) {
            tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN +
                                   static_cast<uint32_t>(g * 64 + h * 32), r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
          uint8_t* rowp = stg + l * 128;
          if (res_tma && h == 0) {
            // the other staging tile's store (item k - 1) has had the TMEM read to drain: fetch the next item's
            // residual into it, then wait for this item's
            if (elected) {
              tma_store_wait_read<0>();
              if (gi + 1 < GPG) res_fetch(w, gi + 1, stage_ctr + 1u);
              else if (w + walk_n < walk_total) res_fetch(w + walk_n, 0, stage_ctr + 1u);
            }
            mbar_wait(&res_bar[eg * 2 + (stage_ctr & 1u)], (stage_ctr >> 1) & 1u);
          }
          if (affine) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              float v = fmaf(__uint_as_float(r[j]), smRed[h * 32 + j], smRed[64 + h * 32 + j]);
              if (has_res && row_ok) {
                const __nv_bfloat16* rb = reinterpret_cast<const __nv_bfloat16*>(&resv[h * 4 + (j >> 3)]);
                v += __bfloat162float(rb[j & 7]);
              }
              if (res_tma) {
                if ((j & 7) == 0) resv[j >> 3] = *reinterpret_cast<const uint4*>(rowp + (((h * 4 + (j >> 3)) ^ (l & 7)) << 4));
                const __nv_bfloat16* rb = reinterpret_cast<const __nv_bfloat16*>(&resv[j >> 3]);
                v += __bfloat162float(rb[j & 7]);
              }
              if (P.ep_relu) v = fmaxf(v, 0.f);
              r[j] = __float_as_uint(v);
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
            v.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
            v.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
            v.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
            if (MODE == MODE_FPROP && !row_ok) v = make_uint4(0u, 0u, 0u, 0u);   // keep the fused statistics clean
            const int chunk = (h * 4 + j) ^ (l & 7);
            *reinterpret_cast<uint4*>(rowp + chunk * 16) = v;
          }
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (elected) {
          // strided outputs (the parity classes of a stride-2 dgrad): the output map walks every o_mul-th pixel
          const int sw = T.q0 * P.o_mul + P.o_off_w, sh = T.p0 * P.o_mul + P.o_off_h;
          if (P.accumulate) tma_reduce_add_4d(&tmO, stg, T.n0 + g * 64, sw, sh, T.i0);
          else tma_store_4d(&tmO, stg, T.n0 + g * 64, sw, sh, T.i0);
          tma_store_commit();
        }
        if (MODE == MODE_FPROP && P.stats != nullptr) {
          // Fused BatchNorm statistics: column sums of the bf16 tile that was just staged.  Thread (warp q, lane)
          // owns columns 2*lane, 2*lane+1 over rows 32q..32q+31 (one conflict-free 4-byte smem read per row) and
          // keeps the sums in registers until the CTA has finished all of its tiles.
#pragma unroll 8
          for (int i = 0; i < 32; ++i) {
            const int row = q * 32 + i;
            const uint32_t w = *reinterpret_cast<const uint32_t*>(stg + row * 128 + (((lane >> 2) ^ (row & 7)) << 4) +
                                                                   ((lane & 3) << 2));
            const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w));
            st_acc[gi][0] += f.x; st_acc[gi][1] = fmaf(f.x, f.x, st_acc[gi][1]);
            st_acc[gi][2] += f.y; st_acc[gi][3] = fmaf(f.y, f.y, st_acc[gi][3]);
          }
        }
        ++stage_ctr;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) mbar_arrive_leader(&tmem_empty[buf]);
        else mbar_arrive(&tmem_empty[buf]);
      }
      continue;
    }
#pragma unroll 1
    for (int c = ALT ? 0 : eg; c < BN / 32; c += ALT ? 1 : 2) {   // 32-column chunks alternate between the two epilogue groups
      uint32_t r[32];
      if (have_work) {
        tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + static_cast<uint32_t>(c * 32), r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0u;
      }
      const int col0 = T.n0 + c * 32;
      if (!row_ok || col0 >= P.ncols) continue;
      if (P.out_fp32) {
        float* p = reinterpret_cast<float*>(P.out) + row_off + col0;
        if (P.vec_ok && col0 + 32 <= P.ncols) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                   __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
            if (P.accumulate) {
              float4 o = reinterpret_cast<float4*>(p)[j];
              v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
            }
            reinterpret_cast<float4*>(p)[j] = v;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < P.ncols) p[j] = (P.accumulate ? p[j] : 0.f) + __uint_as_float(r[j]);
        }
      } else {
        __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(P.out) + row_off + col0;
        if (P.vec_ok && col0 + 32 <= P.ncols) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[8 * j + e]);
            if (P.accumulate) {
              uint4 o = reinterpret_cast<uint4*>(p)[j];
              const __nv_bfloat162* ob = reinterpret_cast<const __nv_bfloat162*>(&o);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float2 t2 = __bfloat1622float2(ob[e]);
                f[2 * e] += t2.x;
                f[2 * e + 1] += t2.y;
              }
            }
            uint4 v;
            v.x = pack_bf16x2(f[0], f[1]);
            v.y = pack_bf16x2(f[2], f[3]);
            v.z = pack_bf16x2(f[4], f[5]);
            v.w = pack_bf16x2(f[6], f[7]);
            reinterpret_cast<uint4*>(p)[j] = v;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < P.ncols) {
              float f = __uint_as_float(r[j]);
              if (P.accumulate) f += __bfloat162float(p[j]);
              p[j] = __float2bfloat16_rn(f);
            }
        }
      }
    }
    // this warp has finished reading the accumulator: hand it back to the MMA issuer
    tc_fence_before();
    __syncwarp();
    if (lane == 0) {
      if (CG == 2) mbar_arrive_leader(&tmem_empty[buf]);
      else mbar_arrive(&tmem_empty[buf]);
    }
  }
  if (MODE != MODE_WGRAD && P.tma_store && ((warp - 2) & 3) == 0 && lane == 0) tma_store_wait<0>();  // smem must outlive the stores
  if (MODE == MODE_FPROP && P.tma_store && P.stats != nullptr) {
    // one row of partial statistics per CTA: warps 1..3 of a group hand their sums to warp 0 through the group's
    // staging tile (idle now: its last store has completed)
    float* smFin = reinterpret_cast<float*>(smStage + eg * NSTG * CONV_STAGING_TILE);
    const int n0 = (blockIdx.x % P.n_tiles) * BN;
    if constexpr (ALT) {
      // both groups hold sums of the same 64 columns (over alternate tiles): the warps park theirs in their own group's
      // staging tile, the eight meet, one warp adds them up in a fixed order
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (n0 < P.ncols) {
        {
          float* r = smFin + (q * 64 + 2 * lane) * 2;
          r[0] = st_acc[0][0]; r[1] = st_acc[0][1]; r[2] = st_acc[0][2]; r[3] = st_acc[0][3];
        }
        asm volatile("bar.sync 3, 256;" ::: "memory");
        if (warp == 2) {
          float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll
          for (int w2 = 0; w2 < 8; ++w2) {
            const float* r = reinterpret_cast<const float*>(smStage + (w2 >> 2) * NSTG * CONV_STAGING_TILE) +
                             ((w2 & 3) * 64 + 2 * lane) * 2;
            a0 += r[0]; a1 += r[1]; b0 += r[2]; b1 += r[3];
          }
          float* dst = P.stats + (static_cast<long long>(blockIdx.x / P.n_tiles) * 2) * P.ncols + n0 + 2 * lane;
          *reinterpret_cast<float2*>(dst) = make_float2(a0, b0);
          *reinterpret_cast<float2*>(dst + P.ncols) = make_float2(a1, b1);
        }
      }
      return;
    }
#pragma unroll
    for (int gi = 0; gi < GPG; ++gi) {
      const int g = eg + 2 * gi;
      if (g >= BN / 64 || n0 + g * 64 >= P.ncols) break;
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (q != 0) {
        float* r = smFin + ((q - 1) * 64 + 2 * lane) * 2;
        r[0] = st_acc[gi][0]; r[1] = st_acc[gi][1]; r[2] = st_acc[gi][2]; r[3] = st_acc[gi][3];
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (q == 0) {
        float a0 = st_acc[gi][0], a1 = st_acc[gi][1], b0 = st_acc[gi][2], b1 = st_acc[gi][3];
#pragma unroll
        for (int w2 = 0; w2 < 3; ++w2) {
          const float* r = smFin + (w2 * 64 + 2 * lane) * 2;
          a0 += r[0]; a1 += r[1]; b0 += r[2]; b1 += r[3];
        }
        // row = blockIdx.x / n_tiles: the n_tiles CTAs of a row cover the column tiles exactly once
        float* dst = P.stats + (static_cast<long long>(blockIdx.x / P.n_tiles) * 2) * P.ncols + n0 + g * 64 + 2 * lane;
        *reinterpret_cast<float2*>(dst) = make_float2(a0, b0);
        *reinterpret_cast<float2*>(dst + P.ncols) = make_float2(a1, b1);
      }
    }
  }
}

// CG = 2 (fprop / dgrad, BN = 256): the CTA PAIR of a cluster shares one M = 256 MMA (tcgen05 cta_group::2).  Each
// CTA keeps its own pixel tile (A, accumulator, epilogue) but loads only HALF of the B tile -- the leader's MMA reads
// both halves -- so a k-block costs 16 + 16 KB of L2->SM traffic per CTA instead of 16 + 32 KB: the 128x256 tile was
// operand-traffic bound.  TMA completions of both CTAs are counted on the leader's `full` barrier, tcgen05.commit is
// multicast to both CTAs' `empty` / `tmem_full` barriers, and both CTAs' epilogue warps release the accumulator on
// the leader's `tmem_empty` barrier.
//
// CG = 3 (fprop / dgrad, BN = 256): ONE CTA works on both pixel tiles of a pair item -- a 256 x 256 output tile as two
// M = 128 MMAs per k-step that share the B tile.  A k-block then brings 2 x 16 + 32 KB for twice the MMA work
// (32 KB per 128x256x64 unit instead of 48): the single-tile kernel is bound by L2 -> SM operand traffic
// (~115 GB/s per SM: 0.43 us per 48 KB k-block against 0.28 us of MMA time).  The two accumulators fill TMEM, so the
// epilogue of an item no longer overlaps the next item's main loop: worth it when the reduction is deep.
template <int BN, int STAGES, int MODE, int CG = 1, int NSTG = 1>
__global__ void __launch_bounds__(CONV_THREADS, 1)
conv_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmO, const __grid_constant__ ConvParams P) {
  constexpr int A_BYTES = MODE == MODE_WGRAD ? 2 * CONV_WGRAD_KR(BN) * 128 : GEMM_BM * 128 * (CG == 3 ? 2 : 1);
  constexpr int B_BYTES = MODE == MODE_WGRAD ? (BN / 64) * CONV_WGRAD_KR(BN) * 128 : BN * 128 / (CG == 2 ? 2 : 1);
  static_assert(CG == 1 || (MODE != MODE_WGRAD && BN == 256), "CG 2 / 3: fprop / dgrad with 256-wide tiles");
  constexpr int CONV_STAGING_BYTES = 2 * NSTG * CONV_STAGING_TILE;
  constexpr int STG_BYTES = MODE == MODE_WGRAD ? 0 : CONV_STAGING_BYTES + CONV_STATS_RED_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;   // two accumulators
  static_assert(2 * BN <= 512, "two accumulators must fit TMEM");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  uint8_t* smA = smem;
  uint8_t* smB = smem + STAGES * A_BYTES;
  uint8_t* smStage = smem + STAGES * (A_BYTES + B_BYTES);   // 1024-aligned: stage sizes are multiples of 1024
  float* smRedBase = reinterpret_cast<float*>(smStage + CONV_STAGING_BYTES);   // per group: [128]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + B_BYTES) + STG_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;    // [2]
  uint64_t* tmem_empty = tmem_full + 2;    // [2]
  uint64_t* res_bar = tmem_empty + 2;      // [4]: residual tile landed, per epilogue group and staging tile
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // persistent walk: CTA b takes tiles b, b + grid, ...; a CTA pair takes pair items c, c + clusters, ... where item
  // w = (column tile w % n_tiles, pixel-tile pair w / n_tiles) and rank r of the pair owns pixel tile 2*pair + r
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
  const int walk_id = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int walk_n = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int walk_total = CG != 1 ? P.total_pairs : P.total_tiles;
  auto tile_of = [&](int w, int r) -> int {
    if (CG == 1) return w;
    return (w % P.n_tiles) + P.n_tiles * (2 * (w / P.n_tiles) + r);
  };
  const uint32_t box_bytes = static_cast<uint32_t>(P.bw * P.bh * P.bn) * 128u;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (MODE != MODE_WGRAD && P.tma_store) tma_prefetch_desc(&tmO);
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], conv_tmem_empty_arrivals<BN, CG>());   // one arrive per epilogue warp (of both CTAs of a pair)
    }
#pragma unroll
    for (int b = 0; b < 4; ++b) mbar_init(&res_bar[b], 1);
    mbar_fence_init();
  }
  if (CG == 2) cluster_sync_all();      // both CTAs' barriers exist before anything remote touches them
  if (warp == 1) {
    if (CG == 2) {
      tmem_alloc_2sm(tmem_slot, TMEM_COLS);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc(tmem_slot, TMEM_COLS);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;          // ring position and phase run continuously across tiles
      uint32_t phase = 0;
      for (int w = walk_id; w < walk_total; w += walk_n) {
        const int tile = tile_of(w, rank);
        const ConvTile T = conv_decode_tile<BN, MODE>(P, tile);
        if (MODE != MODE_WGRAD) {
          const ConvTile T1 = conv_decode_tile<BN, MODE>(P, tile_of(w, 1));   // CG == 3: the item's second pixel tile
          // taps x 64-channel chunks, no per-k-block division; stage index / phase carried incrementally
          for (int t = 0; t < P.ntaps; ++t) {
            const int aw = T.q0 * P.a_mul + P.taps[t].dw, ah = T.p0 * P.a_mul + P.taps[t].dh;
            const int bcol = P.taps[t].bcol;
            for (int cc = 0; cc < P.cchunks; ++cc) {
              mbar_wait(&empty[stage], phase ^ 1);
              uint8_t* a = smA + stage * A_BYTES;
              uint8_t* b = smB + stage * B_BYTES;
              if (CG == 3) {
                mbar_expect_tx(&full[stage], 2 * box_bytes + B_BYTES);
                tma_load_4d(a, &tmA, &full[stage], cc * 64, aw, ah, T.i0);
                tma_load_4d(a + GEMM_BM * 128, &tmA, &full[stage], cc * 64, T1.q0 * P.a_mul + P.taps[t].dw,
                            T1.p0 * P.a_mul + P.taps[t].dh, T1.i0);
                if (MODE == MODE_FPROP) {
                  tma_load_2d(b, &tmB, &full[stage], bcol + cc * 64, T.n0);
                } else {
#pragma unroll
                  for (int h = 0; h < BN / 64; ++h)
                    tma_load_2d(b + h * 8192, &tmB, &full[stage], bcol + T.n0 + h * 64, cc * 64);
                }
              } else if (CG == 2) {
                // bytes of BOTH CTAs land on the leader's barrier; each CTA brings its pixel box and its half of B
                if (rank == 0) mbar_expect_tx(&full[stage], 2 * (box_bytes + B_BYTES));
                tma_load_4d_2sm(a, &tmA, &full[stage], cc * 64, aw, ah, T.i0);
                if (MODE == MODE_FPROP) {
                  tma_load_2d_2sm(b, &tmB, &full[stage], bcol + cc * 64, T.n0 + rank * (BN / 2));  // {64 c, BN/2 cout}
                } else {
#pragma unroll
                  for (int h = 0; h < BN / 128; ++h)
                    tma_load_2d_2sm(b + h * 8192, &tmB, &full[stage], bcol + T.n0 + (rank * (BN / 128) + h) * 64,
                                    cc * 64);
                }
              } else {
              mbar_expect_tx(&full[stage], box_bytes + B_BYTES);
              tma_load_4d(a, &tmA, &full[stage], cc * 64, aw, ah, T.i0);
              if (MODE == MODE_FPROP) {
                tma_load_2d(b, &tmB, &full[stage], bcol + cc * 64, T.n0);  // box {64 c, BN cout}
              } else {
#pragma unroll
                for (int h = 0; h < BN / 64; ++h)
                  tma_load_2d(b + h * 8192, &tmB, &full[stage], bcol + T.n0 + h * 64, cc * 64);  // {64 cin, 64 k}
              }
              }
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
          }
          continue;
        }
        // pixel-box walk (w fastest, then h, then n) carried incrementally from the first box of this split
        int tw = T.kb_begin % P.tiles_w, th = (T.kb_begin / P.tiles_w) % P.tiles_h,
            tn = T.kb_begin / (P.tiles_w * P.tiles_h);
        for (int kb = T.kb_begin; kb < T.kb_end; ++kb) {
          const int s = stage;
          mbar_wait(&empty[s], phase ^ 1);
          uint8_t* a = smA + s * A_BYTES;
          uint8_t* b = smB + s * B_BYTES;
          {
            const int bq = tw * P.bw, bp = th * P.bh, bi = tn * P.bn;
            if (++tw == P.tiles_w) { tw = 0; if (++th == P.tiles_h) { th = 0; ++tn; } }
            if (BN == 128) {
              // one tap per tile (C >= 128): fixed, fully unrolled issue sequence
              const int cin0 = T.n0 - T.tap * P.cdim;
              mbar_expect_tx(&full[s], box_bytes * (2 + BN / 64));
#pragma unroll
              for (int h = 0; h < 2; ++h) tma_load_4d(a + h * box_bytes, &tmA, &full[s], T.m0 + h * 64, bq, bp, bi);
#pragma unroll
              for (int h = 0; h < BN / 64; ++h)
                tma_load_4d(b + h * box_bytes, &tmB, &full[s], cin0 + h * 64, bq * P.x_mul + P.taps[T.tap].dw,
                            bp * P.x_mul + P.taps[T.tap].dh, bi);
            } else {
              // 256-column tile of the packed gradient (C == 64): each 64-column group is its own tap
              const int na = (T.m0 + 64 < P.mrows) ? 2 : 1;          // second 64-cout half only if it exists
              int nb = (P.ncols - T.n0 + 63) / 64;
              if (nb > BN / 64) nb = BN / 64;
              mbar_expect_tx(&full[s], box_bytes * (na + nb));
              for (int h = 0; h < na; ++h) tma_load_4d(a + h * box_bytes, &tmA, &full[s], T.m0 + h * 64, bq, bp, bi);
              if (P.cdim == 64) {
#pragma unroll
                for (int h = 0; h < BN / 64; ++h) {
                  if (h < nb) {
                    const int tp = (T.n0 >> 6) + h;   // C == 64: column group = tap
                    tma_load_4d(b + h * box_bytes, &tmB, &full[s], 0, bq * P.x_mul + P.taps[tp].dw,
                                bp * P.x_mul + P.taps[tp].dh, bi);
                  }
                }
              } else {
                // C is a multiple of 256: the whole 256-column tile lies inside one tap
                const int cin0 = T.n0 - T.tap * P.cdim;
                const int xw = bq * P.x_mul + P.taps[T.tap].dw, xh = bp * P.x_mul + P.taps[T.tap].dh;
#pragma unroll
                for (int h = 0; h < BN / 64; ++h)
                  if (h < nb) tma_load_4d(b + h * box_bytes, &tmB, &full[s], cin0 + h * 64, xw, xh, bi);
              }
            }
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if constexpr (MODE != MODE_WGRAD && CG == 1) {
      // fprop / dgrad, one CTA per tile.  tools/umma_rate_probe.cu: the tensor pipe takes 128 * N / 256 cycles per MMA
      // only for N = 256 (62 cycles at N = 64, 75 at N = 128) and buffers barely two MMAs ahead of the issuing
      // thread, so whatever that thread does between k-blocks beyond ~2 MMA times is added to the tile.  Hence: the
      // ring is walked with the stage index UNROLLED (barrier addresses and descriptors are immediates off one base),
      // the reduction length is the same for every tile (no tile decode here) and the tile boundary is a counter.
      if (elect_one()) {
        constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, false, MODE != MODE_FPROP);
        const int nkb = P.ntaps * P.cchunks;
        int tiles_left = walk_id < walk_total ? (walk_total - walk_id + walk_n - 1) / walk_n : 0;
        const uint64_t a_desc0 = umma_smem_desc(smem_u32(smA), 16, 1024);
        const uint64_t b_desc0 = MODE == MODE_FPROP ? umma_smem_desc(smem_u32(smB), 16, 1024) : umma_smem_desc(smem_u32(smB), 8192, 1024);
        uint32_t phase = 0, it = 0;
        int kb = 0;
        uint32_t d_tmem = tmem_base;
        if (tiles_left > 0) {
          mbar_wait(&tmem_empty[0], 1);
          tc_fence_after();
        }
        while (tiles_left > 0) {
#pragma unroll
          for (int s = 0; s < STAGES; ++s) {
            mbar_wait(&full[s], phase);
            tc_fence_after();
            umma_f16_x4(d_tmem, a_desc0 + static_cast<uint64_t>(s * (A_BYTES >> 4)), b_desc0 + static_cast<uint64_t>(s * (B_BYTES >> 4)),
                        2u, MODE == MODE_FPROP ? 2u : 128u, idesc, kb ? 1u : 0u);
            umma_commit(&empty[s]);
            if (++kb == nkb) {
              umma_commit(&tmem_full[it & 1]);
              kb = 0;
              ++it;
              if (--tiles_left == 0) break;
              mbar_wait(&tmem_empty[it & 1], ((it >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
              tc_fence_after();
              d_tmem = tmem_base + (it & 1) * BN;
            }
          }
          phase ^= 1;
        }
      }
    } else if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM * (CG == 2 ? 2 : 1), BN, MODE == MODE_WGRAD, MODE != MODE_FPROP);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;   // local tile counter
      for (int w = walk_id; w < walk_total; w += walk_n, ++it) {
        const int tile = tile_of(w, rank);
        const ConvTile T = conv_decode_tile<BN, MODE>(P, tile);
        const uint32_t buf = CG == 3 ? 0u : (it & 1), use = CG == 3 ? it : (it >> 1);
        mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);   // epilogue has drained this accumulator
        if (CG == 3) mbar_wait(&tmem_empty[1], (use & 1) ^ 1);   // ... and the second one: an item uses both
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        bool first = true;
        for (int kb = T.kb_begin; kb < T.kb_end; ++kb) {
          const int s = stage;
          mbar_wait(&full[s], phase);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smA + s * A_BYTES);
          const uint32_t b_addr = smem_u32(smB + s * B_BYTES);
          if (MODE != MODE_WGRAD && CG == 1) {
            // one asm block per k-block: four k-steps, descriptors advanced inside (see umma_f16_x4)
            umma_f16_x4(d_tmem, umma_smem_desc(a_addr, 16, 1024),
                        MODE == MODE_FPROP ? umma_smem_desc(b_addr, 16, 1024) : umma_smem_desc(b_addr, 8192, 1024), 2u,
                        MODE == MODE_FPROP ? 2u : 128u, idesc, first ? 0u : 1u);
          } else if (MODE != MODE_WGRAD) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t adesc = umma_smem_desc(a_addr + k * 32, 16, 1024);
              const uint64_t bdesc = MODE == MODE_FPROP ? umma_smem_desc(b_addr + k * 32, 16, 1024)
                                                        : umma_smem_desc(b_addr + k * 2048, 8192, 1024);
              if (CG == 2) umma_f16_2sm(d_tmem, adesc, bdesc, idesc, (!first || k > 0) ? 1u : 0u);
              else umma_f16(d_tmem, adesc, bdesc, idesc, (!first || k > 0) ? 1u : 0u);
              if (CG == 3)   // second pixel tile of the item: same B, its own A rows and accumulator
                umma_f16(d_tmem + BN, umma_smem_desc(a_addr + GEMM_BM * 128 + k * 32, 16, 1024), bdesc, idesc,
                         (!first || k > 0) ? 1u : 0u);
            }
          } else {
            const int ksteps = P.kr >> 4;
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t adesc = umma_smem_desc(a_addr + k * 2048, box_bytes, 1024);
              const uint64_t bdesc = umma_smem_desc(b_addr + k * 2048, box_bytes, 1024);
              umma_f16(d_tmem, adesc, bdesc, idesc, (!first || k > 0) ? 1u : 0u);
            }
          }
          first = false;
          if (CG == 2) umma_commit_2sm(&empty[s]);
          else umma_commit(&empty[s]);
        }
        if (CG == 2) umma_commit_2sm(&tmem_full[buf]);
        else umma_commit(&tmem_full[buf]);
        if (CG == 3) umma_commit(&tmem_full[1]);
      }
    }
    __syncwarp();
  } else {
    conv_epilogue_warps<BN, MODE, CG, NSTG>(tmO, P, smStage, smRedBase, tmem_full, tmem_empty, tmem_base, warp, lane, rank,
                                            walk_id, walk_n, walk_total, res_bar);
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all();     // the peer may still be reading this CTA's shared memory / signalling its barriers
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host side
// bf16 4-D map over an NHWC tensor [N][H][W][C]: box {64, bw*es, bh*es, bn}, traversal stride es along W and H.
inline int make_tmap_bf16_nhwc(CUtensorMap* m, const void* base, int N, int H, int W, int C, int bw, int bh, int bn,
                               int es) {
  EncodeTiledFn enc = tma_encode_fn();
  if (!enc) return -1;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)(bw * es), (cuuint32_t)(bh * es), (cuuint32_t)bn};
  cuuint32_t estr[4] = {1, (cuuint32_t)es, (cuuint32_t)es, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : static_cast<int>(r);
}

struct PixelBox {
  int bw, bh, bn, tiles_w, tiles_h, tiles_n;
};
// Chooses the pixel box {bw,bh,bn} (bw*bh*bn <= max_rows, multiple of `mult`, each TMA box dim <= 256 after the
// element stride) that covers a W x H x N pixel grid with the fewest box rows in total.
inline PixelBox choose_pixel_box(int W, int H, int N, int max_rows, int mult, int es) {
  PixelBox best{1, 1, 1, W, H, N};
  long long best_rows = -1;
  const int wlim = 256 / es;
  for (int bw = 1; bw <= W && bw <= max_rows && bw <= wlim; ++bw) {
    for (int bh = 1; bh <= H && bw * bh <= max_rows && bh <= wlim; ++bh) {
      int bn = max_rows / (bw * bh);
      if (bn > N) bn = N;
      if (bn > 256) bn = 256;
      for (; bn >= 1; --bn)
        if ((bw * bh * bn) % mult == 0) break;
      if (bn < 1) continue;
      const long long tw = (W + bw - 1) / bw, th = (H + bh - 1) / bh, tn = (N + bn - 1) / bn;
      const long long rows = tw * th * tn * (mult == 1 ? (long long)max_rows : (long long)bw * bh * bn);
      if (best_rows < 0 || rows < best_rows || (rows == best_rows && bw > best.bw)) {
        best_rows = rows;
        best = PixelBox{bw, bh, bn, (int)tw, (int)th, (int)tn};
      }
    }
  }
  return best;
}

// Grid of a fprop launch that also produces BN statistics: a multiple of n_tiles (so that tile = blockIdx.x +
// i*gridDim keeps every CTA on one output-column tile) and never more than the number of tiles.  The launch writes
// gridDim / n_tiles rows of partial statistics (CTA b fills its column tile of row b / n_tiles).
inline int conv_stats_grid(int total_tiles, int n_tiles, int num_ctas) {
  int g = total_tiles < num_ctas ? total_tiles : num_ctas;
  g = g / n_tiles * n_tiles;
  return g < n_tiles ? n_tiles : g;
}

template <int BN, int STAGES, int MODE, int CG = 1, int NSTG = 1>
inline int launch_conv_cfg(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const ConvParams& P,
                           int num_ctas, cudaStream_t st) {
  auto kern = conv_sm100_kernel<BN, STAGES, MODE, CG, NSTG>;
  constexpr int smem = conv_smem_bytes<BN, STAGES, MODE, CG, NSTG>();
  static_assert(smem <= 232448, "conv kernel: shared memory budget exceeded");
  static int attr_state[B200_MAX_DEVICES] = {0};
  if (cudaError_t e = ensure_dyn_smem(kern, smem, attr_state)) return 300 + static_cast<int>(e);
  int grid = P.total_tiles < num_ctas ? P.total_tiles : num_ctas;
  if (CG == 3) grid = P.total_pairs < num_ctas ? P.total_pairs : num_ctas;
  if (MODE == MODE_FPROP && P.stats != nullptr) grid = conv_stats_grid(P.total_tiles, P.n_tiles, num_ctas);
  if (CG == 2) {
    // one cluster (CTA pair) per two SMs, persistent over the pair items
    int clusters = P.total_pairs < num_ctas / 2 ? P.total_pairs : num_ctas / 2;
    if (clusters < 1) clusters = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(CONV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e2 = cudaLaunchKernelEx(&cfg, kern, ta, tb, to, P);
    return e2 == cudaSuccess ? 0 : 400 + static_cast<int>(e2);
  }
  kern<<<grid, CONV_THREADS, smem, st>>>(ta, tb, to, P);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : 400 + static_cast<int>(e);
}

}  // namespace b200
