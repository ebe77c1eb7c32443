// 3x3 (R x S, stride 1) convolution fprop / dgrad with the activation HALO TILE RESIDENT in shared memory.
//
// conv_sm100_kernel fetches one pixel box per (filter tap, 64-channel chunk): nine boxes that overlap almost
// completely, so a 3x3 layer pulls its input through the L2 -> SM path nine times.  For the narrow layers (C = 64:
// 128 x 64 tiles, 8 KB of weights per k-block) that traffic is the whole cost: class b of SURVEY 8(a') ran at 0.26 of
// its roofline with the tensor pipe 20 % busy.  Here a tile is `bh` FULL image rows; per channel chunk ONE box
// {64 ch, W + S - 1, bh + R - 1, 1 image} -- the rows plus their halo, padding zero-filled by the TMA unit -- lands in
// shared memory, and the tap (r, s) is a DESCRIPTOR SHIFT: the MMA's A operand starts (r * Wh + s) rows into the
// halo tile (Wh = W + S - 1).  tcgen05 applies the 128-byte swizzle on absolute shared-memory address bits, so an
// operand may start at any 128-byte row of a tile that TMA wrote (tools/umma_shift_probe.cu verifies this on B200).
// The accumulator's M rows then run through the halo tile row-major: row m = ih * Wh + iw, of which iw >= W are junk
// columns that the output TMA store clips (its box is {64, Wh, bh, 1} at w = 0) and the statistics skip.
// Weights still stream per (tap, chunk) through their own ring.  Epilogue = conv_epilogue_warps of conv_sm100.cuh
// with bw := Wh, i.e. unchanged (TMA store / reduce-add, fused BatchNorm statistics, affine + residual + ReLU).
#pragma once
#include "conv_sm100.cuh"

namespace b200 {

constexpr int HALO_ROWS = 256;                  // rows (pixels incl. halo) a halo stage holds: 32 KB
constexpr int HALO_A_BYTES = HALO_ROWS * 128;

struct HaloParams {
  int wh;                    // halo tile width  W + S - 1
  int rows;                  // halo tile rows per box = (bh + R - 1) * wh  (<= HALO_ROWS - what the last tap overshoots)
  int org_w, org_h;          // box origin relative to the tile origin (-pad for fprop)
  int toff[CONV_MAX_TAPS];   // descriptor shift of tap t in rows: (dh - org_h) * wh + (dw - org_w)
};

template <int BN, int BSTAGES, int MODE, int NA>
constexpr int conv_halo_smem_bytes() {
  return NA * HALO_A_BYTES + BSTAGES * BN * 128 + 2 * CONV_STAGING_TILE + CONV_STATS_RED_BYTES + 1024 + 256;
}

template <int BN, int BSTAGES, int MODE, int NA>
__global__ void __launch_bounds__(CONV_THREADS, 1)
conv_halo_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmO, const __grid_constant__ ConvParams P,
                       const __grid_constant__ HaloParams HP) {
  static_assert(MODE == MODE_FPROP || MODE == MODE_DGRAD, "halo kernel: fprop / dgrad");
  constexpr int B_BYTES = BN * 128;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  uint8_t* smA = smem;
  uint8_t* smB = smem + NA * HALO_A_BYTES;
  uint8_t* smStage = smB + BSTAGES * B_BYTES;
  float* smRedBase = reinterpret_cast<float*>(smStage + 2 * CONV_STAGING_TILE);
  uint64_t* b_full = reinterpret_cast<uint64_t*>(smStage + 2 * CONV_STAGING_TILE + CONV_STATS_RED_BYTES);
  uint64_t* b_empty = b_full + BSTAGES;
  uint64_t* a_full = b_empty + BSTAGES;
  uint64_t* a_empty = a_full + NA;
  uint64_t* tmem_full = a_empty + NA;      // [2]
  uint64_t* tmem_empty = tmem_full + 2;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int walk_id = (int)blockIdx.x, walk_n = (int)gridDim.x, walk_total = P.total_tiles;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
    for (int s = 0; s < BSTAGES; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < NA; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], conv_tmem_empty_arrivals<BN, 1>()); }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t halo_bytes = static_cast<uint32_t>(HP.rows) * 128u;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int as = 0, bs = 0;
      uint32_t aph = 0, bph = 0;
      for (int w = walk_id; w < walk_total; w += walk_n) {
        const ConvTile T = conv_decode_tile<BN, MODE>(P, w);
        for (int cc = 0; cc < P.cchunks; ++cc) {
          mbar_wait(&a_empty[as], aph ^ 1);
          mbar_expect_tx(&a_full[as], halo_bytes);
          tma_load_4d(smA + as * HALO_A_BYTES, &tmA, &a_full[as], cc * 64, T.q0 + HP.org_w, T.p0 + HP.org_h, T.i0);
          if (++as == NA) { as = 0; aph ^= 1; }
          for (int t = 0; t < P.ntaps; ++t) {
            mbar_wait(&b_empty[bs], bph ^ 1);
            uint8_t* b = smB + bs * B_BYTES;
            mbar_expect_tx(&b_full[bs], B_BYTES);
            if (MODE == MODE_FPROP) {
              tma_load_2d(b, &tmB, &b_full[bs], P.taps[t].bcol + cc * 64, T.n0);            // box {64 c, BN cout}
            } else {
#pragma unroll
              for (int h = 0; h < BN / 64; ++h)
                tma_load_2d(b + h * 8192, &tmB, &b_full[bs], P.taps[t].bcol + T.n0 + h * 64, cc * 64);  // {64 cin, 64 k}
            }
            if (++bs == BSTAGES) { bs = 0; bph ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, false, MODE != MODE_FPROP);
      int as = 0, bs = 0;
      uint32_t aph = 0, bph = 0, it = 0;
      for (int w = walk_id; w < walk_total; w += walk_n, ++it) {
        const uint32_t buf = it & 1, use = it >> 1;
        mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        bool first = true;
        for (int cc = 0; cc < P.cchunks; ++cc) {
          mbar_wait(&a_full[as], aph);
          const uint32_t halo_addr = smem_u32(smA + as * HALO_A_BYTES);
          for (int t = 0; t < P.ntaps; ++t) {
            mbar_wait(&b_full[bs], bph);
            tc_fence_after();
            const uint32_t a_addr = halo_addr + static_cast<uint32_t>(HP.toff[t]) * 128u;   // the tap: a row shift
            const uint32_t b_addr = smem_u32(smB + bs * B_BYTES);
            umma_f16_x4(d_tmem, umma_smem_desc(a_addr, 16, 1024),
                        MODE == MODE_FPROP ? umma_smem_desc(b_addr, 16, 1024) : umma_smem_desc(b_addr, 8192, 1024), 2u,
                        MODE == MODE_FPROP ? 2u : 128u, idesc, first ? 0u : 1u);
            first = false;
            umma_commit(&b_empty[bs]);
            if (++bs == BSTAGES) { bs = 0; bph ^= 1; }
          }
          umma_commit(&a_empty[as]);          // every tap of this chunk has been issued: the halo tile is free once they finish
          if (++as == NA) { as = 0; aph ^= 1; }
        }
        umma_commit(&tmem_full[buf]);
      }
    }
    __syncwarp();
  } else {
    conv_epilogue_warps<BN, MODE, 1, 1>(tmO, P, smStage, smRedBase, tmem_full, tmem_empty, tmem_base, warp, lane, 0,
                                        walk_id, walk_n, walk_total);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, int BSTAGES, int MODE, int NA>
inline int launch_conv_halo_cfg(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const ConvParams& P,
                                const HaloParams& HP, int num_ctas, cudaStream_t st) {
  auto kern = conv_halo_sm100_kernel<BN, BSTAGES, MODE, NA>;
  constexpr int smem = conv_halo_smem_bytes<BN, BSTAGES, MODE, NA>();
  static_assert(smem <= 232448, "halo conv kernel: shared memory budget exceeded");
  static int attr_state[B200_MAX_DEVICES] = {0};
  if (cudaError_t e = ensure_dyn_smem(kern, smem, attr_state)) return 300 + static_cast<int>(e);
  int grid = P.total_tiles < num_ctas ? P.total_tiles : num_ctas;
  if (MODE == MODE_FPROP && P.stats != nullptr) grid = conv_stats_grid(P.total_tiles, P.n_tiles, num_ctas);
  kern<<<grid, CONV_THREADS, smem, st>>>(ta, tb, to, P, HP);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : 400 + static_cast<int>(e);
}

}  // namespace b200
