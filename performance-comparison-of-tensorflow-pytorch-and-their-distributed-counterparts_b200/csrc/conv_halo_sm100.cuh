// Weights-stationary convolution fprop / dgrad: the CTA's whole weight tile stays in shared memory, only activations
// stream.
//
// conv_sm100_kernel fetches, for every output tile, one activation box AND one weight box per (filter tap, 64-channel
// chunk).  Two things are wasted there (ncu on SURVEY 8(a') class b, 64 -> 64 3x3 at 40x40: 707 MB through the
// L2 -> SM path for 105 MB of tensors, the tensor pipe 26 % busy):
//   * the weight tile is the same for every tile a CTA walks (the launcher makes the grid a multiple of the number of
//     output-column tiles), yet it is re-read from L2 for each of them;
//   * a 3x3 layer's nine activation boxes overlap almost completely.
// Here the weights of the CTA's column tile -- all taps x chunks, at most BSLOTS k-blocks -- are loaded ONCE into
// their own slots (a barrier per slot, so the first tile starts as soon as the first slot lands) and the activation
// ring carries
//   * 1x1 layers: the usual 128-pixel box per channel chunk (A_ROWS = 128, NTAPS = 1), or
//   * "same" 3x3 stride-1 layers: per channel chunk ONE box {64 ch, W + 2, bh + 2, 1 image} -- bh full image rows plus
//     their halo, padding zero-filled by the TMA unit -- and the tap (r, s) is a DESCRIPTOR SHIFT: the MMA's A operand
//     starts (r * Wh + s) rows into the halo tile (Wh = W + 2).  tcgen05 applies the 128-byte swizzle on absolute
//     shared-memory address bits, so an operand may start at any 128-byte row of a tile that TMA wrote
//     (tools/umma_shift_probe.cu verifies this on B200).  The accumulator's M rows then run through the halo tile
//     row-major: row m = ih * Wh + iw, of which iw >= W are junk columns that the output TMA store clips (its box is
//     {64, Wh, bh, 1} at w = 0) and the fused statistics skip.
// The MMA issuer's loop has no weight barriers after the first tile and the taps are unrolled (tools/umma_rate_probe.cu:
// whatever the issuing thread does between MMAs beyond ~2 MMA times is added to the tile).
// Epilogue = conv_epilogue_warps of conv_sm100.cuh, unchanged (TMA store / reduce-add, fused BatchNorm statistics,
// affine + residual + ReLU).
#pragma once
#include "conv_sm100.cuh"

namespace b200 {

constexpr int HALO_ROWS = 256;                  // rows (pixels incl. halo) a 3x3 activation stage holds: 32 KB

struct HaloParams {
  int wh;                    // 3x3: halo tile width W + S - 1
  int rows;                  // rows per activation box (bytes = rows * 128)
  int org_w, org_h;          // box origin relative to the tile origin (-pad for fprop)
  int toff[CONV_MAX_TAPS];   // descriptor shift of tap t in rows: (dh - org_h) * wh + (dw - org_w)
};

template <int BN, int BSLOTS, int NA, int A_ROWS>
constexpr int conv_ws_smem_bytes() {
  return NA * A_ROWS * 128 + BSLOTS * BN * 128 + 2 * CONV_STAGING_TILE + CONV_STATS_RED_BYTES + 1024 + 768;
}

// RESB = false: the weight tile does not fit (3x3 layers with 128+ channels): the weight slots are a ring that streams
// one k-block per (tile, chunk, tap), the activations still come as one halo box per chunk.
template <int BN, int BSLOTS, int MODE, int NA, int A_ROWS, int NTAPS, bool RESB = true>
__global__ void __launch_bounds__(CONV_THREADS, 1)
conv_ws_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmO, const __grid_constant__ ConvParams P,
                     const __grid_constant__ HaloParams HP) {
  static_assert(MODE == MODE_FPROP || MODE == MODE_DGRAD, "weights-stationary kernel: fprop / dgrad");
  constexpr int A_BYTES = A_ROWS * 128, B_BYTES = BN * 128;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  uint8_t* smA = smem;
  uint8_t* smB = smem + NA * A_BYTES;
  uint8_t* smStage = smB + BSLOTS * B_BYTES;
  float* smRedBase = reinterpret_cast<float*>(smStage + 2 * CONV_STAGING_TILE);
  uint64_t* b_full = reinterpret_cast<uint64_t*>(smStage + 2 * CONV_STAGING_TILE + CONV_STATS_RED_BYTES);
  uint64_t* b_empty = b_full + BSLOTS;     // ring mode only
  uint64_t* a_full = b_empty + BSLOTS;
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
    for (int s = 0; s < BSLOTS; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
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
  const uint32_t a_bytes = static_cast<uint32_t>(HP.rows) * 128u;
  const int cchunks = P.cchunks;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one() && walk_id < walk_total) {
      // the weight tile of this CTA's output columns (the same for every tile it walks), slot = chunk * NTAPS + tap
      const int n0 = conv_decode_tile<BN, MODE>(P, walk_id).n0;
      auto load_b = [&](int slot, int cc, int t) {
        uint8_t* b = smB + slot * B_BYTES;
        mbar_expect_tx(&b_full[slot], B_BYTES);
        if (MODE == MODE_FPROP) {
          tma_load_2d(b, &tmB, &b_full[slot], P.taps[t].bcol + cc * 64, n0);               // box {64 c, BN cout}
        } else {
#pragma unroll
          for (int h = 0; h < BN / 64; ++h)
            tma_load_2d(b + h * 8192, &tmB, &b_full[slot], P.taps[t].bcol + n0 + h * 64, cc * 64);   // {64 cin, 64 k}
        }
      };
      if (RESB) {
        for (int cc = 0; cc < cchunks; ++cc)
#pragma unroll
          for (int t = 0; t < NTAPS; ++t) load_b(cc * NTAPS + t, cc, t);
      }
      int as = 0, bs = 0;
      uint32_t aph = 0, bph = 0;
      for (int w = walk_id; w < walk_total; w += walk_n) {
        const ConvTile T = conv_decode_tile<BN, MODE>(P, w);
        for (int cc = 0; cc < cchunks; ++cc) {
          mbar_wait(&a_empty[as], aph ^ 1);
          mbar_expect_tx(&a_full[as], a_bytes);
          tma_load_4d(smA + as * A_BYTES, &tmA, &a_full[as], cc * 64, T.q0 * P.a_mul + HP.org_w, T.p0 * P.a_mul + HP.org_h, T.i0);
          if (++as == NA) { as = 0; aph ^= 1; }
          if (!RESB) {
#pragma unroll 1
            for (int t = 0; t < NTAPS; ++t) {
              mbar_wait(&b_empty[bs], bph ^ 1);
              load_b(bs, cc, t);
              if (++bs == BSLOTS) { bs = 0; bph ^= 1; }
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, false, MODE != MODE_FPROP);
      int tiles_left = walk_id < walk_total ? (walk_total - walk_id + walk_n - 1) / walk_n : 0;
      const uint64_t a_desc0 = umma_smem_desc(smem_u32(smA), 16, 1024);
      const uint64_t b_desc0 = MODE == MODE_FPROP ? umma_smem_desc(smem_u32(smB), 16, 1024) : umma_smem_desc(smem_u32(smB), 8192, 1024);
      uint32_t toff[NTAPS];                 // tap shifts in descriptor units (16 bytes): rows * 128 / 16
#pragma unroll
      for (int t = 0; t < NTAPS; ++t) toff[t] = static_cast<uint32_t>(HP.toff[t]) * 8u;
      int as = 0, bs = 0;
      uint32_t aph = 0, bph = 0, it = 0;
      bool first_tile = true;
      while (tiles_left > 0) {
        mbar_wait(&tmem_empty[it & 1], ((it >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (it & 1) * BN;
        for (int cc = 0; cc < cchunks; ++cc) {
          mbar_wait(&a_full[as], aph);
          tc_fence_after();
          const uint64_t a_desc = a_desc0 + static_cast<uint64_t>(as * (A_BYTES >> 4));
          const uint64_t b_desc = b_desc0 + static_cast<uint64_t>(cc * NTAPS * (B_BYTES >> 4));
#pragma unroll
          for (int t = 0; t < NTAPS; ++t) {
            if (RESB) {
              if (first_tile) {
                mbar_wait(&b_full[cc * NTAPS + t], 0);
                tc_fence_after();
              }
              umma_f16_x4(d_tmem, a_desc + toff[t], b_desc + static_cast<uint64_t>(t * (B_BYTES >> 4)), 2u,
                          MODE == MODE_FPROP ? 2u : 128u, idesc, (cc | t) ? 1u : 0u);
            } else {
              mbar_wait(&b_full[bs], bph);
              tc_fence_after();
              umma_f16_x4(d_tmem, a_desc + toff[t], b_desc0 + static_cast<uint64_t>(bs * (B_BYTES >> 4)), 2u,
                          MODE == MODE_FPROP ? 2u : 128u, idesc, (cc | t) ? 1u : 0u);
              umma_commit(&b_empty[bs]);
              if (++bs == BSLOTS) { bs = 0; bph ^= 1; }
            }
          }
          umma_commit(&a_empty[as]);          // every tap of this chunk has been issued: the stage is free once they finish
          if (++as == NA) { as = 0; aph ^= 1; }
        }
        umma_commit(&tmem_full[it & 1]);
        first_tile = false;
        ++it;
        --tiles_left;
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

// grid: a multiple of the number of output-column tiles (so that a CTA keeps one weight tile), never more than the tiles
template <int BN, int BSLOTS, int MODE, int NA, int A_ROWS, int NTAPS, bool RESB = true>
inline int launch_conv_ws_cfg(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const ConvParams& P,
                              const HaloParams& HP, int num_ctas, cudaStream_t st) {
  auto kern = conv_ws_sm100_kernel<BN, BSLOTS, MODE, NA, A_ROWS, NTAPS, RESB>;
  constexpr int smem = conv_ws_smem_bytes<BN, BSLOTS, NA, A_ROWS>();
  static_assert(smem <= 232448, "weights-stationary conv kernel: shared memory budget exceeded");
  if (P.ntaps != NTAPS || (RESB && P.ntaps * P.cchunks > BSLOTS) || HP.rows > A_ROWS) return 250;
  static int attr_state[B200_MAX_DEVICES] = {0};
  if (cudaError_t e = ensure_dyn_smem(kern, smem, attr_state)) return 300 + static_cast<int>(e);
  const int grid = conv_stats_grid(P.total_tiles, P.n_tiles, num_ctas);
  kern<<<grid, CONV_THREADS, smem, st>>>(ta, tb, to, P, HP);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : 400 + static_cast<int>(e);
}

}  // namespace b200
