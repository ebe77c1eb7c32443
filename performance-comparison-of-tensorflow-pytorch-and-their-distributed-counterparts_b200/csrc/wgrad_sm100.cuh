// Weight gradient of a convolution on tcgen05 for sm_100a, second generation:
//
//     dW[cout][(tap, cin)] = sum over pixels  dY[pixel][cout] * X[pixel shifted by tap][cin]
//
// Why a kernel of its own.  The reduction runs over N*Ho*Wo pixels (up to 1.6 M) while the output is small, so the
// work has to be split along the reduction, and the operands are re-read once per output tile.  Measured with the
// first-generation kernel (128 x 128 tiles, fp32 slabs + a reduce kernel): every layer ran at the L2 -> SM bandwidth
// limit (~6300 B/clk for the whole chip, i.e. ~43 B/clk per SM: a 128x128x128 k-block wants 64 KB per 276 clk), and
// the slabs added up to 40 % more traffic.  This kernel changes both:
//   * 256 (cout) x 256 (tap,cin) output tiles: the two M = 128 halves share the B tile and fill all 512 TMEM columns.
//     Bytes per MMA are halved (64 KB per 1024 clk = 62 B/clk per SM).
//   * split-K over a THREAD-BLOCK CLUSTER: the CTAs of a cluster work on the same output tile and consecutive slices
//     of the pixel range; when their main loops end, each CTA parks its fp32 accumulator in its own shared memory
//     (the operand ring is idle by then) and the cluster reduces through distributed shared memory -- rank r sums
//     column slice r over all ranks in rank order (deterministic) and writes it once.  A layer whose splits fit one
//     cluster writes the OIHW gradient directly: no slabs, no second kernel.  Larger split counts use several
//     clusters per tile (one slab per cluster, reduced by reduce_wgrad_kernel).
// Operands are fetched exactly as in conv_sm100.cuh: 4-D TMA boxes {64 channels, bw, bh, bn pixels}, a filter tap is
// a coordinate shift of the X box, padding / ragged tails are zero-filled by the TMA unit, the conv stride is the
// X map's elementStrides.  Both operands are MN-major (pixel index outermost = the MMA's K dimension).
#pragma once
#include "conv_sm100.cuh"

namespace b200 {

constexpr int WG_KR = 64;                       // pixels per k-block (at most; a multiple of 16)
constexpr int WG_TN = 256;                      // output columns per tile
constexpr int WG_STAGES = 3;
constexpr int WG_BOX_MAX = WG_KR * 128;         // bytes of one {64 ch x 64 px} box
constexpr int WG_STAGE_BYTES = 8 * WG_BOX_MAX;  // 4 dY boxes (256 cout) + 4 X boxes (256 columns)
constexpr int WG_THREADS = 192;
constexpr int WG_SMEM_BYTES = WG_STAGES * WG_STAGE_BYTES + 1024 + 256;
constexpr int WG_MAX_CLUSTER = 8;

struct WgradParams {
  int bw, bh, bn, tiles_w, tiles_h;  // pixel boxes (one box = one k-block), walk order w, h, n
  int kr;                            // pixels per box = bw*bh*bn
  int x_mul;                         // conv stride: X box origin = pixel-box origin * x_mul + tap shift
  ConvTap taps[CONV_MAX_TAPS];
  int cdim;                          // C (input channels): column j <-> tap j / C, channel j % C
  int ncols;                         // R*S*C
  int mrows;                         // K (output channels)
  int n_tiles;                       // column tiles; blockIdx.y = n_tile + n_tiles * m_tile
  int tm;                            // output rows per tile: 128 or 256
  int pb_total, pb_per_split;        // k-blocks in total / per CTA
  int cluster;                       // CTAs per cluster (= splits reduced on chip)
  float* out;                        // direct: OIHW gradient; else slabs [group][K][ldo] (packed columns)
  long long ldo, group_stride;
  int direct;                        // 1: write dw[(k*C + c)*RS + tap], 0: write slab[group][k][tap*C + c]
  int rs;                            // R*S
};

__device__ __forceinline__ float4 ld_dsmem_f32x4(uint32_t local_saddr, uint32_t rank) {
  uint32_t ra;
  float4 v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_saddr), "r"(rank));
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(ra)
               : "memory");
  return v;
}

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_sm100_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                   const __grid_constant__ WgradParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);   // same offset in every CTA of the cluster
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + WG_STAGES * WG_STAGE_BYTES);
  uint64_t* empty = full + WG_STAGES;
  uint64_t* tmem_full = empty + WG_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  // after the main loop: this CTA's accumulator half, [128 rows][256 columns] fp32, 16-byte chunks XOR-swizzled by
  // (row & 7) so that the row-per-thread writes coming out of TMEM spread over all banks
  uint8_t* part = smem;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nt = blockIdx.y % P.n_tiles, mt = blockIdx.y / P.n_tiles;
  const int n0 = nt * WG_TN, m0 = mt * P.tm;
  const int split = blockIdx.x;
  const uint32_t rank = P.cluster > 1 ? cluster_ctarank() : 0u;
  const int group = split / P.cluster;
  const int kb_begin = split * P.pb_per_split;
  int kb_end = kb_begin + P.pb_per_split;
  if (kb_end > P.pb_total) kb_end = P.pb_total;
  const bool have_work = kb_end > kb_begin;
  const uint32_t box_bytes = static_cast<uint32_t>(P.kr) * 128u;
  // boxes of this tile: dY boxes of 64 output channels, X boxes of 64 columns (each with its own tap / channel origin)
  int na = (P.mrows - m0 + 63) / 64;
  if (na > P.tm / 64) na = P.tm / 64;
  int nb = (P.ncols - n0 + 63) / 64;
  if (nb > WG_TN / 64) nb = WG_TN / 64;
  const int halves = na > 2 ? 2 : 1;
  const uint32_t tmem_cols = halves == 2 ? 512u : 256u;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
#pragma unroll
    for (int s = 0; s < WG_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (have_work && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int tw = kb_begin % P.tiles_w, th = (kb_begin / P.tiles_w) % P.tiles_h, tn = kb_begin / (P.tiles_w * P.tiles_h);
      // per X box: channel origin and tap shift (fixed for the whole tile)
      int bc[WG_TN / 64], bdw[WG_TN / 64], bdh[WG_TN / 64];
#pragma unroll
      for (int h = 0; h < WG_TN / 64; ++h) {
        const int col = n0 + h * 64;
        const int tp = col < P.ncols ? col / P.cdim : 0;
        bc[h] = col - tp * P.cdim;
        bdw[h] = P.taps[tp].dw;
        bdh[h] = P.taps[tp].dh;
      }
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* a = smem + stage * WG_STAGE_BYTES;
        uint8_t* b = a + 4 * WG_BOX_MAX;
        const int bq = tw * P.bw, bp = th * P.bh, bi = tn * P.bn;
        if (++tw == P.tiles_w) { tw = 0; if (++th == P.tiles_h) { th = 0; ++tn; } }
        mbar_expect_tx(&full[stage], box_bytes * static_cast<uint32_t>(na + nb));
#pragma unroll
        for (int h = 0; h < 4; ++h)
          if (h < na) tma_load_4d(a + h * box_bytes, &tmDY, &full[stage], m0 + h * 64, bq, bp, bi);
#pragma unroll
        for (int h = 0; h < WG_TN / 64; ++h)
          if (h < nb)
            tma_load_4d(b + h * box_bytes, &tmX, &full[stage], bc[h], bq * P.x_mul + bdw[h], bp * P.x_mul + bdh[h], bi);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (have_work && elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, nb * 64, true, true);
      const int ksteps = P.kr >> 4;
      int stage = 0;
      uint32_t phase = 0;
      bool first = true;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * WG_STAGE_BYTES);
        const uint32_t b_addr = a_addr + 4 * WG_BOX_MAX;
        for (int k = 0; k < ksteps; ++k) {
          const uint64_t bdesc = umma_smem_desc(b_addr + k * 2048, box_bytes, 1024);
          umma_f16(tmem_base, umma_smem_desc(a_addr + k * 2048, box_bytes, 1024), bdesc, idesc, (!first || k > 0) ? 1u : 0u);
          if (halves == 2)
            umma_f16(tmem_base + 256, umma_smem_desc(a_addr + 2 * box_bytes + k * 2048, box_bytes, 1024), bdesc, idesc,
                     (!first || k > 0) ? 1u : 0u);
        }
        first = false;
        umma_commit(&empty[stage]);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(tmem_full);
    }
    __syncwarp();
  }

  // ------------------------------------------------------------ cluster reduction + store (all warps take part in
  // the cluster barriers; warps 2..5 do the work: warp w owns accumulator rows 32*(w&3) .. +31)
  const int q = warp & 3;
  const int l = q * 32 + lane;
  const bool epi = warp >= 2;
  if (epi && have_work) {
    mbar_wait(tmem_full, 0);      // every MMA has completed: accumulators final, operand ring no longer read
    tc_fence_after();
  }
  const int ncol_tile = nb * 64;
  if (P.cluster == 1) {
    // no on-chip reduction: accumulator rows go straight from TMEM to global memory (thread = row, 128 contiguous
    // bytes per tcgen05.ld chunk)
    if (epi) {
      for (int half = 0; half < halves; ++half) {
        const int row = m0 + half * 128 + l;
        for (int c = 0; c < ncol_tile / 32; ++c) {
          uint32_t r[32];
          if (have_work) {
            tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(half * 256 + c * 32), r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
          if (row >= P.mrows) continue;
          const int gcol = n0 + c * 32;
          if (P.direct && P.rs > 1) {
            const int tp = gcol / P.cdim, ch = gcol - tp * P.cdim;     // 32 columns stay inside one tap (C % 64 == 0)
            float* o = P.out + (static_cast<long long>(row) * P.cdim + ch) * P.rs + tp;
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j * P.rs] = __uint_as_float(r[j]);
          } else {
            float* o = P.direct ? P.out + static_cast<long long>(row) * P.cdim + gcol
                                : P.out + static_cast<long long>(group) * P.group_stride + static_cast<long long>(row) * P.ldo + gcol;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              reinterpret_cast<uint4*>(o)[j] = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
          }
        }
      }
    }
  } else {
    // cluster reduction.  log2 of the 16-byte chunks per rank slice: 256 columns / cluster / 4
    const int c16_shift = P.cluster == 8 ? 3 : (P.cluster == 4 ? 4 : 5);
    const int c16_per = 1 << c16_shift;
    for (int half = 0; half < halves; ++half) {
      if (epi) {
        // accumulator half -> shared memory (thread = row l)
        uint8_t* rowp = part + l * 1024;
        for (int c = 0; c < ncol_tile / 32; ++c) {
          uint32_t r[32];
          if (have_work) {
            tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(half * 256 + c * 32), r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(rowp + (((c * 8 + j) ^ (l & 7)) << 4)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        }
      }
      cluster_sync_all();
      {
        // rank r reduces column slice r of all ranks' buffers, in rank order, and stores it.  All six warps take part:
        // thread -> (row, 16-byte chunk) with the chunk index fastest, so a row's slice is read and written as one
        // contiguous segment; two items per thread are in flight (2 x cluster DSMEM loads).
        const int c16_lo = static_cast<int>(rank) << c16_shift;
        const uint32_t base = smem_u32(part);
        const int items = 128 << c16_shift;
        for (int f0 = threadIdx.x; f0 < items; f0 += 2 * WG_THREADS) {
          float4 acc[2];
          int rls[2], c16s[2];
          bool ok[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int f = f0 + u * WG_THREADS;
            rls[u] = f >> c16_shift;
            c16s[u] = c16_lo + (f & (c16_per - 1));
            ok[u] = f < items && c16s[u] * 4 < ncol_tile && (m0 + half * 128 + rls[u]) < P.mrows;
            acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll 2
          for (int rk = 0; rk < P.cluster; ++rk) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              if (!ok[u]) continue;
              const uint32_t off = static_cast<uint32_t>(rls[u] * 1024 + ((c16s[u] ^ (rls[u] & 7)) << 4));
              const float4 v = ld_dsmem_f32x4(base + off, static_cast<uint32_t>(rk));
              acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
            }
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            if (!ok[u]) continue;
            const int row = m0 + half * 128 + rls[u];
            const int gcol = n0 + c16s[u] * 4;         // packed column = tap * C + channel (4 columns stay in one tap)
            if (P.direct && P.rs > 1) {
              const int tp = gcol / P.cdim, ch = gcol - tp * P.cdim;
              float* o = P.out + (static_cast<long long>(row) * P.cdim + ch) * P.rs + tp;
              o[0] = acc[u].x; o[P.rs] = acc[u].y; o[2 * P.rs] = acc[u].z; o[3 * P.rs] = acc[u].w;
            } else {
              float* o = P.direct ? P.out + static_cast<long long>(row) * P.cdim + gcol
                                  : P.out + static_cast<long long>(group) * P.group_stride + static_cast<long long>(row) * P.ldo + gcol;
              *reinterpret_cast<float4*>(o) = acc[u];
            }
          }
        }
      }
      // peers may still be reading this CTA's buffer: nobody overwrites it (next half) or exits before everyone is done
      cluster_sync_all();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

}  // namespace b200
