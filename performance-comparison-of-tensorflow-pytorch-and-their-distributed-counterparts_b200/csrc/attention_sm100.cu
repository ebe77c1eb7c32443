// Fused multi-head attention on tcgen05 for sm_100a (bf16 operands, fp32 accumulation in TMEM), for the
// BertSelfAttention modules of the IMDB model (pytorch_on_language_distr.py:258-261): S in {128, 256}, head dim 64.
//
// One CTA (128 TMEM lanes = 128 rows, 256 threads: warps w and w+4 share a lane quarter and split the columns) per
// (sequence, head, 128-row tile).  All three kernels share one
// skeleton:
//   1. TMA loads the [rows][64] operand tiles of this head straight out of the fused-QKV activation matrix
//      (128-byte rows, 128B swizzle) -- no head split / transpose pass ever runs;
//   2. MMA 1 (one elected thread, tcgen05.mma M=128, N=S, K=64) puts one or two [128 x S] score-like accumulators
//      into TMEM (all of a row's keys at once: no online-softmax streaming is needed at these lengths);
//   3. every thread owns one accumulator row (tcgen05.ld 32x32b): softmax / dropout / dS in registers, result
//      written as bf16 into shared memory in the K-major swizzled layout the next MMA reads as its A operand
//      (over the dead operand tiles of MMA 1);
//   4. MMA 2 (N=64, K=S) with V / K / dO / Q as MN-major B operand -- the very same shared-memory bytes MMA 1 read
//      K-major -- leaves the [128 x 64] output tile(s) in TMEM; the epilogue scales and stores them.
//
//   forward   : acc = Q_i K^T ; P = exp2(acc*c + keybias - rowmax) ; O = (dropout(P) V) / rowsum ; lse saved
//   backward Q: acc1 = Q_i K^T, acc2 = dO_i V^T ; dS = P (acc2*drop - D) ; dQ = scale * dS K      (D = rowsum(dO*O))
//   backward K: acc1 = K_j Q^T, acc2 = V_j dO^T (transposed problem: rows are keys) ; dV = Pd^T dO ; dK = scale dS^T Q
// The probabilities are never written to HBM: backward recomputes them from the saved log-sum-exp.  With dropout the
// forward also leaves the keep mask as one bit per probability (Philox, same indices as the fp32-arithmetic path);
// the backward kernels read bits instead of re-running the generator (the key-side kernel would need one Philox
// call per element, since its rows run across the generator's groups of four).
#include "common.cuh"
#include "philox.cuh"
#include "sm100_ptx.cuh"
#include "gemm_sm100.cuh"

namespace b200 {

enum { ATT_FWD = 0, ATT_BWD_Q = 1, ATT_BWD_KV = 2 };
constexpr int ATC_D = 64;
constexpr int ATC_ROWS = 128;
constexpr int ATC_THREADS = 256;           // two threads per row: each owns half of the row's score columns
constexpr int ATC_TILE = ATC_ROWS * 128;   // bytes of one [128 rows][64 bf16] tile
constexpr float ATC_MASKED = -1.0e30f;     // finite: a fully masked row degrades to uniform weights like the additive
                                           // finfo.min mask of the reference implementation

struct AttParams {
  int S, heads;
  float c;          // softmax scale * log2(e)
  float scale;      // 1/sqrt(D)
  float p, inv_keep;
  uint64_t seed, offset;
  const long long* mask;          // [N][S], 0 = padded key; may be null
  float* lse;                     // [N][heads][S], log2 units
  float* dvec;                    // [N][heads][S]  D = rowsum(dO * O)
  __nv_bfloat16* ctx;             // forward output / backward input [N*S][heads*64]
  const __nv_bfloat16* dctx;
  __nv_bfloat16* dqkv;
  const unsigned long long* ctr;  // optional device-resident call counter added to the Philox offset (<< 40)
  uint32_t* keep_bits;            // [N][heads][S][S/32] dropout keep mask (bit j of word w = key 32w+j); p > 0 only
};

template <int MODE>
__host__ __device__ constexpr int att_tc_smem_bytes(int S) {
  const int a2 = (S / 64) * ATC_TILE;   // one [128][S] bf16 A operand
  int bytes = 0;
  if (MODE == ATT_FWD) bytes = a2 + S * 128;                                              // P | V   (Q,K under P)
  if (MODE == ATT_BWD_Q) bytes = (a2 > 2 * ATC_TILE + S * 128 ? a2 : 2 * ATC_TILE + S * 128) + S * 128;  // dS | K
  if (MODE == ATT_BWD_KV) bytes = 2 * a2 + 2 * S * 128;                                    // PdT dST | Q dO
  return bytes + 2 * S * 4 + (MODE == ATT_BWD_KV ? S * 16 : 0) + 4 * ATC_ROWS * 4 + 64 + 1024;
}

__device__ __forceinline__ void att_store_chunk32(uint8_t* tile_base, int row, int chunk32, const float (&v)[32]) {
  // 32 consecutive K-elements of `row` -> four 16-byte chunks of the 128B-swizzled [128][64] tile
  uint8_t* rowp = tile_base + row * 128;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u;
    u.x = pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
    u.y = pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
    u.z = pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
    u.w = pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
    const int chunk = (chunk32 * 4 + j) ^ (row & 7);
    *reinterpret_cast<uint4*>(rowp + chunk * 16) = u;
  }
}

// store 32 fp32 accumulator columns of this thread's row as bf16 (64 contiguous bytes in global memory)
__device__ __forceinline__ void att_store_row32(__nv_bfloat16* dst, uint32_t taddr, float mul) {
  uint32_t r[32];
  tmem_ld_32x32b_x32(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u;
    u.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]) * mul, __uint_as_float(r[8 * j + 1]) * mul);
    u.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]) * mul, __uint_as_float(r[8 * j + 3]) * mul);
    u.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]) * mul, __uint_as_float(r[8 * j + 5]) * mul);
    u.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]) * mul, __uint_as_float(r[8 * j + 7]) * mul);
    reinterpret_cast<uint4*>(dst)[j] = u;
  }
}

template <int MODE>
__global__ void __launch_bounds__(ATC_THREADS, 1)
att_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
              const __grid_constant__ AttParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  const int S = P.S, HD = P.heads * ATC_D;
  const int ngroups = S / 64;           // 64-wide K groups of the second MMA's A operand
  const int a2 = ngroups * ATC_TILE;
  const int r0 = blockIdx.x * ATC_ROWS, h = blockIdx.y, n = blockIdx.z;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int lrow = tid & (ATC_ROWS - 1);      // row of the tile this thread works on (= its TMEM lane)
  const int half = tid >> 7;                  // which half of the row's score columns it owns
  const int row_base = n * S;

  // shared-memory map (see att_tc_smem_bytes)
  uint8_t *smA2a = smem, *smA2b = nullptr, *smRowA = nullptr, *smRowB = nullptr, *smColA = nullptr, *smColB = nullptr;
  uint8_t* tail;
  if (MODE == ATT_FWD) {
    smRowA = smem;                     // Q_i
    smColA = smem + ATC_TILE;          // K (all keys)
    smColB = smem + a2;                // V
    tail = smColB + S * 128;
  } else if (MODE == ATT_BWD_Q) {
    smRowA = smem;                     // Q_i
    smRowB = smem + ATC_TILE;          // dO_i
    smColB = smem + 2 * ATC_TILE;      // V
    const int reg = a2 > 2 * ATC_TILE + S * 128 ? a2 : 2 * ATC_TILE + S * 128;
    smColA = smem + reg;               // K (survives: B operand of the second MMA)
    tail = smColA + S * 128;
  } else {
    smA2b = smem + a2;                 // dS^T
    smRowA = smem;                     // K_j
    smRowB = smem + ATC_TILE;          // V_j
    smColA = smem + 2 * a2;            // Q (all queries)
    smColB = smColA + S * 128;         // dO (all queries)
    tail = smColB + S * 128;
  }
  float* vecA = reinterpret_cast<float*>(tail);         // fwd / bwd_q: key bias [S]; bwd_kv: lse [S]
  float* vecB = vecA + S;                               // bwd_kv: D [S]
  uint32_t* bitsT = reinterpret_cast<uint32_t*>(vecB + S);  // bwd_kv: keep words [S queries][4] of this key tile
  float* xmax = reinterpret_cast<float*>(bitsT + (MODE == ATT_BWD_KV ? S * 4 : 0));   // [2][128] row-max exchange
  float* xsum = xmax + 2 * ATC_ROWS;                                                  // [2][128] row-sum exchange
  uint64_t* bars = reinterpret_cast<uint64_t*>(xsum + 2 * ATC_ROWS);   // loads, MMA 1, MMA 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const uint32_t tmem_cols = MODE == ATT_FWD ? (uint32_t)S : (uint32_t)(2 * S);
  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    if (MODE != ATT_FWD) tma_prefetch_desc(&tmDO);
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int cq = h * ATC_D, ck = HD + h * ATC_D, cv = 2 * HD + h * ATC_D;
  if (tid == 0) {   // does not wait on anything: no need to park the rest of the warp
    const int ntile = S / ATC_ROWS;
    if (MODE == ATT_FWD) {
      mbar_expect_tx(&bars[0], ATC_TILE + 2 * S * 128);
      tma_load_2d(smRowA, &tmQKV, &bars[0], cq, row_base + r0);
      for (int t = 0; t < ntile; ++t) {
        tma_load_2d(smColA + t * ATC_TILE, &tmQKV, &bars[0], ck, row_base + t * ATC_ROWS);
        tma_load_2d(smColB + t * ATC_TILE, &tmQKV, &bars[0], cv, row_base + t * ATC_ROWS);
      }
    } else if (MODE == ATT_BWD_Q) {
      mbar_expect_tx(&bars[0], 2 * ATC_TILE + 2 * S * 128);
      tma_load_2d(smRowA, &tmQKV, &bars[0], cq, row_base + r0);
      tma_load_2d(smRowB, &tmDO, &bars[0], cq, row_base + r0);
      for (int t = 0; t < ntile; ++t) {
        tma_load_2d(smColA + t * ATC_TILE, &tmQKV, &bars[0], ck, row_base + t * ATC_ROWS);
        tma_load_2d(smColB + t * ATC_TILE, &tmQKV, &bars[0], cv, row_base + t * ATC_ROWS);
      }
    } else {
      mbar_expect_tx(&bars[0], 2 * ATC_TILE + 2 * S * 128);
      tma_load_2d(smRowA, &tmQKV, &bars[0], ck, row_base + r0);
      tma_load_2d(smRowB, &tmQKV, &bars[0], cv, row_base + r0);
      for (int t = 0; t < ntile; ++t) {
        tma_load_2d(smColA + t * ATC_TILE, &tmQKV, &bars[0], cq, row_base + t * ATC_ROWS);
        tma_load_2d(smColB + t * ATC_TILE, &tmDO, &bars[0], cq, row_base + t * ATC_ROWS);
      }
    }
  }

  // per-column vectors -> shared memory; per-row scalars -> registers (overlaps the TMA loads)
  const long long vrow = ((long long)n * P.heads + h) * S;     // base of this head's [S] vectors
  const int row = r0 + lrow;                                   // this thread's row within the sequence
  float row_lse = 0.f, row_D = 0.f, row_bias = 0.f;
  if (MODE == ATT_BWD_KV) {
    for (int i = tid; i < S; i += ATC_THREADS) {
      vecA[i] = P.lse[vrow + i];
      vecB[i] = P.dvec[vrow + i];
    }
    row_bias = (P.mask != nullptr && P.mask[(long long)n * S + row] == 0) ? ATC_MASKED : 0.f;
    if (P.p > 0.f)
      for (int i = tid; i < S * 4; i += ATC_THREADS)
        bitsT[i] = P.keep_bits[(vrow + (i >> 2)) * (S / 32) + (r0 >> 5) + (i & 3)];
  } else {
    for (int i = tid; i < S; i += ATC_THREADS)
      vecA[i] = (P.mask != nullptr && P.mask[(long long)n * S + i] == 0) ? ATC_MASKED : 0.f;
    if (MODE == ATT_BWD_Q) {
      row_lse = P.lse[vrow + row];
      const uint4* od = reinterpret_cast<const uint4*>(P.ctx + ((long long)row_base + row) * HD + cq);
      const uint4* dd = reinterpret_cast<const uint4*>(P.dctx + ((long long)row_base + row) * HD + cq);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 a = od[j], b = dd[j];
        const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
        const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 fa = __bfloat1622float2(pa[e]), fb = __bfloat1622float2(pb[e]);
          row_D += fa.x * fb.x + fa.y * fb.y;
        }
      }
      if (half == 0) P.dvec[vrow + row] = row_D;
    }
  }
  __syncthreads();

  // ---------------------------------------------------------------- MMA 1
  // (lanes 1..31 of warp 0 park at the __syncwarp instead of spinning on the next barrier against their own lane 0)
  if (warp == 0) {
   if (elect_one()) {      // elect.sync: no scalarisation loop around tcgen05.mma (see conv_sm100.cuh)
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(ATC_ROWS, S, false, false);
    const uint32_t a_addr = smem_u32(smRowA), b_addr = smem_u32(smColA);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma_f16(tmem_base, umma_smem_desc(a_addr + k * 32, 16, 1024), umma_smem_desc(b_addr + k * 32, 16, 1024), idesc,
               k > 0 ? 1u : 0u);
    if (MODE != ATT_FWD) {
      const uint32_t a2_addr = smem_u32(smRowB), b2_addr = smem_u32(smColB);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_f16(tmem_base + S, umma_smem_desc(a2_addr + k * 32, 16, 1024),
                 umma_smem_desc(b2_addr + k * 32, 16, 1024), idesc, k > 0 ? 1u : 0u);
    }
    umma_commit(&bars[1]);
   }
   __syncwarp();
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();

  // ---------------------------------------------------------------- row pass
  const uint32_t trow = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const int nch = S / 32, ch0 = half * (nch / 2), ch1 = ch0 + nch / 2;   // this thread's 32-column chunks
  const long long pbase_head = ((long long)n * P.heads + h) * S;   // * S + key gives the flat probability index
  const bool drop = P.p > 0.f;
  const uint64_t rng_offset = P.offset + (P.ctr != nullptr ? *P.ctr << 40 : 0ull);
  float rmax = 0.f, rsum = 0.f;
  if (MODE == ATT_FWD) {
    rmax = -3.0e38f;
    for (int ch = ch0; ch < ch1; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(trow + ch * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) rmax = fmaxf(rmax, fmaf(__uint_as_float(r[j]), P.c, vecA[ch * 32 + j]));
    }
    xmax[half * ATC_ROWS + lrow] = rmax;
    __syncthreads();
    rmax = fmaxf(xmax[lrow], xmax[ATC_ROWS + lrow]);
    const long long pbase = (pbase_head + row) * S;
    for (int ch = ch0; ch < ch1; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(trow + ch * 32, r);
      tmem_ld_wait();
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        // the probability that feeds the second MMA is the bf16-rounded one; sum exactly what is used
        const float pv = exp2f(fmaf(__uint_as_float(r[j]), P.c, vecA[ch * 32 + j]) - rmax);
        v[j] = __bfloat162float(__float2bfloat16_rn(pv));
        rsum += v[j];
      }
      if (drop) {
        uint32_t word = 0u;
        const uint32_t thresh = dropout_thresh16(P.p);
#pragma unroll
        for (int g = 0; g < 4; ++g) word |= dropout_keep8(P.seed, rng_offset, pbase + ch * 32 + g * 8, thresh) << (g * 8);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = ((word >> j) & 1u) ? v[j] : 0.f;
        P.keep_bits[(vrow + row) * (S / 32) + ch] = word;   // the backward kernels read the mask instead of Philox
      }
      att_store_chunk32(smA2a + (ch >> 1) * ATC_TILE, lrow, ch & 1, v);
    }
    xsum[half * ATC_ROWS + lrow] = rsum;      // combined after the barrier that precedes the second MMA
  } else if (MODE == ATT_BWD_Q) {
    for (int ch = ch0; ch < ch1; ++ch) {
      uint32_t r[32], d[32];
      tmem_ld_32x32b_x32(trow + ch * 32, r);
      tmem_ld_32x32b_x32(trow + S + ch * 32, d);
      tmem_ld_wait();
      float v[32];
      const uint32_t word = drop ? P.keep_bits[(vrow + row) * (S / 32) + ch] : 0xFFFFFFFFu;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float keep = ((word >> j) & 1u) ? P.inv_keep : 0.f;
        const float pv = exp2f(fmaf(__uint_as_float(r[j]), P.c, vecA[ch * 32 + j]) - row_lse);
        v[j] = pv * (__uint_as_float(d[j]) * keep - row_D);
      }
      att_store_chunk32(smA2a + (ch >> 1) * ATC_TILE, lrow, ch & 1, v);
    }
  } else {
    // rows are keys, columns are queries: bit `lane` of keep word [q][warp] is this key's mask in query row q
    for (int ch = ch0; ch < ch1; ++ch) {
      uint32_t r[32], d[32];
      tmem_ld_32x32b_x32(trow + ch * 32, r);
      tmem_ld_32x32b_x32(trow + S + ch * 32, d);
      tmem_ld_wait();
      float pd[32], ds[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int q = ch * 32 + j;
        float keep = 1.f;
        if (drop) keep = ((bitsT[q * 4 + (warp & 3)] >> (tid & 31)) & 1u) ? P.inv_keep : 0.f;
        const float pv = exp2f(fmaf(__uint_as_float(r[j]), P.c, row_bias) - vecA[q]);
        pd[j] = pv * keep;
        ds[j] = pv * (__uint_as_float(d[j]) * keep - vecB[q]);
      }
      att_store_chunk32(smA2a + (ch >> 1) * ATC_TILE, lrow, ch & 1, pd);
      att_store_chunk32(smA2b + (ch >> 1) * ATC_TILE, lrow, ch & 1, ds);
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();

  // ---------------------------------------------------------------- MMA 2
  if (warp == 0) {
   if (elect_one()) {      // elect.sync: no scalarisation loop around tcgen05.mma (see conv_sm100.cuh)
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(ATC_ROWS, ATC_D, false, true);
    const uint32_t a_addr = smem_u32(smA2a);
    const uint32_t b_addr = smem_u32(MODE == ATT_FWD ? smColB : (MODE == ATT_BWD_Q ? smColA : smColB));
    for (int kk = 0; kk < S / 16; ++kk)
      umma_f16(tmem_base, umma_smem_desc(a_addr + (kk >> 2) * ATC_TILE + (kk & 3) * 32, 16, 1024),
               umma_smem_desc(b_addr + kk * 2048, 8192, 1024), idesc, kk > 0 ? 1u : 0u);
    if (MODE == ATT_BWD_KV) {
      const uint32_t a2_addr = smem_u32(smA2b), b2_addr = smem_u32(smColA);
      for (int kk = 0; kk < S / 16; ++kk)
        umma_f16(tmem_base + ATC_D, umma_smem_desc(a2_addr + (kk >> 2) * ATC_TILE + (kk & 3) * 32, 16, 1024),
                 umma_smem_desc(b2_addr + kk * 2048, 8192, 1024), idesc, kk > 0 ? 1u : 0u);
    }
    umma_commit(&bars[2]);
   }
   __syncwarp();
  }
  mbar_wait(&bars[2], 0);
  tc_fence_after();

  // ---------------------------------------------------------------- epilogue
  const long long grow = (long long)row_base + row;
  const int c32 = half * 32;                 // each of the row's two threads stores 32 of the 64 output columns
  if (MODE == ATT_FWD) {
    rsum = xsum[lrow] + xsum[ATC_ROWS + lrow];
    att_store_row32(P.ctx + grow * HD + cq + c32, trow + c32, P.inv_keep / rsum);
    if (half == 0) P.lse[vrow + row] = rmax + log2f(rsum);
  } else if (MODE == ATT_BWD_Q) {
    att_store_row32(P.dqkv + grow * 3 * HD + cq + c32, trow + c32, P.scale);
  } else {
    att_store_row32(P.dqkv + grow * 3 * HD + cv + c32, trow + c32, 1.f);
    att_store_row32(P.dqkv + grow * 3 * HD + ck + c32, trow + ATC_D + c32, P.scale);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, tmem_cols);
}

template <int MODE>
static int32_t att_tc_launch(const CUtensorMap& tq, const CUtensorMap& tdo, const AttParams& P, int N,
                             cudaStream_t st) {
  const int smem = att_tc_smem_bytes<MODE>(P.S);
  static int attr_state[B200_MAX_DEVICES] = {0};   // per device, raised once per size (not a stream operation)
  if (cudaError_t e = ensure_dyn_smem(att_tc_kernel<MODE>, smem, attr_state))
    return fail(B200_CUDA_ERROR, "attention: smem attribute: %s", cudaGetErrorString(e));
  att_tc_kernel<MODE><<<dim3(P.S / ATC_ROWS, P.heads, N), ATC_THREADS, smem, st>>>(tq, tdo, P);
  return check_launch("attention_flash");
}

static inline bool att_tc_supported(int S, int D) { return D == ATC_D && (S == 128 || S == 256); }

}  // namespace b200

using namespace b200;

extern "C" {

int32_t b200_attention_flash_supported(int32_t S, int32_t D) { return att_tc_supported(S, D) ? 1 : 0; }

int32_t b200_attention_flash_fwd(const void* qkv, const int64_t* mask, float* lse, uint32_t* keep_bits, void* ctx,
                                 int32_t N, int32_t S, int32_t heads, int32_t D, float p_drop, uint64_t seed,
                                 uint64_t offset, void* stream) {
  if (!qkv || !lse || !ctx || N <= 0 || heads <= 0 || (p_drop > 0.f && !keep_bits))
    return fail(B200_INVALID_ARG, "attention_flash_fwd: bad argument");
  if (!att_tc_supported(S, D) || p_drop < 0.f || p_drop >= 1.f)
    return fail(B200_UNSUPPORTED, "attention_flash_fwd: needs head dim 64 and S in {128, 256} (got D=%d S=%d)", D, S);
  CUtensorMap tq;
  int rc = make_tmap_bf16_2d(&tq, qkv, 3ull * heads * D, (uint64_t)N * S, 3ull * heads * D, 64, ATC_ROWS);
  if (rc) return fail(B200_CUDA_ERROR, "attention_flash_fwd: tensor map failed (%d)", rc);
  AttParams P{};
  P.S = S; P.heads = heads; P.scale = 0.125f; P.c = 0.125f * 1.4426950408889634f;
  P.p = p_drop; P.inv_keep = dropout_inv_keep(p_drop); P.seed = seed; P.offset = offset;
  P.mask = (const long long*)mask; P.lse = lse; P.ctx = (__nv_bfloat16*)ctx; P.keep_bits = keep_bits;
  P.ctr = dropout_ctr();
  return att_tc_launch<ATT_FWD>(tq, tq, P, N, as_stream(stream));
}

int32_t b200_attention_flash_bwd(const void* qkv, const int64_t* mask, const float* lse, const uint32_t* keep_bits,
                                 const void* ctx, const void* dctx, float* dvec, void* dqkv, int32_t N, int32_t S,
                                 int32_t heads, int32_t D, float p_drop, void* stream) {
  if (!qkv || !lse || !ctx || !dctx || !dvec || !dqkv || N <= 0 || heads <= 0 || (p_drop > 0.f && !keep_bits))
    return fail(B200_INVALID_ARG, "attention_flash_bwd: bad argument");
  if (!att_tc_supported(S, D) || p_drop < 0.f || p_drop >= 1.f)
    return fail(B200_UNSUPPORTED, "attention_flash_bwd: needs head dim 64 and S in {128, 256} (got D=%d S=%d)", D, S);
  CUtensorMap tq, tdo;
  int rc = make_tmap_bf16_2d(&tq, qkv, 3ull * heads * D, (uint64_t)N * S, 3ull * heads * D, 64, ATC_ROWS);
  if (rc) return fail(B200_CUDA_ERROR, "attention_flash_bwd: tensor map failed (%d)", rc);
  rc = make_tmap_bf16_2d(&tdo, dctx, (uint64_t)heads * D, (uint64_t)N * S, (uint64_t)heads * D, 64, ATC_ROWS);
  if (rc) return fail(B200_CUDA_ERROR, "attention_flash_bwd: tensor map failed (%d)", rc);
  AttParams P{};
  P.S = S; P.heads = heads; P.scale = 0.125f; P.c = 0.125f * 1.4426950408889634f;
  P.p = p_drop; P.inv_keep = dropout_inv_keep(p_drop);
  P.keep_bits = const_cast<uint32_t*>(keep_bits);
  P.mask = (const long long*)mask; P.lse = const_cast<float*>(lse); P.dvec = dvec;
  P.ctx = (__nv_bfloat16*)const_cast<void*>(ctx); P.dctx = (const __nv_bfloat16*)dctx; P.dqkv = (__nv_bfloat16*)dqkv;
  int32_t r = att_tc_launch<ATT_BWD_Q>(tq, tdo, P, N, as_stream(stream));
  if (r != B200_OK) return r;
  return att_tc_launch<ATT_BWD_KV>(tq, tdo, P, N, as_stream(stream));
}

}  // extern "C"
