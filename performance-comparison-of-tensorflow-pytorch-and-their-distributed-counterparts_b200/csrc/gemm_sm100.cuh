// tcgen05 GEMM for sm_100a: D[M,N] (+)= A * B, bf16 operands staged by TMA (128B swizzle) into a multi-stage
// shared-memory ring, fp32 accumulation in TMEM, warp-specialised (TMA producer / single-thread MMA issuer /
// four epilogue warps reading TMEM with tcgen05.ld).
//
// Operand "major-ness" (which index is contiguous in global memory) is a template parameter per operand:
//   A K-major : A is [M][K] row-major  (activations [pixels][Cin] for a 1x1 fprop, dY for dgrad)
//   A MN-major: A is [K][M] row-major  (dY [pixels][Cout] used as dY^T for wgrad; reduction over pixels)
//   B K-major : B is [N][K] row-major  (weights [Cout][Cin] for fprop)
//   B MN-major: B is [K][N] row-major  (weights [Cout][Cin] used for dgrad; activations for wgrad)
// This is what lets fprop / dgrad / wgrad of a convolution share one mainloop without any transposed copies.
#pragma once
#include "sm100_ptx.cuh"
#include "device_state.cuh"

namespace b200 {

struct GemmOut {
  void* ptr;               // output base ([M][ld] row-major)
  long long ld;            // row stride in elements
  long long split_stride;  // elements between split-K partial slabs (fp32 output only)
  int out_fp32;            // 1 = float output, 0 = bf16 output
  int accumulate;          // 1 = out += result
  int vec_ok;              // base/ld allow 16-byte vector stores
};

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int GEMM_THREADS = 192;

template <int BN, int STAGES>
constexpr int gemm_smem_bytes() {
  return STAGES * (GEMM_BM * GEMM_BK * 2 + BN * GEMM_BK * 2) + 1024 /*align slack*/ + 256 /*barriers*/;
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmOut out,
                  int M, int N, int num_kb, int kb_per_split) {
  static_assert(BN == 64 || BN == 128 || BN == 256, "BN must be 64/128/256");
  constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  constexpr int B_BYTES = BN * GEMM_BK * 2;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  uint8_t* smA = smem;
  uint8_t* smB = smem + STAGES * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + B_BYTES));
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * GEMM_BM;
  const int kb_begin = blockIdx.z * kb_per_split;
  int kb_end = kb_begin + kb_per_split;
  if (kb_end > num_kb) kb_end = num_kb;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
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

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (one lane)
    if (elect_one()) {
      int i = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb, ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
        uint8_t* a = smA + s * A_BYTES;
        uint8_t* b = smB + s * B_BYTES;
        if (!A_MN) {
          tma_load_2d(a, &tmA, &full[s], kb * GEMM_BK, m0);  // box {64 k, 128 m}
        } else {
#pragma unroll
          for (int h = 0; h < GEMM_BM / 64; ++h)
            tma_load_2d(a + h * 8192, &tmA, &full[s], m0 + h * 64, kb * GEMM_BK);  // box {64 m, 64 k}
        }
        if (!B_MN) {
          tma_load_2d(b, &tmB, &full[s], kb * GEMM_BK, n0);  // box {64 k, BN n}
        } else {
#pragma unroll
          for (int h = 0; h < BN / 64; ++h)
            tma_load_2d(b + h * 8192, &tmB, &full[s], n0 + h * 64, kb * GEMM_BK);  // box {64 n, 64 k}
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one lane)
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN, B_MN);
      int i = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb, ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smA + s * A_BYTES);
        const uint32_t b_addr = smem_u32(smB + s * B_BYTES);
#pragma unroll
        for (int k = 0; k < GEMM_BK / 16; ++k) {
          // K-major: step 16 elements (32 B) inside the swizzled 128-byte row.
          // MN-major: step 16 k-rows = two 8-row swizzle atoms (2048 B).
          const uint64_t adesc = A_MN ? umma_smem_desc(a_addr + k * 2048, 8192, 1024)
                                      : umma_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? umma_smem_desc(b_addr + k * 2048, 8192, 1024)
                                      : umma_smem_desc(b_addr + k * 32, 16, 1024);
          umma_f16(tmem_base, adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty[s]);  // frees the smem slot once these MMAs have read it
      }
      umma_commit(tmem_full);  // accumulator complete
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps 2..5 (TMEM lane quarter = warp % 4)
    const int q = warp & 3;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < M;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c * 32), r);
      tmem_ld_wait();
      const int col0 = n0 + c * 32;
      if (!row_ok || col0 >= N) continue;
      if (out.out_fp32) {
        float* p = reinterpret_cast<float*>(out.ptr) + static_cast<long long>(blockIdx.z) * out.split_stride +
                   static_cast<long long>(row) * out.ld + col0;
        if (out.vec_ok && col0 + 32 <= N) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                   __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
            if (out.accumulate) {
              float4 o = reinterpret_cast<float4*>(p)[j];
              v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
            }
            reinterpret_cast<float4*>(p)[j] = v;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < N) p[j] = (out.accumulate ? p[j] : 0.f) + __uint_as_float(r[j]);
        }
      } else {
        __nv_bfloat16* p =
            reinterpret_cast<__nv_bfloat16*>(out.ptr) + static_cast<long long>(row) * out.ld + col0;
        if (out.vec_ok && col0 + 32 <= N) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[8 * j + e]);
            if (out.accumulate) {
              uint4 o = reinterpret_cast<uint4*>(p)[j];
              const __nv_bfloat162* ob = reinterpret_cast<const __nv_bfloat162*>(&o);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float2 t = __bfloat1622float2(ob[e]);
                f[2 * e] += t.x;
                f[2 * e + 1] += t.y;
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
            if (col0 + j < N) {
              float f = __uint_as_float(r[j]);
              if (out.accumulate) f += __bfloat162float(p[j]);
              p[j] = __float2bfloat16_rn(f);
            }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// Resolved through the runtime so the library has no link-time dependency on libcuda (it must load on a CPU box).
inline EncodeTiledFn tma_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 2-D map: `inner` contiguous elements per row, `outer` rows, `row_stride` elements between rows.
inline int make_tmap_bf16_2d(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride,
                             uint32_t box_inner, uint32_t box_outer) {
  EncodeTiledFn enc = tma_encode_fn();
  if (!enc) return -1;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : static_cast<int>(r);
}

struct GemmProblem {
  const void* A;  // bf16
  const void* B;  // bf16
  long long lda;  // elements between rows of A as stored ([M][K] if K-major, [K][M] if MN-major)
  long long ldb;
  int M, N, K;
  bool a_mn, b_mn;
  int splits;  // split-K factor (>=1); >1 requires fp32 output with split_stride
  GemmOut out;
};

template <int BN, int STAGES, bool A_MN, bool B_MN>
inline int launch_gemm_cfg(const GemmProblem& p, cudaStream_t st) {
  CUtensorMap ta, tb;
  int rc;
  if (!A_MN) rc = make_tmap_bf16_2d(&ta, p.A, p.K, p.M, p.lda, GEMM_BK, GEMM_BM);
  else       rc = make_tmap_bf16_2d(&ta, p.A, p.M, p.K, p.lda, 64, GEMM_BK);
  if (rc) return 100 + rc;
  if (!B_MN) rc = make_tmap_bf16_2d(&tb, p.B, p.K, p.N, p.ldb, GEMM_BK, BN);
  else       rc = make_tmap_bf16_2d(&tb, p.B, p.N, p.K, p.ldb, 64, GEMM_BK);
  if (rc) return 200 + rc;
  const int num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
  int splits = p.splits < 1 ? 1 : p.splits;
  if (splits > num_kb) splits = num_kb;
  int kb_per = (num_kb + splits - 1) / splits;
  splits = (num_kb + kb_per - 1) / kb_per;  // every split gets at least one k-block
  auto kern = gemm_sm100_kernel<BN, STAGES, A_MN, B_MN>;
  constexpr int smem = gemm_smem_bytes<BN, STAGES>();
  static int attr_state[B200_MAX_DEVICES] = {0};
  if (cudaError_t e = ensure_dyn_smem(kern, smem, attr_state)) return 300 + static_cast<int>(e);
  dim3 grid((p.N + BN - 1) / BN, (p.M + GEMM_BM - 1) / GEMM_BM, splits);
  kern<<<grid, GEMM_THREADS, smem, st>>>(ta, tb, p.out, p.M, p.N, num_kb, kb_per);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : 400 + static_cast<int>(e);
}

// Returns the number of split-K slabs actually written (callers reduce exactly that many).
inline int gemm_effective_splits(int K, int splits) {
  const int num_kb = (K + GEMM_BK - 1) / GEMM_BK;
  if (splits < 1) splits = 1;
  if (splits > num_kb) splits = num_kb;
  const int kb_per = (num_kb + splits - 1) / splits;
  return (num_kb + kb_per - 1) / kb_per;
}

template <int BN, int STAGES>
inline int launch_gemm_bn(const GemmProblem& p, cudaStream_t st) {
  if (!p.a_mn && !p.b_mn) return launch_gemm_cfg<BN, STAGES, false, false>(p, st);
  if (!p.a_mn && p.b_mn) return launch_gemm_cfg<BN, STAGES, false, true>(p, st);
  if (p.a_mn && !p.b_mn) return launch_gemm_cfg<BN, STAGES, true, false>(p, st);
  return launch_gemm_cfg<BN, STAGES, true, true>(p, st);
}

// Tile-N choice: widest tile that does not waste more than half of itself on the N tail.
inline int launch_gemm_sm100(const GemmProblem& p, cudaStream_t st, int force_bn = 0) {
  int bn = force_bn;
  if (!bn) bn = p.N > 128 ? 256 : (p.N > 64 ? 128 : 64);
  if (bn == 256) return launch_gemm_bn<256, 4>(p, st);
  if (bn == 128) return launch_gemm_bn<128, 6>(p, st);
  return launch_gemm_bn<64, 8>(p, st);
}

}  // namespace b200
