// BatchNorm2d (+ReLU, +residual) over NHWC activations [rows][C]: bandwidth kernels.
//   statistics : vectorised column sums (16-byte loads, four rows in flight per thread), per-block partials in the
//                workspace, fp64 finalize (deterministic: no atomics), running-stat EMA in the same finalize launch.
//                In bf16 training the forward statistics normally come from the convolution epilogue instead
//                (b200_conv_fprop_bnstats) and only the finalize launch runs.
//   apply      : y = act(x*scale + shift (+res)), one pass; optionally emits the ReLU mask as one byte per 16-byte
//                vector so the backward kernels never re-read the activation just to learn its sign
//   backward   : partial sums of g and g*(x-mean) -> finalize (dgamma, dbeta, dx coefficients) -> one dx pass that
//                also emits the residual-branch gradient
#include "common.cuh"

namespace b200 {

constexpr int BN_THREADS = 256;
constexpr int BN_STREAM_U = 4;       // 16-byte vectors of each stream a thread of the elementwise kernels keeps in flight
constexpr int BN_MAX_BLOCKS = 296;  // 2 x 148 SMs (each thread keeps four rows of 16-byte loads in flight)

static inline int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// raw 16-byte vector <-> floats
__device__ __forceinline__ void unpack16(const uint4& v, float (&f)[4]) {
  f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
}
__device__ __forceinline__ void unpack16(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// what the partial-sum kernel accumulates
enum { PM_FWD = 0,       // sum x, sum x^2
       PM_BWD = 1,       // g = dz                      : sum g, sum g*(x-mean)
       PM_BWD_Z = 2,     // g = dz * [z > 0]
       PM_BWD_MASK = 3   // g = dz * mask bit
};

// part layout: [nblk][2][C].  Vector path: a thread owns one 16-byte channel vector and walks rows.
template <typename T, int MODE>
__global__ void __launch_bounds__(BN_THREADS)
bn_partial_vec_kernel(const T* __restrict__ x, const T* __restrict__ dz, const T* __restrict__ z,
                      const uint8_t* __restrict__ mask, const float* __restrict__ mean, long long rows, int C, int cx,
                      long long rows_per_block, float* __restrict__ part, int reverse) {
  constexpr int VEC = VecTraits<T>::W;
  constexpr int U = 4;  // rows in flight per thread (raw 16-byte registers, converted only when consumed)
  __shared__ float red[BN_THREADS][2 * VEC + 1];
  const int nvec = C / VEC;
  const int ry = BN_THREADS / cx;
  const int tx = threadIdx.x % cx, ty = threadIdx.x / cx;
  // reverse: the first blocks to be scheduled take the LAST rows (what the producer kernel touched most recently
  // and L2 still holds); the partial row a block writes follows its rows, so the result does not depend on it
  const int blk = reverse ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
  const long long r0 = blk * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > rows) r1 = rows;
  for (int base = 0; base < nvec; base += cx) {
    const int cv = base + tx;
    const bool active = cv < nvec;
    float s0[VEC], s1[VEC], mu[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) { s0[e] = 0.f; s1[e] = 0.f; mu[e] = 0.f; }
    if (active) {
      if (MODE != PM_FWD) {
#pragma unroll
        for (int e = 0; e < VEC; ++e) mu[e] = mean[cv * VEC + e];
      }
      for (long long r = r0 + ty; r < r1; r += (long long)U * ry) {
        uint4 xr[U], gr[U], zr[U];
        uint32_t mk[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long rr = r + (long long)u * ry;
          ok[u] = rr < r1;
          if (ok[u]) {
            const long long vi = rr * nvec + cv;           // vector index
            xr[u] = reinterpret_cast<const uint4*>(x)[vi];
            if (MODE != PM_FWD) gr[u] = reinterpret_cast<const uint4*>(dz)[vi];
            if (MODE == PM_BWD_Z) zr[u] = reinterpret_cast<const uint4*>(z)[vi];
            if (MODE == PM_BWD_MASK) mk[u] = mask[vi];
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!ok[u]) continue;
          float xv[VEC];
          unpack16(xr[u], xv);
          if (MODE == PM_FWD) {
#pragma unroll
            for (int e = 0; e < VEC; ++e) { s0[e] += xv[e]; s1[e] = fmaf(xv[e], xv[e], s1[e]); }
          } else {
            float gv[VEC], zv[VEC];
            unpack16(gr[u], gv);
            if (MODE == PM_BWD_Z) unpack16(zr[u], zv);
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
              bool on = true;
              if (MODE == PM_BWD_Z) on = zv[e] > 0.f;
              if (MODE == PM_BWD_MASK) on = (mk[u] >> e) & 1u;
              const float g = on ? gv[e] : 0.f;
              s0[e] += g;
              s1[e] = fmaf(g, xv[e] - mu[e], s1[e]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) { red[threadIdx.x][e] = s0[e]; red[threadIdx.x][VEC + e] = s1[e]; }
    __syncthreads();
    if (ty == 0 && active) {
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        float a = 0.f, b = 0.f;
        for (int yy = 0; yy < ry; ++yy) { a += red[yy * cx + tx][e]; b += red[yy * cx + tx][VEC + e]; }
        part[((long long)blk * 2 + 0) * C + cv * VEC + e] = a;
        part[((long long)blk * 2 + 1) * C + cv * VEC + e] = b;
      }
    }
    __syncthreads();
  }
}

// scalar path (C not a multiple of the vector width): a thread owns one channel
template <typename T, int MODE>
__global__ void __launch_bounds__(BN_THREADS)
bn_partial_scalar_kernel(const T* __restrict__ x, const T* __restrict__ dz, const T* __restrict__ z,
                         const float* __restrict__ mean, long long rows, int C, int cx, long long rows_per_block,
                         float* __restrict__ part) {
  __shared__ float red[BN_THREADS][3];
  const int ry = BN_THREADS / cx;
  const int tx = threadIdx.x % cx, ty = threadIdx.x / cx;
  const long long r0 = blockIdx.x * rows_per_block;
  long long r1 = r0 + rows_per_block;
  if (r1 > rows) r1 = rows;
  for (int base = 0; base < C; base += cx) {
    const int c = base + tx;
    const bool active = c < C;
    float s0 = 0.f, s1 = 0.f;
    if (active) {
      const float mu = MODE != PM_FWD ? mean[c] : 0.f;
      for (long long r = r0 + ty; r < r1; r += ry) {
        const long long off = r * C + c;
        const float xv = to_f(x[off]);
        if (MODE == PM_FWD) { s0 += xv; s1 = fmaf(xv, xv, s1); }
        else {
          float g = to_f(dz[off]);
          if (MODE == PM_BWD_Z && !(to_f(z[off]) > 0.f)) g = 0.f;
          s0 += g;
          s1 = fmaf(g, xv - mu, s1);
        }
      }
    }
    red[threadIdx.x][0] = s0;
    red[threadIdx.x][1] = s1;
    __syncthreads();
    if (ty == 0 && active) {
      float a = 0.f, b = 0.f;
      for (int yy = 0; yy < ry; ++yy) { a += red[yy * cx + tx][0]; b += red[yy * cx + tx][1]; }
      part[((long long)blockIdx.x * 2 + 0) * C + c] = a;
      part[((long long)blockIdx.x * 2 + 1) * C + c] = b;
    }
    __syncthreads();
  }
}

// Sums the per-block partials: a block owns 32 channels (coalesced across x) and 32 partial-lanes (y).
constexpr int FIN_CX = 32, FIN_RY = 32;
__device__ __forceinline__ void bn_sum_partials(const float* __restrict__ part, int nblk, int C, int c, double& s,
                                                double& q) {
  __shared__ double red[2][FIN_RY][FIN_CX + 1];
  double a = 0, b = 0;
  if (c < C) {
    // four independent accumulation chains; the eight loads of an iteration are issued together
    double a1 = 0, b1 = 0, a2 = 0, b2 = 0, a3 = 0, b3 = 0;
    int i = threadIdx.y;
    for (; i + 3 * FIN_RY < nblk; i += 4 * FIN_RY) {
      const float x0 = part[((long long)i * 2 + 0) * C + c], y0 = part[((long long)i * 2 + 1) * C + c];
      const float x1 = part[((long long)(i + FIN_RY) * 2 + 0) * C + c], y1 = part[((long long)(i + FIN_RY) * 2 + 1) * C + c];
      const float x2 = part[((long long)(i + 2 * FIN_RY) * 2 + 0) * C + c], y2 = part[((long long)(i + 2 * FIN_RY) * 2 + 1) * C + c];
      const float x3 = part[((long long)(i + 3 * FIN_RY) * 2 + 0) * C + c], y3 = part[((long long)(i + 3 * FIN_RY) * 2 + 1) * C + c];
      a += x0; b += y0; a1 += x1; b1 += y1; a2 += x2; b2 += y2; a3 += x3; b3 += y3;
    }
    for (; i < nblk; i += FIN_RY) {
      a += part[((long long)i * 2 + 0) * C + c];
      b += part[((long long)i * 2 + 1) * C + c];
    }
    a += a1 + a2 + a3;
    b += b1 + b2 + b3;
  }
  red[0][threadIdx.y][threadIdx.x] = a;
  red[1][threadIdx.y][threadIdx.x] = b;
  __syncthreads();
  s = 0;
  q = 0;
  if (threadIdx.y == 0) {
#pragma unroll
    for (int j = 0; j < FIN_RY; ++j) {
      s += red[0][j][threadIdx.x];
      q += red[1][j][threadIdx.x];
    }
  }
}

__global__ void __launch_bounds__(FIN_CX * FIN_RY)
bn_finalize_fwd_kernel(const float* __restrict__ part, int nblk, int C, double n, const float* __restrict__ gamma,
                       const float* __restrict__ beta, float* __restrict__ running_mean,
                       float* __restrict__ running_var, float momentum, float eps, float* __restrict__ save_mean,
                       float* __restrict__ save_invstd, float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * FIN_CX + threadIdx.x;
  double s, q;
  bn_sum_partials(part, nblk, C, c, s, q);
  if (threadIdx.y != 0 || c >= C) return;
  const double mean = s / n;
  double var = q / n - mean * mean;
  if (var < 0) var = 0;
  const float invstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
  save_mean[c] = (float)mean;
  save_invstd[c] = invstd;
  const float sc = g * invstd;
  scale[c] = sc;
  shift[c] = bt - (float)mean * sc;
  if (running_mean) {
    const double unbiased = n > 1 ? var * n / (n - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
  }
}

// coef layout: [3][C] : A = gamma*invstd ; B = gamma*invstd^3*S2/n ; Cc = A*dbeta/n
__global__ void __launch_bounds__(FIN_CX * FIN_RY)
bn_finalize_bwd_kernel(const float* __restrict__ part, int nblk, int C, double n, const float* __restrict__ gamma,
                       const float* __restrict__ invstd, float* __restrict__ dgamma, float* __restrict__ dbeta,
                       float* __restrict__ coef) {
  const int c = blockIdx.x * FIN_CX + threadIdx.x;
  double s, q;
  bn_sum_partials(part, nblk, C, c, s, q);
  if (threadIdx.y != 0 || c >= C) return;
  const double is = invstd[c], g = gamma ? gamma[c] : 1.0;
  if (dgamma) dgamma[c] = (float)(q * is);
  if (dbeta) dbeta[c] = (float)s;
  const double A = g * is;
  coef[c] = (float)A;
  coef[C + c] = (float)(g * is * is * is * q / n);
  coef[2 * C + c] = (float)(A * s / n);
}

// FIXEDC: the grid stride is a multiple of nvec, so a thread always lands on the same channel vector and keeps its
// per-channel constants in registers (the streaming loop is then loads + FMAs + store only).
template <typename T, int VEC, bool FIXEDC, int SU = 2>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, uint8_t* __restrict__ mask,
                long long nvec_total, int nvec, const float* __restrict__ scale, const float* __restrict__ shift,
                int relu, int reverse) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  // reverse: element index = last - i (the walk starts where the producer kernel finished, i.e. at what L2 still
  // holds); the channel of an index changes from i % nvec to nvec-1 - i % nvec, still fixed per thread
  const long long last = nvec_total - 1;
#define B200_IDX(v) (reverse ? last - (v) : (v))
  float sc[VEC], sh[VEC];
  if (FIXEDC) {
    const int c = (int)(B200_IDX(i) % nvec) * VEC;
#pragma unroll
    for (int e = 0; e < VEC; ++e) { sc[e] = scale[c + e]; sh[e] = shift[c + e]; }
  }
  if constexpr (FIXEDC && VEC > 1) {
    // streaming loop: SU raw 16-byte vectors (plus residuals) in flight per thread, unpacked only when consumed.
    // Measured on [409600 x 256] bf16: SU = 2 (63 registers, four blocks per SM) 85 us / 109 us with a residual;
    // SU = 4 (93 registers, two blocks per SM) 97 us / 117 us -- for this kernel more resident threads beat more loads
    // per thread, for the dx pass below it is the other way round.
    for (; i < nvec_total; i += SU * stride) {
      uint4 xr[SU], rr[SU];
      long long ei[SU];
      bool ok[SU];
#pragma unroll
      for (int u = 0; u < SU; ++u) {
        const long long idx = i + u * stride;
        ok[u] = idx < nvec_total;
        ei[u] = ok[u] ? B200_IDX(idx) : 0;
      }
#pragma unroll
      for (int u = 0; u < SU; ++u) {
        if (!ok[u]) continue;
        xr[u] = *reinterpret_cast<const uint4*>(x + ei[u] * VEC);
        if (res) rr[u] = *reinterpret_cast<const uint4*>(res + ei[u] * VEC);
      }
#pragma unroll
      for (int u = 0; u < SU; ++u) {
        if (!ok[u]) continue;
        float xv[VEC], rv[VEC];
        unpack16(xr[u], xv);
        if (res) unpack16(rr[u], rv);
        uint32_t m = 0;
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          float v = fmaf(xv[e], sc[e], sh[e]);
          if (res) v += rv[e];
          if (relu) v = fmaxf(v, 0.f);
          xv[e] = v;
        }
        store_vec(y + ei[u] * VEC, reinterpret_cast<float(&)[VecTraits<T>::W]>(xv));
        if (mask) {
#pragma unroll
          for (int e = 0; e < VEC; ++e) m |= (to_f(from_f<T>(xv[e])) > 0.f ? 1u : 0u) << e;
          mask[ei[u]] = (uint8_t)m;
        }
      }
    }
    return;
  }
  for (; i < nvec_total; i += stride) {
    const long long ei = B200_IDX(i);
    if (!FIXEDC) {
      const int c = (int)(ei % nvec) * VEC;
#pragma unroll
      for (int e = 0; e < VEC; ++e) { sc[e] = scale[c + e]; sh[e] = shift[c + e]; }
    }
    const long long off = ei * VEC;
    float xv[VEC], rv[VEC];
    if constexpr (VEC == 1) { xv[0] = to_f(x[off]); if (res) rv[0] = to_f(res[off]); }
    else {
      load_vec(x + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(xv));
      if (res) load_vec(res + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(rv));
    }
    uint32_t m = 0;
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      float v = fmaf(xv[e], sc[e], sh[e]);
      if (res) v += rv[e];
      if (relu) v = fmaxf(v, 0.f);
      xv[e] = v;
    }
    if constexpr (VEC == 1) y[off] = from_f<T>(xv[0]);
    else {
      store_vec(y + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(xv));
      if (mask) {
        // the bit must describe the STORED value (after rounding to T), which is what "z > 0" meant before
#pragma unroll
        for (int e = 0; e < VEC; ++e) m |= (to_f(from_f<T>(xv[e])) > 0.f ? 1u : 0u) << e;
        mask[ei] = (uint8_t)m;
      }
    }
  }
#undef B200_IDX
}

enum { DM_NONE = 0, DM_Z = 1, DM_MASK = 2 };
template <typename T, int VEC, bool FIXEDC, int DMODE>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_dx_kernel(const T* __restrict__ dz, const T* __restrict__ z, const uint8_t* __restrict__ mask,
                 const T* __restrict__ x, T* __restrict__ dx, T* __restrict__ dres, long long nvec_total, int nvec,
                 int C, const float* __restrict__ mean, const float* __restrict__ coef, int reverse) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long last = nvec_total - 1;      // reverse walk: see bn_apply_kernel
#define B200_IDX(v) (reverse ? last - (v) : (v))
  float cA[VEC], cB[VEC], cC[VEC], mu[VEC];
  if (FIXEDC) {
    const int c = (int)(B200_IDX(i) % nvec) * VEC;
#pragma unroll
    for (int e = 0; e < VEC; ++e) { cA[e] = coef[c + e]; cB[e] = coef[C + c + e]; cC[e] = coef[2 * C + c + e]; mu[e] = mean[c + e]; }
  }
  if constexpr (FIXEDC && VEC > 1) {
    // streaming loop: BN_STREAM_U raw vectors of each input in flight per thread.  With two, ~34 KB per SM were in
    // flight against the ~45 KB that 6.5 TB/s at ~1 us of loaded latency needs (ncu: 4.0 TB/s, issue slots 75 % idle);
    // with four the reduce + dx pair on [409600 x 256] went 245 -> 198 us.
    for (; i < nvec_total; i += BN_STREAM_U * stride) {
      uint4 gr[BN_STREAM_U], xr[BN_STREAM_U], zr[BN_STREAM_U];
      uint32_t mk[BN_STREAM_U];
      long long ei[BN_STREAM_U];
      bool ok[BN_STREAM_U];
#pragma unroll
      for (int u = 0; u < BN_STREAM_U; ++u) {
        const long long idx = i + u * stride;
        ok[u] = idx < nvec_total;
        ei[u] = ok[u] ? B200_IDX(idx) : 0;
      }
#pragma unroll
      for (int u = 0; u < BN_STREAM_U; ++u) {
        mk[u] = 0xffffffffu;
        if (!ok[u]) continue;
        gr[u] = *reinterpret_cast<const uint4*>(dz + ei[u] * VEC);
        xr[u] = *reinterpret_cast<const uint4*>(x + ei[u] * VEC);
        if (DMODE == DM_Z) zr[u] = *reinterpret_cast<const uint4*>(z + ei[u] * VEC);
        if (DMODE == DM_MASK) mk[u] = mask[ei[u]];
      }
#pragma unroll
      for (int u = 0; u < BN_STREAM_U; ++u) {
        if (!ok[u]) continue;
        float gv[VEC], xv[VEC], zv[VEC];
        unpack16(gr[u], gv);
        unpack16(xr[u], xv);
        if (DMODE == DM_Z) unpack16(zr[u], zv);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          bool on = true;
          if (DMODE == DM_Z) on = zv[e] > 0.f;
          if (DMODE == DM_MASK) on = (mk[u] >> e) & 1u;
          const float g = on ? gv[e] : 0.f;
          gv[e] = g;
          xv[e] = cA[e] * g - cC[e] - cB[e] * (xv[e] - mu[e]);
        }
        store_vec(dx + ei[u] * VEC, reinterpret_cast<float(&)[VecTraits<T>::W]>(xv));
        if (dres) store_vec(dres + ei[u] * VEC, reinterpret_cast<float(&)[VecTraits<T>::W]>(gv));
      }
    }
    return;
  }
  for (; i < nvec_total; i += stride) {
    const long long ei = B200_IDX(i);
    if (!FIXEDC) {
      const int c = (int)(ei % nvec) * VEC;
#pragma unroll
      for (int e = 0; e < VEC; ++e) { cA[e] = coef[c + e]; cB[e] = coef[C + c + e]; cC[e] = coef[2 * C + c + e]; mu[e] = mean[c + e]; }
    }
    const long long off = ei * VEC;
    float gv[VEC], zv[VEC], xv[VEC];
    uint32_t m = 0xffffffffu;
    if constexpr (VEC == 1) {
      gv[0] = to_f(dz[off]); xv[0] = to_f(x[off]);
      if (DMODE == DM_Z) zv[0] = to_f(z[off]);
    } else {
      load_vec(dz + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(gv));
      load_vec(x + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(xv));
      if (DMODE == DM_Z) load_vec(z + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(zv));
      if (DMODE == DM_MASK) m = mask[ei];
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      bool on = true;
      if (DMODE == DM_Z) on = zv[e] > 0.f;
      if (DMODE == DM_MASK) on = (m >> e) & 1u;
      const float g = on ? gv[e] : 0.f;
      gv[e] = g;
      xv[e] = cA[e] * g - cC[e] - cB[e] * (xv[e] - mu[e]);
    }
    if constexpr (VEC == 1) { dx[off] = from_f<T>(xv[0]); if (dres) dres[off] = from_f<T>(gv[0]); }
    else {
      store_vec(dx + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(xv));
      if (dres) store_vec(dres + off, reinterpret_cast<float(&)[VecTraits<T>::W]>(gv));
    }
  }
#undef B200_IDX
}

template <typename T>
__global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = from_f<T>(to_f(a[i]) + to_f(b[i]));
}
template <typename T>
__global__ void relu_bwd_kernel(const T* __restrict__ dz, const T* __restrict__ z, T* __restrict__ dx, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = to_f(z[i]) > 0.f ? dz[i] : from_f<T>(0.f);
}

static inline int ew_grid(long long total) {
  long long b = (total + BN_THREADS - 1) / BN_THREADS;
  const long long cap = (long long)num_sms() * 8;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

// grid for the streaming kernels; `fixed` tells whether gridDim*blockDim is a multiple of nvec
static inline int ew_grid_fixed(long long total, int nvec, bool* fixed) {
  int g = ew_grid(total);
  *fixed = false;
  if (nvec <= 0) return g;
  int a = nvec, b = BN_THREADS;
  while (b) { int t = a % b; a = b; b = t; }
  const int per = nvec / a;   // blocks per period: g must be a multiple of it
  if (per <= g) { g = g / per * per; *fixed = true; }
  else if (per <= 4096 && (long long)per * BN_THREADS <= total * 2) { g = per; *fixed = true; }
  return g;
}

struct BnPlan {
  int vec, cx, nblk;
  long long rows_per_block;
};
template <typename T>
static BnPlan bn_plan(long long rows, int C) {
  BnPlan p;
  p.vec = (C % VecTraits<T>::W == 0) ? VecTraits<T>::W : 1;
  const int nvec = C / p.vec;
  p.cx = pow2_ceil(nvec);
  if (p.cx > BN_THREADS) p.cx = BN_THREADS;
  const int ry = BN_THREADS / p.cx;
  // at least 16 rows per row-lane (keeps the partial-sum traffic far below the tensor's own bytes when C is large),
  // at most BN_MAX_BLOCKS blocks
  long long rpb = (long long)ry * 16;
  long long nblk = (rows + rpb - 1) / rpb;
  if (nblk > BN_MAX_BLOCKS) {
    rpb = (rows + BN_MAX_BLOCKS - 1) / BN_MAX_BLOCKS;
    rpb = (rpb + ry - 1) / ry * ry;
    nblk = (rows + rpb - 1) / rpb;
  }
  p.nblk = (int)nblk;
  p.rows_per_block = rpb;
  return p;
}

template <typename T>
static void launch_bn_partial(const BnPlan& p, int mode, const T* x, const T* dz, const T* z, const uint8_t* mask,
                              const float* mean, long long rows, int C, float* part, cudaStream_t st, int reverse = 0) {
  if (p.vec == 1) {
    if (mode == PM_FWD) bn_partial_scalar_kernel<T, PM_FWD><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mean, rows, C, p.cx, p.rows_per_block, part);
    else if (mode == PM_BWD) bn_partial_scalar_kernel<T, PM_BWD><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mean, rows, C, p.cx, p.rows_per_block, part);
    else bn_partial_scalar_kernel<T, PM_BWD_Z><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mean, rows, C, p.cx, p.rows_per_block, part);
    return;
  }
  if (mode == PM_FWD) bn_partial_vec_kernel<T, PM_FWD><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mask, mean, rows, C, p.cx, p.rows_per_block, part, reverse);
  else if (mode == PM_BWD) bn_partial_vec_kernel<T, PM_BWD><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mask, mean, rows, C, p.cx, p.rows_per_block, part, reverse);
  else if (mode == PM_BWD_Z) bn_partial_vec_kernel<T, PM_BWD_Z><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mask, mean, rows, C, p.cx, p.rows_per_block, part, reverse);
  else bn_partial_vec_kernel<T, PM_BWD_MASK><<<p.nblk, BN_THREADS, 0, st>>>(x, dz, z, mask, mean, rows, C, p.cx, p.rows_per_block, part, reverse);
}

template <typename T>
static int32_t bn_stats_impl(const T* x, long long rows, int C, const float* gamma, const float* beta, float* rm,
                             float* rv, float momentum, float eps, float* save_mean, float* save_invstd, float* scale,
                             float* shift, float* part, cudaStream_t st) {
  const BnPlan p = bn_plan<T>(rows, C);
  launch_bn_partial<T>(p, PM_FWD, x, nullptr, nullptr, nullptr, nullptr, rows, C, part, st);
  bn_finalize_fwd_kernel<<<(C + FIN_CX - 1) / FIN_CX, dim3(FIN_CX, FIN_RY), 0, st>>>(part, p.nblk, C, (double)rows, gamma, beta, rm, rv, momentum, eps, save_mean, save_invstd, scale, shift);
  return check_launch("bn_train_stats", 2);
}

template <typename T>
static int32_t bn_apply_impl(const T* x, const T* res, T* y, uint8_t* mask, long long rows, int C, const float* scale,
                             const float* shift, int relu, int reverse, cudaStream_t st) {
  const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(res)) & 15) == 0;
  if (C % VecTraits<T>::W == 0 && aligned) {
    constexpr int V = VecTraits<T>::W;
    bool fixed;
    const int grid = ew_grid_fixed(rows * (C / V), C / V, &fixed);
    if (fixed) bn_apply_kernel<T, V, true, 2><<<grid, BN_THREADS, 0, st>>>(x, res, y, mask, rows * (C / V), C / V, scale, shift, relu, reverse);
    else bn_apply_kernel<T, V, false><<<grid, BN_THREADS, 0, st>>>(x, res, y, mask, rows * (C / V), C / V, scale, shift, relu, reverse);
  } else {
    if (mask) return fail(B200_UNSUPPORTED, "bn_apply: the ReLU mask needs C %% %d == 0 and 16-byte aligned tensors", VecTraits<T>::W);
    bn_apply_kernel<T, 1, false><<<ew_grid(rows * C), BN_THREADS, 0, st>>>(x, res, y, nullptr, rows * C, C, scale, shift, relu, reverse);
  }
  return check_launch("bn_apply");
}

template <typename T>
static int32_t bn_bwd_impl(const T* dz, const T* z, const uint8_t* mask, const T* x, T* dx, T* dres, long long rows,
                           int C, const float* gamma, const float* mean, const float* invstd, int relu, float* dgamma,
                           float* dbeta, float* ws, int walk, cudaStream_t st) {
  const int rev_reduce = walk & 1, reverse = (walk >> 1) & 1;   // walk direction of the reduction pass / the dx pass
  const BnPlan p = bn_plan<T>(rows, C);
  float* part = ws;
  float* coef = ws + (size_t)BN_MAX_BLOCKS * 2 * C;
  const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dz) | reinterpret_cast<uintptr_t>(z) |
                         reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(dres)) & 15) == 0;
  if (mask && (p.vec == 1 || !aligned))
    return fail(B200_UNSUPPORTED, "bn_backward: the ReLU mask needs C %% %d == 0 and 16-byte aligned tensors", VecTraits<T>::W);
  if (p.vec != 1 && !aligned) return fail(B200_INVALID_ARG, "bn_backward: tensors must be 16-byte aligned");
  const int pmode = !relu ? PM_BWD : (mask ? PM_BWD_MASK : PM_BWD_Z);
  launch_bn_partial<T>(p, pmode, x, dz, z, mask, mean, rows, C, part, st, rev_reduce);
  bn_finalize_bwd_kernel<<<(C + FIN_CX - 1) / FIN_CX, dim3(FIN_CX, FIN_RY), 0, st>>>(part, p.nblk, C, (double)rows, gamma, invstd, dgamma, dbeta, coef);
  if (p.vec != 1) {
    constexpr int V = VecTraits<T>::W;
    bool fixed;
    const int grid = ew_grid_fixed(rows * (C / V), C / V, &fixed);
    const long long nv = rows * (C / V);
#define B200_DX(FX, DM) bn_bwd_dx_kernel<T, V, FX, DM><<<grid, BN_THREADS, 0, st>>>(dz, z, mask, x, dx, dres, nv, C / V, C, mean, coef, reverse)
    if (!relu) { if (fixed) B200_DX(true, DM_NONE); else B200_DX(false, DM_NONE); }
    else if (mask) { if (fixed) B200_DX(true, DM_MASK); else B200_DX(false, DM_MASK); }
    else { if (fixed) B200_DX(true, DM_Z); else B200_DX(false, DM_Z); }
#undef B200_DX
  } else {
    if (!relu) bn_bwd_dx_kernel<T, 1, false, DM_NONE><<<ew_grid(rows * C), BN_THREADS, 0, st>>>(dz, z, nullptr, x, dx, dres, rows * C, C, C, mean, coef, reverse);
    else bn_bwd_dx_kernel<T, 1, false, DM_Z><<<ew_grid(rows * C), BN_THREADS, 0, st>>>(dz, z, nullptr, x, dx, dres, rows * C, C, C, mean, coef, reverse);
  }
  return check_launch("bn_backward", 3);
}

__global__ void bn_eval_coeffs_kernel(const float* gamma, const float* beta, const float* rm, const float* rv, float eps,
                                      float* scale, float* shift, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = 1.0f / sqrtf(rv[c] + eps);
  const float sc = (gamma ? gamma[c] : 1.f) * invstd;
  scale[c] = sc;
  shift[c] = (beta ? beta[c] : 0.f) - rm[c] * sc;
}

// for stem_pool.cu (its own reduction pass feeds the same finalize kernel)
int bn_max_partial_blocks() { return BN_MAX_BLOCKS; }
int32_t bn_finalize_bwd_launch(const float* part, int nblk, int C, double n, const float* gamma, const float* invstd,
                               float* dgamma, float* dbeta, float* coef, cudaStream_t st) {
  bn_finalize_bwd_kernel<<<(C + FIN_CX - 1) / FIN_CX, dim3(FIN_CX, FIN_RY), 0, st>>>(part, nblk, C, n, gamma, invstd, dgamma, dbeta, coef);
  return check_launch("bn_finalize_bwd");
}

}  // namespace b200

using namespace b200;

extern "C" int64_t b200_bn_workspace_bytes(int64_t rows, int32_t C) {
  (void)rows;
  return (int64_t)((size_t)BN_MAX_BLOCKS * 2 * C + 3 * (size_t)C) * 4;
}

extern "C" int64_t b200_bn_mask_bytes(int64_t rows, int32_t C, int32_t dtype) {
  const int W = dtype == B200_BF16 ? 8 : 4;
  if (rows <= 0 || C <= 0 || C % W != 0) return 0;
  return rows * (int64_t)(C / W);
}

extern "C" int32_t b200_bn_train_stats(const void* x, int64_t rows, int32_t C, int32_t dtype, const float* gamma,
                                       const float* beta, float* running_mean, float* running_var, float momentum,
                                       float eps, float* save_mean, float* save_invstd, float* scale, float* shift,
                                       void* ws, int64_t ws_bytes, void* stream) {
  if (!x || rows <= 0 || C <= 0 || !save_mean || !save_invstd || !scale || !shift)
    return fail(B200_INVALID_ARG, "bn_train_stats: bad arguments");
  if (!ws || ws_bytes < b200_bn_workspace_bytes(rows, C)) return fail(B200_WORKSPACE_TOO_SMALL, "bn_train_stats: workspace too small");
  if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return fail(B200_INVALID_ARG, "bn_train_stats: x must be 16-byte aligned");
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32)
    return bn_stats_impl<float>((const float*)x, rows, C, gamma, beta, running_mean, running_var, momentum, eps, save_mean, save_invstd, scale, shift, (float*)ws, st);
  if (dtype == B200_BF16)
    return bn_stats_impl<__nv_bfloat16>((const __nv_bfloat16*)x, rows, C, gamma, beta, running_mean, running_var, momentum, eps, save_mean, save_invstd, scale, shift, (float*)ws, st);
  return fail(B200_INVALID_ARG, "bn_train_stats: bad dtype");
}

extern "C" int32_t b200_bn_finalize_stats(const float* partials, int64_t partial_rows, int64_t rows, int32_t C,
                                          const float* gamma, const float* beta, float* running_mean,
                                          float* running_var, float momentum, float eps, float* save_mean,
                                          float* save_invstd, float* scale, float* shift, void* stream) {
  if (!partials || partial_rows <= 0 || rows <= 0 || C <= 0 || !save_mean || !save_invstd || !scale || !shift)
    return fail(B200_INVALID_ARG, "bn_finalize_stats: bad arguments");
  bn_finalize_fwd_kernel<<<(C + FIN_CX - 1) / FIN_CX, dim3(FIN_CX, FIN_RY), 0, as_stream(stream)>>>(
      partials, (int)partial_rows, C, (double)rows, gamma, beta, running_mean, running_var, momentum, eps, save_mean,
      save_invstd, scale, shift);
  return check_launch("bn_finalize_stats");
}

extern "C" int32_t b200_bn_eval_coeffs(const float* gamma, const float* beta, const float* running_mean,
                                       const float* running_var, float eps, float* scale, float* shift, int32_t C,
                                       void* stream) {
  if (!running_mean || !running_var || !scale || !shift || C <= 0) return fail(B200_INVALID_ARG, "bn_eval_coeffs: bad arguments");
  bn_eval_coeffs_kernel<<<(C + 127) / 128, 128, 0, as_stream(stream)>>>(gamma, beta, running_mean, running_var, eps, scale, shift, C);
  return check_launch("bn_eval_coeffs");
}

extern "C" int32_t b200_bn_apply(const void* x, const void* residual, void* y, uint8_t* mask, int64_t rows, int32_t C,
                                 int32_t dtype, const float* scale, const float* shift, int32_t relu, int32_t reverse,
                                 void* stream) {
  if (!x || !y || !scale || !shift || rows <= 0 || C <= 0) return fail(B200_INVALID_ARG, "bn_apply: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32) return bn_apply_impl<float>((const float*)x, (const float*)residual, (float*)y, mask, rows, C, scale, shift, relu, reverse != 0, st);
  if (dtype == B200_BF16)
    return bn_apply_impl<__nv_bfloat16>((const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (__nv_bfloat16*)y, mask, rows, C, scale, shift, relu, reverse != 0, st);
  return fail(B200_INVALID_ARG, "bn_apply: bad dtype");
}

extern "C" int32_t b200_bn_backward(const void* dz, const void* z, const uint8_t* mask, const void* x, void* dx,
                                    void* dres, int64_t rows, int32_t C, int32_t dtype, const float* gamma,
                                    const float* save_mean, const float* save_invstd, int32_t relu, float* dgamma,
                                    float* dbeta, void* ws, int64_t ws_bytes, int32_t walk, void* stream) {
  if (!dz || !x || !dx || !save_mean || !save_invstd || rows <= 0 || C <= 0 || (relu && !z && !mask))
    return fail(B200_INVALID_ARG, "bn_backward: bad arguments");
  if (!ws || ws_bytes < b200_bn_workspace_bytes(rows, C)) return fail(B200_WORKSPACE_TOO_SMALL, "bn_backward: workspace too small");
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32)
    return bn_bwd_impl<float>((const float*)dz, (const float*)z, mask, (const float*)x, (float*)dx, (float*)dres, rows, C, gamma, save_mean, save_invstd, relu, dgamma, dbeta, (float*)ws, walk, st);
  if (dtype == B200_BF16)
    return bn_bwd_impl<__nv_bfloat16>((const __nv_bfloat16*)dz, (const __nv_bfloat16*)z, mask, (const __nv_bfloat16*)x, (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, rows, C, gamma, save_mean, save_invstd, relu, dgamma, dbeta, (float*)ws, walk, st);
  return fail(B200_INVALID_ARG, "bn_backward: bad dtype");
}

extern "C" int32_t b200_add(const void* a, const void* b, void* y, int64_t n, int32_t dtype, void* stream) {
  if (!a || !b || !y || n <= 0) return fail(B200_INVALID_ARG, "add: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32) add_kernel<float><<<ew_grid(n), BN_THREADS, 0, st>>>((const float*)a, (const float*)b, (float*)y, n);
  else if (dtype == B200_BF16) add_kernel<__nv_bfloat16><<<ew_grid(n), BN_THREADS, 0, st>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)y, n);
  else return fail(B200_INVALID_ARG, "add: bad dtype");
  return check_launch("add");
}

extern "C" int32_t b200_relu_bwd(const void* dz, const void* z, void* dx, int64_t n, int32_t dtype, void* stream) {
  if (!dz || !z || !dx || n <= 0) return fail(B200_INVALID_ARG, "relu_bwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32) relu_bwd_kernel<float><<<ew_grid(n), BN_THREADS, 0, st>>>((const float*)dz, (const float*)z, (float*)dx, n);
  else if (dtype == B200_BF16) relu_bwd_kernel<__nv_bfloat16><<<ew_grid(n), BN_THREADS, 0, st>>>((const __nv_bfloat16*)dz, (const __nv_bfloat16*)z, (__nv_bfloat16*)dx, n);
  else return fail(B200_INVALID_ARG, "relu_bwd: bad dtype");
  return check_launch("relu_bwd");
}

// ------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of b200_bn_train_stats.
//
// Some hosts of this library cannot hand us a workspace and cannot allocate one either: a serving process in which
// the framework took all device memory for its own pool when it started, or a latency-bound loop where the
// device-wide synchronisation inside cudaMalloc is not acceptable.  For them the per-block partial sums live in the
// array below, which is part of the library image: each GPU that loads the library carries these 2 MiB until the
// process exits.  Nothing here calls cudaMalloc, cudaFree or any routine that allocates.
//
// `g_bn_noalloc_scratch` names a DEVICE array.  In host code that name is only a handle for the runtime, not an
// address a kernel can use, so the host side asks the runtime for the real device address once per device
// (bn_noalloc_scratch_ptr) and passes that pointer to the kernels.
constexpr size_t BN_NOALLOC_FLOATS = (2u << 20) / sizeof(float);  // 2 MiB
static __device__ __align__(256) float g_bn_noalloc_scratch[BN_NOALLOC_FLOATS];

namespace b200 {

// Device address of the scratch array for the CURRENT device; looked up on the first call per device.
// Do that first call before any stream capture begins.
static int32_t bn_noalloc_scratch_ptr(float** out) {
  constexpr int MAX_DEV = 64;
  static float* ptrs[MAX_DEV] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(B200_CUDA_ERROR, "bn_train_stats_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= MAX_DEV) return fail(B200_UNSUPPORTED, "bn_train_stats_noalloc: device index %d out of range", dev);
  if (!ptrs[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_bn_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return fail(B200_CUDA_ERROR, "bn_train_stats_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    ptrs[dev] = p;  // racing first calls store the same value
  }
  *out = ptrs[dev];
  return B200_OK;
}

template <typename T>
static int32_t bn_stats_noalloc_impl(const T* x, long long rows, int C, const float* gamma, const float* beta,
                                     float* rm, float* rv, float momentum, float eps, float* save_mean,
                                     float* save_invstd, float* scale, float* shift, float* part, cudaStream_t st) {
  // same kernels as the workspace variant; only the number of row blocks is capped so that the
  // [nblk][2][C] partial sums fit the fixed array.  Every slot that the finalize kernel reads is written by this
  // launch, so the contents left behind by an earlier call never matter and nothing needs clearing.
  BnPlan p = bn_plan<T>(rows, C);
  const long long max_blk = (long long)(BN_NOALLOC_FLOATS / (2 * (size_t)C));
  if (p.nblk > max_blk) {
    const int ry = BN_THREADS / p.cx;
    long long rpb = (rows + max_blk - 1) / max_blk;
    rpb = (rpb + ry - 1) / ry * ry;
    p.rows_per_block = rpb;
    p.nblk = (int)((rows + rpb - 1) / rpb);
  }
  launch_bn_partial<T>(p, PM_FWD, x, nullptr, nullptr, nullptr, nullptr, rows, C, part, st);
  bn_finalize_fwd_kernel<<<(C + FIN_CX - 1) / FIN_CX, dim3(FIN_CX, FIN_RY), 0, st>>>(part, p.nblk, C, (double)rows, gamma, beta, rm, rv, momentum, eps, save_mean, save_invstd, scale, shift);
  return check_launch("bn_train_stats_noalloc", 2);
}

}  // namespace b200

extern "C" int32_t b200_bn_train_stats_noalloc(const void* x, int64_t rows, int32_t C, int32_t dtype,
                                               const float* gamma, const float* beta, float* running_mean,
                                               float* running_var, float momentum, float eps, float* save_mean,
                                               float* save_invstd, float* scale, float* shift, void* stream) {
  if (!x || rows <= 0 || C <= 0 || !save_mean || !save_invstd || !scale || !shift)
    return fail(B200_INVALID_ARG, "bn_train_stats_noalloc: bad arguments");
  if (C > B200_BN_NOALLOC_MAX_C)
    return fail(B200_UNSUPPORTED, "bn_train_stats_noalloc: C=%d exceeds the built-in scratch limit of %d channels; "
                "use b200_bn_train_stats with a caller-provided workspace", (int)C, B200_BN_NOALLOC_MAX_C);
  if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return fail(B200_INVALID_ARG, "bn_train_stats_noalloc: x must be 16-byte aligned");
  float* part = nullptr;
  if (int32_t rc = bn_noalloc_scratch_ptr(&part)) return rc;
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32)
    return bn_stats_noalloc_impl<float>((const float*)x, rows, C, gamma, beta, running_mean, running_var, momentum, eps, save_mean, save_invstd, scale, shift, part, st);
  if (dtype == B200_BF16)
    return bn_stats_noalloc_impl<__nv_bfloat16>((const __nv_bfloat16*)x, rows, C, gamma, beta, running_mean, running_var, momentum, eps, save_mean, save_invstd, scale, shift, part, st);
  return fail(B200_INVALID_ARG, "bn_train_stats_noalloc: bad dtype");
}
