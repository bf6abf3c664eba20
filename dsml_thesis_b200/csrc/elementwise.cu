// HBM-bound kernels of the denoising path: fused DDIM update (K13), timestep embedding + small dense layers
// (K12), GroupNorm+SiLU over (virtually concatenated) channels-last maps (K7/K11), LayerNorm (K9), layout
// changes, and the one-launch weight packer for the tensor-core path.
#include <algorithm>
#include <stdlib.h>

#include <cuda_fp16.h>

#include "common.cuh"

namespace ldm {

template <typename T>
struct Store4;
template <>
struct Store4<float> {
  static __device__ __forceinline__ void st(float* p, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
  }
};
template <>
struct Store4<__nv_bfloat16> {
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float a, float b, float c, float d) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&lo);
    u.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = u;
  }
};

template <>
struct Store4<__half> {  // conversions saturate to +-65504 instead of overflowing to inf
  static __device__ __forceinline__ void st(__half* p, float a, float b, float c, float d) {
    uint2 u;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(u.x) : "f"(b), "f"(a));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(u.y) : "f"(d), "f"(c));
    *reinterpret_cast<uint2*>(p) = u;
  }
};
__device__ __forceinline__ void cvt_store(float* p, float v) { *p = v; }
__device__ __forceinline__ void cvt_store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
__device__ __forceinline__ void cvt_store(__half* p, float v) {
  *p = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
}

// launches `call` with T = the C type of an LDMB200_* dtype code
#define LDM_DISPATCH_DTYPE(dt, call)                   \
  do {                                                 \
    if ((dt) == LDMB200_BF16) {                        \
      using T = __nv_bfloat16;                         \
      call;                                            \
    } else if ((dt) == LDMB200_F16) {                  \
      using T = __half;                                \
      call;                                            \
    } else {                                           \
      using T = float;                                 \
      call;                                            \
    }                                                  \
  } while (0)

// ---------------------------------------------------------------------------------------------
// K13 DDIM update.  Each product/sum is a separate rounded fp32 op (no FMA contraction) so the result
// is bit-identical to the reference's chain of elementwise torch ops (ddim.py:177,194-202).
// ---------------------------------------------------------------------------------------------
struct DdimCoef {
  float s1;        // sqrt_one_minus_at
  float sqrt_at;   // sqrtf(a_t)
  float sqrt_ap;   // sqrtf(a_prev)
  float dir;       // sqrtf(1 - a_prev - sigma^2)
  float sigma, temperature, cfg;
};

__device__ __forceinline__ void ddim_one(float x, float ec, float eu, float nz, bool has_u, bool has_n,
                                         const DdimCoef& k, float& xp, float& p0) {
  float e = has_u ? __fadd_rn(eu, __fmul_rn(k.cfg, __fsub_rn(ec, eu))) : ec;
  p0 = __fdiv_rn(__fsub_rn(x, __fmul_rn(k.s1, e)), k.sqrt_at);
  float dir = __fmul_rn(k.dir, e);
  float noise = has_n ? __fmul_rn(__fmul_rn(k.sigma, nz), k.temperature) : 0.0f;
  xp = __fadd_rn(__fadd_rn(__fmul_rn(k.sqrt_ap, p0), dir), noise);
}

__global__ void __launch_bounds__(256) ddim_step_kernel(const float* __restrict__ x, const float* __restrict__ ec,
                                                        const float* __restrict__ eu,
                                                        const float* __restrict__ nz, float* __restrict__ xp,
                                                        float* __restrict__ p0, int64_t n, float a_t, float a_prev,
                                                        float sigma, float s1, float cfg, float temperature) {
  DdimCoef k;
  k.s1 = s1;
  k.sqrt_at = __fsqrt_rn(a_t);
  k.sqrt_ap = __fsqrt_rn(a_prev);
  k.dir = __fsqrt_rn(__fsub_rn(__fsub_rn(1.0f, a_prev), __fmul_rn(sigma, sigma)));
  k.sigma = sigma;
  k.temperature = temperature;
  k.cfg = cfg;
  const bool has_u = eu != nullptr, has_n = nz != nullptr;
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 vx = reinterpret_cast<const float4*>(x)[i];
    float4 ve = reinterpret_cast<const float4*>(ec)[i];
    float4 vu = has_u ? reinterpret_cast<const float4*>(eu)[i] : make_float4(0, 0, 0, 0);
    float4 vn = has_n ? reinterpret_cast<const float4*>(nz)[i] : make_float4(0, 0, 0, 0);
    float4 ox, op;
    ddim_one(vx.x, ve.x, vu.x, vn.x, has_u, has_n, k, ox.x, op.x);
    ddim_one(vx.y, ve.y, vu.y, vn.y, has_u, has_n, k, ox.y, op.y);
    ddim_one(vx.z, ve.z, vu.z, vn.z, has_u, has_n, k, ox.z, op.z);
    ddim_one(vx.w, ve.w, vu.w, vn.w, has_u, has_n, k, ox.w, op.w);
    reinterpret_cast<float4*>(xp)[i] = ox;
    reinterpret_cast<float4*>(p0)[i] = op;
  }
  // scalar tail
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float ox, op;
    ddim_one(x[i], ec[i], has_u ? eu[i] : 0.f, has_n ? nz[i] : 0.f, has_u, has_n, k, ox, op);
    xp[i] = ox;
    p0[i] = op;
  }
}

// ---------------------------------------------------------------------------------------------
// K12 timestep embedding.  The fp32 argument chain matches torch's (scalar*arange, /half); exp and
// cos/sin are evaluated in double and rounded once so the result sits within 1 ulp of torch's fp32 ops.
// ---------------------------------------------------------------------------------------------
__global__ void timestep_embedding_kernel(const int64_t* __restrict__ t, float* __restrict__ out, int batch, int dim,
                                          float neg_log_period) {
  pdl_wait();
  const int half = dim / 2;
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= batch * half) return;
  int b = idx / half, i = idx % half;
  float a = __fdiv_rn(__fmul_rn(neg_log_period, (float)i), (float)half);
  float freq = (float)exp((double)a);
  float arg = __fmul_rn((float)t[b], freq);
  out[(size_t)b * dim + i] = (float)cos((double)arg);
  out[(size_t)b * dim + half + i] = (float)sin((double)arg);
}

// ---------------------------------------------------------------------------------------------
// small dense layers on per-sample vectors (fp32 FMA, exact-parity path): out = act(in) W^T + bias.
// Block = 256 threads, output tile 64 rows x 32 columns, K in chunks of 32 staged in shared memory with the
// next chunk prefetched into registers; thread (ty, tx) owns rows 4ty..4ty+3, columns 2tx..2tx+1.
// One launch serves a whole table of independent layers (all ResBlock emb projections of a UNet eval, or all
// cross-attention value projections): block -> job by the job's first block index.
// (r02s: a warp-per-output-column variant with 64 row accumulators per lane and n/8 blocks per layer was measured and is
// 2x SLOWER — one shared-memory read per FMA; this register tile does 8 FMAs per 2 reads.)
// ---------------------------------------------------------------------------------------------
constexpr int SL_BM = 64, SL_BN = 32, SL_BK = 32, SL_XP = 68, SL_WP = 34;

__device__ __forceinline__ void small_linear_tile(const ldmb200_linear_job& j, int blk) {
  __shared__ __align__(16) float xs[SL_BK][SL_XP];
  __shared__ __align__(16) float ws[SL_BK][SL_WP];
  const int nb = (j.n + SL_BN - 1) / SL_BN;
  const int r0 = (blk / nb) * SL_BM, n0 = (blk % nb) * SL_BN;
  const int tid = threadIdx.x, lk = tid & 31, lr = tid >> 5;  // loader: k = lk, rows / columns lr + 8i
  const int ty = tid >> 4, tx = tid & 15;
  float acc[4][2] = {};
  float xr[8], wr[4];
  auto fetch = [&](int k0) {
    const int k = k0 + lk;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = r0 + lr + 8 * i;
      float v = 0.f;
      if (r < j.rows && k < j.k) {
        v = j.in[(size_t)r * j.in_ld + k];
        if (j.silu_in) v = silu_f(v);
      }
      xr[i] = v;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = n0 + lr + 8 * i;
      wr[i] = (n < j.n && k < j.k) ? j.w[(size_t)n * j.k + k] : 0.f;
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < j.k; k0 += SL_BK) {
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) xs[lk][lr + 8 * i] = xr[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) ws[lk][lr + 8 * i] = wr[i];
    __syncthreads();
    if (k0 + SL_BK < j.k) fetch(k0 + SL_BK);
#pragma unroll
    for (int kk = 0; kk < SL_BK; ++kk) {
      const float4 xv = *reinterpret_cast<const float4*>(&xs[kk][ty * 4]);
      const float2 wv = *reinterpret_cast<const float2*>(&ws[kk][tx * 2]);
      acc[0][0] = fmaf(xv.x, wv.x, acc[0][0]); acc[0][1] = fmaf(xv.x, wv.y, acc[0][1]);
      acc[1][0] = fmaf(xv.y, wv.x, acc[1][0]); acc[1][1] = fmaf(xv.y, wv.y, acc[1][1]);
      acc[2][0] = fmaf(xv.z, wv.x, acc[2][0]); acc[2][1] = fmaf(xv.z, wv.y, acc[2][1]);
      acc[3][0] = fmaf(xv.w, wv.x, acc[3][0]); acc[3][1] = fmaf(xv.w, wv.y, acc[3][1]);
    }
  }
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int r = r0 + ty * 4 + a;
    if (r >= j.rows) continue;
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int n = n0 + tx * 2 + b;
      if (n < j.n) j.out[(size_t)r * j.out_ld + n] = acc[a][b] + (j.bias ? j.bias[n] : 0.f);
    }
  }
}

__global__ void __launch_bounds__(256) small_linear_kernel(const __grid_constant__ ldmb200_linear_job job) {
  pdl_wait();
  small_linear_tile(job, blockIdx.x);
}

__global__ void __launch_bounds__(256) small_linear_batched_kernel(const ldmb200_linear_job* __restrict__ jobs,
                                                                   int n_jobs) {
  pdl_wait();
  int lo = 0;  // last job whose first_block <= blockIdx.x (tables are a few dozen entries)
  for (int i = 1; i < n_jobs; ++i)
    if (jobs[i].first_block <= (int)blockIdx.x) lo = i;
  const ldmb200_linear_job j = jobs[lo];
  small_linear_tile(j, blockIdx.x - j.first_block);
}

// ---------------------------------------------------------------------------------------------
// layout changes
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, T* __restrict__ y, int batch, int C, int HW,
                                    int c_pad) {
  pdl_wait();
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)batch * HW) return;
  int b = idx / HW, p = idx % HW;
  T* dst = y + idx * c_pad;
  for (int c = 0; c < c_pad; ++c) {
    float v = c < C ? x[((size_t)b * C + c) * HW + p] : 0.f;
    cvt_store(dst + c, v);
  }
}

__global__ void nhwc_to_nchw_kernel(const float* __restrict__ y, int ld, float* __restrict__ x, int batch, int C,
                                    int HW) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)batch * HW) return;
  int b = idx / HW, p = idx % HW;
  for (int c = 0; c < C; ++c) x[((size_t)b * C + c) * HW + p] = y[idx * ld + c];
}

// x fp32 [B,H,W,C] -> y [B,4,H/2,W/2,C], plane = (h&1)*2 + (w&1)
template <typename T>
__global__ void space_to_depth2_kernel(const float* __restrict__ x, T* __restrict__ y, int batch, int H, int W,
                                       int C) {
  pdl_wait();
  const int c4n = C >> 2, Ho = H >> 1, Wo = W >> 1;
  int64_t total = (int64_t)batch * H * W * c4n;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    int c4 = idx % c4n;
    int64_t pix = idx / c4n;  // destination pixel index over [B,4,Ho,Wo]
    int wo = pix % Wo;
    int64_t t = pix / Wo;
    int ho = t % Ho;
    t /= Ho;
    int plane = t & 3;
    int b = t >> 2;
    int h = ho * 2 + (plane >> 1), w = wo * 2 + (plane & 1);
    float4 v = *reinterpret_cast<const float4*>(x + (((size_t)b * H + h) * W + w) * C + c4 * 4);
    Store4<T>::st(y + pix * C + c4 * 4, v.x, v.y, v.z, v.w);
  }
}

// x fp32 [B,H,W,C] -> y [B,2H,2W,C] nearest
template <typename T>
__global__ void upsample2x_kernel(const float* __restrict__ x, T* __restrict__ y, int batch, int H, int W, int C) {
  const int c4n = C >> 2, Ho = H * 2, Wo = W * 2;
  int64_t total = (int64_t)batch * Ho * Wo * c4n;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    int c4 = idx % c4n;
    int64_t pix = idx / c4n;
    int wo = pix % Wo;
    int64_t t = pix / Wo;
    int ho = t % Ho;
    int b = t / Ho;
    float4 v = *reinterpret_cast<const float4*>(x + (((size_t)b * H + (ho >> 1)) * W + (wo >> 1)) * C + c4 * 4);
    Store4<T>::st(y + pix * C + c4 * 4, v.x, v.y, v.z, v.w);
  }
}

// x fp32 [n] -> y [n] (16-bit GEMM operand of a map that is consumed as it is: the folded upsample convolution)
template <typename T>
__global__ void cast_rows_kernel(const float* __restrict__ x, T* __restrict__ y, int64_t n4) {
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    Store4<T>::st(y + i * 4, v.x, v.y, v.z, v.w);
  }
}

// fp16 operand range guard: elements at the clamp of the saturating conversion (|v| >= 65504) or not finite
__global__ void count_f16_saturated_kernel(const uint16_t* __restrict__ x, int64_t n, unsigned long long* __restrict__ counter) {
  unsigned int c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    c += (x[i] & 0x7fffu) >= 0x7bffu;
  c = __reduce_add_sync(0xffffffffu, c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(counter, (unsigned long long)c);
}

// ---------------------------------------------------------------------------------------------
// K7 GroupNorm (+SiLU) over channels-last [B, HW, C1 (+C2)].  Two launches:
//   stats: grid (S, B); thread owns one channel quad, sums x and x^2 over its rows; block reduces channel
//          partials to 32 group partials -> ws[b][s][g][2]   (deterministic, no atomics)
//   apply: grid (T, B); combines the S partials in float64, folds gamma/beta into per-channel (scale, shift)
//          registers, streams the slab once: y = silu(x*scale + shift)  (+ raw cast copy)
// ---------------------------------------------------------------------------------------------
constexpr int GN_MAX_SPLITS = 64;
#ifndef LDM_GN_UNR
#define LDM_GN_UNR 4
#endif
constexpr int GN_UNR = LDM_GN_UNR;   // rows per thread in flight in the apply pass

static inline int gn_rows_per_block(int C) {
  int q = C / 4;
  int r = 256 / q;
  return r < 1 ? 1 : r;
}
static inline int gn_splits(int hw, int C) {
  int R = gn_rows_per_block(C);
  int s = hw / (R * 8);
  if (s < 1) s = 1;
  if (s > GN_MAX_SPLITS) s = GN_MAX_SPLITS;
  return s;
}

__device__ __forceinline__ float4 gn_load(const float* __restrict__ x1, int c1, const float* __restrict__ x2, int c2,
                                          int64_t row, int c) {
  if (c < c1) return *reinterpret_cast<const float4*>(x1 + row * c1 + c);
  return *reinterpret_cast<const float4*>(x2 + row * c2 + (c - c1));
}

__global__ void gn_stats_kernel(const float* __restrict__ x1, int c1, const float* __restrict__ x2, int c2, int hw,
                                int groups, int R, float* __restrict__ ws) {
  extern __shared__ float sm[];  // [2][R][C]
  const int C = c1 + c2, q = C >> 2;
  const int cq = threadIdx.x % q, rl = threadIdx.x / q;
  const int S = gridDim.x, s = blockIdx.x, b = blockIdx.y;
  const int rows_per = (hw + S - 1) / S;
  const int r0 = s * rows_per, r1 = min(hw, r0 + rows_per);
  float sx[4] = {0, 0, 0, 0}, sq[4] = {0, 0, 0, 0};
  for (int r = r0 + rl; r < r1; r += R) {
    float4 v = gn_load(x1, c1, x2, c2, (int64_t)b * hw + r, cq * 4);
    sx[0] += v.x; sx[1] += v.y; sx[2] += v.z; sx[3] += v.w;
    sq[0] = fmaf(v.x, v.x, sq[0]); sq[1] = fmaf(v.y, v.y, sq[1]);
    sq[2] = fmaf(v.z, v.z, sq[2]); sq[3] = fmaf(v.w, v.w, sq[3]);
  }
  float* ps = sm;
  float* pq = sm + R * C;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    ps[rl * C + cq * 4 + j] = sx[j];
    pq[rl * C + cq * 4 + j] = sq[j];
  }
  __syncthreads();
  const int cpg = C / groups;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    float a = 0.f, c = 0.f;
    for (int r = 0; r < R; ++r)
      for (int j = 0; j < cpg; ++j) {
        a += ps[r * C + g * cpg + j];
        c += pq[r * C + g * cpg + j];
      }
    float* o = ws + (((size_t)b * S + s) * groups + g) * 2;
    o[0] = a;
    o[1] = c;
  }
}

// Fused-statistics path: the GEMM epilogues that wrote x1 / x2 left {sum, sum of squares} per column per 32-row block
// (ldmb200_gemm_desc.colstats).  One block per (group, sample) folds them in float64 -> ws[b][g] = {mean, rstd}.
__global__ void __launch_bounds__(128) gn_finalize_kernel(const float* __restrict__ cs1, int c1,
                                                          const float* __restrict__ cs2, int c2, int hw, int groups,
                                                          float eps, float* __restrict__ ws) {
  pdl_wait();
  __shared__ double red[2][4];
  const int g = blockIdx.x, b = blockIdx.y;
  const int C = c1 + c2, cpg = C / groups, nrb = hw >> 5;
  double a = 0.0, q = 0.0;
  for (int i = threadIdx.x; i < nrb * cpg; i += blockDim.x) {
    const int rb = i / cpg, c = g * cpg + i % cpg;
    const size_t blk = (size_t)b * nrb + rb;
    const float2 v = (c < c1) ? reinterpret_cast<const float2*>(cs1)[blk * c1 + c]
                              : reinterpret_cast<const float2*>(cs2)[blk * c2 + (c - c1)];
    a += (double)v.x;
    q += (double)v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = a;
    red[1][threadIdx.x >> 5] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = (red[0][0] + red[0][1]) + (red[0][2] + red[0][3]);
    q = (red[1][0] + red[1][1]) + (red[1][2] + red[1][3]);
    const double n = (double)cpg * (double)hw;
    const double mean = a / n;
    double var = q / n - mean * mean;
    if (var < 0.0) var = 0.0;
    ws[((size_t)b * groups + g) * 2] = (float)mean;
    ws[((size_t)b * groups + g) * 2 + 1] = (float)(1.0 / sqrt(var + (double)eps));
  }
}

// MAXT = 256 (every shipped width up to C = 960): capped at 48 registers so that five blocks are resident per SM.
template <typename T, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT == 256 ? (GN_UNR > 4 ? 4 : 5) : (MAXT == 512 ? 2 : 1))
gn_apply_kernel(const float* __restrict__ x1, int c1, const float* __restrict__ x2, int c2,
                                const float* __restrict__ gamma, const float* __restrict__ beta, T* __restrict__ y,
                                T* __restrict__ raw, int hw, int groups, int R, int S, float eps, int silu,
                                const float* __restrict__ ws) {
  pdl_wait();
  __shared__ float s_mean[64], s_rstd[64];
  const int C = c1 + c2, q = C >> 2, cpg = C / groups;
  const int b = blockIdx.y;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    if (S == 0) {  // fused path: gn_finalize_kernel already reduced the producers' column statistics
      s_mean[g] = ws[((size_t)b * groups + g) * 2];
      s_rstd[g] = ws[((size_t)b * groups + g) * 2 + 1];
      continue;
    }
    double a = 0.0, c = 0.0;
    for (int s = 0; s < S; ++s) {
      const float* p = ws + (((size_t)b * S + s) * groups + g) * 2;
      a += (double)p[0];
      c += (double)p[1];
    }
    double n = (double)cpg * (double)hw;
    double mean = a / n;
    double var = c / n - mean * mean;
    if (var < 0.0) var = 0.0;
    s_mean[g] = (float)mean;
    s_rstd[g] = (float)(1.0 / sqrt(var + (double)eps));
  }
  __syncthreads();
  const int cq = threadIdx.x % q, rl = threadIdx.x / q;
  float sc[4], sh[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int c = cq * 4 + j;
    int g = c / cpg;
    float ga = gamma[c] * s_rstd[g];
    sc[j] = ga;
    sh[j] = beta[c] - s_mean[g] * ga;
  }
  const int T_ = gridDim.x;
  const int rows_per = (hw + T_ - 1) / T_;
  const int r0 = blockIdx.x * rows_per, r1 = min(hw, r0 + rows_per);
  // GN_UNR rows per thread in flight: the pass is a pure stream and needs the memory-level parallelism
  int r = r0 + rl;
  for (; r + (GN_UNR - 1) * R < r1; r += GN_UNR * R) {
    float4 v[GN_UNR];
#pragma unroll
    for (int u = 0; u < GN_UNR; ++u) v[u] = gn_load(x1, c1, x2, c2, (int64_t)b * hw + r + u * R, cq * 4);
#pragma unroll
    for (int u = 0; u < GN_UNR; ++u) {
      const int64_t row = (int64_t)b * hw + r + u * R;
      float o0 = fmaf(v[u].x, sc[0], sh[0]), o1 = fmaf(v[u].y, sc[1], sh[1]);
      float o2 = fmaf(v[u].z, sc[2], sh[2]), o3 = fmaf(v[u].w, sc[3], sh[3]);
      if (silu) {
        o0 = silu_f(o0); o1 = silu_f(o1); o2 = silu_f(o2); o3 = silu_f(o3);
      }
      Store4<T>::st(y + row * C + cq * 4, o0, o1, o2, o3);
      if (raw) Store4<T>::st(raw + row * C + cq * 4, v[u].x, v[u].y, v[u].z, v[u].w);
    }
  }
  for (; r < r1; r += R) {
    int64_t row = (int64_t)b * hw + r;
    float4 v = gn_load(x1, c1, x2, c2, row, cq * 4);
    float o0 = fmaf(v.x, sc[0], sh[0]), o1 = fmaf(v.y, sc[1], sh[1]);
    float o2 = fmaf(v.z, sc[2], sh[2]), o3 = fmaf(v.w, sc[3], sh[3]);
    if (silu) {
      o0 = silu_f(o0); o1 = silu_f(o1); o2 = silu_f(o2); o3 = silu_f(o3);
    }
    Store4<T>::st(y + row * C + cq * 4, o0, o1, o2, o3);
    if (raw) Store4<T>::st(raw + row * C + cq * 4, v.x, v.y, v.z, v.w);
  }
}

// ---------------------------------------------------------------------------------------------
// K9 LayerNorm: one warp per row, two-pass statistics on register-resident values.
// ---------------------------------------------------------------------------------------------
template <typename T, int MAXV>
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, T* __restrict__ y,
                                                        int64_t rows, int C, float eps) {
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int q = C >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + row * C);
  float4 v[MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int idx = lane + i * 32;
    if (idx < q) {
      v[i] = xr[idx];
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float mean = warp_sum(s) / (float)C;
  float d = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int idx = lane + i * 32;
    if (idx < q) {
      float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, e = v[i].w - mean;
      d += (a * a + b * b) + (c * c + e * e);
    }
  }
  const float rstd = rsqrtf(warp_sum(d) / (float)C + eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int idx = lane + i * 32;
    if (idx < q) {
      float4 g = reinterpret_cast<const float4*>(gamma)[idx];
      float4 bb = reinterpret_cast<const float4*>(beta)[idx];
      Store4<T>::st(y + row * C + idx * 4, (v[i].x - mean) * rstd * g.x + bb.x, (v[i].y - mean) * rstd * g.y + bb.y,
                    (v[i].z - mean) * rstd * g.z + bb.z, (v[i].w - mean) * rstd * g.w + bb.w);
    }
  }
}

// C = 20 * LPR floats (160 / 320 / 640: every SpatialTransformer width of the shipped UNets): LPR lanes share a row,
// five float4 per lane, so a warp normalises 32 / LPR rows at once with every lane busy and 80 bytes of loads in
// flight per lane (the one-warp-per-row kernel above leaves 24 of 32 lanes idle on its second trip at C = 160 and
// keeps only 640 B in flight per warp).  Statistics: two-pass on the register-resident row, xor-shuffles inside the
// LPR-lane group.
template <typename T, int LPR>
__global__ void __launch_bounds__(256) layernorm_c20_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, T* __restrict__ y,
                                                            int64_t rows, float eps) {
  pdl_wait();
  constexpr int RPW = 32 / LPR, C = 20 * LPR;
  const int lane = threadIdx.x & 31, sub = lane % LPR;
  const int64_t row = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + lane / LPR;
  const bool live = row < rows;
  const float4* xr = reinterpret_cast<const float4*>(x + (live ? row : 0) * C);
  float4 v[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) v[i] = live ? xr[sub + i * LPR] : make_float4(0.f, 0.f, 0.f, 0.f);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 5; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
#pragma unroll
  for (int o = LPR >> 1; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s * (1.0f / (float)C);
  float d = 0.f;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, e = v[i].w - mean;
    d += (a * a + b * b) + (c * c + e * e);
  }
#pragma unroll
  for (int o = LPR >> 1; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  const float rstd = rsqrtf(d * (1.0f / (float)C) + eps);
  if (!live) return;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int idx = sub + i * LPR;
    const float4 g = reinterpret_cast<const float4*>(gamma)[idx];
    const float4 bb = reinterpret_cast<const float4*>(beta)[idx];
    Store4<T>::st(y + row * C + idx * 4, (v[i].x - mean) * rstd * g.x + bb.x, (v[i].y - mean) * rstd * g.y + bb.y,
                  (v[i].z - mean) * rstd * g.z + bb.z, (v[i].w - mean) * rstd * g.w + bb.w);
  }
}

// ---------------------------------------------------------------------------------------------
// weight packer: fp32 nn.Parameter -> bf16 [n_pad, taps, c_pad] (K-major), optional GEGLU row permutation
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int geglu_src_row(int rp, int n_total, int block_n) {
  int t = rp / block_n, rem = rp % block_n;
  int qd = rem >> 5, i = rem & 31;
  int col = t * (block_n >> 1) + qd * 16 + (i & 15);
  return (i < 16) ? col : (n_total >> 1) + col;
}

__global__ void __launch_bounds__(256) pack_weights_kernel(const ldmb200_pack_job* __restrict__ jobs) {
  const ldmb200_pack_job j = jobs[blockIdx.y];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j.kind == 2) {
    float* dst = reinterpret_cast<float*>(j.dst);
    for (int64_t i = i0; i < j.n_pad; i += stride) {
      int src = geglu_src_row((int)i, j.n, j.block_n);
      dst[j.row_offset + i] = src < j.n ? j.src[src] : 0.f;
    }
    return;
  }
  const int taps = j.ksize * j.ksize;
  const int64_t kp = (int64_t)taps * j.c_pad;
  uint16_t* dst = reinterpret_cast<uint16_t*>(j.dst);
  if (j.kind == 3) {
    // data-gradient operand: rows = source input channels (padded to n_pad), row = [tap'][c_pad columns], this job
    // fills columns [row_offset, row_offset + block_n) from source output channels [0, n); taps flipped
    const int ncols = j.block_n;
    const int64_t total3 = (int64_t)j.n_pad * taps * ncols;
    for (int64_t i = i0; i < total3; i += stride) {
      const int nn = (int)(i % ncols);
      const int64_t t = i / ncols;
      const int tapd = (int)(t % taps), cc = (int)(t / taps);
      float v = 0.f;
      if (nn < j.n && cc < j.c) v = j.src[((size_t)nn * j.c + cc) * taps + (taps - 1 - tapd)];
      dst[(size_t)cc * kp + (size_t)tapd * j.c_pad + j.row_offset + nn] =
          j.dst_dtype == LDMB200_F16 ? __half_as_ushort(__float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f)))
                                     : __bfloat16_as_ushort(__float2bfloat16_rn(v));
    }
    return;
  }
  if (j.kind == 4) {
    // K10 folded into the conv: nearest-2x + 3x3 (pad 1) = per output parity (py, px) a 2x2 convolution on the low-res
    // map whose tap (a, b) sums the 3x3 taps that land on the same low-res pixel: rows {py=0: a0 <- kh 0, a1 <- kh 1,2;
    // py=1: a0 <- kh 0,1, a1 <- kh 2}, columns alike.  dst = [4 parities][n_pad][4 taps][c_pad]; sums in fp32.
    const int64_t kp4 = 4 * (int64_t)j.c_pad, per = (int64_t)j.n_pad * kp4;
    for (int64_t i = i0; i < 4 * per; i += stride) {
      const int par = (int)(i / per);
      const int64_t r = i % per;
      const int rp = (int)(r / kp4), rem = (int)(r % kp4), tap = rem / j.c_pad, c = rem % j.c_pad;
      const int py = par >> 1, px = par & 1, a = tap >> 1, b = tap & 1;
      const int kh0 = (py == 0) ? (a == 0 ? 0 : 1) : (a == 0 ? 0 : 2), kh1 = (py == 0) ? (a == 0 ? 0 : 2) : (a == 0 ? 1 : 2);
      const int kw0 = (px == 0) ? (b == 0 ? 0 : 1) : (b == 0 ? 0 : 2), kw1 = (px == 0) ? (b == 0 ? 0 : 2) : (b == 0 ? 1 : 2);
      float v = 0.f;
      if (rp < j.n && c < j.c) {
        const float* w = j.src + ((size_t)rp * j.c + c) * 9;
        for (int kh = kh0; kh <= kh1; ++kh)
          for (int kw = kw0; kw <= kw1; ++kw) v += w[kh * 3 + kw];
      }
      dst[i] = j.dst_dtype == LDMB200_F16 ? __half_as_ushort(__float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f)))
                                          : __bfloat16_as_ushort(__float2bfloat16_rn(v));
    }
    return;
  }
  const int64_t total = (int64_t)j.n_pad * kp;
  for (int64_t i = i0; i < total; i += stride) {
    int rp = (int)(i / kp);
    int rem = (int)(i % kp);
    int tap = rem / j.c_pad, c = rem % j.c_pad;
    int src_row = j.kind == 1 ? geglu_src_row(rp, j.n, j.block_n) : rp;
    float v = 0.f;
    if (src_row < j.n && c < j.c) v = j.src[((size_t)src_row * j.c + c) * taps + tap];
    dst[(size_t)(j.row_offset + rp) * kp + rem] =
        j.dst_dtype == LDMB200_F16 ? __half_as_ushort(__float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f)))
                                   : __bfloat16_as_ushort(__float2bfloat16_rn(v));
  }
}

}  // namespace ldm

using namespace ldm;

// Blocks of the apply pass per sample: ONE wave of resident blocks over the whole batch.  (r01 sized the grid as
// ceil(8 * SMs / batch) while the 64-register kernel fitted 4 blocks per SM: 1216 blocks on 592 slots = 2.05 waves, the
// third nearly empty — the reason the pass sat at 4.1 TB/s next to the LayerNorm's 5.8 TB/s on the same byte pattern.)
template <typename T>
static int gn_apply_launch(const float* x1, int c1, const float* x2, int c2, const float* gamma, const float* beta, void* y,
                           void* raw, int batch, int hw, int groups, int R, int S, float eps, int silu, const float* ws,
                           cudaStream_t st) {
  const int threads = ((c1 + c2) / 4) * R;
  static int occ_small = 0, occ_big = 0;
  int& occ = threads <= 256 ? occ_small : occ_big;
  if (occ == 0) {
    cudaError_t e = threads <= 256
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gn_apply_kernel<T, 256>, 256, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gn_apply_kernel<T, 512>, 512, 0);
    if (e != cudaSuccess || occ < 1) occ = 1;
  }
  static const int per_sm_env = getenv("LDMB200_GN_BLOCKS_PER_SM") ? atoi(getenv("LDMB200_GN_BLOCKS_PER_SM")) : 0;
  const int slots = sm_count() * (per_sm_env > 0 ? per_sm_env : occ);
  int nblk = hw / (R * GN_UNR);
  if (nblk > slots / batch) nblk = slots / batch;
  if (nblk < 1) nblk = 1;
  if (threads <= 256)
    launch_pdl(gn_apply_kernel<T, 256>, dim3(nblk, batch), dim3(threads), 0, st, x1, c1, x2, c2, gamma, beta, (T*)y, (T*)raw, hw,
               groups, R, S, eps, silu, ws);
  else if (threads <= 512)
    launch_pdl(gn_apply_kernel<T, 512>, dim3(nblk, batch), dim3(threads), 0, st, x1, c1, x2, c2, gamma, beta, (T*)y, (T*)raw, hw,
               groups, R, S, eps, silu, ws);
  else  // C > 2048: one block per SM slot
    launch_pdl(gn_apply_kernel<T, 1024>, dim3(nblk, batch), dim3(threads), 0, st, x1, c1, x2, c2, gamma, beta, (T*)y, (T*)raw, hw,
               groups, R, S, eps, silu, ws);
  return 0;
}

extern "C" {

int ldmb200_ddim_step(const float* x, const float* e_cond, const float* e_uncond, const float* noise, float* x_prev,
                      float* pred_x0, int64_t n, float a_t, float a_prev, float sigma_t, float sqrt_one_minus_at,
                      float cfg_scale, float temperature, void* stream) {
  LDM_CHECK_ARG(x && e_cond && x_prev && pred_x0 && n > 0, "ddim_step: null pointer or n <= 0");
  LDM_CHECK_ARG((((uintptr_t)x | (uintptr_t)e_cond | (uintptr_t)e_uncond | (uintptr_t)noise | (uintptr_t)x_prev |
                  (uintptr_t)pred_x0) & 15) == 0, "ddim_step: pointers must be 16-byte aligned");
  int64_t n4 = (n + 3) / 4;
  int blocks = (int)((n4 + 255) / 256);
  int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  ddim_step_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, e_cond, e_uncond, noise, x_prev, pred_x0, n, a_t,
                                                             a_prev, sigma_t, sqrt_one_minus_at, cfg_scale,
                                                             temperature);
  LDM_CHECK_LAUNCH("ddim_step");
  return 0;
}

int ldmb200_timestep_embedding(const int64_t* t, float* out, int batch, int dim, float max_period, void* stream) {
  LDM_CHECK_ARG(t && out && batch > 0 && dim > 0 && dim % 2 == 0, "timestep_embedding: bad arguments");
  int total = batch * (dim / 2);
  // -math.log(max_period) is a python float rounded to fp32 when it multiplies the fp32 arange (util.py:162-164)
  float neg_log = (float)(-log((double)max_period));
  launch_pdl(timestep_embedding_kernel, dim3(ceil_div(total, 128)), dim3(128), 0, (cudaStream_t)stream, t, out, batch, dim, neg_log);
  LDM_CHECK_LAUNCH("timestep_embedding");
  return 0;
}

int ldmb200_small_linear(const float* in, int in_ld, const float* w, const float* bias, float* out, int out_ld,
                         int rows, int k, int n, int silu_in, void* stream) {
  LDM_CHECK_ARG(in && w && out && rows > 0 && k > 0 && n > 0, "small_linear: bad arguments");
  ldmb200_linear_job j;
  j.in = in; j.w = w; j.bias = bias; j.out = out;
  j.in_ld = in_ld; j.out_ld = out_ld; j.rows = rows; j.k = k; j.n = n; j.silu_in = silu_in; j.first_block = 0;
  const int blocks = ceil_div(n, SL_BN) * ceil_div(rows, SL_BM);
  launch_pdl(small_linear_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, j);
  LDM_CHECK_LAUNCH("small_linear");
  return 0;
}

int ldmb200_small_linear_blocks(int rows, int n) { return ceil_div(n, SL_BN) * ceil_div(rows, SL_BM); }

int ldmb200_small_linear_batched(const ldmb200_linear_job* jobs, int n_jobs, int total_blocks, void* stream) {
  LDM_CHECK_ARG(jobs && n_jobs > 0 && total_blocks > 0, "small_linear_batched: bad arguments");
  launch_pdl(small_linear_batched_kernel, dim3(total_blocks), dim3(256), 0, (cudaStream_t)stream, jobs, n_jobs);
  LDM_CHECK_LAUNCH("small_linear_batched");
  return 0;
}

int ldmb200_nchw_to_nhwc(const float* x, void* y, int batch, int c, int hw, int c_pad, int out_dtype, void* stream) {
  LDM_CHECK_ARG(x && y && batch > 0 && c > 0 && hw > 0 && c_pad >= c, "nchw_to_nhwc: bad arguments");
  int64_t total = (int64_t)batch * hw;
  int blocks = (int)ceil_div64(total, 256);
  LDM_DISPATCH_DTYPE(out_dtype, (launch_pdl(nchw_to_nhwc_kernel<T>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, x, (T*)y, batch, c, hw, c_pad)));
  LDM_CHECK_LAUNCH("nchw_to_nhwc");
  return 0;
}

int ldmb200_nhwc_to_nchw(const float* y, int ld, float* x, int batch, int c, int hw, void* stream) {
  LDM_CHECK_ARG(x && y && batch > 0 && c > 0 && hw > 0 && ld >= c, "nhwc_to_nchw: bad arguments");
  int64_t total = (int64_t)batch * hw;
  nhwc_to_nchw_kernel<<<(int)ceil_div64(total, 256), 256, 0, (cudaStream_t)stream>>>(y, ld, x, batch, c, hw);
  LDM_CHECK_LAUNCH("nhwc_to_nchw");
  return 0;
}

int ldmb200_space_to_depth2(const float* x, void* y, int out_dtype, int batch, int h, int w, int c, void* stream) {
  LDM_CHECK_ARG(x && y && batch > 0 && h > 0 && w > 0 && h % 2 == 0 && w % 2 == 0 && c % 4 == 0,
                "space_to_depth2: bad arguments (h, w even; c %% 4 == 0)");
  int64_t total = (int64_t)batch * h * w * (c / 4);
  int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), (int64_t)sm_count() * 16);
  LDM_DISPATCH_DTYPE(out_dtype, (launch_pdl(space_to_depth2_kernel<T>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, x, (T*)y, batch, h, w, c)));
  LDM_CHECK_LAUNCH("space_to_depth2");
  return 0;
}

int ldmb200_upsample2x(const float* x, void* y, int out_dtype, int batch, int h, int w, int c, void* stream) {
  LDM_CHECK_ARG(x && y && batch > 0 && h > 0 && w > 0 && c % 4 == 0, "upsample2x: bad arguments (c %% 4 == 0)");
  int64_t total = (int64_t)batch * h * w * 4 * (c / 4);
  int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), (int64_t)sm_count() * 16);
  LDM_DISPATCH_DTYPE(out_dtype, (upsample2x_kernel<T><<<blocks, 256, 0, (cudaStream_t)stream>>>(x, (T*)y, batch, h, w, c)));
  LDM_CHECK_LAUNCH("upsample2x");
  return 0;
}

int ldmb200_cast_rows(const float* x, void* y, int out_dtype, int64_t n, void* stream) {
  LDM_CHECK_ARG(x && y && n > 0 && n % 4 == 0, "cast_rows: bad arguments (n %% 4 == 0)");
  int blocks = (int)std::min<int64_t>(ceil_div64(n / 4, 256), (int64_t)sm_count() * 16);
  LDM_DISPATCH_DTYPE(out_dtype, (launch_pdl(cast_rows_kernel<T>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, x, (T*)y, n / 4)));
  LDM_CHECK_LAUNCH("cast_rows");
  return 0;
}

int ldmb200_count_f16_saturated(const void* x, int64_t n, void* counter_u64, void* stream) {
  LDM_CHECK_ARG(x && counter_u64 && n > 0, "count_f16_saturated: bad arguments");
  int blocks = (int)std::min<int64_t>(ceil_div64(n, 256 * 8), (int64_t)sm_count() * 8);
  count_f16_saturated_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)x, n, (unsigned long long*)counter_u64);
  LDM_CHECK_LAUNCH("count_f16_saturated");
  return 0;
}

int64_t ldmb200_groupnorm_ws_floats(int batch, int hw, int c) {
  (void)hw;
  (void)c;
  return (int64_t)batch * GN_MAX_SPLITS * 64 * 2;
}

int ldmb200_groupnorm(const float* x1, int c1, const float* x2, int c2, const float* gamma, const float* beta, void* y,
                      void* raw_copy, int out_dtype, int batch, int hw, int groups, float eps, int silu,
                      float* stats_ws, void* stream) {
  const int C = c1 + (x2 ? c2 : 0);
  if (!x2) c2 = 0;
  LDM_CHECK_ARG(x1 && gamma && beta && y && stats_ws && batch > 0 && hw > 0, "groupnorm: null pointer / empty");
  LDM_CHECK_ARG(groups > 0 && groups <= 64 && C % groups == 0, "groupnorm: C=%d not divisible by groups=%d", C, groups);
  LDM_CHECK_ARG(c1 % 4 == 0 && c2 % 4 == 0 && C / 4 <= 1024, "groupnorm: channel counts must be multiples of 4");
  const int R = gn_rows_per_block(C);
  const int S = gn_splits(hw, C);
  const int threads = (C / 4) * R;
  size_t smem = (size_t)2 * R * C * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  gn_stats_kernel<<<dim3(S, batch), threads, smem, st>>>(x1, c1, x2, c2, hw, groups, R, stats_ws);
  LDM_CHECK_LAUNCH("groupnorm stats");
  LDM_DISPATCH_DTYPE(out_dtype, (gn_apply_launch<T>(x1, c1, x2, c2, gamma, beta, y, raw_copy, batch, hw, groups, R, S, eps,
                                                    silu, stats_ws, st)));
  LDM_CHECK_LAUNCH("groupnorm apply");
  return 0;
}

int ldmb200_groupnorm_fused(const float* x1, int c1, const float* colstats1, const float* x2, int c2,
                            const float* colstats2, const float* gamma, const float* beta, void* y, void* raw_copy,
                            int out_dtype, int batch, int hw, int groups, float eps, int silu, float* stats_ws,
                            void* stream) {
  const int C = c1 + (x2 ? c2 : 0);
  if (!x2) c2 = 0;
  LDM_CHECK_ARG(x1 && colstats1 && (!x2 || colstats2) && gamma && beta && y && stats_ws && batch > 0 && hw > 0,
                "groupnorm_fused: null pointer / empty");
  LDM_CHECK_ARG(hw % 32 == 0, "groupnorm_fused: hw=%d must be a multiple of 32 (column statistics are per 32-row block)", hw);
  LDM_CHECK_ARG(groups > 0 && groups <= 64 && C % groups == 0, "groupnorm_fused: C=%d not divisible by groups=%d", C, groups);
  LDM_CHECK_ARG(c1 % 4 == 0 && c2 % 4 == 0 && C / 4 <= 1024, "groupnorm_fused: channel counts must be multiples of 4");
  cudaStream_t st = (cudaStream_t)stream;
  launch_pdl(gn_finalize_kernel, dim3(groups, batch), dim3(128), 0, st, colstats1, c1, colstats2, c2, hw, groups, eps, stats_ws);
  LDM_CHECK_LAUNCH("groupnorm finalize");
  const int R = gn_rows_per_block(C);
  LDM_DISPATCH_DTYPE(out_dtype, (gn_apply_launch<T>(x1, c1, x2, c2, gamma, beta, y, raw_copy, batch, hw, groups, R, 0, eps,
                                                    silu, stats_ws, st)));
  LDM_CHECK_LAUNCH("groupnorm apply");
  return 0;
}

int ldmb200_layernorm(const float* x, const float* gamma, const float* beta, void* y, int out_dtype, int64_t rows,
                      int c, float eps, void* stream) {
  LDM_CHECK_ARG(x && gamma && beta && y && rows > 0, "layernorm: null pointer / empty");
  LDM_CHECK_ARG(c % 4 == 0 && c <= 2048, "layernorm: C=%d must be a multiple of 4 and <= 2048", c);
  const int warps = 8;
  int blocks = (int)ceil_div64(rows, warps);
  cudaStream_t st = (cudaStream_t)stream;
  if (c == 160 || c == 320 || c == 640) {
#define LN20_LAUNCH(T, LPR) \
  launch_pdl(layernorm_c20_kernel<T, LPR>, dim3((int)ceil_div64(rows, warps * (32 / LPR))), dim3(warps * 32), 0, st, x, gamma, beta, (T*)y, rows, eps)
    LDM_DISPATCH_DTYPE(out_dtype, {
      if (c == 160) LN20_LAUNCH(T, 8);
      else if (c == 320) LN20_LAUNCH(T, 16);
      else LN20_LAUNCH(T, 32);
    });
#undef LN20_LAUNCH
    LDM_CHECK_LAUNCH("layernorm");
    return 0;
  }
#define LN_LAUNCH(T, MAXV) launch_pdl(layernorm_kernel<T, MAXV>, dim3(blocks), dim3(warps * 32), 0, st, x, gamma, beta, (T*)y, rows, c, eps)
  const int need = ceil_div(c / 4, 32);
  LDM_DISPATCH_DTYPE(out_dtype, {
    if (need <= 2) LN_LAUNCH(T, 2);
    else if (need <= 5) LN_LAUNCH(T, 5);
    else LN_LAUNCH(T, 16);
  });
#undef LN_LAUNCH
  LDM_CHECK_LAUNCH("layernorm");
  return 0;
}

int ldmb200_pack_weights(const ldmb200_pack_job* jobs, int n_jobs, void* stream) {
  LDM_CHECK_ARG(jobs && n_jobs > 0 && n_jobs <= 65535, "pack_weights: bad job table");
  pack_weights_kernel<<<dim3(48, n_jobs), 256, 0, (cudaStream_t)stream>>>(jobs);
  LDM_CHECK_LAUNCH("pack_weights");
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Allocation-free GroupNorm (ldmb200_groupnorm_static): ldmb200_groupnorm with the statistics workspace in a file-scope
// device array instead of a caller buffer.  The array is part of the module image: every GPU that loads the library
// holds its 2 MiB until exit.  Every call on a device shares it, so calls that overlap on that GPU (two streams, two host
// threads) race on the partial sums; the statistics kernel overwrites every entry the apply kernel reads, so nothing has
// to start at zero and the previous call's contents are harmless.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_gn_static_ws[(size_t)LDMB200_GN_STATIC_MAX_BATCH * GN_MAX_SPLITS * 64 * 2];

// In host code `g_gn_static_ws` is a host-side stand-in, not a device address: resolved once per device, that device current.
static float* gn_static_ws_ptr() {
  constexpr int MAX_DEV = 64;
  static float* ptrs[MAX_DEV] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
    set_error("groupnorm_static: cudaGetDevice failed or device index %d out of range", dev);
    return nullptr;
  }
  if (!ptrs[dev]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_static_ws);
    if (e != cudaSuccess || !p) {
      set_error("groupnorm_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    ptrs[dev] = p;
  }
  return ptrs[dev];
}

extern "C" int ldmb200_groupnorm_static(const float* x1, int c1, const float* x2, int c2, const float* gamma,
                                        const float* beta, void* y, void* raw_copy, int out_dtype, int batch, int hw,
                                        int groups, float eps, int silu, void* stream) {
  const int C = c1 + (x2 ? c2 : 0);
  if (!x2) c2 = 0;
  LDM_CHECK_ARG(x1 && gamma && beta && y && batch > 0 && hw > 0, "groupnorm_static: null pointer / empty");
  LDM_CHECK_ARG(batch <= LDMB200_GN_STATIC_MAX_BATCH,
                "groupnorm_static: batch=%d exceeds the built-in workspace (LDMB200_GN_STATIC_MAX_BATCH = %d); "
                "use ldmb200_groupnorm with a caller workspace", batch, LDMB200_GN_STATIC_MAX_BATCH);
  LDM_CHECK_ARG(groups > 0 && groups <= 64 && C % groups == 0, "groupnorm_static: C=%d not divisible by groups=%d", C, groups);
  LDM_CHECK_ARG(c1 % 4 == 0 && c2 % 4 == 0 && C / 4 <= 1024, "groupnorm_static: channel counts must be multiples of 4");
  float* ws = gn_static_ws_ptr();
  if (!ws) return LDMB200_ERR_CUDA;
  const int R = gn_rows_per_block(C);
  const int S = gn_splits(hw, C);
  const int threads = (C / 4) * R;
  const size_t smem = (size_t)2 * R * C * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  gn_stats_kernel<<<dim3(S, batch), threads, smem, st>>>(x1, c1, x2, c2, hw, groups, R, ws);
  LDM_CHECK_LAUNCH("groupnorm_static stats");
  LDM_DISPATCH_DTYPE(out_dtype, (gn_apply_launch<T>(x1, c1, x2, c2, gamma, beta, y, raw_copy, batch, hw, groups, R, S, eps,
                                                    silu, ws, st)));
  LDM_CHECK_LAUNCH("groupnorm_static apply");
  return 0;
}
