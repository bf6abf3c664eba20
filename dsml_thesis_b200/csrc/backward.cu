// HBM-bound kernels of the training step: q_sample + MSE loss, gradient cast / column sums,
// GroupNorm(+SiLU) / LayerNorm / GEGLU backward, resampling backward, small dense layers backward, fused AdamW + EMA.
// All gradients of the residual stream are fp32; GEMM operands (dy of the next weight / data gradient) are bf16.
#include <algorithm>

#include "common.cuh"

namespace ldm {

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// four consecutive values of a [.., ld] matrix of dtype dt (fp32 | bf16) at element offset off (multiple of 4)
__device__ __forceinline__ float4 load4(const void* p, int dt, int64_t off) {
  if (dt == LDMB200_F32) return *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + off);
  const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(p) + off);
  return make_float4(bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y));
}
__device__ __forceinline__ void store4_bf16(void* p, int64_t off, float4 v) {
  uint2 u;
  u.x = pack_bf16(v.x, v.y);
  u.y = pack_bf16(v.z, v.w);
  *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(p) + off) = u;
}
__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + __expf(-x)); }
// d silu(z) / dz
__device__ __forceinline__ float dsilu_f(float z) {
  const float s = sigmoid_f(z);
  return s * (1.0f + z * (1.0f - s));
}

// ---------------------------------------------------------------------------------------------
// q_sample / loss
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) q_sample_kernel(const float* __restrict__ x0, const float* __restrict__ nz,
                                                       const int64_t* __restrict__ t, const float* __restrict__ sa,
                                                       const float* __restrict__ s1, float* __restrict__ out,
                                                       int64_t per_sample) {
  const int b = blockIdx.y;
  const float a = sa[t[b]], c = s1[t[b]];
  const int64_t base = (int64_t)b * per_sample;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_sample; i += (int64_t)gridDim.x * blockDim.x)
    out[base + i] = __fadd_rn(__fmul_rn(a, x0[base + i]), __fmul_rn(c, nz[base + i]));
}

__global__ void __launch_bounds__(256) mse_loss_grad_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                            float* __restrict__ loss, float* __restrict__ dpred,
                                                            int batch, int64_t per_sample) {
  __shared__ float red[8];
  const int b = blockIdx.x;
  const int64_t base = (int64_t)b * per_sample;
  const float gs = 2.0f / ((float)batch * (float)per_sample);
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < per_sample; i += blockDim.x) {
    const float d = pred[base + i] - tgt[base + i];
    acc = fmaf(d, d, acc);
    dpred[base + i] = d * gs;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += red[i];
    loss[b] = s / (float)per_sample;
  }
}

// ---------------------------------------------------------------------------------------------
// gradient cast + column sums.  grid (chunks, B): a block covers `rows_per_block` rows of one sample; thread
// (rl, cq) strides over rows and column quads; partial column sums go out through atomics.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) grad_cast_kernel(const void* __restrict__ g, int in_dt, void* __restrict__ g16,
                                                        int n, int rows_per_sample, int rows_per_block,
                                                        float* __restrict__ ps, int ps_ld, float* __restrict__ tot,
                                                        float* __restrict__ tot2, int n_valid) {
  extern __shared__ float colsum[];    // [n] block-level column sums: ONE global atomic per column per block
  const int Q = n >> 2;
  const int qt = min(Q, 256);          // column-quad lanes
  const int R = 256 / qt;              // row lanes
  const int cq0 = threadIdx.x % qt, rl = threadIdx.x / qt;
  const int b = blockIdx.y;
  const int r0 = blockIdx.x * rows_per_block, r1 = min(rows_per_sample, r0 + rows_per_block);
  const bool sums = ps || tot || tot2;
  if (sums) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) colsum[i] = 0.f;
    __syncthreads();
  }
  if (rl < R) {
    for (int cq = cq0; cq < Q; cq += qt) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      int r = r0 + rl;
      for (; r + 3 * R < r1; r += 4 * R) {   // four rows in flight: the pass is a pure stream
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = load4(g, in_dt, ((int64_t)b * rows_per_sample + r + u * R) * n + cq * 4);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (g16) store4_bf16(g16, ((int64_t)b * rows_per_sample + r + u * R) * n + cq * 4, v[u]);
          acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w;
        }
      }
      for (; r < r1; r += R) {
        const int64_t off = ((int64_t)b * rows_per_sample + r) * n + cq * 4;
        const float4 v = load4(g, in_dt, off);
        if (g16) store4_bf16(g16, off, v);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      if (sums) {
        atomicAdd(&colsum[cq * 4 + 0], acc.x); atomicAdd(&colsum[cq * 4 + 1], acc.y);
        atomicAdd(&colsum[cq * 4 + 2], acc.z); atomicAdd(&colsum[cq * 4 + 3], acc.w);
      }
    }
  }
  if (!sums) return;
  __syncthreads();
  for (int col = threadIdx.x; col < n_valid; col += blockDim.x) {
    const float a = colsum[col];
    if (ps) atomicAdd(ps + (size_t)b * ps_ld + col, a);
    if (tot) atomicAdd(tot + col, a);
    if (tot2) atomicAdd(tot2 + col, a);
  }
}

// ---------------------------------------------------------------------------------------------
// GroupNorm (+SiLU) backward.  x_hat = (x - mean) rstd, z = x_hat gamma + beta, y = silu(z) | z.
//   dz = dy silu'(z);  dgamma_c = sum dz x_hat;  dbeta_c = sum dz;  dxh = dz gamma
//   dx = rstd (dxh - mean_g(dxh) - x_hat mean_g(dxh x_hat)) + addend
// pass 1: grid (S, B) — per channel partial sums -> atomics into dgamma/dbeta, per group partials -> ws2[b][s][g][2]
// pass 2: grid (T, B) — folds the partials, streams the slab again and writes dx1 / dx2
// ---------------------------------------------------------------------------------------------
constexpr int GNB_MAX_SPLITS = 64;

__device__ __forceinline__ void gn_group_stats(const float* __restrict__ ws, int b, int g, int groups, int S, int cpg,
                                               int hw, float eps, float& mean, float& rstd) {
  if (S == 0) {
    mean = ws[((size_t)b * groups + g) * 2];
    rstd = ws[((size_t)b * groups + g) * 2 + 1];
    return;
  }
  double a = 0.0, c = 0.0;
  for (int s = 0; s < S; ++s) {
    const float* p = ws + (((size_t)b * S + s) * groups + g) * 2;
    a += (double)p[0];
    c += (double)p[1];
  }
  const double n = (double)cpg * (double)hw;
  const double m = a / n;
  double var = c / n - m * m;
  if (var < 0.0) var = 0.0;
  mean = (float)m;
  rstd = (float)(1.0 / sqrt(var + (double)eps));
}

__device__ __forceinline__ float4 gnb_load_x(const float* __restrict__ x1, int c1, const float* __restrict__ x2, int c2,
                                             int64_t row, int c) {
  if (c < c1) return *reinterpret_cast<const float4*>(x1 + row * c1 + c);
  return *reinterpret_cast<const float4*>(x2 + row * c2 + (c - c1));
}

__global__ void gn_bwd_reduce_kernel(const float* __restrict__ x1, int c1, const float* __restrict__ x2, int c2,
                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                     const uint16_t* __restrict__ dy, float* __restrict__ dgamma,
                                     float* __restrict__ dbeta, int hw, int groups, int R, int S_fwd, float eps, int silu,
                                     const float* __restrict__ ws, float* __restrict__ ws2) {
  extern __shared__ float sm[];  // [2][C] channel sums, then [2][R][C] per-row-lane partials (fixed-order fold)
  __shared__ float s_mean[64], s_rstd[64];
  const int C = c1 + c2, q = C >> 2, cpg = C / groups;
  const int cq = threadIdx.x % q, rl = threadIdx.x / q;
  const int S = gridDim.x, s = blockIdx.x, b = blockIdx.y;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) gn_group_stats(ws, b, g, groups, S_fwd, cpg, hw, eps, s_mean[g], s_rstd[g]);
  __syncthreads();
  float mu[4], rs[4], ga[4], be[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = cq * 4 + j, g = c / cpg;
    mu[j] = s_mean[g]; rs[j] = s_rstd[g]; ga[j] = gamma[c]; be[j] = beta[c];
  }
  float* part = sm + 2 * C;  // [2][R][C]
  const int rows_per = (hw + S - 1) / S;
  const int r0 = s * rows_per, r1 = min(hw, r0 + rows_per);
  float s1[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0};
  auto accum = [&](const float4& xv, const float4& dv) {
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w}, ds[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = (xs[j] - mu[j]) * rs[j];
      float dz = ds[j];
      if (silu) dz *= dsilu_f(fmaf(xh, ga[j], be[j]));
      s1[j] += dz;
      s2[j] = fmaf(dz, xh, s2[j]);
    }
  };
  int r = r0 + rl;
  for (; r + 3 * R < r1; r += 4 * R) {   // four rows in flight
    float4 xv[4], dv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t row = (int64_t)b * hw + r + u * R;
      xv[u] = gnb_load_x(x1, c1, x2, c2, row, cq * 4);
      dv[u] = load4(dy, LDMB200_BF16, row * C + cq * 4);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) accum(xv[u], dv[u]);
  }
  for (; r < r1; r += R) {
    const int64_t row = (int64_t)b * hw + r;
    accum(gnb_load_x(x1, c1, x2, c2, row, cq * 4), load4(dy, LDMB200_BF16, row * C + cq * 4));
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    part[rl * C + cq * 4 + j] = s1[j];
    part[(R + rl) * C + cq * 4 + j] = s2[j];
  }
  __syncthreads();
  // fixed-order fold over the row lanes: the group sums that enter the data gradient are run-to-run deterministic
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f, b2 = 0.f;
    for (int r = 0; r < R; ++r) {
      a += part[r * C + c];
      b2 += part[(R + r) * C + c];
    }
    sm[c] = a;
    sm[C + c] = b2;
    atomicAdd(dbeta + c, a);
    atomicAdd(dgamma + c, b2);
  }
  __syncthreads();
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    float a = 0.f, c = 0.f;
    for (int j = 0; j < cpg; ++j) {
      a = fmaf(gamma[g * cpg + j], sm[g * cpg + j], a);
      c = fmaf(gamma[g * cpg + j], sm[C + g * cpg + j], c);
    }
    float* o = ws2 + (((size_t)b * S + s) * groups + g) * 2;
    o[0] = a;
    o[1] = c;
  }
}

__global__ void gn_bwd_apply_kernel(const float* __restrict__ x1, int c1, const float* __restrict__ x2, int c2,
                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                    const uint16_t* __restrict__ dy, const void* __restrict__ addend, int add_dt,
                                    float* __restrict__ dx1, int acc1, float* __restrict__ dx2, int acc2, int hw,
                                    int groups, int R, int S_fwd, int S2, float eps, int silu,
                                    const float* __restrict__ ws, const float* __restrict__ ws2) {
  __shared__ float s_mean[64], s_rstd[64], s_m1[64], s_m2[64];
  const int C = c1 + c2, q = C >> 2, cpg = C / groups;
  const int b = blockIdx.y;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    gn_group_stats(ws, b, g, groups, S_fwd, cpg, hw, eps, s_mean[g], s_rstd[g]);
    double a = 0.0, c = 0.0;
    for (int s = 0; s < S2; ++s) {
      const float* p = ws2 + (((size_t)b * S2 + s) * groups + g) * 2;
      a += (double)p[0];
      c += (double)p[1];
    }
    const double n = (double)cpg * (double)hw;
    s_m1[g] = (float)(a / n);
    s_m2[g] = (float)(c / n);
  }
  __syncthreads();
  const int cq = threadIdx.x % q, rl = threadIdx.x / q;
  float mu[4], rs[4], ga[4], be[4], m1[4], m2[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = cq * 4 + j, g = c / cpg;
    mu[j] = s_mean[g]; rs[j] = s_rstd[g]; ga[j] = gamma[c]; be[j] = beta[c]; m1[j] = s_m1[g]; m2[j] = s_m2[g];
  }
  const int T_ = gridDim.x;
  const int rows_per = (hw + T_ - 1) / T_;
  const int r0 = blockIdx.x * rows_per, r1 = min(hw, r0 + rows_per);
  const int c0 = cq * 4;
  auto finish = [&](int64_t row, const float4& xv, const float4& dv, const float4& av) {
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w}, ds[4] = {dv.x, dv.y, dv.z, dv.w}, as[4] = {av.x, av.y, av.z, av.w};
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = (xs[j] - mu[j]) * rs[j];
      float dz = ds[j];
      if (silu) dz *= dsilu_f(fmaf(xh, ga[j], be[j]));
      o[j] = rs[j] * (dz * ga[j] - m1[j] - xh * m2[j]) + as[j];
    }
    float* dst;
    int acc;
    if (c0 < c1) {
      dst = dx1 + row * c1 + c0;
      acc = acc1;
    } else {
      dst = dx2 + row * c2 + (c0 - c1);
      acc = acc2;
    }
    if (acc) {
      const float4 old = *reinterpret_cast<const float4*>(dst);
      o[0] += old.x; o[1] += old.y; o[2] += old.z; o[3] += old.w;
    }
    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
  };
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  int r = r0 + rl;
  for (; r + 3 * R < r1; r += 4 * R) {   // four rows in flight: the pass is a pure stream
    float4 xv[4], dv[4], av[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t row = (int64_t)b * hw + r + u * R;
      xv[u] = gnb_load_x(x1, c1, x2, c2, row, c0);
      dv[u] = load4(dy, LDMB200_BF16, row * C + c0);
      av[u] = addend ? load4(addend, add_dt, row * C + c0) : zero4;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) finish((int64_t)b * hw + r + u * R, xv[u], dv[u], av[u]);
  }
  for (; r < r1; r += R) {
    const int64_t row = (int64_t)b * hw + r;
    finish(row, gnb_load_x(x1, c1, x2, c2, row, c0), load4(dy, LDMB200_BF16, row * C + c0),
           addend ? load4(addend, add_dt, row * C + c0) : zero4);
  }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm backward: one warp per row (persistent over rows), dgamma/dbeta partials in registers -> smem -> atomics
// ---------------------------------------------------------------------------------------------
template <int MAXV>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const uint16_t* __restrict__ dy,
                                                            const float* __restrict__ addend, float* __restrict__ dx,
                                                            float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                            int64_t rows, int C, float eps) {
  extern __shared__ float sm[];  // [2][C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int q = C >> 2;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  float4 g[MAXV], dg[MAXV], db[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int idx = lane + i * 32;
    g[i] = idx < q ? reinterpret_cast<const float4*>(gamma)[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
    dg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    db[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int64_t row = (int64_t)blockIdx.x * nw + warp; row < rows; row += (int64_t)gridDim.x * nw) {
    const float4* xr = reinterpret_cast<const float4*>(x + row * C);
    float4 v[MAXV], d[MAXV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + i * 32;
      if (idx < q) {
        v[i] = xr[idx];
        d[i] = load4(dy, LDMB200_BF16, row * C + idx * 4);
        s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
      }
    }
    const float mean = warp_sum(s) / (float)C;
    float var = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + i * 32;
      if (idx < q) {
        v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
        var += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
      }
    }
    const float rstd = rsqrtf(warp_sum(var) / (float)C + eps);
    float a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + i * 32;
      if (idx < q) {
        v[i].x *= rstd; v[i].y *= rstd; v[i].z *= rstd; v[i].w *= rstd;   // x_hat
        db[i].x += d[i].x; db[i].y += d[i].y; db[i].z += d[i].z; db[i].w += d[i].w;
        dg[i].x = fmaf(d[i].x, v[i].x, dg[i].x); dg[i].y = fmaf(d[i].y, v[i].y, dg[i].y);
        dg[i].z = fmaf(d[i].z, v[i].z, dg[i].z); dg[i].w = fmaf(d[i].w, v[i].w, dg[i].w);
        d[i].x *= g[i].x; d[i].y *= g[i].y; d[i].z *= g[i].z; d[i].w *= g[i].w;   // dxh
        a1 += (d[i].x + d[i].y) + (d[i].z + d[i].w);
        a2 += (d[i].x * v[i].x + d[i].y * v[i].y) + (d[i].z * v[i].z + d[i].w * v[i].w);
      }
    }
    const float m1 = warp_sum(a1) / (float)C, m2 = warp_sum(a2) / (float)C;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + i * 32;
      if (idx < q) {
        float4 o;
        o.x = rstd * (d[i].x - m1 - v[i].x * m2); o.y = rstd * (d[i].y - m1 - v[i].y * m2);
        o.z = rstd * (d[i].z - m1 - v[i].z * m2); o.w = rstd * (d[i].w - m1 - v[i].w * m2);
        if (addend) {
          const float4 a = reinterpret_cast<const float4*>(addend + row * C)[idx];
          o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
        }
        reinterpret_cast<float4*>(dx + row * C)[idx] = o;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int idx = lane + i * 32;
    if (idx < q) {
      atomicAdd(&sm[idx * 4 + 0], dg[i].x); atomicAdd(&sm[idx * 4 + 1], dg[i].y);
      atomicAdd(&sm[idx * 4 + 2], dg[i].z); atomicAdd(&sm[idx * 4 + 3], dg[i].w);
      atomicAdd(&sm[C + idx * 4 + 0], db[i].x); atomicAdd(&sm[C + idx * 4 + 1], db[i].y);
      atomicAdd(&sm[C + idx * 4 + 2], db[i].z); atomicAdd(&sm[C + idx * 4 + 3], db[i].w);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    atomicAdd(dgamma + c, sm[c]);
    atomicAdd(dbeta + c, sm[C + c]);
  }
}

// ---------------------------------------------------------------------------------------------
// GEGLU forward / backward on the stored projection u = [value | gate]
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float gelu_grad_f(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
  return fmaf(x, pdf, cdf);
}

__global__ void __launch_bounds__(256) geglu_fwd_kernel(const uint16_t* __restrict__ u, uint16_t* __restrict__ out,
                                                        int64_t rows, int inner) {
  const int q = inner >> 2;
  const int64_t total = rows * q;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / q;
    const int c = (int)(i % q) * 4;
    const float4 a = load4(u, LDMB200_BF16, r * 2 * inner + c);
    const float4 g = load4(u, LDMB200_BF16, r * 2 * inner + inner + c);
    store4_bf16(out, r * inner + c,
                make_float4(a.x * gelu_erf_f(g.x), a.y * gelu_erf_f(g.y), a.z * gelu_erf_f(g.z), a.w * gelu_erf_f(g.w)));
  }
}

__global__ void __launch_bounds__(256) geglu_bwd_kernel(const uint16_t* __restrict__ u, const uint16_t* __restrict__ dout,
                                                        uint16_t* __restrict__ du, int64_t rows, int inner) {
  const int q = inner >> 2;
  const int64_t total = rows * q;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / q;
    const int c = (int)(i % q) * 4;
    const float4 a = load4(u, LDMB200_BF16, r * 2 * inner + c);
    const float4 g = load4(u, LDMB200_BF16, r * 2 * inner + inner + c);
    const float4 d = load4(dout, LDMB200_BF16, r * inner + c);
    store4_bf16(du, r * 2 * inner + c,
                make_float4(d.x * gelu_erf_f(g.x), d.y * gelu_erf_f(g.y), d.z * gelu_erf_f(g.z), d.w * gelu_erf_f(g.w)));
    store4_bf16(du, r * 2 * inner + inner + c,
                make_float4(d.x * a.x * gelu_grad_f(g.x), d.y * a.y * gelu_grad_f(g.y), d.z * a.z * gelu_grad_f(g.z),
                            d.w * a.w * gelu_grad_f(g.w)));
  }
}

// ---------------------------------------------------------------------------------------------
// resampling backward
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) zero_insert2x_kernel(const uint16_t* __restrict__ dy, uint16_t* __restrict__ z,
                                                            int batch, int h, int w, int C) {
  const int q = C >> 3;  // 8 bf16 = 16 bytes
  const int64_t total = (int64_t)batch * 2 * h * 2 * w * q;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % q);
    int64_t t = i / q;
    const int ow = (int)(t % (2 * w)); t /= 2 * w;
    const int oh = (int)(t % (2 * h));
    const int b = (int)(t / (2 * h));
    uint4 v = make_uint4(0, 0, 0, 0);
    if (!(oh & 1) && !(ow & 1))
      v = reinterpret_cast<const uint4*>(dy)[(((int64_t)b * h + (oh >> 1)) * w + (ow >> 1)) * q + c];
    reinterpret_cast<uint4*>(z)[i] = v;
  }
}

__global__ void __launch_bounds__(256) downsum2x_kernel(const float* __restrict__ du, float* __restrict__ dx, int batch,
                                                        int h, int w, int C) {
  const int q = C >> 2;
  const int64_t total = (int64_t)batch * h * w * q;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % q);
    int64_t t = i / q;
    const int x = (int)(t % w); t /= w;
    const int y = (int)(t % h);
    const int b = (int)(t / h);
    const float4* src = reinterpret_cast<const float4*>(du);
    const int64_t r0 = (((int64_t)b * 2 * h + 2 * y) * 2 * w + 2 * x) * q + c;
    const int64_t r1 = r0 + (int64_t)2 * w * q;
    const float4 a = src[r0], bb = src[r0 + q], cc = src[r1], d = src[r1 + q];
    reinterpret_cast<float4*>(dx)[i] =
        make_float4((a.x + bb.x) + (cc.x + d.x), (a.y + bb.y) + (cc.y + d.y), (a.z + bb.z) + (cc.z + d.z), (a.w + bb.w) + (cc.w + d.w));
  }
}

// ---------------------------------------------------------------------------------------------
// small dense layers backward (rows = batch: per-sample vectors)
// ---------------------------------------------------------------------------------------------
// The same for a whole table of independent layers in ONE launch (grid.y = job).  Thread <-> input column kk of a
// 256-wide k tile; the block walks its share of the output columns n in 64-wide tiles of dout staged in smem, so every
// weight element is read once and feeds 32 row FMAs (din) while act(in) sits in registers (dw).  din is accumulated with
// atomics: several jobs / n-parts feed the same input gradient (all 17 ResBlock emb projections read emb).
constexpr int SLB_ROWS = 32, SLB_NT = 64, SLB_PARTS = 4;

__device__ __forceinline__ void slb_body(const ldmb200_linear_bwd_job& j) {
  __shared__ float sd[SLB_ROWS][SLB_NT + 1];
  const int ktiles = (j.k + 255) / 256;
  const int kt = blockIdx.x % ktiles, part = blockIdx.x / ktiles;
  if (blockIdx.x >= ktiles * SLB_PARTS) return;
  const int kk = kt * 256 + threadIdx.x;
  const bool kok = kk < j.k;
  const int n_per = (j.n + SLB_PARTS - 1) / SLB_PARTS;
  const int n_lo = part * n_per, n_hi = min(j.n, n_lo + n_per);
  for (int r0 = 0; r0 < j.rows; r0 += SLB_ROWS) {
    const int nr = min(SLB_ROWS, j.rows - r0);
    float a[SLB_ROWS], acc[SLB_ROWS], raw[SLB_ROWS];
#pragma unroll
    for (int r = 0; r < SLB_ROWS; ++r) {
      raw[r] = (kok && r < nr) ? j.in[(size_t)(r0 + r) * j.in_ld + kk] : 0.f;
      a[r] = j.silu_in ? silu_f(raw[r]) : raw[r];
      acc[r] = 0.f;
    }
    for (int nt = n_lo; nt < n_hi; nt += SLB_NT) {
      const int nn_max = min(SLB_NT, n_hi - nt);
      __syncthreads();
      for (int i = threadIdx.x; i < SLB_ROWS * SLB_NT; i += 256) {
        const int r = i / SLB_NT, c = i % SLB_NT;
        sd[r][c] = (r < nr && c < nn_max) ? j.dout[(size_t)(r0 + r) * j.dout_ld + nt + c] : 0.f;
      }
      __syncthreads();
      if (j.dbias && kt == 0 && threadIdx.x < nn_max) {
        float bsum = 0.f;
#pragma unroll
        for (int r = 0; r < SLB_ROWS; ++r) bsum += sd[r][threadIdx.x];
        j.dbias[nt + threadIdx.x] += bsum;       // one block owns (k tile 0, this n range): no race; row chunks add up
      }
      if (kok) {
        for (int c = 0; c < nn_max; ++c) {
          const float w = j.din ? j.w[(size_t)(nt + c) * j.k + kk] : 0.f;
          float s_ = 0.f;
#pragma unroll
          for (int r = 0; r < SLB_ROWS; ++r) {
            const float d = sd[r][c];
            s_ = fmaf(d, a[r], s_);
            acc[r] = fmaf(d, w, acc[r]);
          }
          if (j.dw) j.dw[(size_t)(nt + c) * j.k + kk] += s_;
        }
      }
    }
    if (j.din && kok) {
#pragma unroll
      for (int r = 0; r < SLB_ROWS; ++r)
        if (r < nr) atomicAdd(j.din + (size_t)(r0 + r) * j.din_ld + kk, j.silu_in ? acc[r] * dsilu_f(raw[r]) : acc[r]);
    }
  }
}

__global__ void __launch_bounds__(256) slb_batched_kernel(const ldmb200_linear_bwd_job* __restrict__ jobs) {
  const ldmb200_linear_bwd_job j = jobs[blockIdx.y];
  slb_body(j);
}
__global__ void __launch_bounds__(256) slb_single_kernel(const __grid_constant__ ldmb200_linear_bwd_job j) { slb_body(j); }

// ---------------------------------------------------------------------------------------------
// fused AdamW + EMA over a flat parameter arena
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) adamw_ema_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v,
                                                        float* __restrict__ ema, int64_t n, float lr, float b1, float b2,
                                                        float eps, float wd, float bc1, float bc2, float omd) {
  const float step = lr / bc1, isb2 = rsqrtf(bc2);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    float pi = p[i] * (1.0f - lr * wd);
    const float mi = m[i] + (1.0f - b1) * (gi - m[i]);       // lerp, like torch's exp_avg.lerp_(grad, 1 - beta1)
    const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    pi -= step * mi / (sqrtf(vi) * isb2 + eps);
    p[i] = pi;
    if (ema) ema[i] -= omd * (ema[i] - pi);
  }
}

}  // namespace ldm

using namespace ldm;

static inline int grid_for(int64_t work, int per_block = 256, int waves = 16) {
  int64_t b = (work + per_block - 1) / per_block;
  int64_t cap = (int64_t)sm_count() * waves;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

extern "C" {

int ldmb200_q_sample(const float* x0, const float* noise, const int64_t* t, const float* sqrt_ac,
                     const float* sqrt_1mac, float* out, int batch, int64_t per_sample, void* stream) {
  LDM_CHECK_ARG(x0 && noise && t && sqrt_ac && sqrt_1mac && out && batch > 0 && per_sample > 0, "q_sample: bad arguments");
  dim3 grid((unsigned)std::min<int64_t>((per_sample + 255) / 256, 64), batch);
  q_sample_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x0, noise, t, sqrt_ac, sqrt_1mac, out, per_sample);
  LDM_CHECK_LAUNCH("q_sample");
  return 0;
}

int ldmb200_mse_loss_grad(const float* pred, const float* target, float* loss_per_sample, float* dpred, int batch,
                          int64_t per_sample, void* stream) {
  LDM_CHECK_ARG(pred && target && loss_per_sample && dpred && batch > 0 && per_sample > 0, "mse_loss_grad: bad arguments");
  mse_loss_grad_kernel<<<batch, 256, 0, (cudaStream_t)stream>>>(pred, target, loss_per_sample, dpred, batch, per_sample);
  LDM_CHECK_LAUNCH("mse_loss_grad");
  return 0;
}

int ldmb200_grad_cast(const void* g, int in_dtype, void* g16, int64_t rows, int n, int rows_per_sample,
                      float* per_sample_sum, int ps_ld, float* total_sum, float* total_sum2, int n_valid, void* stream) {
  LDM_CHECK_ARG(g && rows > 0 && n > 0 && n % 4 == 0, "grad_cast: bad arguments (n %% 4 == 0)");
  LDM_CHECK_ARG(in_dtype == LDMB200_F32 || in_dtype == LDMB200_BF16, "grad_cast: in_dtype must be fp32 or bf16");
  LDM_CHECK_ARG(rows_per_sample > 0 && rows % rows_per_sample == 0, "grad_cast: rows must be a multiple of rows_per_sample");
  const int batch = (int)(rows / rows_per_sample);
  LDM_CHECK_ARG(batch <= 65535, "grad_cast: batch exceeds grid limits");
  if (n_valid <= 0 || n_valid > n) n_valid = n;
  int chunks = ceil_div(sm_count() * 4, batch);
  int rpb = std::max(16, ceil_div(rows_per_sample, chunks));
  chunks = ceil_div(rows_per_sample, rpb);
  LDM_CHECK_ARG(n <= 12288, "grad_cast: n=%d too wide for the shared-memory column sums", n);
  grad_cast_kernel<<<dim3(chunks, batch), 256, (size_t)n * sizeof(float), (cudaStream_t)stream>>>(g, in_dtype, g16, n, rows_per_sample, rpb,
                                                                           per_sample_sum, ps_ld, total_sum, total_sum2,
                                                                           n_valid);
  LDM_CHECK_LAUNCH("grad_cast");
  return 0;
}

int ldmb200_groupnorm_bwd(const float* x1, int c1, const float* x2, int c2, const float* gamma, const float* beta,
                          const void* dy, const void* addend, int add_dtype, float* dx1, int acc1, float* dx2,
                          int acc2, float* dgamma, float* dbeta, int batch, int hw, int groups, float eps, int silu,
                          const float* stats_ws, int fused, float* ws2, void* stream) {
  const int C = c1 + (x2 ? c2 : 0);
  if (!x2) c2 = 0;
  LDM_CHECK_ARG(x1 && gamma && beta && dy && dx1 && (!x2 || dx2) && dgamma && dbeta && stats_ws && ws2 && batch > 0 && hw > 0,
                "groupnorm_bwd: null pointer / empty");
  LDM_CHECK_ARG(groups > 0 && groups <= 64 && C % groups == 0, "groupnorm_bwd: C=%d not divisible by groups=%d", C, groups);
  LDM_CHECK_ARG(c1 % 4 == 0 && c2 % 4 == 0 && C / 4 <= 1024, "groupnorm_bwd: channel counts must be multiples of 4");
  LDM_CHECK_ARG(!addend || add_dtype == LDMB200_F32 || add_dtype == LDMB200_BF16, "groupnorm_bwd: bad add_dtype");
  // the forward's split count (must mirror ldmb200_groupnorm: elementwise.cu gn_rows_per_block / gn_splits)
  int q = C / 4;
  int R = 256 / q;
  if (R < 1) R = 1;
  int S_fwd = hw / (R * 8);
  if (S_fwd < 1) S_fwd = 1;
  if (S_fwd > GNB_MAX_SPLITS) S_fwd = GNB_MAX_SPLITS;
  if (fused) S_fwd = 0;
  int S2 = hw / (R * 8);
  if (S2 < 1) S2 = 1;
  if (S2 > GNB_MAX_SPLITS) S2 = GNB_MAX_SPLITS;
  const int threads = q * R;
  cudaStream_t st = (cudaStream_t)stream;
  gn_bwd_reduce_kernel<<<dim3(S2, batch), threads, (size_t)(2 + 2 * R) * C * sizeof(float), st>>>(
      x1, c1, x2, c2, gamma, beta, (const uint16_t*)dy, dgamma, dbeta, hw, groups, R, S_fwd, eps, silu, stats_ws, ws2);
  LDM_CHECK_LAUNCH("groupnorm_bwd reduce");
  int T = hw / (R * 4);
  int want = ceil_div(sm_count() * 4, batch);
  if (T > want) T = want;
  if (T < 1) T = 1;
  gn_bwd_apply_kernel<<<dim3(T, batch), threads, 0, st>>>(x1, c1, x2, c2, gamma, beta, (const uint16_t*)dy, addend,
                                                          add_dtype, dx1, acc1, dx2, acc2, hw, groups, R, S_fwd, S2, eps,
                                                          silu, stats_ws, ws2);
  LDM_CHECK_LAUNCH("groupnorm_bwd apply");
  return 0;
}

int ldmb200_layernorm_bwd(const float* x, const float* gamma, const void* dy, const float* addend, float* dx,
                          float* dgamma, float* dbeta, int64_t rows, int c, float eps, void* stream) {
  LDM_CHECK_ARG(x && gamma && dy && dx && dgamma && dbeta && rows > 0, "layernorm_bwd: null pointer / empty");
  LDM_CHECK_ARG(c % 4 == 0 && c <= 2048, "layernorm_bwd: C=%d must be a multiple of 4 and <= 2048", c);
  const int warps = 8;
  int blocks = (int)std::min<int64_t>(ceil_div64(rows, warps), (int64_t)sm_count() * 4);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = 2 * (size_t)c * sizeof(float);
  const int need = ceil_div(c / 4, 32);
#define LNB_LAUNCH(MAXV) \
  layernorm_bwd_kernel<MAXV><<<blocks, warps * 32, smem, st>>>(x, gamma, (const uint16_t*)dy, addend, dx, dgamma, dbeta, rows, c, eps)
  if (need <= 2) LNB_LAUNCH(2);
  else if (need <= 5) LNB_LAUNCH(5);
  else LNB_LAUNCH(16);
#undef LNB_LAUNCH
  LDM_CHECK_LAUNCH("layernorm_bwd");
  return 0;
}

int ldmb200_geglu_fwd(const void* u, void* out, int64_t rows, int inner, void* stream) {
  LDM_CHECK_ARG(u && out && rows > 0 && inner > 0 && inner % 4 == 0, "geglu_fwd: bad arguments");
  geglu_fwd_kernel<<<grid_for(rows * (inner / 4)), 256, 0, (cudaStream_t)stream>>>((const uint16_t*)u, (uint16_t*)out, rows, inner);
  LDM_CHECK_LAUNCH("geglu_fwd");
  return 0;
}

int ldmb200_geglu_bwd(const void* u, const void* dout, void* du, int64_t rows, int inner, void* stream) {
  LDM_CHECK_ARG(u && dout && du && rows > 0 && inner > 0 && inner % 4 == 0, "geglu_bwd: bad arguments");
  geglu_bwd_kernel<<<grid_for(rows * (inner / 4)), 256, 0, (cudaStream_t)stream>>>((const uint16_t*)u, (const uint16_t*)dout,
                                                                                  (uint16_t*)du, rows, inner);
  LDM_CHECK_LAUNCH("geglu_bwd");
  return 0;
}

int ldmb200_zero_insert2x(const void* dy, void* z, int batch, int h, int w, int c, void* stream) {
  LDM_CHECK_ARG(dy && z && batch > 0 && h > 0 && w > 0 && c % 8 == 0, "zero_insert2x: bad arguments (c %% 8 == 0)");
  zero_insert2x_kernel<<<grid_for((int64_t)batch * 4 * h * w * (c / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const uint16_t*)dy, (uint16_t*)z, batch, h, w, c);
  LDM_CHECK_LAUNCH("zero_insert2x");
  return 0;
}

int ldmb200_downsum2x(const float* du, float* dx, int batch, int h, int w, int c, void* stream) {
  LDM_CHECK_ARG(du && dx && batch > 0 && h > 0 && w > 0 && c % 4 == 0, "downsum2x: bad arguments (c %% 4 == 0)");
  downsum2x_kernel<<<grid_for((int64_t)batch * h * w * (c / 4)), 256, 0, (cudaStream_t)stream>>>(du, dx, batch, h, w, c);
  LDM_CHECK_LAUNCH("downsum2x");
  return 0;
}

int ldmb200_small_linear_bwd(const float* in, int in_ld, const float* w, const float* dout, int dout_ld, float* dw,
                             float* dbias, float* din, int din_ld, int acc_din, int rows, int k, int n, int silu_in,
                             void* stream) {
  LDM_CHECK_ARG(in && dout && rows > 0 && k > 0 && n > 0, "small_linear_bwd: bad arguments");
  LDM_CHECK_ARG(!din || w, "small_linear_bwd: din needs w");
  LDM_CHECK_ARG(k <= 4096, "small_linear_bwd: k=%d exceeds the 16 k tiles of the launch grid", k);
  cudaStream_t st = (cudaStream_t)stream;
  if (din && !acc_din) {   // the kernel accumulates (its n parts add up): "write" = zero first
    cudaError_t e = cudaMemset2DAsync(din, (size_t)din_ld * 4, 0, (size_t)k * 4, rows, st);
    if (e != cudaSuccess) {
      set_error("small_linear_bwd: memset failed: %s", cudaGetErrorString(e));
      return LDMB200_ERR_CUDA;
    }
  }
  ldmb200_linear_bwd_job j;
  j.in = in; j.w = w; j.dout = dout; j.dw = dw; j.dbias = dbias; j.din = din;
  j.in_ld = in_ld; j.dout_ld = dout_ld; j.din_ld = din_ld; j.rows = rows; j.k = k; j.n = n; j.silu_in = silu_in; j.pad_ = 0;
  slb_single_kernel<<<dim3(ceil_div(k, 256) * SLB_PARTS, 1), 256, 0, st>>>(j);
  LDM_CHECK_LAUNCH("small_linear_bwd");
  return 0;
}

int ldmb200_small_linear_bwd_batched(const ldmb200_linear_bwd_job* jobs, int n_jobs, void* stream) {
  LDM_CHECK_ARG(jobs && n_jobs > 0 && n_jobs <= 65535, "small_linear_bwd_batched: bad job table");
  // grid.x covers the widest job (k <= 4096: 16 k tiles x 4 n parts); narrower jobs leave their extra blocks idle
  slb_batched_kernel<<<dim3(16 * SLB_PARTS, n_jobs), 256, 0, (cudaStream_t)stream>>>(jobs);
  LDM_CHECK_LAUNCH("small_linear_bwd_batched");
  return 0;
}

int ldmb200_adamw_ema(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* ema, int64_t n,
                      float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1,
                      float bias_corr2, float ema_one_minus_decay, void* stream) {
  LDM_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && n > 0, "adamw_ema: bad arguments");
  adamw_ema_kernel<<<grid_for(n, 256, 32), 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, ema, n, lr, beta1,
                                                                          beta2, eps, weight_decay, bias_corr1, bias_corr2,
                                                                          ema_one_minus_decay);
  LDM_CHECK_LAUNCH("adamw_ema");
  return 0;
}

}  // extern "C"
