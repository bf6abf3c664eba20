// fp32 SIMT implicit-GEMM for convolutions / linears (precision mode "fp32": fp32 FMA accumulation, weights
// read in their native nn.Parameter layout, no packing).  Covers every A-operand index map of the path:
// 1x1 / 3x3 (pad 1), stride 1 / 2, nearest-2x upsample folded into the gather (K10), two-pointer virtual
// channel concat (K11), and the fused epilogue (bias, per-sample row vector (K8), residual, GEGLU, NCHW store).
#include "common.cuh"

namespace ldm {

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16, SG_THREADS = 256;

struct SimtGeom {
  int c1, c2, ctot, taps, ks, stride, upsample, pad_asym;
  int h_in, w_in, h_eff, w_eff, h_out, w_out;
  int64_t M;
  int K;
};

template <typename TA>
__device__ __forceinline__ float lda(const TA* p, size_t i) {
  return (float)p[i];
}
template <>
__device__ __forceinline__ float lda<__nv_bfloat16>(const __nv_bfloat16* p, size_t i) {
  return __bfloat162float(p[i]);
}

template <typename TA, typename TO, bool GEGLU>
__global__ void __launch_bounds__(SG_THREADS) igemm_simt_kernel(const TA* __restrict__ a, const TA* __restrict__ a2,
                                                                const float* __restrict__ w, SimtGeom g, int N,
                                                                const float* __restrict__ bias,
                                                                const float* __restrict__ rowvec, int rowvec_ld,
                                                                const float* __restrict__ residual, int residual_ld,
                                                                TO* __restrict__ out, int out_ld, int out_nchw,
                                                                int n_valid) {
  __shared__ float As[SG_BK][SG_BM + 4];
  __shared__ float Bs[GEGLU ? 2 : 1][SG_BK][SG_BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * SG_BM;
  const int n0 = blockIdx.y * SG_BN;
  const int n_cols = GEGLU ? N / 2 : N;  // output columns of the GEMM as seen by the tile grid

  // loader mapping: element (row = ty + 16*i, k = tx)
  int lb[4], loh[4], low[4];
  bool lvalid[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + ty + 16 * i;
    lvalid[i] = m < g.M;
    int64_t mm = lvalid[i] ? m : 0;
    int hw = g.h_out * g.w_out;
    lb[i] = (int)(mm / hw);
    int p = (int)(mm % hw);
    loh[i] = p / g.w_out;
    low[i] = p % g.w_out;
  }
  const int pad = (g.ks == 3 && !g.pad_asym) ? 1 : 0;

  float acc[4][4], accg[GEGLU ? 4 : 1][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  if (GEGLU) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) accg[GEGLU ? i : 0][j] = 0.f;
  }

  for (int k0 = 0; k0 < g.K; k0 += SG_BK) {
    const int k = k0 + tx;
    const bool kvalid = k < g.K;
    const int tap = kvalid ? k / g.ctot : 0;
    const int c = kvalid ? k % g.ctot : 0;
    const int kh = tap / g.ks, kw = tap % g.ks;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float v = 0.f;
      if (kvalid && lvalid[i]) {
        int ih = loh[i] * g.stride + kh - pad, iw = low[i] * g.stride + kw - pad;
        if (ih >= 0 && ih < g.h_eff && iw >= 0 && iw < g.w_eff) {
          if (g.upsample) {
            ih >>= 1;
            iw >>= 1;
          }
          size_t pix = ((size_t)lb[i] * g.h_in + ih) * g.w_in + iw;
          v = c < g.c1 ? lda<TA>(a, pix * g.c1 + c) : lda<TA>(a2, pix * g.c2 + (c - g.c1));
        }
      }
      As[tx][ty + 16 * i] = v;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int n = n0 + ty + 16 * i;
      float v = 0.f, vg = 0.f;
      if (kvalid && n < n_cols) {
        v = w[((size_t)n * g.ctot + c) * g.taps + tap];
        if (GEGLU) vg = w[((size_t)(n + n_cols) * g.ctot + c) * g.taps + tap];
      }
      Bs[0][tx][ty + 16 * i] = v;
      if (GEGLU) Bs[GEGLU ? 1 : 0][tx][ty + 16 * i] = vg;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SG_BK; ++kk) {
      float av[4], bv[4], bg[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        bv[j] = Bs[0][kk][tx * 4 + j];
        if (GEGLU) bg[j] = Bs[GEGLU ? 1 : 0][kk][tx * 4 + j];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
          if (GEGLU) accg[GEGLU ? i : 0][j] = fmaf(av[i], bg[j], accg[GEGLU ? i : 0][j]);
        }
    }
    __syncthreads();
  }

  const int hw_out = g.h_out * g.w_out;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + ty * 4 + i;
    if (m >= g.M) continue;
    int b = (int)(m / hw_out);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= n_cols) continue;
      float v = acc[i][j];
      if (bias) v += bias[n];
      if (GEGLU) {
        float gt = accg[GEGLU ? i : 0][j];
        if (bias) gt += bias[n + n_cols];
        v = v * gelu_erf_f(gt);
      }
      if (rowvec) v += rowvec[(size_t)b * rowvec_ld + n];
      if (residual) v += residual[(size_t)m * residual_ld + n];
      if (n < n_valid) {
        if (out_nchw)
          out[((size_t)b * n_valid + n) * hw_out + (m % hw_out)] = (TO)v;
        else
          out[(size_t)m * out_ld + n] = (TO)v;
      }
    }
  }
}

}  // namespace ldm

using namespace ldm;

extern "C" int ldmb200_igemm_simt(const ldmb200_gemm_desc* d, void* stream) {
  LDM_CHECK_ARG(d && d->a && d->w && d->out, "igemm_simt: null pointer");
  LDM_CHECK_ARG(d->ksize == 1 || d->ksize == 3, "igemm_simt: ksize must be 1 or 3");
  LDM_CHECK_ARG(d->stride == 1 || d->stride == 2, "igemm_simt: stride must be 1 or 2");
  LDM_CHECK_ARG(d->pad_mode == 0 || (d->pad_mode == 1 && d->ksize == 3 && d->stride == 2 && !d->upsample),
                "igemm_simt: pad_mode 1 is the (0,1,0,1)-padded 3x3 stride-2 convolution");
  LDM_CHECK_ARG(d->batch > 0 && d->h_in > 0 && d->w_in > 0 && d->c1 > 0 && d->n > 0, "igemm_simt: empty problem");
  LDM_CHECK_ARG(!(d->geglu && (d->n % 2)), "igemm_simt: geglu needs even n");
  SimtGeom g;
  g.c1 = d->c1;
  g.c2 = d->a2 ? d->c2 : 0;
  g.ctot = g.c1 + g.c2;
  g.ks = d->ksize;
  g.taps = g.ks * g.ks;
  g.stride = d->stride;
  g.upsample = d->upsample;
  g.pad_asym = d->pad_mode == 1;
  g.h_in = d->h_in;
  g.w_in = d->w_in;
  g.h_eff = d->upsample ? 2 * d->h_in : d->h_in;
  g.w_eff = d->upsample ? 2 * d->w_in : d->w_in;
  g.h_out = d->stride == 2 ? (g.h_eff + 1) / 2 : g.h_eff;
  g.w_out = d->stride == 2 ? (g.w_eff + 1) / 2 : g.w_eff;
  g.M = (int64_t)d->batch * g.h_out * g.w_out;
  g.K = g.taps * g.ctot;
  const int n_cols = d->geglu ? d->n / 2 : d->n;
  const int n_valid = d->n_valid > 0 ? d->n_valid : n_cols;
  dim3 grid((unsigned)ceil_div64(g.M, SG_BM), (unsigned)ceil_div(n_cols, SG_BN));
  cudaStream_t st = (cudaStream_t)stream;
#define SG_LAUNCH(TA, TO, GG)                                                                                     \
  igemm_simt_kernel<TA, TO, GG><<<grid, SG_THREADS, 0, st>>>(                                                     \
      (const TA*)d->a, (const TA*)d->a2, (const float*)d->w, g, d->n, d->bias, d->rowvec, d->rowvec_ld, d->residual, \
      d->residual_ld, (TO*)d->out, d->out_ld, d->out_nchw, n_valid)
  const bool abf = d->a_dtype == LDMB200_BF16, obf = d->out_dtype == LDMB200_BF16;
  LDM_CHECK_ARG(!(obf && d->out_nchw), "igemm_simt: NCHW output is fp32 only");
  if (d->geglu) {
    if (abf && obf) SG_LAUNCH(__nv_bfloat16, __nv_bfloat16, true);
    else if (abf) SG_LAUNCH(__nv_bfloat16, float, true);
    else if (obf) SG_LAUNCH(float, __nv_bfloat16, true);
    else SG_LAUNCH(float, float, true);
  } else {
    if (abf && obf) SG_LAUNCH(__nv_bfloat16, __nv_bfloat16, false);
    else if (abf) SG_LAUNCH(__nv_bfloat16, float, false);
    else if (obf) SG_LAUNCH(float, __nv_bfloat16, false);
    else SG_LAUNCH(float, float, false);
  }
#undef SG_LAUNCH
  LDM_CHECK_LAUNCH("igemm_simt");
  return 0;
}
