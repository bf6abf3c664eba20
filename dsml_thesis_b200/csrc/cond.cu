// Audio conditioning encoder of the talking-face model: Conv1DTemporalAttention.forward,
// talking_face/ldm/modules/encoders/modules.py:76-113 — five Conv1d(k=3, pad=1) + LeakyReLU(0.02) layers squeeze the
// [seq_len, 768] window of audio features to one score per frame, Linear(seq_len, seq_len) + softmax turns the scores
// into weights, and the output is the weighted sum of the window's feature vectors.  Per-frame, latency-bound work on
// the critical path of progressive sampling (progressive_sampling_difftalk.py:286-291): one launch, one CTA per sample,
// every intermediate in shared memory.
#include <algorithm>

#include "common.cuh"

namespace ldm {

constexpr int CTA_THREADS = 256;
constexpr int CTA_MAX_L = 32;     // seq_len bound (17 in the shipped config)
constexpr int CTA_MAX_LAYERS = 8;

struct CtaParams {
  const float* x;          // [B, L, D]
  float* out;              // [B, D]
  const float* w[CTA_MAX_LAYERS];   // Conv1d weights [cout, cin, 3]
  const float* b[CTA_MAX_LAYERS];
  int cin[CTA_MAX_LAYERS], cout[CTA_MAX_LAYERS];
  int n_layers;
  const float* lin_w;      // [L, L]
  const float* lin_b;      // [L]
  int L, D;
  float slope;
};

// out[o][l] = leaky(b[o] + sum_{c,k} w[o][c][k] * in[c][l + k - 1]); in / out are [channels][L] in shared memory
// (layer 0 reads the transposed window xs[l][c] instead).  Warp per output channel, lanes over the (c, k) index so
// that the weight row is read coalesced; L partial sums per lane, reduced by shuffles.
template <bool FIRST>
__device__ __forceinline__ void conv1d_layer(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ w,
                                             const float* __restrict__ b, int cin, int cout, int L, int D, float slope) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = CTA_THREADS / 32;
  const int K = cin * 3;
  for (int o = warp; o < cout; o += nwarps) {
    float acc[CTA_MAX_L];
#pragma unroll
    for (int l = 0; l < CTA_MAX_L; ++l) acc[l] = 0.f;
    const float* wr = w + (size_t)o * K;
    for (int j = lane; j < K; j += 32) {
      const float wv = wr[j];
      const int c = j / 3, k = j - 3 * c;
#pragma unroll
      for (int l = 0; l < CTA_MAX_L; ++l) {
        const int src = l + k - 1;
        if (l < L && src >= 0 && src < L) acc[l] = fmaf(wv, FIRST ? in[src * D + c] : in[c * L + src], acc[l]);
      }
    }
#pragma unroll
    for (int l = 0; l < CTA_MAX_L; ++l) {
      if (l < L) {
        float v = warp_sum(acc[l]);
        if (lane == 0) {
          v += b[o];
          out[o * L + l] = v > 0.f ? v : v * slope;
        }
      }
    }
  }
}

__global__ void __launch_bounds__(CTA_THREADS) conv1d_temporal_attention_kernel(CtaParams p) {
  extern __shared__ float smem[];
  const int L = p.L, D = p.D;
  float* xs = smem;                       // [L][D]
  int max_c = 0;
  for (int i = 0; i < p.n_layers; ++i) max_c = max(max_c, p.cout[i]);
  float* buf0 = xs + L * D;               // [max_c][L]
  float* buf1 = buf0 + max_c * L;
  float* att = buf1 + max_c * L;          // [L]
  const float* x = p.x + (size_t)blockIdx.x * L * D;
  for (int i = threadIdx.x; i < L * D; i += CTA_THREADS) xs[i] = x[i];
  __syncthreads();
  const float* cur = xs;
  float* nxt = buf0;
  for (int i = 0; i < p.n_layers; ++i) {
    if (i == 0)
      conv1d_layer<true>(cur, nxt, p.w[i], p.b[i], p.cin[i], p.cout[i], L, D, p.slope);
    else
      conv1d_layer<false>(cur, nxt, p.w[i], p.b[i], p.cin[i], p.cout[i], L, D, p.slope);
    __syncthreads();
    cur = nxt;
    nxt = (nxt == buf0) ? buf1 : buf0;
  }
  // attentionNet: softmax(Linear(L, L)(scores)) — the last conv layer has one channel, so cur = scores[L]
  if (threadIdx.x < 32) {
    const int l = threadIdx.x;
    float v = -INFINITY;
    if (l < L) {
      v = p.lin_b[l];
      for (int j = 0; j < L; ++j) v = fmaf(p.lin_w[l * L + j], cur[j], v);
    }
    const float m = warp_max(v);
    const float e = l < L ? __expf(v - m) : 0.f;
    const float s = warp_sum(e);
    if (l < L) att[l] = e / s;
  }
  __syncthreads();
  // torch.bmm(x^T, attention): out[d] = sum_l x[l][d] * att[l]
  float* out = p.out + (size_t)blockIdx.x * D;
  for (int d = threadIdx.x; d < D; d += CTA_THREADS) {
    float v = 0.f;
    for (int l = 0; l < L; ++l) v = fmaf(xs[l * D + d], att[l], v);
    out[d] = v;
  }
}

}  // namespace ldm

using namespace ldm;

extern "C" int ldmb200_conv1d_temporal_attention(const float* x, float* out, const float* const* conv_w,
                                                 const float* const* conv_b, const int* channels, int n_layers,
                                                 const float* lin_w, const float* lin_b, int batch, int seq_len,
                                                 float leaky_slope, void* stream) {
  LDM_CHECK_ARG(x && out && conv_w && conv_b && channels && lin_w && lin_b && batch > 0, "conv1d_temporal_attention: null pointer / empty");
  LDM_CHECK_ARG(n_layers >= 1 && n_layers <= CTA_MAX_LAYERS, "conv1d_temporal_attention: 1..%d conv layers", CTA_MAX_LAYERS);
  LDM_CHECK_ARG(seq_len >= 1 && seq_len <= CTA_MAX_L, "conv1d_temporal_attention: seq_len %d (max %d)", seq_len, CTA_MAX_L);
  LDM_CHECK_ARG(channels[n_layers] == 1, "conv1d_temporal_attention: the last conv layer must have one output channel");
  CtaParams p;
  memset(&p, 0, sizeof(p));
  p.x = x;
  p.out = out;
  int max_c = 0;
  for (int i = 0; i < n_layers; ++i) {
    LDM_CHECK_ARG(conv_w[i] && conv_b[i] && channels[i] > 0 && channels[i + 1] > 0, "conv1d_temporal_attention: layer %d", i);
    p.w[i] = conv_w[i];
    p.b[i] = conv_b[i];
    p.cin[i] = channels[i];
    p.cout[i] = channels[i + 1];
    max_c = std::max(max_c, channels[i + 1]);
  }
  p.n_layers = n_layers;
  p.lin_w = lin_w;
  p.lin_b = lin_b;
  p.L = seq_len;
  p.D = channels[0];
  p.slope = leaky_slope;
  const size_t smem = ((size_t)seq_len * p.D + 2 * (size_t)max_c * seq_len + seq_len) * sizeof(float);
  LDM_CHECK_ARG(smem <= 200 * 1024, "conv1d_temporal_attention: window of %d x %d floats does not fit shared memory", seq_len, p.D);
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(conv1d_temporal_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  conv1d_temporal_attention_kernel<<<batch, CTA_THREADS, smem, (cudaStream_t)stream>>>(p);
  LDM_CHECK_LAUNCH("conv1d_temporal_attention");
  return 0;
}
