// fp32 SIMT attention (precision mode "fp32", and the general cross-attention path with a short context).
// One thread per query row (d_head = 32 lives in registers), K/V tiles of 64 keys staged in shared memory,
// online softmax in chunks of 16 keys.  CrossAttention.forward, attention.py:170-193.
#include "common.cuh"

namespace ldm {

constexpr int AS_D = 32, AS_TQ = 128, AS_TK = 64, AS_CH = 16;

__global__ void __launch_bounds__(AS_TQ) attention_simt_kernel(const float* __restrict__ q, int q_ld, int q_off,
                                                               const float* __restrict__ k, int k_off,
                                                               const float* __restrict__ v, int v_off, int kv_ld,
                                                               float* __restrict__ out, int out_ld, int nq, int nkv,
                                                               float scale) {
  __shared__ float Ks[AS_TK][AS_D];
  __shared__ float Vs[AS_TK][AS_D];
  const int b = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * AS_TQ + threadIdx.x;
  const bool active = qi < nq;
  float qr[AS_D], o[AS_D];
  {
    const float* qp = q + ((size_t)b * nq + (active ? qi : 0)) * q_ld + q_off + h * AS_D;
#pragma unroll
    for (int d = 0; d < AS_D; d += 4) {
      float4 t = *reinterpret_cast<const float4*>(qp + d);
      qr[d] = t.x; qr[d + 1] = t.y; qr[d + 2] = t.z; qr[d + 3] = t.w;
    }
#pragma unroll
    for (int d = 0; d < AS_D; ++d) o[d] = 0.f;
  }
  float m = -INFINITY, l = 0.f;
  for (int j0 = 0; j0 < nkv; j0 += AS_TK) {
    __syncthreads();
    for (int e = threadIdx.x; e < AS_TK * (AS_D / 4); e += AS_TQ) {
      int j = e / (AS_D / 4), d4 = e % (AS_D / 4);
      float4 kk = make_float4(0, 0, 0, 0), vv = make_float4(0, 0, 0, 0);
      if (j0 + j < nkv) {
        size_t row = (size_t)b * nkv + j0 + j;
        kk = *reinterpret_cast<const float4*>(k + row * kv_ld + k_off + h * AS_D + d4 * 4);
        vv = *reinterpret_cast<const float4*>(v + row * kv_ld + v_off + h * AS_D + d4 * 4);
      }
      *reinterpret_cast<float4*>(&Ks[j][d4 * 4]) = kk;
      *reinterpret_cast<float4*>(&Vs[j][d4 * 4]) = vv;
    }
    __syncthreads();
    const int jn = min(AS_TK, nkv - j0);
    for (int c0 = 0; c0 < jn; c0 += AS_CH) {
      float s[AS_CH];
      float cmax = -INFINITY;
#pragma unroll
      for (int j = 0; j < AS_CH; ++j) {
        float acc = 0.f;
#pragma unroll
        for (int d = 0; d < AS_D; ++d) acc = fmaf(qr[d], Ks[c0 + j][d], acc);
        s[j] = (c0 + j < jn) ? acc * scale : -INFINITY;
        cmax = fmaxf(cmax, s[j]);
      }
      const float m_new = fmaxf(m, cmax);
      const float corr = (m == -INFINITY) ? 0.f : expf(m - m_new);
      l *= corr;
#pragma unroll
      for (int d = 0; d < AS_D; ++d) o[d] *= corr;
#pragma unroll
      for (int j = 0; j < AS_CH; ++j) {
        float p = (c0 + j < jn) ? expf(s[j] - m_new) : 0.f;
        l += p;
#pragma unroll
        for (int d = 0; d < AS_D; ++d) o[d] = fmaf(p, Vs[c0 + j][d], o[d]);
      }
      m = m_new;
    }
  }
  if (active) {
    const float inv = 1.0f / l;
    float* op = out + ((size_t)b * nq + qi) * out_ld + h * AS_D;
#pragma unroll
    for (int d = 0; d < AS_D; d += 4)
      *reinterpret_cast<float4*>(op + d) = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
  }
}

}  // namespace ldm

using namespace ldm;

extern "C" int ldmb200_attention_simt(const float* q, int q_ld, int q_off, const float* k, int k_off, const float* v,
                                      int v_off, int kv_ld, float* out, int out_ld, int batch, int heads, int nq,
                                      int nkv, float scale, void* stream) {
  LDM_CHECK_ARG(q && k && v && out && batch > 0 && heads > 0 && nq > 0 && nkv > 0, "attention_simt: bad arguments");
  LDM_CHECK_ARG(q_ld % 4 == 0 && kv_ld % 4 == 0 && out_ld % 4 == 0 && q_off % 4 == 0 && k_off % 4 == 0 &&
                    v_off % 4 == 0,
                "attention_simt: leading dims / offsets must be multiples of 4 floats");
  LDM_CHECK_ARG(batch <= 65535 && heads <= 65535, "attention_simt: batch/heads exceed grid limits");
  dim3 grid(ceil_div(nq, AS_TQ), heads, batch);
  attention_simt_kernel<<<grid, AS_TQ, 0, (cudaStream_t)stream>>>(q, q_ld, q_off, k, k_off, v, v_off, kv_ld, out,
                                                                  out_ld, nq, nkv, scale);
  LDM_CHECK_LAUNCH("attention_simt");
  return 0;
}
