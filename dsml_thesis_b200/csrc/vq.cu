// First-stage decode helpers: the VectorQuantizer2 nearest-code lookup fused with post_quant_conv
// and the channels-last staging of the Decoder input, and the row softmax of the Decoder's single-head AttnBlock
// (whose two contractions run on the implicit-GEMM kernels with per-sample activations as the "weight" operand).
#include <algorithm>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace ldm {

#define VQ_DISPATCH_DTYPE(dt, call)    \
  do {                                 \
    if ((dt) == LDMB200_BF16) {        \
      using T = __nv_bfloat16;         \
      call;                            \
    } else if ((dt) == LDMB200_F16) {  \
      using T = __half;                \
      call;                            \
    } else {                           \
      using T = float;                 \
      call;                            \
    }                                  \
  } while (0)

template <typename T>
__device__ __forceinline__ T vq_cvt(float v);
template <>
__device__ __forceinline__ float vq_cvt<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 vq_cvt<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ __half vq_cvt<__half>(float v) { return __float2half_rn(v); }

constexpr int VQ_MAX_E = 8;       // embed_dim / z_channels bound (3 in every shipped config)
constexpr int VQ_CHUNK = 1024;    // codes staged in shared memory per pass
constexpr int VQ_THREADS = 256;

// 32 latent pixels per CTA (one per lane); the CTA's 8 warps each scan one eighth of every codebook chunk staged in
// shared memory and the per-warp winners are merged (smaller distance, then smaller index = torch.argmin's first
// minimum).  Distances follow quantize.py:292-294 term by term in fp32:
//   d = (sum z^2 + sum e^2) - 2 * (z . e)
// z_q = z + (e - z) (the straight-through form of quantize.py:309, kept because it is not exactly e in fp32).
// EC = 3 is the shipped embed_dim: codes are staged as {e0, e1, e2, |e|^2} quads (one 16-byte shared-memory load and
// seven fp32 instructions per pixel-code pair); EC = 0 is the generic embed_dim <= 8 path.
template <typename T, int EC>
__global__ void __launch_bounds__(VQ_THREADS)
vq_lookup_postconv_kernel(const float* __restrict__ z, const float* __restrict__ codebook, int n_e, int E,
                          const float* __restrict__ w, const float* __restrict__ bias, int zc, T* __restrict__ y,
                          int c_pad, int64_t* __restrict__ indices, float* __restrict__ zq_out, int batch, int hw) {
  __shared__ __align__(16) float s_code[VQ_CHUNK * VQ_MAX_E];
  __shared__ float s_ee[VQ_CHUNK];
  __shared__ float s_best[VQ_THREADS];
  __shared__ int s_besti[VQ_THREADS];
  constexpr int NW = VQ_THREADS / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t pix = (int64_t)blockIdx.x * 32 + lane;
  const int64_t total = (int64_t)batch * hw;
  const bool live = pix < total;
  const int b = live ? (int)(pix / hw) : 0, p = live ? (int)(pix % hw) : 0;
  float zv[VQ_MAX_E];
#pragma unroll
  for (int e = 0; e < VQ_MAX_E; ++e) zv[e] = (live && e < E) ? z[((int64_t)b * E + e) * hw + p] : 0.f;
  float q[VQ_MAX_E];
#pragma unroll
  for (int e = 0; e < VQ_MAX_E; ++e) q[e] = zv[e];
  if (codebook) {
    float zz = 0.f;
#pragma unroll
    for (int e = 0; e < VQ_MAX_E; ++e)
      if (e < E) zz = e == 0 ? __fmul_rn(zv[0], zv[0]) : __fadd_rn(zz, __fmul_rn(zv[e], zv[e]));
    float best = INFINITY;
    int best_i = 0;
    for (int c0 = 0; c0 < n_e; c0 += VQ_CHUNK) {
      const int nc = min(VQ_CHUNK, n_e - c0);
      __syncthreads();
      if constexpr (EC == 3) {
        float4* c4 = reinterpret_cast<float4*>(s_code);
        for (int i = threadIdx.x; i < nc; i += VQ_THREADS) {
          const float* src = codebook + (int64_t)(c0 + i) * 3;
          const float e0 = src[0], e1 = src[1], e2 = src[2];
          c4[i] = make_float4(e0, e1, e2, __fadd_rn(__fadd_rn(__fmul_rn(e0, e0), __fmul_rn(e1, e1)), __fmul_rn(e2, e2)));
        }
      } else {
        for (int i = threadIdx.x; i < nc * E; i += VQ_THREADS) s_code[i] = codebook[(int64_t)c0 * E + i];
        __syncthreads();
        for (int i = threadIdx.x; i < nc; i += VQ_THREADS) {
          float ee = __fmul_rn(s_code[i * E], s_code[i * E]);
          for (int e = 1; e < E; ++e) ee = __fadd_rn(ee, __fmul_rn(s_code[i * E + e], s_code[i * E + e]));
          s_ee[i] = ee;
        }
      }
      __syncthreads();
      const int per = (nc + NW - 1) / NW, i0 = warp * per, i1 = min(nc, i0 + per);
      if constexpr (EC == 3) {
        const float4* c4 = reinterpret_cast<const float4*>(s_code);
#pragma unroll 4
        for (int i = i0; i < i1; ++i) {
          const float4 c = c4[i];
          const float dot = fmaf(zv[2], c.z, fmaf(zv[1], c.y, __fmul_rn(zv[0], c.x)));
          const float d = fmaf(-2.0f, dot, __fadd_rn(zz, c.w));   // == (zz + ee) - 2*dot: the product by 2 is exact
          if (d < best) {
            best = d;
            best_i = c0 + i;
          }
        }
      } else {
        for (int i = i0; i < i1; ++i) {
          const float* ce = s_code + i * E;
          float dot = __fmul_rn(zv[0], ce[0]);
#pragma unroll
          for (int e = 1; e < VQ_MAX_E; ++e)
            if (e < E) dot = fmaf(zv[e], ce[e], dot);
          const float d = __fsub_rn(__fadd_rn(zz, s_ee[i]), 2.0f * dot);
          if (d < best) {   // ascending scan inside the warp's slice: the first minimum of the slice
            best = d;
            best_i = c0 + i;
          }
        }
      }
    }
    s_best[threadIdx.x] = best;
    s_besti[threadIdx.x] = best_i;
    __syncthreads();
    if (warp == 0) {
#pragma unroll
      for (int ww = 1; ww < NW; ++ww) {
        const float d = s_best[ww * 32 + lane];
        const int di = s_besti[ww * 32 + lane];
        if (d < best || (d == best && di < best_i)) {
          best = d;
          best_i = di;
        }
      }
      if (live) {
#pragma unroll
        for (int e = 0; e < VQ_MAX_E; ++e)
          if (e < E) q[e] = __fadd_rn(zv[e], __fsub_rn(codebook[(int64_t)best_i * E + e], zv[e]));
        if (indices) indices[pix] = best_i;
      }
    }
  }
  if (!live || warp != 0) return;
  if (zq_out) {
#pragma unroll
    for (int e = 0; e < VQ_MAX_E; ++e)
      if (e < E) zq_out[((int64_t)b * E + e) * hw + p] = q[e];
  }
  T* row = y + pix * c_pad;
  for (int c = 0; c < c_pad; ++c) {
    float v = 0.f;
    if (c < zc) {
      v = bias ? bias[c] : 0.f;
#pragma unroll
      for (int e = 0; e < VQ_MAX_E; ++e)
        if (e < E) v = fmaf(w[c * E + e], q[e], v);
    }
    row[c] = vq_cvt<T>(v);
  }
}

// softmax(scale * s) over each row (model.py:186-187).  One block per row, the row staged in shared memory.
template <typename T>
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ s, int64_t s_ld, T* __restrict__ p, int64_t p_ld, int n, float scale) {
  extern __shared__ float s_row[];
  __shared__ float s_red[8];
  const float* src = s + (int64_t)blockIdx.x * s_ld;
  T* dst = p + (int64_t)blockIdx.x * p_ld;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float m = -INFINITY;
  for (int i = threadIdx.x; i < n; i += 256) {
    const float v = src[i] * scale;
    s_row[i] = v;
    m = fmaxf(m, v);
  }
  m = warp_max(m);
  if (lane == 0) s_red[warp] = m;
  __syncthreads();
  m = s_red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) m = fmaxf(m, s_red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) {
    const float e = __expf(s_row[i] - m);
    s_row[i] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  if (lane == 0) s_red[warp] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) sum += s_red[i];
  const float inv = 1.0f / sum;
  for (int i = threadIdx.x; i < n; i += 256) dst[i] = vq_cvt<T>(s_row[i] * inv);
}

}  // namespace ldm

using namespace ldm;

extern "C" {

int ldmb200_vq_lookup_postconv(const float* z, const float* codebook, int n_e, int e_dim, const float* w,
                               const float* bias, int z_channels, void* y, int out_dtype, int c_pad, int64_t* indices,
                               float* zq, int batch, int hw, void* stream) {
  LDM_CHECK_ARG(z && w && y && batch > 0 && hw > 0, "vq_lookup_postconv: null pointer / empty");
  LDM_CHECK_ARG(e_dim >= 1 && e_dim <= VQ_MAX_E && z_channels >= 1 && z_channels <= c_pad,
                "vq_lookup_postconv: embed_dim %d (max %d) / z_channels %d / c_pad %d", e_dim, VQ_MAX_E, z_channels, c_pad);
  LDM_CHECK_ARG(!codebook || n_e > 0, "vq_lookup_postconv: empty codebook");
  const int64_t total = (int64_t)batch * hw;
  const int blocks = (int)ceil_div64(total, 32);
  if (e_dim == 3) {
    VQ_DISPATCH_DTYPE(out_dtype, (vq_lookup_postconv_kernel<T, 3><<<blocks, VQ_THREADS, 0, (cudaStream_t)stream>>>(
                                     z, codebook, n_e, e_dim, w, bias, z_channels, (T*)y, c_pad, indices, zq, batch, hw)));
  } else {
    VQ_DISPATCH_DTYPE(out_dtype, (vq_lookup_postconv_kernel<T, 0><<<blocks, VQ_THREADS, 0, (cudaStream_t)stream>>>(
                                     z, codebook, n_e, e_dim, w, bias, z_channels, (T*)y, c_pad, indices, zq, batch, hw)));
  }
  LDM_CHECK_LAUNCH("vq_lookup_postconv");
  return 0;
}

int ldmb200_softmax_rows(const float* s, int64_t s_ld, void* p, int out_dtype, int64_t p_ld, int rows, int n, float scale,
                         void* stream) {
  LDM_CHECK_ARG(s && p && rows > 0 && n > 0 && s_ld >= n && p_ld >= n, "softmax_rows: bad arguments");
  LDM_CHECK_ARG(n <= 16384, "softmax_rows: rows of up to 16384 scores are staged in shared memory, got %d", n);
  const size_t smem = (size_t)n * sizeof(float);
  if (smem > 48 * 1024) {
    VQ_DISPATCH_DTYPE(out_dtype, cudaFuncSetAttribute(softmax_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                      (int)smem));
  }
  VQ_DISPATCH_DTYPE(out_dtype, (softmax_rows_kernel<T><<<rows, 256, smem, (cudaStream_t)stream>>>(s, s_ld, (T*)p, p_ld, n,
                                                                                                 scale)));
  LDM_CHECK_LAUNCH("softmax_rows");
  return 0;
}

}  // extern "C"
