// Self-attention backward for d_head = 32 (training step; autograd of attention.py:178-192).
// Flash-style: P is recomputed from the forward's log-sum-exp, never stored.
//   S = q k^T,  P = exp2(S c - lse2),  dP = dO v^T,  D_i = sum_d dO_id O_id,  dS = P (dP - D) scale
//   dV = P^T dO,  dK = dS^T q,  dQ = dS k
// Round-1 implementation on the warp-level tensor-core path (nvcuda::wmma, fp32 accumulate): two kernels without
// atomics — one CTA per 64-key block accumulates dK/dV over all query blocks, one CTA per 64-query block accumulates
// dQ over all key blocks (S and dP are recomputed in both; 7 instead of 5 small GEMMs per tile pair).
#include <stdlib.h>

#include <mma.h>

#include "common.cuh"

namespace ldm {

using namespace nvcuda;

constexpr int AB_T = 64;          // tile: 64 queries x 64 keys
constexpr int AB_PD = 40;         // smem pitch of the [64][32] bf16 tiles (elements)
constexpr int AB_PS = 68;         // pitch of the [64][64] fp32 tiles
constexpr int AB_PP = 72;         // pitch of the [64][64] bf16 tiles
constexpr int AB_THREADS = 128;
constexpr int AB_TILE16 = AB_T * AB_PD * 2;
constexpr int AB_SMEM = 4 * AB_TILE16 + 2 * AB_T * AB_PS * 4 + 2 * AB_T * AB_PP * 2 + 2 * AB_T * 4;

struct AbParams {
  const uint16_t* qkv;
  const uint16_t* o;
  const uint16_t* dout;
  const float* lse2;
  uint16_t* dqkv;
  int ld, q_off, k_off, v_off, o_ld;
  int n_tok, heads;
  float scale, scale_log2e;
};

__device__ __forceinline__ float ab_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float ab_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// [64 rows][32 cols] bf16 tile of a row-major matrix (row pitch ld elements) -> smem pitch AB_PD; rows >= n_tok zeroed.
// 128 threads: 256 16-byte chunks, two per thread.
__device__ __forceinline__ void ab_load_tile(uint16_t* dst, const uint16_t* src, int ld, int row0, int n_tok) {
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int i = threadIdx.x + u * AB_THREADS;
    const int r = i >> 2, ch = i & 3;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (row0 + r < n_tok) v = *reinterpret_cast<const uint4*>(src + (size_t)(row0 + r) * ld + ch * 8);
    *reinterpret_cast<uint4*>(dst + r * AB_PD + ch * 8) = v;
  }
}

// D_i = sum_d dO_id O_id for the 64 rows of a query tile -> sD; also loads lse2 -> sL (rows >= n_tok: D = 0, lse = 0)
__device__ __forceinline__ void ab_load_rowstats(float* sD, float* sL, const uint16_t* o, const uint16_t* dout, int o_ld,
                                                 const float* lse, int row0, int n_tok) {
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int i = threadIdx.x + u * AB_THREADS;
    const int r = i >> 2, ch = i & 3;
    float acc = 0.f;
    if (row0 + r < n_tok) {
      const uint4 a = *reinterpret_cast<const uint4*>(o + (size_t)(row0 + r) * o_ld + ch * 8);
      const uint4 b = *reinterpret_cast<const uint4*>(dout + (size_t)(row0 + r) * o_ld + ch * 8);
      acc = ab_lo(a.x) * ab_lo(b.x) + ab_hi(a.x) * ab_hi(b.x) + ab_lo(a.y) * ab_lo(b.y) + ab_hi(a.y) * ab_hi(b.y) +
            ab_lo(a.z) * ab_lo(b.z) + ab_hi(a.z) * ab_hi(b.z) + ab_lo(a.w) * ab_lo(b.w) + ab_hi(a.w) * ab_hi(b.w);
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (ch == 0) {
      sD[r] = acc;
      sL[r] = (row0 + r < n_tok) ? lse[row0 + r] : 0.f;
    }
  }
}

// S = Q K^T and dP = dO V^T for the warp's 16 query rows x 64 keys -> sS / sdP (fp32), then
// P = exp2(S c - lse) and dS = P (dP - D) scale -> sP / sdS (bf16).  Keys >= n_tok and queries >= n_tok give 0.
__device__ __forceinline__ void ab_scores(const uint16_t* sQ, const uint16_t* sK, const uint16_t* sV, const uint16_t* sdO,
                                          float* sS, float* sdP, uint16_t* sP, uint16_t* sdS, const float* sD,
                                          const float* sL, int q0, int k0, int n_tok, float scale, float c_) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r0 = warp * 16;
  wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> fq[2], fo[2];
#pragma unroll
  for (int kk = 0; kk < 2; ++kk) {
    wmma::load_matrix_sync(fq[kk], reinterpret_cast<const __nv_bfloat16*>(sQ) + r0 * AB_PD + kk * 16, AB_PD);
    wmma::load_matrix_sync(fo[kk], reinterpret_cast<const __nv_bfloat16*>(sdO) + r0 * AB_PD + kk * 16, AB_PD);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    wmma::fragment<wmma::accumulator, 16, 16, 16, float> s, dp;
    wmma::fill_fragment(s, 0.f);
    wmma::fill_fragment(dp, 0.f);
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
      wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> fk, fv;
      wmma::load_matrix_sync(fk, reinterpret_cast<const __nv_bfloat16*>(sK) + (j * 16) * AB_PD + kk * 16, AB_PD);
      wmma::load_matrix_sync(fv, reinterpret_cast<const __nv_bfloat16*>(sV) + (j * 16) * AB_PD + kk * 16, AB_PD);
      wmma::mma_sync(s, fq[kk], fk, s);
      wmma::mma_sync(dp, fo[kk], fv, dp);
    }
    wmma::store_matrix_sync(sS + r0 * AB_PS + j * 16, s, AB_PS, wmma::mem_row_major);
    wmma::store_matrix_sync(sdP + r0 * AB_PS + j * 16, dp, AB_PS, wmma::mem_row_major);
  }
  __syncwarp();
  // 16 rows x 64 columns per warp: lane owns column pair (2*lane, 2*lane+1) of every row
#pragma unroll 4
  for (int r = 0; r < 16; ++r) {
    const int row = r0 + r;
    const float2 sv = *reinterpret_cast<const float2*>(sS + row * AB_PS + 2 * lane);
    const float2 dv = *reinterpret_cast<const float2*>(sdP + row * AB_PS + 2 * lane);
    const bool qok = q0 + row < n_tok;
    const float lse = sL[row], dd = sD[row];
    float p0 = (qok && k0 + 2 * lane < n_tok) ? exp2f(fmaf(sv.x, c_, -lse)) : 0.f;
    float p1 = (qok && k0 + 2 * lane + 1 < n_tok) ? exp2f(fmaf(sv.y, c_, -lse)) : 0.f;
    const float d0 = p0 * (dv.x - dd) * scale, d1 = p1 * (dv.y - dd) * scale;
    uint32_t pp, ds;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pp) : "f"(p1), "f"(p0));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(ds) : "f"(d1), "f"(d0));
    *reinterpret_cast<uint32_t*>(sP + row * AB_PP + 2 * lane) = pp;
    *reinterpret_cast<uint32_t*>(sdS + row * AB_PP + 2 * lane) = ds;
  }
}

struct AbSmem {
  uint16_t *sQ, *sK, *sV, *sdO, *sP, *sdS;
  float *sS, *sdP, *sD, *sL;
};
__device__ __forceinline__ AbSmem ab_carve(uint8_t* smem) {
  AbSmem m;
  m.sQ = reinterpret_cast<uint16_t*>(smem);
  m.sK = m.sQ + AB_T * AB_PD;
  m.sV = m.sK + AB_T * AB_PD;
  m.sdO = m.sV + AB_T * AB_PD;
  m.sS = reinterpret_cast<float*>(m.sdO + AB_T * AB_PD);
  m.sdP = m.sS + AB_T * AB_PS;
  m.sP = reinterpret_cast<uint16_t*>(m.sdP + AB_T * AB_PS);
  m.sdS = m.sP + AB_T * AB_PP;
  m.sD = reinterpret_cast<float*>(m.sdS + AB_T * AB_PP);
  m.sL = m.sD + AB_T;
  return m;
}

// fp32 [64][32] (pitch AB_PS, in sS) -> bf16 rows of a global matrix
__device__ __forceinline__ void ab_store_tile(const float* sT, uint16_t* dst, int ld, int row0, int n_tok) {
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int i = threadIdx.x + u * AB_THREADS;
    const int r = i >> 2, ch = i & 3;
    if (row0 + r >= n_tok) continue;
    const float* s = sT + r * AB_PS + ch * 8;
    uint4 v;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v.x) : "f"(s[1]), "f"(s[0]));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v.y) : "f"(s[3]), "f"(s[2]));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v.z) : "f"(s[5]), "f"(s[4]));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v.w) : "f"(s[7]), "f"(s[6]));
    *reinterpret_cast<uint4*>(dst + (size_t)(row0 + r) * ld + ch * 8) = v;
  }
}

// ---- dK, dV: one CTA per (64-key block, head, sample) ----
__global__ void __launch_bounds__(AB_THREADS) attention_bwd_kv_kernel(const AbParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const AbSmem m = ab_carve(smem);
  const int kb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const size_t row_base = (size_t)b * p.n_tok;
  const uint16_t* qkv = p.qkv + row_base * p.ld + h * 32;
  const uint16_t* o = p.o + row_base * p.o_ld + h * 32;
  const uint16_t* dout = p.dout + row_base * p.o_ld + h * 32;
  const float* lse = p.lse2 + ((size_t)b * p.heads + h) * p.n_tok;
  const int k0 = kb * AB_T;
  ab_load_tile(m.sK, qkv + p.k_off, p.ld, k0, p.n_tok);
  ab_load_tile(m.sV, qkv + p.v_off, p.ld, k0, p.n_tok);
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> dk[2], dv[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    wmma::fill_fragment(dk[j], 0.f);
    wmma::fill_fragment(dv[j], 0.f);
  }
  const int nqb = (p.n_tok + AB_T - 1) / AB_T;
  for (int qb = 0; qb < nqb; ++qb) {
    const int q0 = qb * AB_T;
    __syncthreads();  // previous iteration's readers of sQ / sdO / sP / sdS are done
    ab_load_tile(m.sQ, qkv + p.q_off, p.ld, q0, p.n_tok);
    ab_load_tile(m.sdO, dout, p.o_ld, q0, p.n_tok);
    ab_load_rowstats(m.sD, m.sL, o, dout, p.o_ld, lse, q0, p.n_tok);
    __syncthreads();
    ab_scores(m.sQ, m.sK, m.sV, m.sdO, m.sS, m.sdP, m.sP, m.sdS, m.sD, m.sL, q0, k0, p.n_tok, p.scale, p.scale_log2e);
    __syncthreads();
    // dV[key, d] += sum_q P[q, key] dO[q, d];  dK[key, d] += sum_q dS[q, key] Q[q, d]   (warp: 16 keys x 32 d)
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> fp, fs;
      wmma::load_matrix_sync(fp, reinterpret_cast<const __nv_bfloat16*>(m.sP) + (kk * 16) * AB_PP + warp * 16, AB_PP);
      wmma::load_matrix_sync(fs, reinterpret_cast<const __nv_bfloat16*>(m.sdS) + (kk * 16) * AB_PP + warp * 16, AB_PP);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> fo, fq;
        wmma::load_matrix_sync(fo, reinterpret_cast<const __nv_bfloat16*>(m.sdO) + (kk * 16) * AB_PD + j * 16, AB_PD);
        wmma::load_matrix_sync(fq, reinterpret_cast<const __nv_bfloat16*>(m.sQ) + (kk * 16) * AB_PD + j * 16, AB_PD);
        wmma::mma_sync(dv[j], fp, fo, dv[j]);
        wmma::mma_sync(dk[j], fs, fq, dk[j]);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    wmma::store_matrix_sync(m.sS + (warp * 16) * AB_PS + j * 16, dk[j], AB_PS, wmma::mem_row_major);
    wmma::store_matrix_sync(m.sdP + (warp * 16) * AB_PS + j * 16, dv[j], AB_PS, wmma::mem_row_major);
  }
  __syncthreads();
  uint16_t* dq = p.dqkv + row_base * p.ld + h * 32;
  ab_store_tile(m.sS, dq + p.k_off, p.ld, k0, p.n_tok);
  ab_store_tile(m.sdP, dq + p.v_off, p.ld, k0, p.n_tok);
}

// ---- dQ: one CTA per (64-query block, head, sample) ----
__global__ void __launch_bounds__(AB_THREADS) attention_bwd_q_kernel(const AbParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const AbSmem m = ab_carve(smem);
  const int qb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const size_t row_base = (size_t)b * p.n_tok;
  const uint16_t* qkv = p.qkv + row_base * p.ld + h * 32;
  const uint16_t* o = p.o + row_base * p.o_ld + h * 32;
  const uint16_t* dout = p.dout + row_base * p.o_ld + h * 32;
  const float* lse = p.lse2 + ((size_t)b * p.heads + h) * p.n_tok;
  const int q0 = qb * AB_T;
  ab_load_tile(m.sQ, qkv + p.q_off, p.ld, q0, p.n_tok);
  ab_load_tile(m.sdO, dout, p.o_ld, q0, p.n_tok);
  ab_load_rowstats(m.sD, m.sL, o, dout, p.o_ld, lse, q0, p.n_tok);
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> dq[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) wmma::fill_fragment(dq[j], 0.f);
  const int nkb = (p.n_tok + AB_T - 1) / AB_T;
  for (int kb = 0; kb < nkb; ++kb) {
    const int k0 = kb * AB_T;
    __syncthreads();
    ab_load_tile(m.sK, qkv + p.k_off, p.ld, k0, p.n_tok);
    ab_load_tile(m.sV, qkv + p.v_off, p.ld, k0, p.n_tok);
    __syncthreads();
    ab_scores(m.sQ, m.sK, m.sV, m.sdO, m.sS, m.sdP, m.sP, m.sdS, m.sD, m.sL, q0, k0, p.n_tok, p.scale, p.scale_log2e);
    __syncwarp();  // dS rows of this warp are produced and consumed by the same warp
    // dQ[q, d] += sum_key dS[q, key] K[key, d]   (warp: 16 queries x 32 d)
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> fs;
      wmma::load_matrix_sync(fs, reinterpret_cast<const __nv_bfloat16*>(m.sdS) + (warp * 16) * AB_PP + kk * 16, AB_PP);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> fk;
        wmma::load_matrix_sync(fk, reinterpret_cast<const __nv_bfloat16*>(m.sK) + (kk * 16) * AB_PD + j * 16, AB_PD);
        wmma::mma_sync(dq[j], fs, fk, dq[j]);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 2; ++j)
    wmma::store_matrix_sync(m.sS + (warp * 16) * AB_PS + j * 16, dq[j], AB_PS, wmma::mem_row_major);
  __syncthreads();
  ab_store_tile(m.sS, p.dqkv + row_base * p.ld + h * 32 + p.q_off, p.ld, q0, p.n_tok);
}

int attention_bwd_tc_launch(const void* qkv, int ld, int q_off, int k_off, int v_off, const void* out, const void* dout,
                            int o_ld, const float* lse2, void* dqkv, float* ws, int batch, int heads, int n_tokens,
                            float scale, cudaStream_t st);  // attention_bwd_tc.cu

}  // namespace ldm

using namespace ldm;

extern "C" int64_t ldmb200_attention_bwd_ws_floats(int batch, int heads, int n_tokens) {
  return (int64_t)batch * heads * ((n_tokens + 63) / 64 * 64) * 2;
}

extern "C" int ldmb200_attention_bwd(const void* qkv, int ld, int q_off, int k_off, int v_off, const void* out,
                                     const void* dout, int o_ld, const float* lse2, void* dqkv, float* ws, int batch,
                                     int heads, int n_tokens, float scale, void* stream) {
  LDM_CHECK_ARG(qkv && out && dout && lse2 && dqkv && batch > 0 && heads > 0 && n_tokens > 0, "attention_bwd: bad arguments");
  LDM_CHECK_ARG(ld % 8 == 0 && o_ld % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
                "attention_bwd: leading dims / offsets must be multiples of 8 elements (16 bytes)");
  LDM_CHECK_ARG(batch <= 65535 && heads <= 65535, "attention_bwd: batch/heads exceed grid limits");
  // tcgen05 kernels when the caller provides the {lse2, D} workspace; LDMB200_ATTN_BWD_MMA=1 forces the warp-level kernels
  static const bool force_mma = getenv("LDMB200_ATTN_BWD_MMA") && atoi(getenv("LDMB200_ATTN_BWD_MMA")) != 0;
  if (ws && !force_mma)
    return attention_bwd_tc_launch(qkv, ld, q_off, k_off, v_off, out, dout, o_ld, lse2, dqkv, ws, batch, heads, n_tokens, scale,
                                   (cudaStream_t)stream);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e1 = cudaFuncSetAttribute(attention_bwd_kv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AB_SMEM);
    cudaError_t e2 = cudaFuncSetAttribute(attention_bwd_q_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AB_SMEM);
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
      set_error("attention_bwd: cudaFuncSetAttribute failed");
      return LDMB200_ERR_CUDA;
    }
    attr_set = true;
  }
  AbParams p;
  p.qkv = (const uint16_t*)qkv; p.o = (const uint16_t*)out; p.dout = (const uint16_t*)dout; p.lse2 = lse2;
  p.dqkv = (uint16_t*)dqkv;
  p.ld = ld; p.q_off = q_off; p.k_off = k_off; p.v_off = v_off; p.o_ld = o_ld;
  p.n_tok = n_tokens; p.heads = heads;
  p.scale = scale; p.scale_log2e = scale * 1.4426950408889634f;
  dim3 grid(ceil_div(n_tokens, AB_T), heads, batch);
  cudaStream_t st = (cudaStream_t)stream;
  attention_bwd_kv_kernel<<<grid, AB_THREADS, AB_SMEM, st>>>(p);
  LDM_CHECK_LAUNCH("attention_bwd kv");
  attention_bwd_q_kernel<<<grid, AB_THREADS, AB_SMEM, st>>>(p);
  LDM_CHECK_LAUNCH("attention_bwd q");
  return 0;
}
