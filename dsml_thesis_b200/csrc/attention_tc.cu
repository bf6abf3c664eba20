// Fused flash-style self-attention for d_head = 32 on tcgen05 (K5): softmax(q k^T * scale) v per (sample, head)
// without materialising the [N, N] score matrix the reference builds (attention.py:180-191).
//
// One CTA per (128-query tile, head, sample); 2 CTAs per SM so one CTA's softmax (MUFU-bound: one exp per
// score vs 128 tensor flops) hides the other's MMA/TMA latency.
//   warp 4 : TMA producer  — Q tile once, then K/V tiles (128 keys x 32) double-buffered, all from the fused
//            QKV matrix [B*N, 3C] through one 2-D tensor map (64B-swizzled K-major rows of 32 bf16)
//   warp 5 : MMA issuer    — S = Q K^T (M128 N128 K32) into TMEM cols [0,128); O_j = P V (M128 N32 K128, V as
//            MN-major B operand straight from the row-major tile) into TMEM cols [128,160)
//   warps 0-3 : softmax    — thread r owns query row r: two passes over S in TMEM (row max, then exp2 + row
//            sum), P written as bf16 into a 128B-swizzled K-major smem tile for the PV MMA; the running
//            output lives in registers (32 fp32) and is rescaled online.
#include <algorithm>

#include "common.cuh"

namespace ldm {

constexpr int AT_THREADS = 192;
constexpr int AT_TILE_BYTES = 128 * 32 * 2;  // 8 KiB: 128 rows x 32 bf16
constexpr int AT_P_BYTES = 2 * 128 * 128;    // two K-major chunks of [128 x 64] bf16, 128B swizzle
constexpr int AT_SMEM = AT_TILE_BYTES * 5 + AT_P_BYTES + 16 * 8 + 16 + 1024;
constexpr int AT_TMEM_COLS = 256;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// MN-major B operand (V tile: rows = keys (K dim), 32 contiguous d = N dim), 64B swizzle:
// 8-key groups are SBO = 512 B apart; N = 32 spans a single 64 B atom so LBO is unused.
__device__ __forceinline__ uint64_t umma_desc_mnmajor_sw64(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;
  return d;
}

__global__ void __launch_bounds__(AT_THREADS, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out, int out_ld, int n_tok,
                    int q_off, int k_off, int v_off, float scale_log2e) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + AT_TILE_BYTES;       // 2 stages
  uint8_t* sV = sK + 2 * AT_TILE_BYTES;   // 2 stages
  uint8_t* sP = sV + 2 * AT_TILE_BYTES;   // 32 KiB, 1024-aligned (5 * 8 KiB offset)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + AT_P_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;   // [2]
  uint64_t* kv_empty = bars + 3;  // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int n_kv_tiles = (n_tok + 127) / 128;
  const int row_base = b * n_tok;  // first row of this sample in the [B*N, ld] matrix

  if (warp == 4 && lane == 0) prefetch_tmap(&tm);
  if (warp == 5) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int s = 0; s < 2; ++s) {
        mbar_init(&kv_full[s], 1);
        mbar_init(&kv_empty[s], 1);
      }
      mbar_init(s_full, 1);
      mbar_init(p_full, 128);
      mbar_init(o_full, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, AT_TMEM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_O = tmem_base + 128;

  if (warp == 4) {
    if (lane == 0) {
      mbar_expect_tx(q_full, AT_TILE_BYTES);
      tma_load_2d(sQ, &tm, q_full, q_off + h * 32, row_base + qt * 128);
      for (int j = 0; j < n_kv_tiles; ++j) {
        const int s = j & 1;
        mbar_wait(&kv_empty[s], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&kv_full[s], 2 * AT_TILE_BYTES);
        tma_load_2d(sK + s * AT_TILE_BYTES, &tm, &kv_full[s], k_off + h * 32, row_base + j * 128);
        tma_load_2d(sV + s * AT_TILE_BYTES, &tm, &kv_full[s], v_off + h * 32, row_base + j * 128);
      }
    }
  } else if (warp == 5) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 32, 0, 1);  // B = V is MN-major
      mbar_wait(q_full, 0);
      const uint32_t q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
      for (int j = 0; j < n_kv_tiles; ++j) {
        const int s = j & 1;
        mbar_wait(&kv_full[s], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + s * AT_TILE_BYTES), v_addr = smem_u32(sV + s * AT_TILE_BYTES);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
          umma_bf16_ss(tmem_S, umma_desc_kmajor(q_addr + kk * 32, 64), umma_desc_kmajor(k_addr + kk * 32, 64), idesc_qk,
                       kk > 0);
        umma_commit(s_full);
        mbar_wait(p_full, j & 1);
        tc_fence_after();
#pragma unroll
        for (int t = 0; t < 8; ++t)
          umma_bf16_ss(tmem_O, umma_desc_kmajor(p_addr + (t >> 2) * (128 * 128) + (t & 3) * 32, 128),
                       umma_desc_mnmajor_sw64(v_addr + t * 1024), idesc_pv, t > 0);
        umma_commit(&kv_empty[s]);
        if (j == n_kv_tiles - 1) umma_commit(o_full);
      }
    }
  } else {
    // ===================== softmax warps: thread r <-> query row r =====================
    const int r = threadIdx.x;  // 0..127 == TMEM lane
    const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;
    float o[32];
#pragma unroll
    for (int d = 0; d < 32; ++d) o[d] = 0.f;
    float m = -INFINITY, l = 0.f;
    const bool ragged = (n_tok & 127) != 0;
    for (int j = 0; j < n_kv_tiles; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      float tmp[32];
      if (j > 0) {  // S_j complete implies P V of tile j-1 complete (in-order tensor pipe): fetch it
        tmem_ld_32x32(tmem_O + lane_sel, tmp);
      }
      const int valid = ragged ? min(128, n_tok - j * 128) : 128;
      // ---- pass 1: row max
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld_32x32(tmem_S + lane_sel + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) mx = fmaxf(mx, (c * 32 + i < valid) ? v[i] : -INFINITY);
      }
      const float m_new = fmaxf(m, mx);
      const float corr = ex2_approx((m - m_new) * scale_log2e);  // m = -inf on the first tile -> 0
      const float mb = m_new * scale_log2e;
      l *= corr;
      if (j > 0) {
        tmem_ld_wait();
#pragma unroll
        for (int d = 0; d < 32; ++d) o[d] = (o[d] + tmp[d]) * corr;
      }
      // ---- pass 2: p = exp2(s*c - m*c), row sum, bf16 P tile (K-major, 128B swizzle)
      float lsum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld_32x32(tmem_S + lane_sel + c * 32, v);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = (c * 32 + i < valid) ? ex2_approx(fmaf(v[i], scale_log2e, -mb)) : 0.f;
          float p1 = (c * 32 + i + 1 < valid) ? ex2_approx(fmaf(v[i + 1], scale_log2e, -mb)) : 0.f;
          lsum += p0 + p1;
          __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
          pk[i >> 1] = *reinterpret_cast<uint32_t*>(&t);
        }
        // 32 keys = 64 B = 4 units of 16 B; chunk (c>>1) of 64 keys, units (c&1)*4 .. +3 of row r
        uint8_t* rowp = sP + (c >> 1) * (128 * 128) + r * 128;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int unit = (c & 1) * 4 + u;
          *reinterpret_cast<uint4*>(rowp + ((unit ^ (r & 7)) << 4)) =
              make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
        }
      }
      l += lsum;
      m = m_new;
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the UMMA (async proxy)
      tc_fence_before();
      mbar_arrive(p_full);
    }
    mbar_wait(o_full, 0);
    tc_fence_after();
    {
      float tmp[32];
      tmem_ld_32x32(tmem_O + lane_sel, tmp);
      tmem_ld_wait();
      const float inv = 1.0f / l;
      const int qi = qt * 128 + r;
      if (qi < n_tok) {
        __nv_bfloat16* op = out + (size_t)(row_base + qi) * out_ld + h * 32;
#pragma unroll
        for (int d = 0; d < 32; d += 8) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            __nv_bfloat162 t = __floats2bfloat162_rn((o[d + 2 * e] + tmp[d + 2 * e]) * inv,
                                                     (o[d + 2 * e + 1] + tmp[d + 2 * e + 1]) * inv);
            w[e] = *reinterpret_cast<uint32_t*>(&t);
          }
          *reinterpret_cast<uint4*>(op + d) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AT_TMEM_COLS);
  }
}

}  // namespace ldm

using namespace ldm;

extern "C" int ldmb200_attention_tc(const void* qkv, int ld, int q_off, int k_off, int v_off, void* out, int out_ld,
                                    int batch, int heads, int n_tokens, float scale, void* stream) {
  LDM_CHECK_ARG(qkv && out && batch > 0 && heads > 0 && n_tokens > 0, "attention_tc: bad arguments");
  LDM_CHECK_ARG(ld % 8 == 0 && out_ld % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
                "attention_tc: leading dims / offsets must be multiples of 8 bf16 (16 bytes)");
  LDM_CHECK_ARG(batch <= 65535 && heads <= 65535, "attention_tc: batch/heads exceed grid limits");
  LDM_CHECK_ARG((int64_t)batch * n_tokens < (1ll << 31), "attention_tc: too many rows");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
    if (e != cudaSuccess) {
      set_error("attention_tc: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
      return LDMB200_ERR_CUDA;
    }
    attr_set = true;
  }
  CUtensorMap tm;
  uint64_t dims[2] = {(uint64_t)ld, (uint64_t)batch * n_tokens};
  uint64_t str[1] = {(uint64_t)ld * 2};
  uint32_t box[2] = {32, 128};
  int rc = make_tensor_map_bf16(&tm, qkv, 2, dims, str, box, 64);
  if (rc) return rc;
  dim3 grid(ceil_div(n_tokens, 128), heads, batch);
  const float scale_log2e = scale * 1.4426950408889634f;
  attention_tc_kernel<<<grid, AT_THREADS, AT_SMEM, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, out_ld, n_tokens,
                                                                           q_off, k_off, v_off, scale_log2e);
  LDM_CHECK_LAUNCH("attention_tc");
  return 0;
}
