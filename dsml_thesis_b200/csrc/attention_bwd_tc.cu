// Self-attention backward for d_head = 32 on tcgen05 / TMEM / TMA (training step).
//
//   S = q k^T, P = exp2(S c - lse2), dP = dO v^T, dS = P (dP - D) scale, dV = P^T dO, dK = dS^T q, dQ = dS k
//
// One templated kernel in two roles (no atomics, P recomputed in both):
//   KV role: CTA = (128-key block, head, sample); rows = keys;    streams 64-query slabs of (q, dO, {lse2, D});
//            T1 = k q^T = S^T, T2 = v dO^T = dP^T;  dV += P^T dO,  dK += dS^T q
//   Q  role: CTA = (128-query block, head, sample); rows = queries; streams 64-key slabs of (k, v);
//            T1 = q k^T = S,   T2 = dO v^T = dP;    dQ += dS k
// Per slab the tensor pipe writes T1 / T2 (M128 N64 K32) into a double-buffered TMEM pair; eight softmax warps
// (thread <-> row <-> TMEM lane, two warpgroups splitting the 64 columns) read them once, apply lse / D / scale and
// write the bf16 operand tiles (P^T and dS^T, or dS) K-major with the 128-byte swizzle; the MMA warp then accumulates
// the outputs in TMEM with the streamed slab as the MN-major B operand (like V in the forward kernel), always one slab
// ahead on T1 / T2.
//   warps 0-15: softmax, two sets of eight — set u owns the TMEM / operand buffer pair u, i.e. every other slab, so two
//   slabs are in the softmax stage at any time (the stage is latency-bound: ex2, two packs and two swizzled 16-byte
//   stores per column pair on one thread per row);   warp 16: MMA issuer   warp 17: TMA producer (+ barrier init,
//   TMEM allocation)
#include <algorithm>
#include <type_traits>

#include "common.cuh"

namespace ldm {

constexpr int AX_ST = 4;                       // streamed-slab ring depth
constexpr int AX_RES = 128 * 32 * 2;           // resident tile: 128 rows x 32 (8 KiB)
constexpr int AX_SLAB = 64 * 32 * 2;           // streamed tile: 64 rows x 32 (4 KiB)
constexpr int AX_LD = 64 * 8;                  // {lse2, D} of the slab's 64 queries (KV role)
constexpr int AX_STAGE = 2 * AX_SLAB + AX_LD;  // 8.5 KiB, keeps 512-byte alignment of the slabs
constexpr int AX_OP = 128 * 128;               // operand tile: 128 rows x 64 bf16 (16 KiB, 128B-swizzled K-major)
constexpr int AX_SMEM = 2 * AX_RES + AX_ST * AX_STAGE + 4 * AX_OP + (1 + 2 * AX_ST + 6) * 8 + 16 + 1024;
#ifndef LDM_AX_SETS
#define LDM_AX_SETS 1   // measured r01t at 64x64, B=32: one set 21.8 ms, two sets 22.1 ms (the stage is TMEM-read / issue bound, not latency bound)
#endif
constexpr int AX_SETS = LDM_AX_SETS;           // softmax warp sets of 8 (set u owns buffer pair u when there are two)
constexpr int AX_MMA_WARP = 8 * AX_SETS, AX_TMA_WARP = 8 * AX_SETS + 1;
constexpr int AX_THREADS = 32 * (8 * AX_SETS + 2);
constexpr int AX_ACC = 256;                    // TMEM: T1_u at u*128, T2_u at u*128 + 64, accA at 256, accB at 288

struct AxParams {
  uint16_t* dqkv;
  const float2* ld;     // [B, heads, n_pad] x 2 floats: {-lse2, -D scale}, pair-interleaved (attention_bwd_prep_kernel)
  int ld_qkv, q_off, k_off, v_off;
  int n_tok, n_pad, heads;
  float scale, scale_log2e;
};

__device__ __forceinline__ float ax_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t ax_pack(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void ax_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
#define AX_R8(v, o) "+f"(v[o]), "+f"(v[o + 1]), "+f"(v[o + 2]), "+f"(v[o + 3]), "+f"(v[o + 4]), "+f"(v[o + 5]), "+f"(v[o + 6]), "+f"(v[o + 7])
__device__ __forceinline__ void ax_wait_regs(float (&a)[32], float (&b)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : AX_R8(a, 0), AX_R8(a, 8), AX_R8(a, 16), AX_R8(a, 24), AX_R8(b, 0), AX_R8(b, 8), AX_R8(b, 16), AX_R8(b, 24)::"memory");
}

template <bool KV>
__global__ void __launch_bounds__(AX_THREADS, 1)
attention_bwd_tc_kernel(const __grid_constant__ CUtensorMap tm_qkv128, const __grid_constant__ CUtensorMap tm_qkv64,
                        const __grid_constant__ CUtensorMap tm_do128, const __grid_constant__ CUtensorMap tm_do64,
                        const AxParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // (offset, not uintptr_t: keeps the shared address space)
  uint8_t* sR1 = smem;
  uint8_t* sR2 = sR1 + AX_RES;
  uint8_t* sOp = sR2 + AX_RES;                 // [2 buffers][Ds | Pw] 4 x 16 KiB, 1024-aligned
  uint8_t* sSt = sOp + 4 * AX_OP;              // ring: [S1 | S2 | LD] per stage
  uint64_t* bars = reinterpret_cast<uint64_t*>(sSt + AX_ST * AX_STAGE);
  uint64_t* r_full = bars;
  uint64_t* st_full = bars + 1;
  uint64_t* st_empty = st_full + AX_ST;
  uint64_t* t_full = st_empty + AX_ST;          // [2]
  uint64_t* op_full = t_full + 2;               // [2]
  uint64_t* op_free = op_full + 2;              // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(op_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int row_base = b * p.n_tok;
  const int row0 = rb * 128;
  const int n_slabs = (p.n_tok + 63) / 64;

  if (warp == AX_TMA_WARP) {
    if (lane == 0) {
      prefetch_tmap(&tm_qkv128);
      prefetch_tmap(&tm_qkv64);
      prefetch_tmap(KV ? &tm_do64 : &tm_do128);
      mbar_init(r_full, 1);
      for (int s = 0; s < AX_ST; ++s) {
        mbar_init(&st_full[s], 1);
        mbar_init(&st_empty[s], 1);
      }
      for (int u = 0; u < 2; ++u) {
        mbar_init(&t_full[u], 1);
        mbar_init(&op_full[u], 256);
        mbar_init(&op_free[u], 1);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == AX_TMA_WARP) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mbar_expect_tx(r_full, 2 * AX_RES);
      if (KV) {
        tma_load_2d(sR1, &tm_qkv128, r_full, p.k_off + h * 32, row_base + row0);
        tma_load_2d(sR2, &tm_qkv128, r_full, p.v_off + h * 32, row_base + row0);
      } else {
        tma_load_2d(sR1, &tm_qkv128, r_full, p.q_off + h * 32, row_base + row0);
        tma_load_2d(sR2, &tm_do128, r_full, h * 32, row_base + row0);
      }
    }
    __syncwarp();
    const float2* ldrow = p.ld + ((size_t)b * p.heads + h) * p.n_pad;
    for (int i = 0; i < n_slabs; ++i) {
      const int s = i % AX_ST;
      mbar_wait(&st_empty[s], ((i / AX_ST) & 1) ^ 1);
      if (elect_one()) {
        uint8_t* st = sSt + s * AX_STAGE;
        mbar_expect_tx(&st_full[s], KV ? AX_STAGE : 2 * AX_SLAB);
        if (KV) {
          tma_load_2d(st, &tm_qkv64, &st_full[s], p.q_off + h * 32, row_base + i * 64);
          tma_load_2d(st + AX_SLAB, &tm_do64, &st_full[s], h * 32, row_base + i * 64);
          ax_bulk_g2s(st + 2 * AX_SLAB, ldrow + i * 64, AX_LD, &st_full[s]);
        } else {
          tma_load_2d(st, &tm_qkv64, &st_full[s], p.k_off + h * 32, row_base + i * 64);
          tma_load_2d(st + AX_SLAB, &tm_qkv64, &st_full[s], p.v_off + h * 32, row_base + i * 64);
        }
      }
      __syncwarp();
    }
  } else if (warp == AX_MMA_WARP) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_t = umma_idesc_bf16(128, 64, 0, 0);          // T = R . S^T, both K-major
    constexpr uint32_t idesc_acc = umma_idesc_bf16(128, 32, 0, 1);        // acc += Op . S, S as MN-major B
    constexpr uint32_t hi_sw64 = (512u >> 4) | (1u << 14) | (4u << 29);
    constexpr uint32_t hi_sw128 = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t r1_lo = ((smem_u32(sR1) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t r2_lo = ((smem_u32(sR2) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t op_lo = ((smem_u32(sOp) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t st_lo = ((smem_u32(sSt) & 0x3FFFFu) >> 4) | (1u << 16);
    auto issue_t = [&](int i) {     // T1_u = R1 . S1^T, T2_u = R2 . S2^T  (K = 32: two MMAs each)
      const int u = i & 1, s = i % AX_ST;
      const uint32_t s1 = st_lo + s * (AX_STAGE >> 4), s2 = s1 + (AX_SLAB >> 4);
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        umma_bf16_ss(tmem_base + u * 128, ((uint64_t)hi_sw64 << 32) | (r1_lo + kk * 2), ((uint64_t)hi_sw64 << 32) | (s1 + kk * 2),
                     idesc_t, kk);
        umma_bf16_ss(tmem_base + u * 128 + 64, ((uint64_t)hi_sw64 << 32) | (r2_lo + kk * 2),
                     ((uint64_t)hi_sw64 << 32) | (s2 + kk * 2), idesc_t, kk);
      }
    };
    auto issue_acc = [&](int i) {   // accA += Ds_u . S1 ; KV: accB += Pw_u . S2   (K = 64 slab rows: four MMAs each)
      const int u = i & 1, s = i % AX_ST;
      const uint32_t s1 = st_lo + s * (AX_STAGE >> 4), s2 = s1 + (AX_SLAB >> 4);
      const uint32_t ds = op_lo + u * (2 * AX_OP >> 4), pw = ds + (AX_OP >> 4);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const uint32_t accf = (i > 0 || t > 0) ? 1u : 0u;
        umma_bf16_ss(tmem_base + AX_ACC, ((uint64_t)hi_sw128 << 32) | (ds + t * 2),
                     ((uint64_t)hi_sw64 << 32) | (s1 + t * (1024 >> 4)), idesc_acc, accf);
        if (KV)
          umma_bf16_ss(tmem_base + AX_ACC + 32, ((uint64_t)hi_sw128 << 32) | (pw + t * 2),
                       ((uint64_t)hi_sw64 << 32) | (s2 + t * (1024 >> 4)), idesc_acc, accf);
      }
    };
    mbar_wait(r_full, 0);
    mbar_wait(&st_full[0], 0);
    tc_fence_after();
    if (elect_one()) {
      issue_t(0);
      umma_commit(&t_full[0]);
    }
    __syncwarp();
    for (int i = 0; i < n_slabs; ++i) {
      const int u = i & 1;
      if (i + 1 < n_slabs) {   // T(i+1) into the other buffer: its previous reader (slab i-1) has arrived on op_full
        const int s1 = (i + 1) % AX_ST;
        mbar_wait(&st_full[s1], ((i + 1) / AX_ST) & 1);
        tc_fence_after();
        if (elect_one()) {
          issue_t(i + 1);
          umma_commit(&t_full[(i + 1) & 1]);
        }
        __syncwarp();
      }
      mbar_wait(&op_full[u], (i >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        issue_acc(i);
        umma_commit(&st_empty[i % AX_ST]);   // slab i fully consumed
        umma_commit(&op_free[u]);            // operand buffers u reusable; after the last slab: accumulators final
      }
      __syncwarp();
    }
  } else {
    // ===================== softmax warps: thread <-> row (TMEM lane), warpgroup <-> 32-column half ===============
    const int set = warp >> 3;                      // owns buffer pair u == set (slabs i with i & 1 == set) if AX_SETS == 2
    const int wg = (warp >> 2) & 1;                 // 0 | 1: 32-column half of the slab
    const int r = (warp & 3) * 32 + lane;           // row of the 128-row block == TMEM lane
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const int grow = row0 + r;                      // row index inside the sample
    float lse_r = 0.f, d_r = 0.f;
    if (!KV && grow < p.n_tok) {  // (already negated: the stage only adds)
      const float* lp = reinterpret_cast<const float*>(p.ld + ((size_t)b * p.heads + h) * p.n_pad) + (size_t)(grow >> 1) * 4 + (grow & 1);
      lse_r = lp[0];
      d_r = lp[2];
    }
    const f32x2 c2 = pk2(p.scale_log2e, p.scale_log2e), sc2 = pk2(p.scale, p.scale);
    for (int i = set; i < n_slabs; i += AX_SETS) {
      const int u = i & 1, s = i % AX_ST;
      if (KV) mbar_wait(&st_full[s], (i / AX_ST) & 1);     // the slab's {lse2, D}
      mbar_wait(&t_full[u], (i >> 1) & 1);
      tc_fence_after();
      float t1[32], t2[32];
      tmem_ld_32x32(tmem_base + lane_sel + u * 128 + wg * 32, t1);
      tmem_ld_32x32(tmem_base + lane_sel + u * 128 + 64 + wg * 32, t2);
      if (i >= 2) mbar_wait(&op_free[u], ((i >> 1) + 1) & 1);   // MMAs of slab i-2 no longer read the operand tiles
      ax_wait_regs(t1, t2);
      const float2* ldp = reinterpret_cast<const float2*>(sSt + s * AX_STAGE + 2 * AX_SLAB) + wg * 32;
      uint8_t* ds_row = sOp + u * (2 * AX_OP) + r * 128;
      uint8_t* pw_row = ds_row + AX_OP;
      const int col0 = i * 64 + wg * 32;            // first streamed index of this thread's 32 columns
      // {lse2, D * scale} per streamed index (KV role) or per row (Q role): dS = P (dP scale - D scale)
      auto columns = [&](auto check_tag) {
        constexpr bool CHECK = decltype(check_tag)::value;   // only the sample's last, partial slab needs column masks
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {            // 8 columns -> one 16-byte unit of each operand row
          uint32_t pk[4], dk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {  // one column pair on the packed fp32 pipe: 2 FFMA2 + 2 ex2 + FMUL2 + packs
            const int j = q4 * 8 + e * 2;
            f32x2 nl2 = pk2(lse_r, lse_r), nd2 = pk2(d_r, d_r);
            if (KV) {   // {-lse_j, -lse_j+1 | -D_j, -D_j+1}: the two halves of one 16-byte word (broadcast read)
              const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(ldp + j);
              nl2 = v.x;
              nd2 = v.y;
            }
            float p0, p1, x0, x1;
            unpk2(fma2(pk2(t1[j], t1[j + 1]), c2, nl2), x0, x1);
            p0 = ax_ex2(x0);
            p1 = ax_ex2(x1);
            if (CHECK) {
              if (col0 + j >= p.n_tok) p0 = 0.f;
              if (col0 + j + 1 >= p.n_tok) p1 = 0.f;
            }
            float s0, s1;
            unpk2(mul2(pk2(p0, p1), fma2(pk2(t2[j], t2[j + 1]), sc2, nd2)), s0, s1);
            dk[e] = ax_pack(s0, s1);
            if (KV) pk[e] = ax_pack(p0, p1);
          }
          const int unit = wg * 4 + q4;
          *reinterpret_cast<uint4*>(ds_row + ((unit ^ (r & 7)) << 4)) = make_uint4(dk[0], dk[1], dk[2], dk[3]);
          if (KV) *reinterpret_cast<uint4*>(pw_row + ((unit ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      };
      if (col0 + 32 <= p.n_tok) columns(std::false_type{});
      else columns(std::true_type{});
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(&op_full[u]);
    }
    // ---- epilogue: accumulators -> bf16 rows of dqkv
    // (by the set that owns the last slab's buffers: only its threads have followed every phase of that op_free barrier)
    const int last = n_slabs - 1;
    if (AX_SETS == 1 || set == (last & 1)) {
    mbar_wait(&op_free[last & 1], (last >> 1) & 1);
    tc_fence_after();
    if (wg == 0 || KV) {                            // Q role: one accumulator, warpgroup 0 stores it
      float o[32];
      tmem_ld_32x32(tmem_base + lane_sel + AX_ACC + wg * 32, o);
      tmem_ld_wait();
      if (grow < p.n_tok) {
        // accA = dK (KV) | dQ (Q);  accB = dV
        const int off = KV ? (wg == 0 ? p.k_off : p.v_off) : p.q_off;
        uint16_t* op = p.dqkv + (size_t)(row_base + grow) * p.ld_qkv + off + h * 32;
#pragma unroll
        for (int d = 0; d < 32; d += 8)
          *reinterpret_cast<uint4*>(op + d) = make_uint4(ax_pack(o[d], o[d + 1]), ax_pack(o[d + 2], o[d + 3]),
                                                         ax_pack(o[d + 4], o[d + 5]), ax_pack(o[d + 6], o[d + 7]));
      }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == AX_TMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// D_i = sum_d dO_id O_id and the forward's lse2 -> {-lse2, -D * scale} of row i, stored PAIR-INTERLEAVED per (b, h):
// floats [(i >> 1) * 4 + (i & 1)] = -lse2_i and [(i >> 1) * 4 + 2 + (i & 1)] = -D_i * scale, so that the softmax stage reads
// the two 64-bit halves of one 16-byte word as ready-made packed-fp32 addends of a column pair (thread per (row, head))
__global__ void __launch_bounds__(256) attention_bwd_prep_kernel(const uint16_t* __restrict__ o, const uint16_t* __restrict__ dout,
                                                                 int o_ld, const float* __restrict__ lse2,
                                                                 float2* __restrict__ ld, int batch, int heads, int n_tok,
                                                                 int n_pad, float scale) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)batch * n_pad * heads;
  if (idx >= total) return;
  const int h = (int)(idx % heads);
  const int64_t t = idx / heads;
  const int i = (int)(t % n_pad), b = (int)(t / n_pad);
  float2 out = make_float2(0.f, 0.f);
  if (i < n_tok) {
    const size_t row = (size_t)b * n_tok + i;
    const uint4* a = reinterpret_cast<const uint4*>(o + row * o_ld + h * 32);
    const uint4* g = reinterpret_cast<const uint4*>(dout + row * o_ld + h * 32);
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint4 x = a[q], y = g[q];
      const uint32_t xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        acc = fmaf(__uint_as_float(xs[e] << 16), __uint_as_float(ys[e] << 16), acc);
        acc = fmaf(__uint_as_float(xs[e] & 0xffff0000u), __uint_as_float(ys[e] & 0xffff0000u), acc);
      }
    }
    out = make_float2(-lse2[((size_t)b * heads + h) * n_tok + i], -acc * scale);
  }
  float* dst = reinterpret_cast<float*>(ld + ((size_t)b * heads + h) * n_pad) + (size_t)(i >> 1) * 4 + (i & 1);
  dst[0] = out.x;
  dst[2] = out.y;
}

int attention_bwd_tc_launch(const void* qkv, int ld, int q_off, int k_off, int v_off, const void* out, const void* dout,
                            int o_ld, const float* lse2, void* dqkv, float* ws, int batch, int heads, int n_tokens,
                            float scale, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e1 = cudaFuncSetAttribute(attention_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AX_SMEM);
    cudaError_t e2 = cudaFuncSetAttribute(attention_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AX_SMEM);
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
      set_error("attention_bwd_tc: cudaFuncSetAttribute failed");
      return LDMB200_ERR_CUDA;
    }
    attr_set = true;
  }
  const int n_pad = (n_tokens + 63) / 64 * 64;
  float2* ldbuf = reinterpret_cast<float2*>(ws);
  const int64_t total = (int64_t)batch * n_pad * heads;
  attention_bwd_prep_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, st>>>((const uint16_t*)out, (const uint16_t*)dout, o_ld,
                                                                             lse2, ldbuf, batch, heads, n_tokens, n_pad, scale);
  CUtensorMap t128, t64, d128, d64;
  {
    uint64_t dims[2] = {(uint64_t)ld, (uint64_t)batch * n_tokens};
    uint64_t str[1] = {(uint64_t)ld * 2};
    uint32_t box128[2] = {32, 128}, box64[2] = {32, 64};
    int rc = make_tensor_map_bf16(&t128, qkv, 2, dims, str, box128, 64);
    if (rc) return rc;
    rc = make_tensor_map_bf16(&t64, qkv, 2, dims, str, box64, 64);
    if (rc) return rc;
    uint64_t ddims[2] = {(uint64_t)o_ld, (uint64_t)batch * n_tokens};
    uint64_t dstr[1] = {(uint64_t)o_ld * 2};
    rc = make_tensor_map_bf16(&d128, dout, 2, ddims, dstr, box128, 64);
    if (rc) return rc;
    rc = make_tensor_map_bf16(&d64, dout, 2, ddims, dstr, box64, 64);
    if (rc) return rc;
  }
  AxParams p;
  p.dqkv = (uint16_t*)dqkv;
  p.ld = ldbuf;
  p.ld_qkv = ld; p.q_off = q_off; p.k_off = k_off; p.v_off = v_off;
  p.n_tok = n_tokens; p.n_pad = n_pad; p.heads = heads;
  p.scale = scale; p.scale_log2e = scale * 1.4426950408889634f;
  dim3 grid(ceil_div(n_tokens, 128), heads, batch);
  attention_bwd_tc_kernel<true><<<grid, AX_THREADS, AX_SMEM, st>>>(t128, t64, d128, d64, p);
  attention_bwd_tc_kernel<false><<<grid, AX_THREADS, AX_SMEM, st>>>(t128, t64, d128, d64, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("attention_bwd_tc: launch failed: %s", cudaGetErrorString(e));
    return LDMB200_ERR_CUDA;
  }
  return 0;
}

}  // namespace ldm
