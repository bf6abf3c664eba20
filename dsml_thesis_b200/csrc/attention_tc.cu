// Fused flash-style self-attention for d_head = 32 on tcgen05 (K5): softmax(q k^T * scale) v per (sample, head)
// without materialising the [N, N] score matrix the reference builds (attention.py:180-191).
//
// The op is MUFU / TMEM-read bound, not tensor-bound: every score needs one ex2 (16/clk/SM) and 4 bytes of TMEM
// read (64 B/clk/SM) but only 128 tensor flops, so the design goal is to keep the exponent pipe busy all the time.
// One CTA per (384-query block, head, sample), one CTA per SM, three 128-query tiles per CTA (three softmax
// warpgroups = 3 warps per scheduler to hide the ex2 / TMEM latencies); keys advance in 64-key steps with S and P
// double-buffered per query tile, so the tensor pipe always works one step ahead of every softmax warpgroup.
//   warps 0-11  : softmax warpgroup q = warp / 4   (thread r <-> query row r <-> TMEM lane r)
//   warps 12-14 : MMA issuer of query tile q = warp - 12 (one elected lane) — one per scheduler, so no scheduler
//                 carries two service warps:  S_q(i) = Q_q K(i)^T (M128 N64 K32) into TMEM cols q*128 + (i&1)*64;
//                 O_q += P_q(i) V(i) (M128 N32 K64, V as MN-major B operand straight from the row-major tile) into
//                 TMEM cols 384 + 32q.  Per step i: wait p_full_q(i) -> issue P.V_q(i), S_q(i+2), commit s_full_q.
//   warp  15    : TMA producer — the Q tiles once, then K/V tiles (128 keys x 32) through a 3-stage ring, all from
//                 the fused QKV matrix [B*N, 3C] through one 2-D tensor map (64B-swizzled K-major rows of 32 x 16
//                 bit); also owns barrier initialisation and the TMEM allocation
// Softmax: ONE pass over S in TMEM per KV tile (TMEM read bandwidth, ~64 B/clk/SM, and the ex2 pipe bound the
// kernel): p = exp2(s*c - ref) against a per-row reference that is only re-based when a tile's row max outgrows it
// by more than 2^8 (any finite reference is exact after normalisation); the running output accumulates in TMEM
// across KV tiles (no per-tile read-back / rescale) and is re-based in place (tcgen05.ld/st) on that rare path.
// P is written as bf16 into a 128B-swizzled K-major smem tile for the PV MMA.
#include <algorithm>

#include "common.cuh"

namespace ldm {

constexpr int AT_NQ = 3;                      // 128-query tiles per CTA (softmax warpgroups)
constexpr int AT_THREADS = 32 * (5 * AT_NQ + 1);  // 12 softmax warps, one MMA warp per query tile, one TMA warp
constexpr int AT_MMA_WARP0 = 4 * AT_NQ, AT_TMA_WARP = 5 * AT_NQ;
constexpr int AT_O_COL = AT_NQ * 128;          // TMEM: S_q buffers at q*128 + u*64, O_q at AT_O_COL + q*32
constexpr int AT_STAGES = 3;
constexpr int AT_TILE_BYTES = 128 * 32 * 2;  // 8 KiB: 128 rows x 32 16-bit values
constexpr int AT_P_BYTES = 2 * 128 * 128;    // two K-major chunks of [128 x 64] bf16, 128B swizzle
constexpr int AT_SMEM = AT_NQ * AT_TILE_BYTES + 2 * AT_STAGES * AT_TILE_BYTES + AT_NQ * AT_P_BYTES + 32 * 8 + 16 + 1024;
constexpr int AT_TMEM_COLS = 512;
#ifndef LDM_AT_POLY_PAIRS
#define LDM_AT_POLY_PAIRS 0
#endif
// 0..4: share (x 25 %) of the exponentials emulated on the FMA pipe.  Measured r01s (tools/attn_sweep.py, B=64 N=4096):
// 0 -> 1.766 ms, 1 -> 1.831 ms, 2 -> 2.048 ms: the softmax warps are latency-bound (ex2 unit 70 % busy, issue slots
// 36 %), the extra ~8 FMA-pipe instructions per emulated exponential only lengthen the critical path, so the default is 0.
constexpr int AT_POLY_PAIRS = LDM_AT_POLY_PAIRS;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// exp2 on the FMA pipe (FlashAttention-4 style): the ex2 unit (16/clk/SM) is the co-bound of this kernel with the TMEM
// read port, so a fraction of the exponentials is emulated — round-to-nearest split x = n + f, |f| <= 0.5 (magic-number
// add), degree-3 minimax 2^f (max rel. error 7.5e-5, far below the bf16 rounding of P), exponent add of n.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -125.0f);
  const float t = x + 12582912.0f;                 // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const float f = x - (t - 12582912.0f);
  float p = fmaf(0.05517146f, f, 0.24261086f);
  p = fmaf(p, f, 0.69326099f);
  p = fmaf(p, f, 0.99992809f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float y;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(y) : "f"(a), "f"(b), "f"(c));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack_out16(float lo, float hi, int dt) {
  uint32_t r;
  if (dt == LDMB200_F16)
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  else
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// row max of one 32-column chunk on 4 independent chains; MASK: columns >= valid are ignored
template <bool MASK>
__device__ __forceinline__ void chunk_max(const float (&v)[32], float (&mx)[4], int col0, int valid) {
  if (MASK) {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (col0 + i < valid) mx[i & 3] = fmaxf(mx[i & 3], v[i]);
  } else {
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      mx[0] = fmax3(mx[0], v[i], v[i + 1]);
      mx[1] = fmax3(mx[1], v[i + 2], v[i + 3]);
      mx[2] = fmax3(mx[2], v[i + 4], v[i + 5]);
      mx[3] = fmax3(mx[3], v[i + 6], v[i + 7]);
    }
  }
}

// p = exp2(s*c - mb) for one 32-column chunk, row sum on 4 chains, bf16 pack and swizzled store of the
// thread's 64 bytes: chunk c of the row -> 64-key half (c>>1), 16-byte units (c&1)*4 .. +3
template <bool MASK>
__device__ __forceinline__ void chunk_exp_store(const float (&v)[32], float (&ls)[4], float c_, float mb, int c,
                                                int valid, uint8_t* sP, int r) {
  uint8_t* rowp = sP + (c >> 1) * (128 * 128) + r * 128;
#pragma unroll
  for (int u = 0; u < 4; ++u) {  // 8 columns -> one 16-byte unit
    uint32_t pk[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = u * 8 + e * 2;
      // AT_POLY_PAIRS of every 4 column pairs go to the FMA pipe, the rest to the ex2 unit
      float p0, p1;
      if (e >= 4 - AT_POLY_PAIRS) {
        p0 = ex2_poly(fmaf(v[i], c_, -mb));
        p1 = ex2_poly(fmaf(v[i + 1], c_, -mb));
      } else {
        p0 = ex2_approx(fmaf(v[i], c_, -mb));
        p1 = ex2_approx(fmaf(v[i + 1], c_, -mb));
      }
      if (MASK) {
        if (c * 32 + i >= valid) p0 = 0.f;
        if (c * 32 + i + 1 >= valid) p1 = 0.f;
      }
      ls[e] += p0 + p1;
      pk[e] = pack_bf16x2(p0, p1);
    }
    const int unit = (c & 1) * 4 + u;
    *reinterpret_cast<uint4*>(rowp + ((unit ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

// tcgen05.wait::ld that also names the destination registers of the load it completes, so the compiler cannot
// schedule a use of them above the wait (the asynchronous write is invisible to it otherwise)
#define LDM_R8(v, o) "+f"(v[o]), "+f"(v[o + 1]), "+f"(v[o + 2]), "+f"(v[o + 3]), "+f"(v[o + 4]), "+f"(v[o + 5]), "+f"(v[o + 6]), "+f"(v[o + 7])
__device__ __forceinline__ void tmem_wait_regs(float (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" : LDM_R8(v, 0), LDM_R8(v, 8), LDM_R8(v, 16), LDM_R8(v, 24)::"memory");
}
__device__ __forceinline__ void reg_fence(float (&v)[32]) {
  asm volatile("" : LDM_R8(v, 0), LDM_R8(v, 8), LDM_R8(v, 16), LDM_R8(v, 24)::"memory");
}

// One 64-key step for query row r, ONE pass over S (64 fp32 columns at tS; TMEM reads are the scarce resource):
// p = exp2(s*c - mb) against the row's current reference mb (= m_ref*c), the step's row max and row sum on the
// side, P (bf16) into the swizzled smem tile.  FIRST: the reference is not known yet — take the max of the
// first 32 keys (any finite reference is exact after normalisation; the caller re-references on overflow risk).
template <bool MASK, bool FIRST>
__device__ __forceinline__ void score_pass(uint32_t tS, int valid, float c_, float& mb, float& tmax, float& lsum,
                                           uint8_t* sP, int r) {
  float a[32], b[32];
  float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
  float ls[4] = {0.f, 0.f, 0.f, 0.f};
  tmem_ld_32x32(tS, a);
  tmem_ld_32x32(tS + 32, b);
  tmem_wait_regs(a);
  reg_fence(b);
  chunk_max<MASK>(a, mx, 0, valid);
  if (FIRST) mb = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * c_;
  chunk_exp_store<MASK>(a, ls, c_, mb, 0, valid, sP, r);
  chunk_max<MASK>(b, mx, 32, valid);
  chunk_exp_store<MASK>(b, ls, c_, mb, 1, valid, sP, r);
  tmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
  lsum = (ls[0] + ls[1]) + (ls[2] + ls[3]);
}

// 32 fp32 columns of this thread's TMEM lane <- registers (the rare re-referencing of the running output)
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// how far (in exp2 units) a tile's row max may exceed the row's reference before the row is re-referenced:
// p <= 2^8 per score, sums stay far inside the fp32 / bf16 exponent range
constexpr float AT_REREF = 8.0f;

__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm, uint16_t* __restrict__ out, int out_ld, int out_dtype,
                    int n_tok, int q_off, int k_off, int v_off, float scale_log2e, float* __restrict__ lse2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;                               // 2 tiles (256 rows), one TMA box
  uint8_t* sK = sQ + AT_NQ * AT_TILE_BYTES;             // AT_STAGES tiles
  uint8_t* sV = sK + AT_STAGES * AT_TILE_BYTES;     // AT_STAGES tiles
  uint8_t* sP = sV + AT_STAGES * AT_TILE_BYTES;     // 2 x 32 KiB, 1024-aligned (8 x 8 KiB offset)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + AT_NQ * AT_P_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;                     // [AT_STAGES]
  uint64_t* kv_empty = kv_full + AT_STAGES;         // [AT_STAGES]
  uint64_t* s_full = kv_empty + AT_STAGES;          // [2 q][2 buf]  S_q(i) ready (and every MMA issued before it)
  uint64_t* p_full = s_full + 2 * AT_NQ;                    // [2 q][2 buf]  P_q(i) written, S_q(i) consumed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_full + 2 * AT_NQ);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int n_kv = (n_tok + 127) / 128;      // 128-key K/V tiles (TMA stages)
  const int n_steps = (n_tok + 63) / 64;     // 64-key softmax / MMA steps: S and P are double-buffered per query tile
  const int row_base = b * n_tok;  // first row of this sample in the [B*N, ld] matrix

  // query tiles of this CTA that hold real rows (the last CTA of a sample may have fewer than AT_NQ)
  const int nq = min(AT_NQ, (n_tok - qb * (AT_NQ * 128) + 127) / 128);
  if (warp == AT_TMA_WARP) {
    if (lane == 0) {
      prefetch_tmap(&tm);
      mbar_init(q_full, 1);
      for (int s = 0; s < AT_STAGES; ++s) {
        mbar_init(&kv_full[s], 1);
        mbar_init(&kv_empty[s], nq);  // one tcgen05.commit per active query tile
      }
      for (int q = 0; q < 2 * AT_NQ; ++q) {
        mbar_init(&s_full[q], 1);
        mbar_init(&p_full[q], 128);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, AT_TMEM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == AT_TMA_WARP) {
    // ===================== TMA producer (warp converged, one elected lane issues) =====================
    if (elect_one()) {
      mbar_expect_tx(q_full, nq * AT_TILE_BYTES);
      for (int q = 0; q < nq; ++q)
        tma_load_2d(sQ + q * AT_TILE_BYTES, &tm, q_full, q_off + h * 32, row_base + qb * (AT_NQ * 128) + q * 128);
    }
    __syncwarp();
    for (int j = 0; j < n_kv; ++j) {
      const int s = j % AT_STAGES;
      mbar_wait(&kv_empty[s], ((j / AT_STAGES) & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(&kv_full[s], 2 * AT_TILE_BYTES);
        tma_load_2d(sK + s * AT_TILE_BYTES, &tm, &kv_full[s], k_off + h * 32, row_base + j * 128);
        tma_load_2d(sV + s * AT_TILE_BYTES, &tm, &kv_full[s], v_off + h * 32, row_base + j * 128);
      }
      __syncwarp();
    }
  } else if (warp >= AT_MMA_WARP0) {
    // ===================== MMA issuers: warp AT_MMA_WARP0 + q drives query tile q (converged, one elected lane) ===
    // One warp per tile: a single MMA warp serialised the three tiles behind its in-order waits and, sharing its
    // scheduler with three busy softmax warps, became the bottleneck (profiles/ r01g); with one issuer per
    // scheduler the load is balanced.  Smem descriptors: constant high word (SBO, version, swizzle mode), low
    // word = (address >> 4) | LBO.
    const int q = warp - AT_MMA_WARP0;
    if (q < nq) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 64, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 32, 0, 1);          // B = V is MN-major
      constexpr uint32_t hi_sw64 = (512u >> 4) | (1u << 14) | (4u << 29);    // Q, K (K-major) and V (MN-major) tiles
      constexpr uint32_t hi_sw128 = (1024u >> 4) | (1u << 14) | (2u << 29);  // P tile
      const uint32_t q_lo = (((smem_u32(sQ) & 0x3FFFFu) >> 4) | (1u << 16)) + q * (AT_TILE_BYTES >> 4);
      const uint32_t k_lo = ((smem_u32(sK) & 0x3FFFFu) >> 4) | (1u << 16);
      const uint32_t v_lo = ((smem_u32(sV) & 0x3FFFFu) >> 4) | (1u << 16);
      const uint32_t p_lo = (((smem_u32(sP) & 0x3FFFFu) >> 4) | (1u << 16)) + q * (AT_P_BYTES >> 4);
      const uint32_t dS = tmem_base + q * 128, dO = tmem_base + AT_O_COL + q * 32;
      // S_q(i) = Q_q . K(keys 64 i ..)^T -> TMEM columns q*128 + (i&1)*64;  st = smem stage of K/V tile i >> 1
      auto issue_s = [&](int i, int st) {
        const int hf = i & 1;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
          umma_bf16_ss(dS + hf * 64, ((uint64_t)hi_sw64 << 32) | (q_lo + kk * 2),
                       ((uint64_t)hi_sw64 << 32) | (k_lo + st * (AT_TILE_BYTES >> 4) + hf * (4096 >> 4) + kk * 2),
                       idesc_qk, kk > 0);
      };
      // O_q += P_q(i) . V(keys 64 i ..)   (the first step overwrites)
      auto issue_pv = [&](int i, int st) {
        const int hf = i & 1;
#pragma unroll
        for (int t = 0; t < 4; ++t)
          umma_bf16_ss(dO, ((uint64_t)hi_sw128 << 32) | (p_lo + hf * ((128 * 128) >> 4) + t * 2),
                       ((uint64_t)hi_sw64 << 32) | (v_lo + st * (AT_TILE_BYTES >> 4) + (hf * 4 + t) * (1024 >> 4)),
                       idesc_pv, (i > 0 || t > 0) ? 1u : 0u);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      if (elect_one()) {
        issue_s(0, 0);
        umma_commit(&s_full[q * 2 + 0]);
        if (n_steps > 1) {
          issue_s(1, 0);
          umma_commit(&s_full[q * 2 + 1]);
        }
      }
      __syncwarp();
      int st = 0;  // smem stage of K/V tile i >> 1
      for (int i = 0; i < n_steps; ++i) {
        const int u = i & 1;
        const int st_next = (st + 1 == AT_STAGES) ? 0 : st + 1;
        const bool more = i + 2 < n_steps;
        mbar_wait(&p_full[q * 2 + u], (i >> 1) & 1);  // all 128 rows of P_q(i) are in smem, S_q(i) is consumed
        tc_fence_after();
        if (elect_one()) issue_pv(i, st);
        __syncwarp();
        if (more && u == 0) {  // S_q(i+2) is the first use of K/V tile (i + 2) >> 1
          const int jn = (i + 2) >> 1;
          mbar_wait(&kv_full[st_next], (jn / AT_STAGES) & 1);
          tc_fence_after();
        }
        if (elect_one()) {
          if (u == 1 || i == n_steps - 1) umma_commit(&kv_empty[st]);  // last read of K/V tile i >> 1 by this query tile
          if (more) issue_s(i + 2, st_next);
          umma_commit(&s_full[q * 2 + u]);  // always: it also tells the softmax warps that P.V(i) has landed
        }
        __syncwarp();
        if (u == 1) st = st_next;
      }
    }
  } else {
    // ===================== softmax warpgroups: thread r <-> query row r of tile q =====================
    const int q = warp >> 2;
    const int r = threadIdx.x & 127;  // == TMEM lane
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + lane_sel + q * 128, tO = tmem_base + lane_sel + AT_O_COL + q * 32;
    uint8_t* myP = sP + q * AT_P_BYTES;
    float mb = 0.f, l = 0.f;  // mb = (row reference) * c: the running output and l are relative to 2^mb
    const int last_valid = n_tok - (n_steps - 1) * 64;  // keys in the last step (1..64)
    const int my_steps = q < nq ? n_steps : 0;  // a query tile past the end of the sample does nothing
    for (int i = 0; i < my_steps; ++i) {
      const int u = i & 1;
      const bool masked = (i == n_steps - 1) && last_valid < 64;
      const uint32_t tSu = tS + u * 64;
      uint8_t* Pu = myP + u * (128 * 128);
      mbar_wait(&s_full[q * 2 + u], (i >> 1) & 1);  // S_q(i) complete (and P_q.V(i-2): the P buffer is free again)
      tc_fence_after();
      float tmax, lsum;
      if (i == 0) {
        if (masked) score_pass<true, true>(tSu, last_valid, scale_log2e, mb, tmax, lsum, Pu, r);
        else score_pass<false, true>(tSu, 64, scale_log2e, mb, tmax, lsum, Pu, r);
      } else {
        if (masked) score_pass<true, false>(tSu, last_valid, scale_log2e, mb, tmax, lsum, Pu, r);
        else score_pass<false, false>(tSu, 64, scale_log2e, mb, tmax, lsum, Pu, r);
      }
      // lazy re-referencing: only when some row's scores outgrew its reference by more than 2^AT_REREF
      const bool need = fmaf(tmax, scale_log2e, -mb) > AT_REREF;
      if (__any_sync(0xffffffffu, need)) {
        const float mb_new = need ? tmax * scale_log2e : mb;
        const float f = ex2_approx(mb - mb_new);  // 1 for the rows that keep their reference
        if (i > 0) {
          // the output is rescaled in place: P.V(i-1) must have landed first (its commit is s_full of step i+1,
          // issued even when there is no such step); P.V(i) cannot start before this thread arrives on p_full
          mbar_wait(&s_full[q * 2 + (u ^ 1)], (((i - 1) >> 1) + 1) & 1);
          tc_fence_after();
          float t[32];
          tmem_ld_32x32(tO, t);
          tmem_wait_regs(t);
#pragma unroll
          for (int d = 0; d < 32; ++d) t[d] *= f;
          tmem_st_32x32(tO, t);
        }
        l *= f;
        mb = mb_new;
        if (masked) score_pass<true, false>(tSu, last_valid, scale_log2e, mb, tmax, lsum, Pu, r);
        else score_pass<false, false>(tSu, 64, scale_log2e, mb, tmax, lsum, Pu, r);
      }
      l += lsum;
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the UMMA (async proxy)
      tc_fence_before();
      mbar_arrive(&p_full[q * 2 + u]);
    }
    if (q < nq) {
      mbar_wait(&s_full[q * 2 + ((n_steps - 1) & 1)], (((n_steps - 1) >> 1) + 1) & 1);  // the commit after the last P.V
      tc_fence_after();
      float o[32];
      tmem_ld_32x32(tO, o);
      tmem_wait_regs(o);
      const float inv = 1.0f / l;
      const int qi = qb * (AT_NQ * 128) + q * 128 + r;
      if (qi < n_tok) {
        // training: log2 of the softmax denominator in the exp2 domain (scores scaled by scale*log2 e)
        if (lse2) lse2[((size_t)b * gridDim.y + h) * n_tok + qi] = mb + log2f(l);
        uint16_t* op = out + (size_t)(row_base + qi) * out_ld + h * 32;
#pragma unroll
        for (int d = 0; d < 32; d += 8) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) w[e] = pack_out16(o[d + 2 * e] * inv, o[d + 2 * e + 1] * inv, out_dtype);
          *reinterpret_cast<uint4*>(op + d) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == AT_TMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AT_TMEM_COLS);
  }
}

}  // namespace ldm

using namespace ldm;

static int attention_tc_launch(const void* qkv, int ld, int q_off, int k_off, int v_off, void* out, int out_ld,
                               int out_dtype, float* lse2, int batch, int heads, int n_tokens, float scale, void* stream) {
  LDM_CHECK_ARG(qkv && out && batch > 0 && heads > 0 && n_tokens > 0, "attention_tc: bad arguments");
  LDM_CHECK_ARG(ld % 8 == 0 && out_ld % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
                "attention_tc: leading dims / offsets must be multiples of 8 elements (16 bytes)");
  LDM_CHECK_ARG(out_dtype == LDMB200_BF16 || out_dtype == LDMB200_F16, "attention_tc: out_dtype must be bf16 or f16");
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
  dim3 grid(ceil_div(n_tokens, AT_NQ * 128), heads, batch);
  const float scale_log2e = scale * 1.4426950408889634f;
  attention_tc_kernel<<<grid, AT_THREADS, AT_SMEM, (cudaStream_t)stream>>>(tm, (uint16_t*)out, out_ld, out_dtype,
                                                                           n_tokens, q_off, k_off, v_off, scale_log2e, lse2);
  LDM_CHECK_LAUNCH("attention_tc");
  return 0;
}

extern "C" int ldmb200_attention_tc(const void* qkv, int ld, int q_off, int k_off, int v_off, void* out, int out_ld,
                                    int out_dtype, int batch, int heads, int n_tokens, float scale, void* stream) {
  return attention_tc_launch(qkv, ld, q_off, k_off, v_off, out, out_ld, out_dtype, nullptr, batch, heads, n_tokens, scale,
                             stream);
}

extern "C" int ldmb200_attention_tc_lse(const void* qkv, int ld, int q_off, int k_off, int v_off, void* out, int out_ld,
                                        int out_dtype, float* lse2, int batch, int heads, int n_tokens, float scale,
                                        void* stream) {
  LDM_CHECK_ARG(lse2, "attention_tc_lse: lse2 is NULL");
  return attention_tc_launch(qkv, ld, q_off, k_off, v_off, out, out_ld, out_dtype, lse2, batch, heads, n_tokens, scale,
                             stream);
}
