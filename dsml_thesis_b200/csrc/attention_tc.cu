// Fused flash-style self-attention for d_head = 32 on tcgen05 (K5): softmax(q k^T * scale) v per (sample, head)
// without materialising the [N, N] score matrix the reference builds (attention.py:180-191).
//
// The op is MUFU / issue-slot bound, not tensor-bound: every score needs one ex2 (16/clk/SM) but only 128 tensor
// flops, so the design goal is the fewest issue slots per score next to the exponential.
// One CTA per (384-query block, head, sample), one CTA per SM, three 128-query tiles per CTA (three softmax
// warpgroups = 3 warps per scheduler to hide the ex2 / TMEM latencies); keys advance in 64-key steps with S
// double-buffered per query tile, so the tensor pipe always works one step ahead of every softmax warpgroup.
//   warps 0-11  : softmax warpgroup q = warp / 4   (thread r <-> query row r <-> TMEM lane r)
//   warps 12-14 : MMA issuer of query tile q = warp - 12 (one elected lane) — one per scheduler, so no scheduler
//                 carries two service warps:  S_q(i) = Q_q K(i)^T (M128 N64 K32) into TMEM cols q*128 + (i&1)*64;
//                 O_q += P_q(i) V(i) (M128 N32 K64; A = P straight from TENSOR MEMORY, V as MN-major B operand from
//                 the row-major smem tile) into TMEM cols 384 + 32q.  Per step i: wait p_full_q(i) -> issue
//                 P.V_q(i), S_q(i+2), commit s_full_q.
//   warp  15    : TMA producer — the Q tiles once, then K/V tiles (128 keys x 32) through a 3-stage ring, all from
//                 the fused QKV matrix [B*N, 3C] through one 2-D tensor map (64B-swizzled K-major rows of 32 x 16
//                 bit); also owns barrier initialisation and the TMEM allocation
// Softmax: ONE pass over S in TMEM per 64-key step: p = exp2(s*c - ref) against a per-row reference that is only
// re-based when a p outgrows 2^8 (any finite reference is exact after normalisation); the running output
// accumulates in TMEM across KV tiles (no per-tile read-back / rescale) and is re-based in place (tcgen05.ld/st) on
// that rare path.  Issue-slot diet of the inner loop (r02): s*c - ref and the row sum run on the packed fp32 pipe
// (FFMA2 / FADD2: two scores per instruction); the per-step row max is NOT computed on the fast path — a step whose
// row sum stays <= 2^8 cannot hold a p above 2^8, so the exact max is only taken when that gate trips; P (bf16) is
// written back over the first 32 columns of its own S buffer with one tcgen05.st and consumed by the PV MMA as its
// TMEM A operand, so there is no shared-memory restage, no swizzle arithmetic and no async-proxy fence.
#include <algorithm>

#include "common.cuh"

namespace ldm {

constexpr int AT_NQ = 3;                      // 128-query tiles per CTA (softmax warpgroups)
constexpr int AT_THREADS = 32 * (5 * AT_NQ + 1);  // 12 softmax warps, one MMA warp per query tile, one TMA warp
constexpr int AT_MMA_WARP0 = 4 * AT_NQ, AT_TMA_WARP = 5 * AT_NQ;
constexpr int AT_O_COL = AT_NQ * 128;          // TMEM: S_q buffers at q*128 + u*64 (P_q(i) over cols 0..31), O_q at AT_O_COL + q*32
#ifndef LDM_AT_STAGES
#define LDM_AT_STAGES 4
#endif
constexpr int AT_STAGES = LDM_AT_STAGES;   // K/V tile ring depth (128 keys x 32 each)
constexpr int AT_TILE_BYTES = 128 * 32 * 2;  // 8 KiB: 128 rows x 32 16-bit values
constexpr int AT_SMEM = AT_NQ * AT_TILE_BYTES + 2 * AT_STAGES * AT_TILE_BYTES + 32 * 8 + 16 + 1024;
constexpr int AT_TMEM_COLS = 512;
#ifndef LDM_AT_POLY
#define LDM_AT_POLY 2
#endif
// 0..8: of every 8 score pairs, this many take their exponentials on the FMA pipe (packed degree-3 polynomial)
// instead of the ex2 unit.  Measured r02a (tools/attn_sweep.py, B=64 N=4096 h=5): 0 -> 1.606 ms, 2 -> 1.393, 3 -> 1.425,
// 4 -> 1.634 (r01 kernel: 1.767).  A software-pipelined variant (next step's tcgen05.ld issued before the P store) was
// measured r02c and rejected: 1.635 ms at 2/8 — the loop-carried register arrays push the kernel into spills at the
// 128-register cap of a 512-thread CTA, and setmaxnreg re-distribution made ptxas spill more, not less.
constexpr int AT_POLY = LDM_AT_POLY;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
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

// exp2 of a score pair on the FMA pipe (FlashAttention-4 style): the ex2 unit (16/clk/SM) bounds this kernel, so a
// share of the exponentials is emulated — round-to-nearest split x = n + f, |f| <= 0.5 (magic-number add), degree-3
// minimax 2^f (max rel. error 7.5e-5, far below the bf16 rounding of P), exponent add of n.  x is clamped from below
// (2^-125 ~ 0); the caller guards the upper side (x > 126 would wrap the exponent) through `gmax`.
__device__ __forceinline__ void ex2_poly2(f32x2 x2, float& p0, float& p1, float& gmax) {
  float x0, x1;
  unpk2(x2, x0, x1);
  x0 = fmaxf(x0, -125.0f);
  x1 = fmaxf(x1, -125.0f);
  gmax = fmax3(gmax, x0, x1);
  x2 = pk2(x0, x1);
  const f32x2 magic = pk2(12582912.0f, 12582912.0f);   // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const f32x2 t2 = add2(x2, magic);
  const f32x2 f2 = sub2(x2, sub2(t2, magic));
  f32x2 q2 = fma2(pk2(0.05517146f, 0.05517146f), f2, pk2(0.24261086f, 0.24261086f));
  q2 = fma2(q2, f2, pk2(0.69326099f, 0.69326099f));
  q2 = fma2(q2, f2, pk2(0.99992809f, 0.99992809f));
  float q0, q1, t0, t1;
  unpk2(q2, q0, q1);
  unpk2(t2, t0, t1);
  p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
  p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
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

// p = exp2(s*c - mb) for one 32-column chunk: packed FFMA2 for the argument, ex2 unit (or FMA-pipe polynomial for
// AT_POLY of every 8 pairs), packed row sum on two chains, bf16 pack into 16 registers.  MASK: columns >= valid -> 0.
template <bool MASK>
__device__ __forceinline__ void chunk_exp(const float (&v)[32], f32x2 (&ls)[2], f32x2 c2, f32x2 nmb2, int col0, int valid,
                                          uint32_t* pk, float& gmax) {
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const f32x2 x2 = fma2(pk2(v[2 * j], v[2 * j + 1]), c2, nmb2);
    float p0, p1;
    if ((j & 7) < AT_POLY) {
      ex2_poly2(x2, p0, p1, gmax);
    } else {
      float x0, x1;
      unpk2(x2, x0, x1);
      p0 = ex2_approx(x0);
      p1 = ex2_approx(x1);
    }
    if (MASK) {
      if (col0 + 2 * j >= valid) p0 = 0.f;
      if (col0 + 2 * j + 1 >= valid) p1 = 0.f;
    }
    ls[j & 1] = add2(ls[j & 1], pk2(p0, p1));
    pk[j] = pack_bf16x2(p0, p1);
  }
}

// exact row max of one 64-key step (the slow path and the first step's reference); MASK: columns >= valid ignored
template <bool MASK>
__device__ __forceinline__ float chunk_max32(const float (&v)[32], int col0, int valid) {
  float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
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
  return fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
}
template <bool MASK>
__device__ __noinline__ float step_rowmax(uint32_t tS, int valid) {
  float a[32], b[32];
  tmem_ld_32x32(tS, a);
  tmem_ld_32x32(tS + 32, b);
  tmem_wait_regs(a);
  reg_fence(b);
  return fmaxf(chunk_max32<MASK>(a, 0, valid), chunk_max32<MASK>(b, 32, valid));
}

// One 64-key step for query row r, ONE pass over S (64 fp32 columns at tS): p = exp2(s*c - mb) against the row's
// current reference mb (= m_ref*c), the step's row sum on the side, P (bf16 pairs) into pk[32] (stored to TMEM by the
// caller once the overflow gate has passed).  FIRST: the reference is not known yet — take the max of the first 32
// keys (any finite reference is exact after normalisation; the caller re-references on overflow risk).
// gmax: largest exponent argument seen by the FMA-pipe exponentials (their upper-range guard).
template <bool MASK, bool FIRST>
__device__ __forceinline__ void score_pass(uint32_t tS, int valid, float c_, float& mb, float& lsum, float& gmax,
                                           uint32_t (&pk)[32]) {
  float a[32], b[32];
  tmem_ld_32x32(tS, a);
  tmem_ld_32x32(tS + 32, b);
  tmem_wait_regs(a);
  reg_fence(b);
  if (FIRST) mb = chunk_max32<MASK>(a, 0, valid) * c_;
  const f32x2 c2 = pk2(c_, c_), nmb2 = pk2(-mb, -mb);
  f32x2 ls[2] = {0ull, 0ull};
  gmax = -INFINITY;
  chunk_exp<MASK>(a, ls, c2, nmb2, 0, valid, pk, gmax);
  chunk_exp<MASK>(b, ls, c2, nmb2, 32, valid, pk + 16, gmax);
  float s0, s1;
  unpk2(add2(ls[0], ls[1]), s0, s1);
  lsum = s0 + s1;
}

// 32 fp32 columns of this thread's TMEM lane <- registers (the rare re-referencing of the running output)
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
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

// D[tmem] (+)= A[tmem] * B[smem desc]: A = bf16 pairs, lane = row, one 32-bit column per two K elements (K-major)
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// how far (in exp2 units) a tile's row max may exceed the row's reference before the row is re-referenced:
// p <= 2^8 per score, sums stay far inside the fp32 / bf16 exponent range
constexpr float AT_REREF = 8.0f;
constexpr float AT_GATE = 256.0f;           // 2^AT_REREF

__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm, uint16_t* __restrict__ out, int out_ld, int out_dtype,
                    int n_tok, int q_off, int k_off, int v_off, float scale_log2e, float* __restrict__ lse2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // (offset, not uintptr_t: keeps the shared address space)
  uint8_t* sQ = smem;                               // AT_NQ tiles of 128 query rows
  uint8_t* sK = sQ + AT_NQ * AT_TILE_BYTES;             // AT_STAGES tiles
  uint8_t* sV = sK + AT_STAGES * AT_TILE_BYTES;     // AT_STAGES tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + AT_STAGES * AT_TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;                     // [AT_STAGES]
  uint64_t* kv_empty = kv_full + AT_STAGES;         // [AT_STAGES]
  uint64_t* s_full = kv_empty + AT_STAGES;          // [q][2 buf]  S_q(i) ready (and every MMA issued before it)
  uint64_t* p_full = s_full + 2 * AT_NQ;                    // [q][2 buf]  P_q(i) stored over S_q(i) in TMEM
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
  pdl_wait();  // q / k / v of the predecessor are visible from here (barriers and TMEM were set up under its tail)

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
      const uint32_t q_lo = (((smem_u32(sQ) & 0x3FFFFu) >> 4) | (1u << 16)) + q * (AT_TILE_BYTES >> 4);
      const uint32_t k_lo = ((smem_u32(sK) & 0x3FFFFu) >> 4) | (1u << 16);
      const uint32_t v_lo = ((smem_u32(sV) & 0x3FFFFu) >> 4) | (1u << 16);
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
      // O_q += P_q(i) . V(keys 64 i ..)   (the first step overwrites); P_q(i) = bf16 pairs in the first 32 columns of
      // the S buffer it was computed from: 16 keys (one K step) = 8 columns
      auto issue_pv = [&](int i, int st) {
        const int hf = i & 1;
#pragma unroll
        for (int t = 0; t < 4; ++t)
          umma_bf16_ts(dO, dS + hf * 64 + t * 8,
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
        mbar_wait(&p_full[q * 2 + u], (i >> 1) & 1);  // all 128 rows of P_q(i) are in TMEM, S_q(i) is consumed
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
    float mb = 0.f, l = 0.f;  // mb = (row reference) * c: the running output and l are relative to 2^mb
    const int last_valid = n_tok - (n_steps - 1) * 64;  // keys in the last step (1..64)
    const int my_steps = q < nq ? n_steps : 0;  // a query tile past the end of the sample does nothing
    for (int i = 0; i < my_steps; ++i) {
      const int u = i & 1;
      const bool masked = (i == n_steps - 1) && last_valid < 64;
      const uint32_t tSu = tS + u * 64;
      mbar_wait(&s_full[q * 2 + u], (i >> 1) & 1);  // S_q(i) complete (and P_q.V(i-2): the buffer's P half is free again)
      tc_fence_after();
      float lsum, gmax;
      uint32_t pk[32];
      if (i == 0) {
        if (masked) score_pass<true, true>(tSu, last_valid, scale_log2e, mb, lsum, gmax, pk);
        else score_pass<false, true>(tSu, 64, scale_log2e, mb, lsum, gmax, pk);
      } else {
        if (masked) score_pass<true, false>(tSu, last_valid, scale_log2e, mb, lsum, gmax, pk);
        else score_pass<false, false>(tSu, 64, scale_log2e, mb, lsum, gmax, pk);
      }
      // overflow gate: every p is positive, so a row sum <= 2^AT_REREF bounds every p of the step; the comparison is
      // written so that inf / NaN sums trip it.  FMA-pipe exponentials wrap above 2^126: their largest argument is
      // tracked separately.  Only when the gate trips is the exact row max taken (S is still intact in TMEM).
      const bool gate = !(lsum <= AT_GATE) || (AT_POLY > 0 && gmax > 100.0f);
      if (__any_sync(0xffffffffu, gate)) {
        const float tmax = masked ? step_rowmax<true>(tSu, last_valid) : step_rowmax<false>(tSu, 64);
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
            tmem_st_32x32(tO, reinterpret_cast<const uint32_t*>(t));
          }
          l *= f;
          mb = mb_new;
          if (masked) score_pass<true, false>(tSu, last_valid, scale_log2e, mb, lsum, gmax, pk);
          else score_pass<false, false>(tSu, 64, scale_log2e, mb, lsum, gmax, pk);
        }
      }
      l += lsum;
      tmem_st_32x32(tSu, pk);  // P_q(i) over the first 32 columns of S_q(i) (incl. tcgen05.wait::st)
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

  pdl_launch_dependents();
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
  cudaError_t le = launch_pdl(attention_tc_kernel, grid, dim3(AT_THREADS), AT_SMEM, (cudaStream_t)stream, tm, (uint16_t*)out, out_ld,
                              out_dtype, n_tokens, q_off, k_off, v_off, scale_log2e, lse2);
  if (le != cudaSuccess) {
    set_error("attention_tc: launch failed: %s", cudaGetErrorString(le));
    return LDMB200_ERR_CUDA;
  }
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
