// tcgen05 / TMEM / TMA implicit GEMM for the convolutions and linears of the UNet (K1-K4, K8, K11).
//
//   acc[m, n] = sum_{tap, c} A(m, tap, c) * W[n, tap, c]       bf16 x bf16 -> fp32 in tensor memory
//
// * persistent CTAs (one per SM), static round-robin over (m-tile, n-tile); tile = 128 x BN
// * warp 0: TMA producer   — A tiles come straight from the channels-last activation map through a 5-D
//           tensor map {C, W, H, plane, B}: a 3x3 tap is a coordinate shift, the padding halo is TMA
//           out-of-bounds zero fill, stride 2 reads the space-to-depth planes, the virtual channel concat
//           (K11) is a second tensor map; B tiles from the packed K-major weight matrix
// * warp 1: MMA issuer     — tcgen05.mma cta_group::1 kind::f16, M=128, N=BN, K=16 per instruction,
//           accumulators double-buffered in TMEM (2 x BN columns) so the epilogue overlaps the next tile
// * warps 2-5: epilogue    — tcgen05.ld 32x32b, fused bias / per-sample row vector (K8) / residual / GEGLU,
//           transposed through shared memory so global loads and stores are 128-byte coalesced
// smem operand tiles are K-major with the hardware 64B (BK=32) or 128B (BK=64) swizzle, written by TMA and
// read by the UMMA descriptors (cute/arch/mma_sm100_desc.hpp bit layout).
#include <algorithm>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

#ifndef LDM_TC_RES_PREFETCH
#define LDM_TC_RES_PREFETCH 1
#endif
#ifndef LDM_TC_RUNCNT
#define LDM_TC_RUNCNT 1   // (r02s: -1.7 % on the family, -3..-6 % on the deep convolutions: the per-chunk instruction count of the two service warps is a limiter) running stage / phase counters in the producer and MMA loops instead of it % STAGES, it / STAGES
#endif
#ifndef LDM_TC_SPLITP
#define LDM_TC_SPLITP 0   // CTA-pair kernels: a second producer warp issues the B (weight) loads, warp 0 the A loads only.  Measured r02s
                          // (same run, after the running counters): implicit GEMMs 12.47 vs 12.49 ms per eval - the concat convolutions gain
                          // 2-4 %, the rest is flat: off
#endif
#ifndef LDM_TC_GG_PIPE
#define LDM_TC_GG_PIPE 1   // software-pipelined GEGLU epilogue (halves of the chunk; see the epilogue branch)
#endif

namespace ldm {

struct TcParams {
  int M;
  int tiles_m, tiles_n;
  int taps, ks, stride, is_conv;
  int pad_asym;  // stride 2 over an input padded (0,1,0,1): input row = 2*oh + kh instead of 2*oh + kh - 1
  int c1_chunks, c2_chunks, ctot, c1;
  int c1_tail_k, c2_tail_k;  // MMAs (K = 16 each) of the last chunk of each channel range
  int hw_out, w_out, rows_per_sample;
  const float* bias;
  const float* rowvec;
  int rowvec_ld;
  int rowvec_uniform;  // rows_per_sample % 128 == 0: every row of a tile belongs to one sample
  const float* residual;
  int residual_ld;
  int geglu;
  void* out;
  int out_dtype, out_ld, out_nchw, n_valid;
  float* colstats;  // optional [M/32][n_valid][2]: per 32-row block column sums / sums of squares of the stored values
  // K10 folded into the conv (nearest-2x upsample + 3x3 = four 2x2 convolutions on the LOW-RES map, one per output
  // parity, with tap-summed weights): the GEMM's M space is the low-res pixel grid, row m = (n, y, x) is stored at
  // output pixel (n, 2y + up_py, 2x + up_px).  up_w == 0: off.
  int up_w, up_hw, up_py, up_px;
  int b_row0;  // first row of this launch's weight set inside the B tensor map
  // split-K for grids that would leave most SMs idle (small M, deep K): `splits` CTAs share one output tile, each reduces
  // kiters_per_split of the tile's taps x channel-chunk iterations and leaves its fp32 partial in sk_part; the CTA that
  // arrives last at the tile's counter folds the partials in split order (deterministic) and runs the fused epilogue.
  int splits, kiters_per_split;
  float* sk_part;          // [tile][split][chunk][CH/4][128 rows] float4
  unsigned int* sk_cnt;    // [tile] arrival counters: zero before the launch, reset to zero by the last arrival
  // weights of a LATER launch to pull into L2 while this one runs (ldmb200_gemm_desc.prefetch): at batch 1-4 a GEMM's
  // 2-15 MB of weights are read once, cold, by 8-32 CTAs whose 5 TMA stages cannot cover the DRAM latency
  const char* pf_ptr;
  unsigned long long pf_bytes;
  uint32_t idesc;  // tcgen05 instruction descriptor (operand format bf16 or f16, M = 128, N = BN)
};

// Epilogue warps: EW / 4 per TMEM lane quarter, round-robin over the tile's 32-column chunks.  8 for the plain
// epilogues; 16 for GEGLU, whose per-element work (erf GELU) makes the epilogue the bound: 128-wide tiles = 4 chunks
// = exactly one chunk per warp, 4 warps per scheduler to hide the MUFU / TMEM / smem latencies.

// RS: the TMA-store epilogue with an fp32 residual (igemm_tc_kernel<..., TS, RS>): 12 KB of residual / output staging
// tiles per epilogue warp instead of the 4.5 KB transpose tile, paid for with two operand stages (K <= 640 linears only).
// C2: CTA pair (cta_group::2, M = 256 per MMA): each CTA stages its own A rows and HALF of the B tile -> 26 KB instead of
// 36 KB per stage through each SM's L2 port, and seven stages instead of five.
template <int BN, int BK, int EW, bool RS = false, bool C2 = false>
struct TcCfg {
  static constexpr int THREADS = 64 + EW * 32 + ((C2 && LDM_TC_SPLITP) ? 32 : 0);
  static constexpr int BM = 128;
  static constexpr int SW = BK * 2;  // swizzle span in bytes == one K-major row
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int BN_CTA = C2 ? BN / 2 : BN;          // B rows staged by one CTA
  static constexpr int B_BYTES = ((BN_CTA * BK * 2 + 1023) / 1024) * 1024;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = C2 ? (EW > 8 ? 6 : 7) : RS ? 3 : (BN >= 128) ? (EW > 8 ? 4 : 5) : 8;
  static constexpr int TX_BYTES = (BM * BK * 2 + BN_CTA * BK * 2) * (C2 ? 2 : 1);   // C2: both CTAs' bytes land on rank 0's barrier
  static constexpr int EPI_PITCH = 36;  // floats; 144 B row pitch: conflict-free 16 B row-owner writes + row reads
  static constexpr int EPI_BYTES = RS ? EW * 12288 + 256 : EW * 32 * EPI_PITCH * 4;
  static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + (2 * STAGES + 4) * 8 + 16 + 1024;
  static constexpr int CH = BN >= 32 ? 32 : 16;  // epilogue column chunk
};

// two fp32 -> one packed pair of 16-bit floats (lo in the low half); f16 saturates instead of overflowing to inf
__device__ __forceinline__ uint32_t pack16(float lo, float hi, int dt) {
  uint32_t r;
  if (dt == LDMB200_F16)
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  else
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// ---- epilogue, coalesced phase --------------------------------------------------------------------
// A warp owns 32 accumulator rows.  tcgen05.ld hands thread i row i; the chunk is transposed through a
// warp-private smem tile so that global traffic is row-contiguous: LPR lanes cover one row's COLS columns
// (CPL = 4 fp32 or 8 16-bit columns per lane), 32/LPR rows per instruction.
template <bool OUT16>
__device__ __forceinline__ void epi_load_residual(const TcParams& p, float4 (&res)[8], int m0, int col0, int lane) {
  constexpr int CPL = OUT16 ? 8 : 4, LPR = 32 / CPL, RPI = 32 / LPR;
  const int lr = lane / LPR, lc = (lane % LPR) * CPL;
  const bool col_ok = col0 + lc < p.n_valid;
#pragma unroll
  for (int it = 0; it < LPR; ++it) {
    const int m = m0 + it * RPI + lr;
    const float* src = p.residual + (size_t)m * p.residual_ld + col0 + lc;
#pragma unroll
    for (int q = 0; q < CPL / 4; ++q)
      res[it * (CPL / 4) + q] = (col_ok && m < p.M) ? *reinterpret_cast<const float4*>(src + 4 * q)
                                                   : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// bias (+ the per-sample row vector when a tile lies inside one sample) of this lane's columns of a chunk: loaded by
// the caller BEFORE it waits for the accumulator so that the L1/L2 latency hides under the TMEM read
template <bool OUT16, int COLS>
__device__ __forceinline__ void epi_load_add(const TcParams& p, float4 (&add)[2], int col0, int lane, bool add_bias,
                                             int tile_sample) {
  constexpr int CPL = OUT16 ? 8 : 4, LPR = COLS / CPL, NQ = CPL / 4;
  const int col = col0 + (lane % LPR) * CPL;
  add[0] = add[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (col >= p.n_valid) return;
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    if (add_bias && p.bias) add[q] = *reinterpret_cast<const float4*>(p.bias + col + 4 * q);
    if (p.rowvec && p.rowvec_uniform) {
      const float4 rv = *reinterpret_cast<const float4*>(p.rowvec + (size_t)tile_sample * p.rowvec_ld + col + 4 * q);
      add[q].x += rv.x; add[q].y += rv.y; add[q].z += rv.z; add[q].w += rv.w;
    }
  }
}

template <bool OUT16, int COLS, bool STATS = false, bool UP = false>
__device__ __forceinline__ void epi_store(const TcParams& p, const float* st, int pitch, int m0, int col0, int lane,
                                          const float4 (&res)[8], bool has_res, const float4 (&add)[2]) {
  static_assert(!(STATS && OUT16), "column statistics are taken of fp32 outputs only");
  float sx[4] = {0.f, 0.f, 0.f, 0.f}, sq[4] = {0.f, 0.f, 0.f, 0.f};
  constexpr int CPL = OUT16 ? 8 : 4, LPR = COLS / CPL, RPI = 32 / LPR, NQ = CPL / 4;
  const int lr = lane / LPR, lc = (lane % LPR) * CPL;
  const int col = col0 + lc;
  if (!STATS && col >= p.n_valid) return;  // (STATS: n_valid is a multiple of the chunk width, checked on the host)
#pragma unroll
  for (int it = 0; it < LPR; ++it) {
    const int r = it * RPI + lr;
    const int m = m0 + r;
    if (m >= p.M) continue;
    float4 t[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      t[q] = *reinterpret_cast<const float4*>(&st[r * pitch + lc + 4 * q]);
      t[q].x += add[q].x; t[q].y += add[q].y; t[q].z += add[q].z; t[q].w += add[q].w;
    }
    if (p.rowvec && !p.rowvec_uniform) {
      const float* rvp = p.rowvec + (size_t)(m / p.rows_per_sample) * p.rowvec_ld + col;
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const float4 rv = *reinterpret_cast<const float4*>(rvp + 4 * q);
        t[q].x += rv.x; t[q].y += rv.y; t[q].z += rv.z; t[q].w += rv.w;
      }
    }
    if (has_res) {
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const float4 rs = res[it * NQ + q];
        t[q].x += rs.x; t[q].y += rs.y; t[q].z += rs.z; t[q].w += rs.w;
      }
    }
    size_t mo = (size_t)m;
    if (UP) {  // folded upsample: low-res pixel -> its output pixel of this launch's parity (compile-time: the r02g build
               // carried this as a runtime branch in every instantiation and lost 15 % on all 1x1 GEMMs to code size)
      const int n = m / p.up_hw, rem = m - n * p.up_hw, y = rem / p.up_w, x = rem - y * p.up_w;
      mo = (size_t)n * 4 * p.up_hw + (size_t)(2 * y + p.up_py) * (2 * p.up_w) + 2 * x + p.up_px;
    }
    if (OUT16) {
      uint4 u;
      u.x = pack16(t[0].x, t[0].y, p.out_dtype);
      u.y = pack16(t[0].z, t[0].w, p.out_dtype);
      u.z = pack16(t[NQ - 1].x, t[NQ - 1].y, p.out_dtype);
      u.w = pack16(t[NQ - 1].z, t[NQ - 1].w, p.out_dtype);
      *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out) + mo * p.out_ld + col) = u;
    } else {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + mo * p.out_ld + col) = t[0];
      if (STATS) {
        sx[0] += t[0].x; sx[1] += t[0].y; sx[2] += t[0].z; sx[3] += t[0].w;
        sq[0] = fmaf(t[0].x, t[0].x, sq[0]); sq[1] = fmaf(t[0].y, t[0].y, sq[1]);
        sq[2] = fmaf(t[0].z, t[0].z, sq[2]); sq[3] = fmaf(t[0].w, t[0].w, sq[3]);
      }
    }
  }
  if (STATS) {
    // GroupNorm statistics of the tensor being written (K7 fused): the lanes that share a column quad differ in
    // lane bits >= log2(LPR); fold them, one float2 {sum, sum of squares} per column per 32-row block
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sx[j] += __shfl_xor_sync(0xffffffffu, sx[j], o);
        sq[j] += __shfl_xor_sync(0xffffffffu, sq[j], o);
      }
    }
    if (lr == 0 && m0 < p.M) {  // (M % 32 == 0: a 32-row block is entirely inside or outside the matrix)
      size_t blk = (size_t)(m0 >> 5);
      if (UP) {  // folded upsample: the output sample's blocks are [4 parities][low-res blocks] (any order sums alike)
        const int n = m0 / p.up_hw;
        blk = (size_t)n * (p.up_hw >> 3) + (size_t)(p.up_py * 2 + p.up_px) * (p.up_hw >> 5) + ((m0 - n * p.up_hw) >> 5);
      }
      float4* dst = reinterpret_cast<float4*>(p.colstats + (blk * p.n_valid + col) * 2);
      dst[0] = make_float4(sx[0], sq[0], sx[1], sq[1]);
      dst[1] = make_float4(sx[2], sq[2], sx[3], sq[3]);
    }
  }
}

// GG: the GEGLU-epilogue specialisation (no residual / NCHW / statistics code in it: the erf-GELU epilogue is the bound of
// those GEMMs and needs the registers); GG == false handles every other epilogue.
// O16: the output is a 16-bit type (bf16 / f16) — a template parameter, like GG, so that each instantiation carries one
// store path only (the kernel was losing 8 % of its issue slots to instruction fetch, ncu r01v).
// SK: split-K specialisation (see TcParams::splits); plain epilogues only.
// TS: nothing but a bias / per-tile row vector to add and no statistics (fused QKV, proj_in, 1x1 skip convolutions): each
// thread finishes its row of the 32 x 32 chunk in registers, writes it into a swizzled shared-memory tile, and a TMA store
// (tmO) moves the tile — no transposing read-back through shared memory, no per-lane global stores.
// RS (with TS): + an fp32 residual, TMA-loaded into a per-warp ring of swizzled tiles one chunk ahead (tmR); every thread adds
// its own row in shared memory and the tile goes out through the TMA store — attn1.to_out, ff.net.2 (no statistics).
// C2: CTA-pair main loop (launched as 2-clusters; plain epilogue only): unit = a pair of vertically adjacent M tiles.
template <int BN, int BK, int EW, bool GG, bool O16, bool UP = false, bool SK = false, bool TS = false, bool RS = false,
          bool C2 = false>
__global__ void __launch_bounds__(TcCfg<BN, BK, EW, RS, C2>::THREADS, 1)
igemm_tc_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
                const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                const __grid_constant__ CUtensorMap tmR, const TcParams p) {
  using Cfg = TcCfg<BN, BK, EW, RS, C2>;
  static_assert(!RS || TS, "RS extends the TMA-store epilogue");
  static_assert(!C2 || (!UP && !SK && !RS && BN % 32 == 0 && (GG ? TS : !TS)), "CTA pair: plain epilogue, or GEGLU with the TMA store");
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment as an OFFSET from the shared array (not through uintptr_t): the compiler keeps the shared address
  // space and emits STS / LDS for the epilogue's staging tile instead of generic ST / LD
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + Cfg::STAGES * Cfg::A_BYTES;
  float* epi = reinterpret_cast<float*>(sB + Cfg::STAGES * Cfg::B_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(epi) + Cfg::EPI_BYTES);
  uint64_t* empty = full + Cfg::STAGES;
  uint64_t* tfull = empty + Cfg::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  static_assert(!SK || (!GG && !UP), "split-K serves the plain epilogues");
  static_assert(!TS || (!UP && !SK && BN % 32 == 0 && (GG ? O16 && BN == 128 : BN / 32 >= EW / 4)), "TMA-store epilogue: plain or GEGLU output");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // work units of the persistent loop: output tiles, or (tile, split) pairs with the split index fastest
  // (C2: units are PAIRS of M tiles — tiles_m is even — walked by the cluster; rank r takes the pair's r-th tile)
  const int ntiles = C2 ? (p.tiles_m >> 1) * p.tiles_n : p.tiles_m * p.tiles_n * (SK ? p.splits : 1);
  const int crank = C2 ? (int)cluster_ctarank() : 0;
  const int unit0 = C2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, ustep = C2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  auto pair_tile = [&](int unit) { return (2 * (unit / p.tiles_n) + crank) * p.tiles_n + unit % p.tiles_n; };

  if (p.pf_bytes && warp == 2 && lane == 0) {
    // constant data, nothing of the predecessor is touched: issued before the dependency wait.  CTA b takes slice b.
    const unsigned long long per = (((p.pf_bytes + gridDim.x - 1) / gridDim.x) + 4095ull) & ~4095ull;
    const unsigned long long b0 = blockIdx.x * per, b1 = b0 + per < p.pf_bytes ? b0 + per : p.pf_bytes;
    for (unsigned long long o = b0; o < b1; o += 16384ull) {
      const unsigned int n = (unsigned int)(b1 - o < 16384ull ? b1 - o : 16384ull);
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.pf_ptr + o), "r"(n) : "memory");
    }
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA1);
    prefetch_tmap(&tmB);
    if (p.c2_chunks) prefetch_tmap(&tmA2);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < Cfg::STAGES; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], 1);
      }
      for (int b = 0; b < 2; ++b) {
        mbar_init(&tfull[b], 1);
        mbar_init(&tempty[b], C2 ? 2 * EW : EW);   // C2: rank 0's barrier collects both CTAs' epilogue warps
      }
      fence_mbar_init();
    }
    __syncwarp();
    if constexpr (C2) tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS);
    else tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (C2) cluster_sync_all();   // the peer's barriers exist before anything is signalled across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // the predecessor's writes (activations, packed weights, residual, statistics buffers) are visible from here

  // K loop: taps x channel chunks of BK; the last chunk of each channel range may be partial (multiple of 16):
  // TMA zero-fills the out-of-range channels of the A box, and only `valid/16` MMAs are issued for it.
  const int nchunks = p.c1_chunks + p.c2_chunks;
  if (warp == 0) {
    // ===================== TMA producer (whole warp converged, one elected lane issues) =====================
    // (a single-thread producer / MMA loop with running stage counters was measured r02l and is SLOWER: big-K convolutions
    // 0.753 -> 0.934 ms at K = 5760, implicit GEMMs 13.6 -> 14.9 ms per eval; the converged-warp form stays)
    uint32_t it = 0;
    [[maybe_unused]] int rs_ = 0;
    [[maybe_unused]] uint32_t rp_ = 0;
    [[maybe_unused]] const uint32_t full0 = C2 ? mapa_u32(smem_u32(full), 0) : 0u;   // rank 0's full[] barriers
    for (int unit = unit0; unit < ntiles; unit += ustep) {
      const int tile = C2 ? pair_tile(unit) : SK ? unit / p.splits : unit;
      const int tm = tile / p.tiles_n, tn = tile % p.tiles_n;
      const int m0 = tm * Cfg::BM;
      // this unit's range [k0, k1) of the flattened (tap, chunk) iterations
      const int k0 = SK ? (unit - tile * p.splits) * p.kiters_per_split : 0;
      const int k1 = SK ? min(p.taps * nchunks, k0 + p.kiters_per_split) : p.taps * nchunks;
      int n_img = 0, h0 = 0, w0 = 0;
      if (p.is_conv) {
        n_img = m0 / p.hw_out;
        h0 = (m0 % p.hw_out) / p.w_out;
        w0 = m0 % p.w_out;  // non-zero only for maps wider than one 128-pixel tile (first-stage decoder at 256x256)
      }
      for (int tap = SK ? k0 / nchunks : 0; tap < (SK ? (k1 + nchunks - 1) / nchunks : p.taps); ++tap) {
        int dh = 0, dw = 0, plane = 0;
        if (p.ks == 2) {  // folded upsample: tap (a, b) reads low-res pixel (y + a - 1 + py, x + b - 1 + px)
          dh = (tap >> 1) - 1 + p.up_py;
          dw = (tap & 1) - 1 + p.up_px;
        } else if (p.ks == 3) {
          const int kh = tap / 3, kw = tap % 3;
          if (p.stride == 1) {
            dh = kh - 1;
            dw = kw - 1;
          } else {  // stride 2 on space-to-depth planes: row 2*oh+kh-1 -> (parity, index)
            int ph, pw;
            if (p.pad_asym) {  // row 2*oh+kh: kh 0 / 2 are even rows (oh, oh + 1), kh 1 the odd row oh
              ph = (kh == 1) ? 1 : 0;
              pw = (kw == 1) ? 1 : 0;
              dh = (kh == 2) ? 1 : 0;
              dw = (kw == 2) ? 1 : 0;
            } else {
              ph = (kh == 1) ? 0 : 1;
              pw = (kw == 1) ? 0 : 1;
              dh = (kh == 0) ? -1 : 0;
              dw = (kw == 0) ? -1 : 0;
            }
            plane = ph * 2 + pw;
          }
        }
        for (int j = SK ? max(k0 - tap * nchunks, 0) : 0; j < (SK ? min(k1 - tap * nchunks, nchunks) : nchunks); ++j, ++it) {
#if LDM_TC_RUNCNT
          const int s = rs_;
          const uint32_t ph_ = rp_;
          if (++rs_ == Cfg::STAGES) { rs_ = 0; rp_ ^= 1u; }
#else
          const int s = it % Cfg::STAGES;
          const uint32_t ph_ = (it / Cfg::STAGES) & 1;
#endif
          mbar_wait(&empty[s], ph_ ^ 1);
          if (elect_one()) {
            const bool second = j >= p.c1_chunks;
            const CUtensorMap* tm_a = second ? &tmA2 : &tmA1;
            const int c0 = (second ? j - p.c1_chunks : j) * BK;
            if constexpr (C2) {
              if (crank == 0) mbar_expect_tx(&full[s], Cfg::TX_BYTES);
              const uint32_t bar = full0 + s * 8;
              if (p.is_conv)
                tma_load_5d_pair(sA + s * Cfg::A_BYTES, tm_a, bar, c0, w0 + dw, h0 + dh, plane, n_img);
              else
                tma_load_5d_pair(sA + s * Cfg::A_BYTES, tm_a, bar, c0, 0, m0, 0, 0);
              if (!LDM_TC_SPLITP)
                tma_load_2d_pair(sB + s * Cfg::B_BYTES, &tmB, bar, tap * p.ctot + (second ? p.c1 : 0) + c0,
                                 p.b_row0 + tn * BN + crank * (BN / 2));
            } else {
            mbar_expect_tx(&full[s], Cfg::TX_BYTES);
            if (p.is_conv)
              tma_load_5d(sA + s * Cfg::A_BYTES, tm_a, &full[s], c0, w0 + dw, h0 + dh, plane, n_img);
            else
              tma_load_5d(sA + s * Cfg::A_BYTES, tm_a, &full[s], c0, 0, m0, 0, 0);
            tma_load_2d(sB + s * Cfg::B_BYTES, &tmB, &full[s], tap * p.ctot + (second ? p.c1 : 0) + c0, p.b_row0 + tn * BN);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (C2 && LDM_TC_SPLITP && warp == 2 + EW) {
    // ===================== second producer of the CTA-pair kernels: this CTA's half of the weight tile =====================
    // (the one-lane instruction stream of a producer that issues A and B — ~80 SASS instructions per 64-deep chunk, most of
    // them register-to-uniform moves for the TMA coordinates — took as long as the chunk's MMAs: ncu r02t)
    const uint32_t full0 = mapa_u32(smem_u32(full), 0);
    int rs_ = 0;
    uint32_t rp_ = 0;
    for (int unit = unit0; unit < ntiles; unit += ustep) {
      const int brow = p.b_row0 + (unit % p.tiles_n) * BN + crank * (BN / 2);
      for (int tap = 0; tap < p.taps; ++tap) {
        const int kcol = tap * p.ctot;
        for (int j = 0; j < nchunks; ++j) {
          const int s = rs_;
          const uint32_t ph_ = rp_;
          if (++rs_ == Cfg::STAGES) { rs_ = 0; rp_ ^= 1u; }
          mbar_wait(&empty[s], ph_ ^ 1);
          if (elect_one())
            tma_load_2d_pair(sB + s * Cfg::B_BYTES, &tmB, full0 + s * 8,
                             kcol + (j >= p.c1_chunks ? p.c1 + (j - p.c1_chunks) * BK : j * BK), brow);
          __syncwarp();
        }
      }
    }
  } else if (warp == 1 && (!C2 || crank == 0)) {
    // ===================== MMA issuer (whole warp converged, one elected lane issues; C2: rank 0 for the pair) ==========
    // descriptors: the high word is constant, the low word is (address >> 4) | LBO; K advances by 32 B = +2
    constexpr uint32_t desc_hi = (uint32_t)((8u * Cfg::SW) >> 4) | (1u << 14) | ((Cfg::SW == 128 ? 2u : 4u) << 29);
    const uint32_t a_lo0 = ((smem_u32(sA) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t b_lo0 = ((smem_u32(sB) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t idesc = p.idesc;
    const int kmax1 = p.c1_tail_k, kmax2 = p.c2_tail_k;
    uint32_t it = 0, t_iter = 0;
    [[maybe_unused]] int rs_ = 0;
    [[maybe_unused]] uint32_t rp_ = 0;
    for (int unit = unit0; unit < ntiles; unit += ustep, ++t_iter) {
      const uint32_t buf = t_iter & 1;
      mbar_wait(&tempty[buf], ((t_iter >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * BN;
      uint32_t acc = 0;
      const int k0 = SK ? (unit % p.splits) * p.kiters_per_split : 0;
      const int k1 = SK ? min(p.taps * nchunks, k0 + p.kiters_per_split) : p.taps * nchunks;
      for (int tap = SK ? k0 / nchunks : 0; tap < (SK ? (k1 + nchunks - 1) / nchunks : p.taps); ++tap) {
        for (int j = SK ? max(k0 - tap * nchunks, 0) : 0; j < (SK ? min(k1 - tap * nchunks, nchunks) : nchunks); ++j, ++it) {
#if LDM_TC_RUNCNT
          const int s = rs_;
          const uint32_t ph_ = rp_;
          if (++rs_ == Cfg::STAGES) { rs_ = 0; rp_ ^= 1u; }
#else
          const int s = it % Cfg::STAGES;
          const uint32_t ph_ = (it / Cfg::STAGES) & 1;
#endif
          const int kmax = (j == p.c1_chunks - 1) ? kmax1 : (j == nchunks - 1) ? kmax2 : BK / 16;
          mbar_wait(&full[s], ph_);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_lo = a_lo0 + s * (Cfg::A_BYTES >> 4), b_lo = b_lo0 + s * (Cfg::B_BYTES >> 4);
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk) {
              if (kk < kmax) {
                if constexpr (C2)
                  umma_bf16_ss_pair(d_tmem, ((uint64_t)desc_hi << 32) | (a_lo + 2 * kk),
                                    ((uint64_t)desc_hi << 32) | (b_lo + 2 * kk), idesc, acc);
                else
                umma_bf16_ss(d_tmem, ((uint64_t)desc_hi << 32) | (a_lo + 2 * kk), ((uint64_t)desc_hi << 32) | (b_lo + 2 * kk),
                             idesc, acc);
                acc = 1;
              }
            }
            if constexpr (C2) umma_commit_pair(&empty[s]);  // frees the stage in BOTH CTAs
            else umma_commit(&empty[s]);  // frees the smem stage once these MMAs retire
          }
          __syncwarp();
          acc = 1;
        }
      }
      if (elect_one()) {  // accumulator complete -> epilogue (C2: of both CTAs)
        if constexpr (C2) umma_commit_pair(&tfull[buf]);
        else umma_commit(&tfull[buf]);
      }
      __syncwarp();
    }
  } else if (warp >= 2 && warp < 2 + EW) {
    // ===================== epilogue (EW warps) =====================
    const int ew = warp - 2;
    const int qw = warp & 3;          // TMEM lane quarter accessible to this warp (hardware: warp id % 4)
    const int par = ew >> 2;          // which of the quarter's EW/4 warps: takes chunks cc % (EW/4) == par
    constexpr int CSTEP = EW / 4;
    constexpr int NCH = BN / Cfg::CH;
    constexpr int PITCH = Cfg::EPI_PITCH;
    float* st = epi + ew * 32 * PITCH;
    constexpr bool out16 = O16;
    const bool has_res = !GG && p.residual != nullptr;
    uint32_t t_iter = 0;
    [[maybe_unused]] const uint32_t tempty0 = C2 ? mapa_u32(smem_u32(tempty), 0) : 0u;   // rank 0's tempty[] barriers
    if constexpr (RS) {
      // ---- fp32 residual + bias / uniform row vector: TMA-load / TMA-store epilogue.  Per warp 12 KB: fp32 output = three
      // 4 KB tiles used IN PLACE (residual lands, every thread adds its row, the same tile is stored); 16-bit output = two
      // residual tiles + two 2 KB output tiles.  Work items of the warp = its chunks, tile after tile; item i + 1's residual
      // is requested when item i starts (it does not depend on the accumulator), so the load has a whole chunk to land. ----
      constexpr int NRB = O16 ? 2 : 3;                                  // residual (fp32: residual == output) tiles
      uint8_t* stb = reinterpret_cast<uint8_t*>(epi) + ew * 12288;      // 1024-byte aligned
      uint8_t* outb = stb + 8192;                                       // 16-bit output tiles (2 x 2 KB)
      uint64_t* rfull = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(epi) + EW * 12288) + ew * 3;
      if (lane == 0) {
        for (int b = 0; b < NRB; ++b) mbar_init(&rfull[b], 1);
        fence_mbar_init();
        prefetch_tmap(&tmO);
        prefetch_tmap(&tmR);
      }
      __syncwarp();
      const int sw = lane & 7, sw16 = (lane >> 1) & 3;
      const int odt = p.out_dtype;
      uint32_t item = 0;
      auto request = [&](uint32_t it, int tile_, int cc_) {  // lane 0: residual chunk (tile_, cc_) -> ring slot it % NRB
        const int tm_ = tile_ / p.tiles_n, tn_ = tile_ - tm_ * p.tiles_n;
        uint64_t* bar = &rfull[it % NRB];
        mbar_expect_tx(bar, 4096);
        tma_load_2d(stb + (it % NRB) * 4096, &tmR, bar, tn_ * BN + cc_ * 32, tm_ * Cfg::BM + qw * 32);
      };
      if (par < NCH && (int)blockIdx.x < ntiles && lane == 0) request(0, blockIdx.x, par);
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t_iter) {
        const int tm = tile / p.tiles_n, tn = tile % p.tiles_n;
        const int m0 = tm * Cfg::BM + qw * 32;
        const int n0 = tn * BN;
        const uint32_t buf = t_iter & 1;
        const float* rvp = p.rowvec ? p.rowvec + (size_t)((tm * Cfg::BM) / p.rows_per_sample) * p.rowvec_ld : nullptr;
        float4 bq[8];
        auto load_add = [&](float4 (&b)[8], int cc) {
          const int col = n0 + cc * 32;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            b[j] = p.bias ? __ldg(reinterpret_cast<const float4*>(p.bias + col) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (rvp) {
              const float4 r = __ldg(reinterpret_cast<const float4*>(rvp + col) + j);
              b[j].x += r.x; b[j].y += r.y; b[j].z += r.z; b[j].w += r.w;
            }
          }
        };
        if (par < NCH) load_add(bq, par);
        mbar_wait(&tfull[buf], (t_iter >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int cc = par; cc < NCH; cc += CSTEP, ++item) {
          const bool more = cc + CSTEP < NCH;
          if (lane == 0) {  // next item's residual
            const int ntile = more ? tile : tile + (int)gridDim.x, ncc = more ? cc + CSTEP : par;
            if (ntile < ntiles) {
              // fp32: the slot was item - 2's tile; its store (one of the last two groups) has finished READING it
              if constexpr (!O16) tma_store_wait_read<1>();
              request(item + 1, ntile, ncc);
            }
          }
          float v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(qw * 32) << 16) + buf * BN + cc * 32, v);
          float4 bn[8];
          if (more) load_add(bn, cc + CSTEP);
          tmem_ld_wait();
          if (!more) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[buf]);
          }
          mbar_wait(&rfull[item % NRB], (item / NRB) & 1);
          uint8_t* rrow = stb + (item % NRB) * 4096 + lane * 128;
          if constexpr (O16) {
            if (lane == 0) tma_store_wait_read<1>();    // the output tile used two items ago has been read
            __syncwarp();
            uint8_t* tile_o = outb + (item & 1) * 2048;
            uint8_t* orow = tile_o + lane * 64;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const float4 r0 = *reinterpret_cast<const float4*>(rrow + (((2 * c) ^ sw) << 4));
              const float4 r1 = *reinterpret_cast<const float4*>(rrow + (((2 * c + 1) ^ sw) << 4));
              const float4 b0 = bq[2 * c], b1 = bq[2 * c + 1];
              uint4 o;
              o.x = pack16(v[8 * c] + b0.x + r0.x, v[8 * c + 1] + b0.y + r0.y, odt);
              o.y = pack16(v[8 * c + 2] + b0.z + r0.z, v[8 * c + 3] + b0.w + r0.w, odt);
              o.z = pack16(v[8 * c + 4] + b1.x + r1.x, v[8 * c + 5] + b1.y + r1.y, odt);
              o.w = pack16(v[8 * c + 6] + b1.z + r1.z, v[8 * c + 7] + b1.w + r1.w, odt);
              *reinterpret_cast<uint4*>(orow + ((c ^ sw16) << 4)) = o;
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmO, tile_o, n0 + cc * 32, m0);
              tma_store_commit();
            }
          } else {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              float4* q = reinterpret_cast<float4*>(rrow + ((c ^ sw) << 4));
              const float4 r = *q;
              *q = make_float4(v[4 * c] + bq[c].x + r.x, v[4 * c + 1] + bq[c].y + r.y, v[4 * c + 2] + bq[c].z + r.z,
                               v[4 * c + 3] + bq[c].w + r.w);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmO, stb + (item % NRB) * 4096, n0 + cc * 32, m0);
              tma_store_commit();
            }
          }
          if (more) {
#pragma unroll
            for (int j = 0; j < 8; ++j) bq[j] = bn[j];
          }
        }
      }
      if (lane == 0) tma_store_wait_all();
    } else
    if constexpr (TS && !GG) {
      // ---- bias / uniform row vector only: TMA-store epilogue.  Staging per warp: 16-bit output = two 2 KB tiles (32 rows
      // x 64 B, 64B swizzle), fp32 output = one 4 KB tile (32 rows x 128 B, 128B swizzle); the 16-byte piece c of a row sits
      // at piece c ^ (bits 7.. of the row's shared address), so the row owners' 16-byte writes are bank-conflict-free ----
      constexpr int ROWB = O16 ? 64 : 128, NBUF = O16 ? 2 : 1;
      uint8_t* stb = reinterpret_cast<uint8_t*>(epi) + ew * 4096;      // 1024-byte aligned (the staging area is)
      if (warp == 2 && lane == 0) prefetch_tmap(&tmO);
      uint32_t nstore = 0;
      const int sw = O16 ? ((lane >> 1) & 3) : (lane & 7);
      const int odt = p.out_dtype;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t_iter) {
        const int tm = tile / p.tiles_n, tn = tile % p.tiles_n;
        const int m0 = tm * Cfg::BM + qw * 32;
        const int n0 = tn * BN;
        const uint32_t buf = t_iter & 1;
        const float* rvp = p.rowvec ? p.rowvec + (size_t)((tm * Cfg::BM) / p.rows_per_sample) * p.rowvec_ld : nullptr;
        float4 bq[8];
        auto load_add = [&](float4 (&b)[8], int cc) {  // the same 32 column addends for every row of the tile
          const int col = n0 + cc * 32;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            b[j] = p.bias ? __ldg(reinterpret_cast<const float4*>(p.bias + col) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (rvp) {
              const float4 r = __ldg(reinterpret_cast<const float4*>(rvp + col) + j);
              b[j].x += r.x; b[j].y += r.y; b[j].z += r.z; b[j].w += r.w;
            }
          }
        };
        load_add(bq, par);
        mbar_wait(&tfull[buf], (t_iter >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int cc = par; cc < NCH; cc += CSTEP) {
          float v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(qw * 32) << 16) + buf * BN + cc * 32, v);
          const bool more = cc + CSTEP < NCH;
          float4 bn[8];
          if (more) load_add(bn, cc + CSTEP);
          tmem_ld_wait();
          if (!more) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[buf]);
          }
          // the store that last read this staging tile has finished reading it
          if (lane == 0) tma_store_wait_read<NBUF - 1>();
          __syncwarp();
          uint8_t* tile_s = stb + (nstore % NBUF) * 2048;
          uint8_t* row_s = tile_s + lane * ROWB;
          if constexpr (O16) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const float4 b0 = bq[2 * c], b1 = bq[2 * c + 1];
              uint4 o;
              o.x = pack16(v[8 * c] + b0.x, v[8 * c + 1] + b0.y, odt);
              o.y = pack16(v[8 * c + 2] + b0.z, v[8 * c + 3] + b0.w, odt);
              o.z = pack16(v[8 * c + 4] + b1.x, v[8 * c + 5] + b1.y, odt);
              o.w = pack16(v[8 * c + 6] + b1.z, v[8 * c + 7] + b1.w, odt);
              *reinterpret_cast<uint4*>(row_s + ((c ^ sw) << 4)) = o;
            }
          } else {
#pragma unroll
            for (int c = 0; c < 8; ++c)
              *reinterpret_cast<float4*>(row_s + ((c ^ sw) << 4)) =
                  make_float4(v[4 * c] + bq[c].x, v[4 * c + 1] + bq[c].y, v[4 * c + 2] + bq[c].z, v[4 * c + 3] + bq[c].w);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmO, tile_s, n0 + cc * 32, m0);
            tma_store_commit();
          }
          ++nstore;
          if (more) {
#pragma unroll
            for (int j = 0; j < 8; ++j) bq[j] = bn[j];
          }
        }
      }
      if (lane == 0) tma_store_wait_all();
    } else
    if constexpr (GG && Cfg::CH == 32 && LDM_TC_GG_PIPE) {
      // ---- GEGLU epilogue, software-pipelined over the two 16-column halves of the warp's chunk ----
      // All 16 epilogue warps wait for the same accumulator and have identical work, so they ran in lock step: first
      // everyone reads tensor memory (a 128x128 fp32 tile = 64 KB through a 64 B/clk read port = ~1000 clk with nothing
      // else to do), then everyone evaluates GELUs (issue-bound, the read port idle) — ~3100 clk per tile of which the two
      // phases explain 2500.  Here the second half's tcgen05.ld is in flight while the first half is evaluated, and the
      // NEXT tile's first half is fetched (other accumulator buffer, complete long ago) while the second half is evaluated.
      // Measured on top of this and rejected (r02w / r02x, same-run A/B/A/B on the N = 1280 GEMMs): the biases fetched a whole
      // tile ahead + no zero-add in the store (1.079 -> 1.135 ms), and a rows-fastest lane map that removes the 2-way bank
      // conflicts of the 16-column staging read (1.07 -> 1.17 ms: the stores lose their lane-pair-contiguous 32-byte sectors).
      static_assert(!GG || NCH <= CSTEP, "one chunk per warp and tile");
      const int cc = par;
      if (cc < NCH) {
        float va[16], vb[16];  // half h: [8 value | 8 gate] columns 8h..8h+7 and 16+8h..16+8h+7 of the chunk
        const uint32_t lane_base = tmem_base + ((uint32_t)(qw * 32) << 16) + cc * 32;
        // TS: each half is exactly one 16-byte piece (8 x 16-bit) of the thread's output row: two 1 KB tiles (32 rows x 32 B,
        // 32B swizzle: piece ^= bit 2 of the row) per warp, one TMA store per chunk instead of the transposing read-back
        uint8_t* gtile = reinterpret_cast<uint8_t*>(st);
        [[maybe_unused]] const int gsw = (lane >> 2) & 1;
        [[maybe_unused]] const int godt = p.out_dtype;
        if (TS && warp == 2 && lane == 0) prefetch_tmap(&tmO);
        int unit = unit0;
        if (unit < ntiles) {
          mbar_wait(&tfull[0], 0);
          tc_fence_after();
          tmem_ld_32x8(lane_base, va);
          tmem_ld_32x8(lane_base + 16, va + 8);
        }
        for (; unit < ntiles; unit += ustep, ++t_iter) {
          const int tile = C2 ? pair_tile(unit) : unit;
          const int tm = tile / p.tiles_n, tn = tile - tm * p.tiles_n;
          const int m0 = tm * Cfg::BM + qw * 32;
          const int n0 = tn * BN;
          const uint32_t buf = t_iter & 1;
          const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + cc * 32);
          float4 b4[4] = {bp[0], bp[1], bp[4], bp[5]};          // half 0: value biases 0..7, gate biases 16..23
          tmem_ld_wait16(va);
          tmem_ld_32x8(lane_base + buf * BN + 8, vb);           // half 1 in flight while half 0 is evaluated
          tmem_ld_32x8(lane_base + buf * BN + 24, vb + 8);
          auto half = [&](const float (&v)[16], const float4 (&b)[4], int h) {
            const float bb[16] = {b[0].x, b[0].y, b[0].z, b[0].w, b[1].x, b[1].y, b[1].z, b[1].w,
                                  b[2].x, b[2].y, b[2].z, b[2].w, b[3].x, b[3].y, b[3].z, b[3].w};
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
              const f32x2 val2 = add2(pk2(v[j], v[j + 1]), pk2(bb[j], bb[j + 1]));
              const f32x2 g2 = add2(pk2(v[8 + j], v[9 + j]), pk2(bb[8 + j], bb[9 + j]));
              unpk2(mul2(val2, gelu_erf_fast2(g2)), o[j], o[j + 1]);
            }
            if constexpr (TS) {
              uint4 u;
              u.x = pack16(o[0], o[1], godt); u.y = pack16(o[2], o[3], godt);
              u.z = pack16(o[4], o[5], godt); u.w = pack16(o[6], o[7], godt);
              *reinterpret_cast<uint4*>(gtile + (t_iter & 1) * 1024 + lane * 32 + ((h ^ gsw) << 4)) = u;
            } else {
              *reinterpret_cast<float4*>(&st[lane * PITCH + 8 * h]) = make_float4(o[0], o[1], o[2], o[3]);
              *reinterpret_cast<float4*>(&st[lane * PITCH + 8 * h + 4]) = make_float4(o[4], o[5], o[6], o[7]);
            }
          };
          if constexpr (TS) {  // the store that read this staging tile two tiles ago has finished reading it
            if (lane == 0) tma_store_wait_read<1>();
            __syncwarp();
          }
          half(va, b4, 0);
          const float4 c4[4] = {bp[2], bp[3], bp[6], bp[7]};    // half 1: value biases 8..15, gate biases 24..31
          tmem_ld_wait16(vb);
          tc_fence_before();                                    // the whole chunk has left tensor memory
          __syncwarp();
          if (lane == 0) { if constexpr (C2) mbar_arrive_cluster(tempty0 + buf * 8); else mbar_arrive(&tempty[buf]); }
          if (unit + ustep < ntiles) {                          // next tile's half 0 (other buffer) during half 1's GELUs
            mbar_wait(&tfull[buf ^ 1], ((t_iter + 1) >> 1) & 1);
            tc_fence_after();
            tmem_ld_32x8(lane_base + (buf ^ 1) * BN, va);
            tmem_ld_32x8(lane_base + (buf ^ 1) * BN + 16, va + 8);
          }
          half(vb, c4, 1);
          const int col0 = (n0 >> 1) + cc * 16;
          if constexpr (TS) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmO, gtile + (t_iter & 1) * 1024, col0, m0);
              tma_store_commit();
            }
          } else {
            __syncwarp();
            const float4 zero_add[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
            float4 res0[8];
            if (out16) epi_store<true, 16>(p, st, PITCH, m0, col0, lane, res0, false, zero_add);
            else epi_store<false, 16>(p, st, PITCH, m0, col0, lane, res0, false, zero_add);
            __syncwarp();
          }
        }
        if (TS && lane == 0) tma_store_wait_all();
      } else {  // a warp without a chunk (BN == 32): it still hands every accumulator buffer back
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t_iter) {
          const uint32_t buf = t_iter & 1;
          mbar_wait(&tfull[buf], (t_iter >> 1) & 1);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[buf]);
        }
      }
    } else
    for (int unit = unit0; unit < ntiles; unit += ustep, ++t_iter) {
      const int tile = C2 ? pair_tile(unit) : SK ? unit / p.splits : unit;
      const int tm = tile / p.tiles_n, tn = tile % p.tiles_n;
      const int m0 = tm * Cfg::BM + qw * 32;
      const int n0 = tn * BN;
      const uint32_t buf = t_iter & 1;
      const int tile_sample = (tm * Cfg::BM) / p.rows_per_sample;
      float4 res[8];
      const int cc_first = par;
      // partials of this tile: [split][chunk][CH/4][128 rows] float4 — a warp's store of one float4 index is 512 contiguous bytes
      const float4* sk_tile = nullptr;
      if constexpr (SK) {
        sk_tile = reinterpret_cast<const float4*>(p.sk_part) + (size_t)tile * p.splits * (NCH * (Cfg::CH / 4) * 128);
        float4* mine = const_cast<float4*>(sk_tile) + (size_t)(unit - tile * p.splits) * (NCH * (Cfg::CH / 4) * 128) + qw * 32 + lane;
        mbar_wait(&tfull[buf], (t_iter >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int cc = cc_first; cc < NCH; cc += CSTEP) {
          float v[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16) + buf * BN + cc * Cfg::CH;
          if (Cfg::CH == 32) tmem_ld_32x32(taddr, v);
          else tmem_ld_32x16(taddr, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < Cfg::CH / 4; ++j)
            __stcg(mine + (cc * (Cfg::CH / 4) + j) * 128, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[buf]);
        // every epilogue thread's partial is visible device-wide before the tile's counter moves (threadFenceReduction pattern)
        __threadfence();
        asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");
        if (ew == 0 && lane == 0) {
          const unsigned int prev = atomicAdd(p.sk_cnt + tile, 1u);
          const bool last = prev == (unsigned int)p.splits - 1u;
          if (last) p.sk_cnt[tile] = 0u;  // every split of the tile has arrived: the counter is ready for the next launch
          tmem_slot[1] = last ? 1u : 0u;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");
        const bool last = reinterpret_cast<volatile uint32_t*>(tmem_slot)[1] != 0u;
        asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");  // (the flag is rewritten by the next unit)
        if (!last) continue;
        __threadfence();
      }
#if LDM_TC_RES_PREFETCH
      // The +residual GEMMs with K <= 640 are bound by the latency of the residual read (ncu r01s: 80 KB of fp32
      // residual per tile with at most one 4 KB chunk per warp in flight = 7.4 us per tile): pull the NEXT tile's
      // residual rows into L2 now, a whole tile time ahead; fire-and-forget, no registers held.
      if (!SK && has_res) {
        const int tile_n = C2 ? pair_tile(unit + ustep) : tile + (int)gridDim.x;
        if (unit + ustep < ntiles) {
          const int m = (tile_n / p.tiles_n) * Cfg::BM + qw * 32 + lane;
          if (m < p.M) {
            const char* rp = reinterpret_cast<const char*>(p.residual + (size_t)m * p.residual_ld + (tile_n % p.tiles_n) * BN);
#pragma unroll
            for (int l = par; l < (BN * 4 + 127) / 128; l += CSTEP)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + l * 128));
          }
        }
      }
#endif
      // the residual tile does not depend on the accumulator: fetch it while the MMAs are still running
      // (a second chunk of lookahead was measured r01u: slower by 15-45 % on the K <= 640 GEMMs - register spills)
      if (has_res && cc_first < NCH) {
        if (out16) epi_load_residual<true>(p, res, m0, n0 + cc_first * Cfg::CH, lane);
        else epi_load_residual<false>(p, res, m0, n0 + cc_first * Cfg::CH, lane);
      }
      // bias / per-sample row vector of the first chunk: fetched before the accumulator wait, and one chunk ahead from then
      // on (ncu r02j: the FADD that folds the row vector into the bias stalled on these loads — 8 % of the epilogue warps'
      // samples on the K = 160 GEMMs — because they were issued in the chunk that consumes them)
      float4 addv[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
      // GEGLU: every warp owns at most one 32-column chunk per tile; its 32 biases ([16 value | 16 gate], packed order) are
      // fetched here, before the accumulator wait (they used to be loaded after it: an exposed L1/L2 round trip per chunk,
      // the long-scoreboard share of the r02e profile of this specialisation)
      float bv[32];
      if constexpr (GG) {
        static_assert(NCH <= CSTEP, "the GEGLU epilogue prefetches the bias of ONE chunk per warp and tile");
        if (Cfg::CH == 32 && cc_first < NCH) {
          const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + cc_first * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 t = bp[j];
            bv[4 * j] = t.x; bv[4 * j + 1] = t.y; bv[4 * j + 2] = t.z; bv[4 * j + 3] = t.w;
          }
        }
      }
      if constexpr (!GG) {
        if (!p.out_nchw && cc_first < NCH) {
          if (out16) epi_load_add<true, Cfg::CH>(p, addv, n0 + cc_first * Cfg::CH, lane, true, tile_sample);
          else epi_load_add<false, Cfg::CH>(p, addv, n0 + cc_first * Cfg::CH, lane, true, tile_sample);
        }
      }
      if constexpr (!SK) {
        mbar_wait(&tfull[buf], (t_iter >> 1) & 1);
        tc_fence_after();
      }
#pragma unroll 1
      for (int cc = cc_first; cc < NCH; cc += CSTEP) {
        float v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16) + buf * BN + cc * Cfg::CH;
        if constexpr (SK) {  // accumulator = the splits' partials, folded in split order whichever CTA arrived last
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
          const float4* src = sk_tile + cc * (Cfg::CH / 4) * 128 + qw * 32 + lane;
#pragma unroll 1
          for (int sp = 0; sp < p.splits; ++sp, src += NCH * (Cfg::CH / 4) * 128) {
            float4 t[Cfg::CH / 4];
#pragma unroll
            for (int j = 0; j < Cfg::CH / 4; ++j) t[j] = __ldcg(src + j * 128);
#pragma unroll
            for (int j = 0; j < Cfg::CH / 4; ++j) {
              v[4 * j] += t[j].x; v[4 * j + 1] += t[j].y; v[4 * j + 2] += t[j].z; v[4 * j + 3] += t[j].w;
            }
          }
        } else if (Cfg::CH == 32)
          tmem_ld_32x32(taddr, v);
        else
          tmem_ld_32x16(taddr, v);
        float4 res_next[8];
        const bool more = cc + CSTEP < NCH;
        if (has_res && more) {
          if (out16) epi_load_residual<true>(p, res_next, m0, n0 + (cc + CSTEP) * Cfg::CH, lane);
          else epi_load_residual<false>(p, res_next, m0, n0 + (cc + CSTEP) * Cfg::CH, lane);
        }
        float4 addv_next[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
        if constexpr (!GG) {
          if (!p.out_nchw && more) {
            if (out16) epi_load_add<true, Cfg::CH>(p, addv_next, n0 + (cc + CSTEP) * Cfg::CH, lane, true, tile_sample);
            else epi_load_add<false, Cfg::CH>(p, addv_next, n0 + (cc + CSTEP) * Cfg::CH, lane, true, tile_sample);
          }
        }
        if constexpr (!SK) {
          tmem_ld_wait();
          if (!more) {  // last TMEM read of this warp for this tile: hand the accumulator buffer back early
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if constexpr (C2) mbar_arrive_cluster(tempty0 + buf * 8); else mbar_arrive(&tempty[buf]); }
          }
        }
        if (!GG && !O16 && p.out_nchw) {
          // UNetModel.out conv: thread owns pixel m, writes the n_valid real channels as fp32 NCHW
          const int m = m0 + lane;
          if (m < p.M) {
            const int b = m / p.hw_out, pix = m % p.hw_out;
            float* o = reinterpret_cast<float*>(p.out);
#pragma unroll
            for (int j = 0; j < Cfg::CH; ++j) {
              const int n = n0 + cc * Cfg::CH + j;
              if (n < p.n_valid) o[((size_t)b * p.n_valid + n) * p.hw_out + pix] = v[j] + (p.bias ? p.bias[n] : 0.f);
            }
          }
          continue;
        }
        if constexpr (GG) {  // chunk = [16 value | 16 gate] columns (packed order), biases in the same order
          if (Cfg::CH == 32) {
#pragma unroll
            for (int j = 0; j < 16; j += 2) {  // packed fp32 pairs: half the issue slots of the epilogue's bound
              const f32x2 val2 = add2(pk2(v[j], v[j + 1]), pk2(bv[j], bv[j + 1]));
              const f32x2 g2 = add2(pk2(v[16 + j], v[17 + j]), pk2(bv[16 + j], bv[17 + j]));
              unpk2(mul2(val2, gelu_erf_fast2(g2)), v[j], v[j + 1]);
            }
            // (storing each thread's 16 outputs = one 32-byte sector straight from registers, without this transpose, cuts the
            // chunk from ~360 to ~230 instructions and was measured r02t: 1.149 -> 1.204 ms on the N = 1280 GEMMs — the
            // row-scattered stores cost more than the instructions save; the epilogue is not issue-bound)
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<float4*>(&st[lane * PITCH + j * 4]) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
            __syncwarp();
            const int col0 = (n0 >> 1) + cc * 16;
            const float4 zero_add[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
            if (out16) epi_store<true, 16>(p, st, PITCH, m0, col0, lane, res, false, zero_add);
            else epi_store<false, 16>(p, st, PITCH, m0, col0, lane, res, false, zero_add);
            __syncwarp();
          }
          continue;
        }
        if constexpr (!GG) {
#pragma unroll
        for (int j = 0; j < Cfg::CH / 4; ++j)
          *reinterpret_cast<float4*>(&st[lane * PITCH + j * 4]) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
        __syncwarp();
        {
          const int col0 = n0 + cc * Cfg::CH;
          if (out16) epi_store<true, Cfg::CH>(p, st, PITCH, m0, col0, lane, res, has_res, addv);
          else if (p.colstats) epi_store<false, Cfg::CH, true, UP>(p, st, PITCH, m0, col0, lane, res, has_res, addv);
          else epi_store<false, Cfg::CH, false, UP>(p, st, PITCH, m0, col0, lane, res, has_res, addv);
        }
        __syncwarp();
        if (has_res && more) {
#pragma unroll
          for (int j = 0; j < 8; ++j) res[j] = res_next[j];
        }
        addv[0] = addv_next[0];
        addv[1] = addv_next[1];
        }  // !GG
      }
      if (!SK && cc_first >= NCH) {  // this warp had no chunk in the tile (BN == CH): still release the buffer
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if constexpr (C2) mbar_arrive_cluster(tempty0 + buf * 8); else mbar_arrive(&tempty[buf]); }
      }
    }
  }

  pdl_launch_dependents();  // PDL (opt-in): this CTA's work is issued; the successor may start being scheduled
  tc_fence_before();
  __syncthreads();
  if constexpr (C2) cluster_sync_all();   // the peer has finished reading this CTA's operands and signalling its barriers
  if (warp == 1) {
    tc_fence_after();
    if constexpr (C2) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BN, int BK, int EW, bool GG, bool O16, bool UP = false, bool SK = false, bool TS = false, bool RS = false,
          bool C2 = false>
static int launch_tc_impl(const CUtensorMap& a1, const CUtensorMap& a2, const CUtensorMap& b, const TcParams& p,
                     cudaStream_t st, const CUtensorMap* o = nullptr, const CUtensorMap* r = nullptr) {
  using Cfg = TcCfg<BN, BK, EW, RS, C2>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(igemm_tc_kernel<BN, BK, EW, GG, O16, UP, SK, TS, RS, C2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("igemm_tc: cudaFuncSetAttribute(smem=%d) failed: %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
      return LDMB200_ERR_CUDA;
    }
    attr_set = true;
  }
  int grid = std::min(p.tiles_m * p.tiles_n * (SK ? p.splits : 1), sm_count());
  cudaError_t le;
  if constexpr (C2) {  // 2-clusters: one per pair of M tiles, at most one per pair of SMs
    grid = 2 * std::min((p.tiles_m / 2) * p.tiles_n, sm_count() / 2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(Cfg::THREADS);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    le = cudaLaunchKernelEx(&cfg, igemm_tc_kernel<BN, BK, EW, GG, O16, UP, SK, TS, RS, C2>, a1, a2, b, o ? *o : b, b, p);
  } else
  le = launch_pdl(igemm_tc_kernel<BN, BK, EW, GG, O16, UP, SK, TS, RS>, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, st, a1,
                              a2, b, o ? *o : b, r ? *r : b, p);
  if (le != cudaSuccess) {
    set_error("igemm_tc: launch failed: %s", cudaGetErrorString(le));
    return LDMB200_ERR_CUDA;
  }
  LDM_CHECK_LAUNCH("igemm_tc");
  return 0;
}

template <int BN, int BK, int EW, bool GG = false>
static int launch_tc(const CUtensorMap& a1, const CUtensorMap& a2, const CUtensorMap& b, const TcParams& p,
                     cudaStream_t st) {
  if (p.out_dtype != LDMB200_F32) return launch_tc_impl<BN, BK, EW, GG, true>(a1, a2, b, p, st);
  return launch_tc_impl<BN, BK, EW, GG, false>(a1, a2, b, p, st);
}

static bool is_pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

}  // namespace ldm

using namespace ldm;

extern "C" int ldmb200_igemm_tc(const ldmb200_gemm_desc* d, void* stream) {
  LDM_CHECK_ARG(d && d->a && d->w && d->out, "igemm_tc: null pointer");
  LDM_CHECK_ARG(d->a_dtype == LDMB200_BF16 || d->a_dtype == LDMB200_F16,
                "igemm_tc: A operand must be bf16 or f16 (and the packed weights of the same type)");
  const bool fold = d->upsample != 0;
  LDM_CHECK_ARG(!fold || d->out_dtype == LDMB200_F32, "igemm_tc: the folded upsample convolution writes fp32");
  LDM_CHECK_ARG(!fold || (d->ksize == 3 && d->stride == 1 && !d->a2 && !d->residual && !d->rowvec && !d->geglu &&
                          !d->out_nchw && d->pad_mode == 0),
                "igemm_tc: upsample=1 is the folded nearest-2x + 3x3 convolution (one input, bias only, row-major output); "
                "w must be the four tap-summed 2x2 weight sets of ldmb200_pack_weights kind 4");
  LDM_CHECK_ARG(d->ksize == 1 || d->ksize == 3, "igemm_tc: ksize must be 1 or 3");
  LDM_CHECK_ARG(d->stride == 1 || (d->stride == 2 && d->ksize == 3), "igemm_tc: stride 2 only for 3x3");
  const int c1 = d->c1, c2 = d->a2 ? d->c2 : 0;
  LDM_CHECK_ARG(c1 > 0 && c1 % 16 == 0 && c2 % 16 == 0, "igemm_tc: channel counts must be multiples of 16 (c1=%d c2=%d)", c1, c2);
  LDM_CHECK_ARG(d->n > 0 && d->n % 16 == 0, "igemm_tc: n=%d must be a multiple of 16 (pad the packed weights)", d->n);
  const int h_out = d->stride == 2 ? d->h_in / 2 : d->h_in, w_out = d->stride == 2 ? d->w_in / 2 : d->w_in;
  const int hw_out = h_out * w_out;
  TcParams p;
  memset(&p, 0, sizeof(p));
  LDM_CHECK_ARG((int64_t)d->batch * hw_out > 0 && (int64_t)d->batch * hw_out < (1ll << 31), "igemm_tc: M out of range");
  p.M = d->batch * hw_out;          // (folded upsample: h_in x w_in IS the low-res grid the GEMM runs on)
  p.ks = fold ? 2 : d->ksize;
  p.taps = p.ks * p.ks;
  if (fold) {
    p.up_w = d->w_in;
    p.up_hw = hw_out;
    LDM_CHECK_ARG(!d->colstats || hw_out % 32 == 0, "igemm_tc: folded upsample with colstats needs h_in*w_in %% 32 == 0");
  }
  p.stride = d->stride;
  p.is_conv = d->ksize == 3;
  LDM_CHECK_ARG(d->pad_mode == 0 || (d->pad_mode == 1 && d->ksize == 3 && d->stride == 2),
                "igemm_tc: pad_mode 1 is the (0,1,0,1)-padded 3x3 stride-2 convolution");
  p.pad_asym = d->pad_mode == 1;
  p.ctot = c1 + c2;
  p.hw_out = hw_out;
  p.w_out = w_out;
  p.rows_per_sample = hw_out;
  p.bias = d->bias;
  p.rowvec = d->rowvec;
  p.rowvec_ld = d->rowvec_ld;
  p.residual = d->residual;
  p.residual_ld = d->residual_ld;
  p.geglu = d->geglu;
  p.out = d->out;
  p.out_dtype = d->out_dtype;
  p.rowvec_uniform = hw_out % 128 == 0;
  p.out_ld = d->out_ld;
  p.out_nchw = d->out_nchw;
  p.colstats = d->colstats;
  if (d->prefetch && d->prefetch_bytes > 0) {
    LDM_CHECK_ARG(((uintptr_t)d->prefetch & 15) == 0 && d->prefetch_bytes % 16 == 0,
                  "igemm_tc: prefetch range must be 16-byte aligned and a multiple of 16 bytes");
    p.pf_ptr = reinterpret_cast<const char*>(d->prefetch);
    p.pf_bytes = (unsigned long long)d->prefetch_bytes;
  }
  LDM_CHECK_ARG(!d->colstats || (d->out_dtype == LDMB200_F32 && !d->geglu && !d->out_nchw && p.M % 32 == 0 &&
                                 d->n % 32 == 0 && (d->n_valid <= 0 || d->n_valid == d->n)),
                "igemm_tc: colstats needs an fp32 row-major output with M %% 32 == 0 and n %% 32 == 0, unpadded");
  const int n_cols = d->geglu ? d->n / 2 : d->n;
  p.n_valid = d->n_valid > 0 ? d->n_valid : n_cols;
  LDM_CHECK_ARG(d->out_dtype == LDMB200_F32 || d->out_dtype == LDMB200_BF16 || d->out_dtype == LDMB200_F16,
                "igemm_tc: bad out_dtype");
  const bool out16 = d->out_dtype != LDMB200_F32;
  LDM_CHECK_ARG(d->out_nchw || p.n_valid % (out16 ? 8 : 4) == 0,
                "igemm_tc: n_valid must be a multiple of 4 (fp32 out) / 8 (16-bit out) unless out_nchw");
  LDM_CHECK_ARG(d->out_nchw || d->out_ld % (out16 ? 8 : 4) == 0, "igemm_tc: out_ld must keep rows 16-byte aligned");
  LDM_CHECK_ARG(!d->residual || d->residual_ld % 4 == 0, "igemm_tc: residual_ld must be a multiple of 4");
  LDM_CHECK_ARG(!d->rowvec || d->rowvec_ld % 4 == 0, "igemm_tc: rowvec_ld must be a multiple of 4");
  LDM_CHECK_ARG(!(d->out_nchw && (out16 || d->geglu || d->rowvec || d->residual)), "igemm_tc: out_nchw is a plain fp32 store");
  LDM_CHECK_ARG(!(d->geglu && (!d->bias || d->rowvec || d->residual)), "igemm_tc: geglu epilogue takes the packed bias only");

  // BK = 64 (128-byte swizzled rows) for every shape: a channel range that is not a multiple of 64 ends in a
  // partial chunk (TMA zero-fills the missing channels, the MMA warp issues only valid/16 instructions for it)
  constexpr int BK = 64;
  // 128-wide tiles also serve the plain GEMMs whose n is a multiple of 128 but not of 160 (the first-stage decoder's
  // 128/256/512-channel convolutions and its n x n attention scores); no UNet shape falls in that class
  int BN = d->geglu ? ((d->n % 128 == 0) ? 128 : 32)
                    : (d->n % 160 == 0) ? 160 : (d->n % 128 == 0) ? 128 : (d->n % 32 == 0) ? 32 : 16;
  LDM_CHECK_ARG(BN != 16 || d->n == 16, "igemm_tc: n=%d unsupported (need multiple of 32, or exactly 16)", d->n);
  LDM_CHECK_ARG(!(d->geglu && BN == 16), "igemm_tc: geglu needs n %% 32 == 0");
  p.c1 = c1;
  p.c1_chunks = ceil_div(c1, BK);
  p.c2_chunks = ceil_div(c2, BK);
  p.c1_tail_k = (c1 - (p.c1_chunks - 1) * BK) / 16;
  p.c2_tail_k = c2 ? (c2 - (p.c2_chunks - 1) * BK) / 16 : 0;
  p.tiles_m = ceil_div(p.M, 128);
  p.tiles_n = d->n / BN;
  p.idesc = umma_idesc_f16(128, BN, d->a_dtype == LDMB200_F16 ? 0 : 1);

  // ---- A tensor maps: {C, W, H, plane, B}, box {BK, bw, bh, 1, bn} with bw*bh*bn = 128 rows
  CUtensorMap tA1, tA2, tB;
  const int planes = d->stride == 2 ? 4 : 1;
  for (int which = 0; which < (c2 ? 2 : 1); ++which) {
    const void* base = which ? d->a2 : d->a;
    const uint64_t C = which ? c2 : c1;
    uint64_t dims[5], str[4];
    uint32_t box[5];
    if (p.is_conv) {
      LDM_CHECK_ARG(is_pow2(w_out) && is_pow2(h_out) && (w_out <= 128 || d->stride == 1),
                    "igemm_tc: 3x3 conv needs power-of-two output H,W (W <= 128 for stride 2) (got %dx%d)", h_out, w_out);
      const int bw = std::min(w_out, 128), bh = std::min(h_out, 128 / bw), bn = 128 / (bw * bh);
      dims[0] = C; dims[1] = w_out; dims[2] = h_out; dims[3] = planes; dims[4] = d->batch;
      str[0] = C * 2; str[1] = (uint64_t)w_out * C * 2; str[2] = (uint64_t)hw_out * C * 2; str[3] = (uint64_t)planes * hw_out * C * 2;
      box[0] = BK; box[1] = bw; box[2] = bh; box[3] = 1; box[4] = bn;
    } else {
      dims[0] = C; dims[1] = 1; dims[2] = (uint64_t)p.M; dims[3] = 1; dims[4] = 1;
      str[0] = C * 2; str[1] = C * 2; str[2] = (uint64_t)p.M * C * 2; str[3] = (uint64_t)p.M * C * 2;
      box[0] = BK; box[1] = 1; box[2] = 128; box[3] = 1; box[4] = 1;
    }
    int rc = make_tensor_map_bf16(which ? &tA2 : &tA1, base, 5, dims, str, box, BK * 2);
    if (rc) return rc;
  }
  if (!c2) tA2 = tA1;
  {
    uint64_t dims[2] = {(uint64_t)p.taps * p.ctot, (uint64_t)d->n * (fold ? 4 : 1)};
    uint64_t str[1] = {(uint64_t)p.taps * p.ctot * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)BN};
    int rc = make_tensor_map_bf16(&tB, d->w, 2, dims, str, box, BK * 2);
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (fold) {  // four launches, one per output parity; each reads its own weight set and writes its own pixels
    for (int par = 0; par < 4; ++par) {
      p.up_py = par >> 1;
      p.up_px = par & 1;
      p.b_row0 = par * d->n;
      int rc = (BN == 160) ? launch_tc_impl<160, 64, 8, false, false, true>(tA1, tA2, tB, p, st)
               : (BN == 128) ? launch_tc_impl<128, 64, 16, false, false, true>(tA1, tA2, tB, p, st)
               : (BN == 32) ? launch_tc_impl<32, 64, 8, false, false, true>(tA1, tA2, tB, p, st)
                            : launch_tc_impl<16, 64, 8, false, false, true>(tA1, tA2, tB, p, st);
      if (rc) return rc;
    }
    return 0;
  }
  // ---- split-K (opt-in: the caller passes a workspace): a grid that would occupy at most half of the SMs with a deep
  // reduction per tile (batch 1-4 sampling: M = 256 ... 4096 rows against K = 2560 ... 11520).  S ~ sqrt(0.6 kiters) splits,
  // at most 8, at most SMs / tiles, at least 4 iterations each, K >= 2560 (40 chunks of 64).  Measured r02 (ncu, batch 1,
  // 64x64 latents): M256 N640 K5760 32.0 -> 29.4 us with 7 splits, K11520 54 -> 33 us with 8; K <= 2880 LOSES (18 -> 23 us).
  // The main loops do shrink to ~8 us, but the CTA that arrives last then pulls S x 80 KB of partials through ONE SM's
  // load path in S dependent rounds (~20 us at S = 7: sm__cycles_active shows 48 of the 56 CTAs done after 15 K of the
  // 56 K cycles).  Whole eval under CUDA-graph replay: 3.98 -> 3.89 ms at batch 1, 6.85 -> 6.65 ms at batch 1 / 128x128
  // (tools/eval_time.py) — a 2-3 % gain, hence not the default; a cluster / DSMEM reduce-scatter fold is the next step.
  if (d->splitk_ws && BN == 160 && !d->geglu && !d->out_nchw) {
    const int kiters = p.taps * (p.c1_chunks + p.c2_chunks), tiles = p.tiles_m * p.tiles_n, sms = sm_count();
    static int tune_mink = 40, tune_max = 8;
    static double tune_coef = 0.6;
    static bool tune_read = false;
    if (!tune_read) {  // LDMB200_SPLITK_TUNE="min_kiters,coef,max_splits" (measurement knob)
      if (const char* e = getenv("LDMB200_SPLITK_TUNE")) sscanf(e, "%d,%lf,%d", &tune_mink, &tune_coef, &tune_max);
      tune_read = true;
    }
    if (tiles * 2 <= sms && kiters >= tune_mink) {
      const int64_t tile_bytes = (int64_t)128 * BN * 4;
      const int64_t cap = (d->splitk_ws_bytes - LDMB200_SPLITK_COUNTER_BYTES) / (tile_bytes * tiles);
      int s = (int)(sqrt(tune_coef * (double)kiters) + 0.5);
      s = std::min(std::min(s, tune_max), std::min(sms / tiles, kiters / 4));
      s = (int)std::min<int64_t>(s, cap);
      if (s >= 2 && tiles * 4 <= LDMB200_SPLITK_COUNTER_BYTES) {
        LDM_CHECK_ARG(((uintptr_t)d->splitk_ws & 15) == 0, "igemm_tc: splitk_ws must be 16-byte aligned");
        p.kiters_per_split = ceil_div(kiters, s);
        p.splits = ceil_div(kiters, p.kiters_per_split);
        p.sk_cnt = reinterpret_cast<unsigned int*>(d->splitk_ws);
        p.sk_part = reinterpret_cast<float*>(reinterpret_cast<char*>(d->splitk_ws) + LDMB200_SPLITK_COUNTER_BYTES);
        if (p.splits >= 2)
          return out16 ? launch_tc_impl<160, 64, 8, false, true, false, true>(tA1, tA2, tB, p, st)
                       : launch_tc_impl<160, 64, 8, false, false, false, true>(tA1, tA2, tB, p, st);
        p.splits = 0;
      }
    }
  }
  // ---- TMA-store epilogue: bias / per-tile row vector only, no statistics (fused QKV, proj_in, 1x1 skip convolutions)
  static const bool tma_store_on = !(getenv("LDMB200_TMA_STORE") && atoi(getenv("LDMB200_TMA_STORE")) == 0);
  if (tma_store_on && BN == 160 && !d->geglu && !d->residual && (!d->rowvec || p.rowvec_uniform) && !d->colstats &&
      !d->out_nchw && p.n_valid == d->n && d->out_ld % 8 == 0 && ((uintptr_t)d->out & 15) == 0) {
    CUtensorMap tO;
    const int eb = out16 ? 2 : 4;
    uint64_t dims[2] = {(uint64_t)d->n, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)d->out_ld * eb};
    uint32_t box[2] = {32u, 32u};
    int rc = make_tensor_map_bf16(&tO, d->out, 2, dims, str, box, out16 ? 64 : 128, eb);
    if (rc) return rc;
    return out16 ? launch_tc_impl<160, 64, 8, false, true, false, false, true>(tA1, tA2, tB, p, st, &tO)
                 : launch_tc_impl<160, 64, 8, false, false, false, false, true>(tA1, tA2, tB, p, st, &tO);
  }
  // ---- the same with an fp32 residual (attn1.to_out, ff.net.2 of the 160-channel level), K <= 640 only: the 96 KB of
  // staging leave three operand stages.  Measured r02s, same run: M262144 N160 K160 0.926 -> 0.853 ms (10 launches),
  // N160 K640 0.559 -> 0.529, M65536 N320 K320 0.511 -> 0.486, M16384 N640 K640 0.426 -> 0.412; K = 1280 / 2560 LOSE on
  // three stages (0.350 -> 0.401, 0.383 -> 0.469 ms) and stay on the per-lane path.
  static const bool tma_res_on = !(getenv("LDMB200_TMA_RES") && atoi(getenv("LDMB200_TMA_RES")) == 0);
  if (tma_store_on && tma_res_on && BN == 160 && d->ksize == 1 && d->residual && !d->geglu && (!d->rowvec || p.rowvec_uniform) &&
      !d->colstats && !d->out_nchw && p.n_valid == d->n && d->out_ld % 8 == 0 && ((uintptr_t)d->out & 15) == 0 &&
      ((uintptr_t)d->residual & 15) == 0 && d->residual_ld % 4 == 0 && p.ctot <= 640) {
    CUtensorMap tO, tR;
    const int eb = out16 ? 2 : 4;
    uint64_t dims[2] = {(uint64_t)d->n, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)d->out_ld * eb};
    uint32_t box[2] = {32u, 32u};
    int rc = make_tensor_map_bf16(&tO, d->out, 2, dims, str, box, out16 ? 64 : 128, eb);
    if (rc) return rc;
    uint64_t rstr[1] = {(uint64_t)d->residual_ld * 4};
    rc = make_tensor_map_bf16(&tR, d->residual, 2, dims, rstr, box, 128, 4);
    if (rc) return rc;
    return out16 ? launch_tc_impl<160, 64, 8, false, true, false, false, true, true>(tA1, tA2, tB, p, st, &tO, &tR)
                 : launch_tc_impl<160, 64, 8, false, false, false, false, true, true>(tA1, tA2, tB, p, st, &tO, &tR);
  }
  // ---- GEGLU with a 16-bit output: TMA store of each warp's 32 x 16 chunk
  static const bool tma_gg_on = !(getenv("LDMB200_TMA_GEGLU") && atoi(getenv("LDMB200_TMA_GEGLU")) == 0);
  if (tma_store_on && tma_gg_on && BN == 128 && d->geglu && out16 && p.n_valid == d->n / 2 && d->out_ld % 8 == 0 &&
      ((uintptr_t)d->out & 15) == 0) {
    CUtensorMap tO;
    uint64_t dims[2] = {(uint64_t)(d->n / 2), (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)d->out_ld * 2};
    uint32_t box[2] = {16u, 32u};
    int rc = make_tensor_map_bf16(&tO, d->out, 2, dims, str, box, 32, 2);
    if (rc) return rc;
    // CTA pair for the GEGLU GEMMs (24 KB instead of 32 KB per 64-deep chunk and CTA): measured r02s and NOT taken by
    // default — N5120 K640 0.655 -> 0.680 ms (6 launches), N2560 K320 0.673 -> 0.692: these GEMMs are bound by the erf-GELU
    // epilogue, not by the operand feed, and the pair only adds its synchronisation.  LDMB200_C2_GG_MINK=320 turns it on.
    static const int c2g_on = getenv("LDMB200_C2") ? atoi(getenv("LDMB200_C2")) : 1;
    static const int c2g_mink = getenv("LDMB200_C2_GG_MINK") ? atoi(getenv("LDMB200_C2_GG_MINK")) : (1 << 30);
    if (c2g_on && p.tiles_m % 2 == 0 && p.ctot >= c2g_mink && sm_count() >= 2 && p.tiles_m * p.tiles_n >= 2 * sm_count()) {
      CUtensorMap tBh;
      uint64_t bdims[2] = {(uint64_t)p.taps * p.ctot, (uint64_t)d->n};
      uint64_t bstr[1] = {(uint64_t)p.taps * p.ctot * 2};
      uint32_t bbox[2] = {(uint32_t)BK, 64u};
      rc = make_tensor_map_bf16(&tBh, d->w, 2, bdims, bstr, bbox, BK * 2);
      if (rc) return rc;
      p.idesc = umma_idesc_f16(256, 128, d->a_dtype == LDMB200_F16 ? 0 : 1);
      return launch_tc_impl<128, 64, 16, true, true, false, false, true, false, true>(tA1, tA2, tBh, p, st, &tO);
    }
    return launch_tc_impl<128, 64, 16, true, true, false, false, true>(tA1, tA2, tB, p, st, &tO);
  }
  // ---- CTA pair (cta_group::2, M = 256): the deep reductions (3x3 convolutions, K >= 1280 linears) on 160-wide tiles
  static const int c2_on = getenv("LDMB200_C2") ? atoi(getenv("LDMB200_C2")) : 1;
  // Measured r02s (same run, one eval at batch 64): 25.41 -> 24.91 ms with K >= 1280, 24.78 with K >= 640 (every shape gains
  // 3-12 %: N=160 K=1440 757 -> 854 TFLOP/s).  Below two waves of tiles the pairs lose (batch 1-8: +2 % per eval: half as
  // many independent schedulable units and a cluster barrier at both ends), so small grids stay on single CTAs.
  static const int c2_mink = getenv("LDMB200_C2_MINK") ? atoi(getenv("LDMB200_C2_MINK")) : 640;
  static const int c2_waves = getenv("LDMB200_C2_WAVES") ? atoi(getenv("LDMB200_C2_WAVES")) : 2;
  if (c2_on && BN == 160 && !d->geglu && !d->out_nchw && p.tiles_m % 2 == 0 && p.taps * p.ctot >= c2_mink && sm_count() >= 2 &&
      p.tiles_m * p.tiles_n >= c2_waves * sm_count()) {
    CUtensorMap tBh;  // half of the B tile per CTA
    uint64_t dims[2] = {(uint64_t)p.taps * p.ctot, (uint64_t)d->n};
    uint64_t str[1] = {(uint64_t)p.taps * p.ctot * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)(BN / 2)};
    int rc = make_tensor_map_bf16(&tBh, d->w, 2, dims, str, box, BK * 2);
    if (rc) return rc;
    static bool c2_unavailable = false;  // a device / partition that cannot co-schedule 2-clusters: single CTAs from then on
    if (!c2_unavailable) {
      TcParams p2 = p;
      p2.idesc = umma_idesc_f16(256, BN, d->a_dtype == LDMB200_F16 ? 0 : 1);
      rc = out16 ? launch_tc_impl<160, 64, 8, false, true, false, false, false, false, true>(tA1, tA2, tBh, p2, st)
                 : launch_tc_impl<160, 64, 8, false, false, false, false, false, false, true>(tA1, tA2, tBh, p2, st);
      if (rc == 0) return 0;
      cudaGetLastError();   // the launch was refused (nothing ran): clear it and take the single-CTA kernel below
      c2_unavailable = true;
    }
  }
  if (BN == 128) return d->geglu ? launch_tc<128, 64, 16, true>(tA1, tA2, tB, p, st) : launch_tc<128, 64, 16>(tA1, tA2, tB, p, st);
  LDM_CHECK_ARG(!d->geglu || d->n % 32 == 0, "igemm_tc: geglu needs n %% 32 == 0 (got n=%d)", d->n);
  if (d->geglu) return launch_tc<32, 64, 8, true>(tA1, tA2, tB, p, st);
  // (16 epilogue warps on the plain 160-wide tiles were measured r01n and are SLOWER, 32.1 vs 30.0 ms per UNet eval —
  // 5 chunks over 4 warps per quarter is unbalanced and the ring loses a stage — so that instantiation is not built)
  if (BN == 160) return launch_tc<160, 64, 8>(tA1, tA2, tB, p, st);
  if (BN == 32) return launch_tc<32, 64, 8>(tA1, tA2, tB, p, st);
  return launch_tc<16, 64, 8>(tA1, tA2, tB, p, st);
}
