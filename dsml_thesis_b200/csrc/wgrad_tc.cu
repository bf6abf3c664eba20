// Weight gradient on tcgen05 / TMEM / TMA (training step):
//   dw[n, c, tap] += sum_m dy[m, n] * X(m, tap, c)        bf16 x bf16 -> fp32 in tensor memory, split over the pixels
//
// Both operands are consumed in their natural [pixel][channel] layouts as MN-MAJOR UMMA operands: a TMA box of
// {64 channels, 64 pixels} with the 128-byte swizzle IS the canonical MN-major SWIZZLE_128B atom stack
// (cute/atom/mma_traits_sm100.hpp: ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units), so dy^T (A, M = 128 output
// channels = two atoms) and the tap-shifted X (B, N <= 256 input channels = up to four atoms) need no transposes.
// The 3x3 taps are coordinate shifts of the same 5-D tensor map the forward uses ({C, W, H, plane, B}; halo = TMA zero
// fill; stride 2 = space-to-depth planes).
//   CTA = one (<=256-wide n tile = two M = 128 accumulators sharing the X slabs, <=256-wide c tile, tap, pixel split);
//   warp 0 TMA producer, warp 1 MMA issuer (K = 16 pixels per tcgen05.mma, 4 per 64-pixel stage and accumulator),
//   warps 2-5 epilogue: tcgen05.ld -> transpose through smem -> row-contiguous fp32 reductions into dw.
#include <algorithm>

#include "common.cuh"

namespace ldm {

constexpr int WT_KP = 64, WT_STAGES = 3;
constexpr int WT_ATOM = 64 * WT_KP * 2;          // 8 KiB: 64 channels x 64 pixels
constexpr int WT_A_BYTES = 4 * WT_ATOM;          // up to 256 output channels = two M = 128 accumulator tiles
constexpr int WT_B_BYTES = 4 * WT_ATOM;          // up to 256 input channels
constexpr int WT_STAGE = WT_A_BYTES + WT_B_BYTES;
constexpr int WT_SMEM = WT_STAGES * WT_STAGE + (2 * WT_STAGES + 1) * 8 + 16 + 1024;
constexpr int WT_THREADS = 192;
constexpr int WT_TMEM_COLS = 512;           // two accumulator tiles of up to 256 fp32 columns

struct WtParams {
  float* dw;
  int64_t n_stride, c_stride, tap_stride;
  int n, c, c_span;   // c_span: channels covered by the c tiles (c_ld rounded to 32)
  int M, hw, w, is_conv, ks, stride, taps;
  int wide;           // w >= 64: a 64-pixel slab is part of one image row
  int tiles_c, splits, chunks, chunks_per_split;
};

__global__ void __launch_bounds__(WT_THREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX, const WtParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // (offset, not uintptr_t: keeps the shared address space)
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + WT_STAGES * WT_STAGE);
  uint64_t* empty = full + WT_STAGES;
  uint64_t* tfull = empty + WT_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int unit = blockIdx.x;
  const int split = unit % p.splits; unit /= p.splits;
  const int tap = unit % p.taps; unit /= p.taps;
  const int tc = unit % p.tiles_c, tn = unit / p.tiles_c;
  const int n0 = tn * 256, c0 = tc * 256;
  const int n_mt = (p.n - n0 > 128) ? 2 : 1;  // accumulator tiles (128 output channels each)
  const int cw = min(256, p.c_span - c0);     // multiple of 32
  const int n_atoms_b = (cw + 63) >> 6;
  const int chunk0 = split * p.chunks_per_split, chunk1 = min(p.chunks, chunk0 + p.chunks_per_split);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmDY);
    prefetch_tmap(&tmX);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < WT_STAGES; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], 1);
      }
      mbar_init(tfull, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, WT_TMEM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    int dh = 0, dw_ = 0, plane = 0;
    if (p.ks == 3) {
      const int kh = tap / 3, kw = tap % 3;
      if (p.stride == 1) {
        dh = kh - 1;
        dw_ = kw - 1;
      } else {
        const int ph = (kh == 1) ? 0 : 1, pw = (kw == 1) ? 0 : 1;
        dh = (kh == 0) ? -1 : 0;
        dw_ = (kw == 0) ? -1 : 0;
        plane = ph * 2 + pw;
      }
    }
    const uint32_t tx = (2 * n_mt + n_atoms_b) * WT_ATOM;
    for (int ch = chunk0, it = 0; ch < chunk1; ++ch, ++it) {
      const int s = it % WT_STAGES;
      mbar_wait(&empty[s], ((it / WT_STAGES) & 1) ^ 1);
      if (elect_one()) {
        uint8_t* sa = smem + s * WT_STAGE;
        uint8_t* sb = sa + WT_A_BYTES;
        const int m0 = ch * WT_KP;
        mbar_expect_tx(&full[s], tx);
        for (int a = 0; a < 2 * n_mt; ++a) tma_load_2d(sa + a * WT_ATOM, &tmDY, &full[s], n0 + a * 64, m0);
        if (p.is_conv) {
          const int n_img = m0 / p.hw, rem = m0 % p.hw;
          const int h0 = rem / p.w, w0 = p.wide ? rem % p.w : 0;
          for (int a = 0; a < n_atoms_b; ++a)
            tma_load_5d(sb + a * WT_ATOM, &tmX, &full[s], c0 + a * 64, w0 + dw_, h0 + dh, plane, n_img);
        } else {
          for (int a = 0; a < n_atoms_b; ++a)
            tma_load_5d(sb + a * WT_ATOM, &tmX, &full[s], c0 + a * 64, 0, m0, 0, 0);
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // MN-major SWIZZLE_128B descriptors: SBO = 1024 B (8 pixel rows), LBO = one 8 KiB atom (next 64 channels)
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t lbo = (uint32_t)(WT_ATOM >> 4) << 16;
    const uint32_t idesc = umma_idesc_f16(128, cw, 1, 1, 1);
    uint32_t acc = 0;
    for (int ch = chunk0, it = 0; ch < chunk1; ++ch, ++it) {
      const int s = it % WT_STAGES;
      mbar_wait(&full[s], (it / WT_STAGES) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_lo = ((smem_u32(smem + s * WT_STAGE) & 0x3FFFFu) >> 4) | lbo;
        const uint32_t b_lo = ((smem_u32(smem + s * WT_STAGE + WT_A_BYTES) & 0x3FFFFu) >> 4) | lbo;
        for (int mt = 0; mt < n_mt; ++mt) {
#pragma unroll
          for (int kk = 0; kk < WT_KP / 16; ++kk)   // 16 pixels = 16 rows x 128 B = +128 in 16-byte units
            umma_bf16_ss(tmem_base + mt * 256, ((uint64_t)desc_hi << 32) | (a_lo + mt * (2 * WT_ATOM >> 4) + kk * 128),
                         ((uint64_t)desc_hi << 32) | (b_lo + kk * 128), idesc, (acc | kk) ? 1u : 0u);
        }
        umma_commit(&empty[s]);
      }
      __syncwarp();
      acc = 1;
    }
    if (elect_one()) umma_commit(tfull);
    __syncwarp();
  } else {
    // ===================== epilogue: TMEM lane quarter -> smem transpose -> row-contiguous reductions ==========
    const int qw = warp & 3;
    mbar_wait(tfull, 0);   // every MMA has retired: the operand ring is free and serves as the transpose buffer
    tc_fence_after();
    float* tr = reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 33);
    for (int mt = 0; mt < n_mt; ++mt) {
      const int nbase = n0 + mt * 128 + qw * 32;
      for (int cc = 0; cc * 32 < cw; ++cc) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(qw * 32) << 16) + mt * 256 + cc * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) tr[lane * 33 + j] = v[j];
        __syncwarp();
        const int ch = c0 + cc * 32 + lane;
        if (ch < p.c) {
          float* col = p.dw + (size_t)ch * p.c_stride + (size_t)tap * p.tap_stride;
          const int rmax = min(32, p.n - nbase);
          for (int r = 0; r < rmax; ++r) atomicAdd(col + (size_t)(nbase + r) * p.n_stride, tr[r * 33 + lane]);
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, WT_TMEM_COLS);
  }
}

static bool wt_pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

// returns 1 when the shape is not supported by the tensor-memory kernel (the caller falls back), 0 on success
int wgrad_tc_launch(const ldmb200_wgrad_desc* d, cudaStream_t st) {
  const bool conv = d->ksize == 3;
  if (conv && !(wt_pow2(d->w) && wt_pow2(d->h))) return 1;
  if (d->c_ld % 32 != 0 && d->c_ld != 16) return 1;
  const int64_t M = (int64_t)d->batch * d->h * d->w;
  WtParams p;
  memset(&p, 0, sizeof(p));
  p.dw = d->dw;
  p.n_stride = d->n_stride; p.c_stride = d->c_stride; p.tap_stride = d->tap_stride;
  p.n = d->n; p.c = d->c;
  p.c_span = (std::min(d->c_ld, (d->c + 31) / 32 * 32) + 31) / 32 * 32;
  p.M = (int)M; p.hw = d->h * d->w; p.w = d->w; p.is_conv = conv; p.ks = d->ksize; p.stride = d->stride;
  p.taps = d->ksize * d->ksize;
  p.wide = d->w >= 64;
  const int tiles_n = ceil_div(d->n, 256);
  p.tiles_c = ceil_div(p.c_span, 256);
  p.chunks = ceil_div((int)M, WT_KP);
  const int base_units = tiles_n * p.tiles_c * p.taps;
  int splits = ceil_div(sm_count(), base_units);
  splits = std::max(1, std::min(splits, ceil_div(p.chunks, 4)));
  p.chunks_per_split = ceil_div(p.chunks, splits);
  p.splits = ceil_div(p.chunks, p.chunks_per_split);

  CUtensorMap tDY, tX;
  {
    // columns >= n are out of bounds (zero fill): the pointer may be a column offset into a wider matrix (fused QKV)
    uint64_t dims[2] = {(uint64_t)d->n, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)d->dy_ld * 2};
    uint32_t box[2] = {64, WT_KP};
    int rc = make_tensor_map_bf16(&tDY, d->dy, 2, dims, str, box, 128);
    if (rc) return rc;
  }
  {
    const uint64_t C = d->c_ld;
    uint64_t dims[5], str[4];
    uint32_t box[5];
    if (conv) {
      const int planes = d->stride == 2 ? 4 : 1;
      const int bw = std::min(d->w, WT_KP), bh = std::min(d->h, WT_KP / bw), bn = WT_KP / (bw * bh);
      dims[0] = C; dims[1] = d->w; dims[2] = d->h; dims[3] = planes; dims[4] = d->batch;
      str[0] = C * 2; str[1] = (uint64_t)d->w * C * 2; str[2] = (uint64_t)p.hw * C * 2; str[3] = (uint64_t)planes * p.hw * C * 2;
      box[0] = 64; box[1] = bw; box[2] = bh; box[3] = 1; box[4] = bn;
    } else {
      dims[0] = C; dims[1] = 1; dims[2] = (uint64_t)M; dims[3] = 1; dims[4] = 1;
      str[0] = C * 2; str[1] = C * 2; str[2] = (uint64_t)M * C * 2; str[3] = (uint64_t)M * C * 2;
      box[0] = 64; box[1] = 1; box[2] = WT_KP; box[3] = 1; box[4] = 1;
    }
    int rc = make_tensor_map_bf16(&tX, d->x, 5, dims, str, box, 128);
    if (rc) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WT_SMEM);
    if (e != cudaSuccess) {
      set_error("wgrad_tc: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
      return LDMB200_ERR_CUDA;
    }
    attr_set = true;
  }
  const int64_t grid = (int64_t)base_units * p.splits;
  wgrad_tc_kernel<<<(unsigned)grid, WT_THREADS, WT_SMEM, st>>>(tDY, tX, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("wgrad_tc: launch failed: %s", cudaGetErrorString(e));
    return LDMB200_ERR_CUDA;
  }
  return 0;
}

}  // namespace ldm
