// Weight gradient of the convolutions / linears of the UNet (training step):
//   dw[n, c, tap] += sum_m dy[m, n] * X(m, tap, c)          bf16 x bf16 -> fp32, reduction over the B*H*W pixels
// Round-1 implementation on the warp-level tensor-core path (mma.sync through nvcuda::wmma; both operands are read
// in their natural [pixel][channel] layout: dy^T is a col-major A, X a row-major B).  A CTA owns one
// (160-wide n tile, 64-wide c tile, tap, pixel split) unit, streams 32-pixel slabs of dy and of the tap-shifted X
// through a 3-stage cp.async ring (zero-filled halo = the conv padding) and adds its partial tile into dw with
// fp32 atomics.  The tcgen05 version (MN-major UMMA operands straight from the TMA boxes) is the next step.
#include <algorithm>
#include <stdlib.h>

#include <mma.h>

#include "common.cuh"

namespace ldm {

using namespace nvcuda;

constexpr int WG_BN = 160, WG_BC = 64, WG_KP = 32, WG_STAGES = 3;
constexpr int WG_WARPS_N = 5, WG_WARPS_C = 2, WG_THREADS = 32 * WG_WARPS_N * WG_WARPS_C;
constexpr int WG_PA = WG_BN + 8, WG_PB = WG_BC + 8;  // smem row pitches (elements): 16-byte multiples, bank-skewed
constexpr int WG_A_BYTES = WG_KP * WG_PA * 2, WG_B_BYTES = WG_KP * WG_PB * 2;
constexpr int WG_STAGE_BYTES = WG_A_BYTES + WG_B_BYTES;
constexpr int WG_EPI_PITCH = WG_BC + 4;
constexpr int WG_SMEM = std::max(WG_STAGES * WG_STAGE_BYTES, WG_BN * WG_EPI_PITCH * 4);

struct WgParams {
  const uint16_t* dy;
  const uint16_t* x;
  float* dw;
  int64_t n_stride, c_stride, tap_stride;
  int dy_ld, n, c_ld, c;
  int batch, h, w, M;
  int ks, stride, taps;
  int tiles_n, tiles_c, splits, chunks, chunks_per_split;
};

__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem, bool valid) {
  const int sz = valid ? 16 : 0;  // src-size 0: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem)), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(WG_THREADS) wgrad_mma_kernel(const WgParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int wn = warp % WG_WARPS_N, wc = warp / WG_WARPS_N;
  int unit = blockIdx.x;
  const int split = unit % p.splits; unit /= p.splits;
  const int tap = unit % p.taps; unit /= p.taps;
  const int tc = unit % p.tiles_c, tn = unit / p.tiles_c;
  const int n0 = tn * WG_BN, c0 = tc * WG_BC;
  const int kh = p.ks == 3 ? tap / 3 : 1, kw = p.ks == 3 ? tap % 3 : 1;
  const int chunk0 = split * p.chunks_per_split, chunk1 = min(p.chunks, chunk0 + p.chunks_per_split);
  const int hw = p.h * p.w;

  auto load_stage = [&](int chunk, int s) {
    uint8_t* sa = smem + s * WG_STAGE_BYTES;
    uint8_t* sb = sa + WG_A_BYTES;
    const int m0 = chunk * WG_KP;
    // dy slab: WG_KP rows x 160 columns = 20 16-byte chunks per row
    for (int i = threadIdx.x; i < WG_KP * (WG_BN / 8); i += WG_THREADS) {
      const int r = i / (WG_BN / 8), ch = i % (WG_BN / 8);
      const int m = m0 + r, col = n0 + ch * 8;
      const bool ok = m < p.M && col < p.dy_ld;
      cp_async_16(sa + (r * WG_PA + ch * 8) * 2, ok ? p.dy + (size_t)m * p.dy_ld + col : p.dy, ok);
    }
    // X slab: WG_KP rows x 64 channels = 8 chunks per row, shifted by the tap (zero-filled outside the image)
    for (int i = threadIdx.x; i < WG_KP * (WG_BC / 8); i += WG_THREADS) {
      const int r = i / (WG_BC / 8), ch = i % (WG_BC / 8);
      const int m = m0 + r, col = c0 + ch * 8;
      bool ok = m < p.M && col < p.c_ld;
      const uint16_t* src = p.x;
      if (ok) {
        if (p.ks == 1) {
          src = p.x + (size_t)m * p.c_ld + col;
        } else {
          const int b = m / hw, rem = m % hw;
          const int oh = rem / p.w, ow = rem % p.w;
          if (p.stride == 1) {
            const int ih = oh + kh - 1, iw = ow + kw - 1;
            ok = ih >= 0 && ih < p.h && iw >= 0 && iw < p.w;
            if (ok) src = p.x + (((size_t)b * p.h + ih) * p.w + iw) * p.c_ld + col;
          } else {
            const int ih = 2 * oh + kh - 1, iw = 2 * ow + kw - 1;
            ok = ih >= 0 && iw >= 0;
            if (ok) {
              const int plane = (ih & 1) * 2 + (iw & 1);
              src = p.x + ((((size_t)b * 4 + plane) * p.h + (ih >> 1)) * p.w + (iw >> 1)) * p.c_ld + col;
            }
          }
        }
      }
      cp_async_16(sb + (r * WG_PB + ch * 8) * 2, src, ok);
    }
  };

  wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) wmma::fill_fragment(acc[i][j], 0.f);
  const bool warp_active = n0 + wn * 32 < p.n && c0 + wc * 32 < p.c;

  const int nchunks = chunk1 - chunk0;
#pragma unroll
  for (int s = 0; s < WG_STAGES - 1; ++s) {
    if (s < nchunks) load_stage(chunk0 + s, s);
    cp_async_commit();
  }
  for (int it = 0; it < nchunks; ++it) {
    cp_async_wait<WG_STAGES - 2>();
    __syncthreads();
    {  // prefetch the slab WG_STAGES-1 ahead into the stage that was consumed in the previous iteration
      const int nx = it + WG_STAGES - 1;
      if (nx < nchunks) load_stage(chunk0 + nx, nx % WG_STAGES);
      cp_async_commit();
    }
    if (warp_active) {
      const __nv_bfloat16* sa = reinterpret_cast<const __nv_bfloat16*>(smem + (it % WG_STAGES) * WG_STAGE_BYTES);
      const __nv_bfloat16* sb = reinterpret_cast<const __nv_bfloat16*>(smem + (it % WG_STAGES) * WG_STAGE_BYTES + WG_A_BYTES);
#pragma unroll
      for (int k = 0; k < WG_KP; k += 16) {
        wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> fa[2];
        wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> fb[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) wmma::load_matrix_sync(fa[i], sa + k * WG_PA + wn * 32 + i * 16, WG_PA);
#pragma unroll
        for (int j = 0; j < 2; ++j) wmma::load_matrix_sync(fb[j], sb + k * WG_PB + wc * 32 + j * 16, WG_PB);
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int j = 0; j < 2; ++j) wmma::mma_sync(acc[i][j], fa[i], fb[j], acc[i][j]);
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();
  // epilogue: tile -> smem (fp32) -> atomics into dw
  float* epi = reinterpret_cast<float*>(smem);
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
      wmma::store_matrix_sync(epi + (wn * 32 + i * 16) * WG_EPI_PITCH + wc * 32 + j * 16, acc[i][j], WG_EPI_PITCH,
                              wmma::mem_row_major);
  __syncthreads();
  for (int i = threadIdx.x; i < WG_BN * WG_BC; i += WG_THREADS) {
    const int r = i / WG_BC, cc = i % WG_BC;
    const int nn = n0 + r, ch = c0 + cc;
    if (nn < p.n && ch < p.c)
      atomicAdd(p.dw + (size_t)nn * p.n_stride + (size_t)ch * p.c_stride + (size_t)tap * p.tap_stride,
                epi[r * WG_EPI_PITCH + cc]);
  }
}

}  // namespace ldm

namespace ldm {
int wgrad_tc_launch(const ldmb200_wgrad_desc* d, cudaStream_t st);  // wgrad_tc.cu
}
using namespace ldm;

extern "C" int ldmb200_wgrad(const ldmb200_wgrad_desc* d, void* stream) {
  LDM_CHECK_ARG(d && d->dy && d->x && d->dw, "wgrad: null pointer");
  LDM_CHECK_ARG(d->ksize == 1 || d->ksize == 3, "wgrad: ksize must be 1 or 3");
  LDM_CHECK_ARG(d->stride == 1 || (d->stride == 2 && d->ksize == 3), "wgrad: stride 2 only for 3x3");
  LDM_CHECK_ARG(d->dy_ld % 8 == 0 && d->c_ld % 8 == 0, "wgrad: row pitches must be multiples of 8 elements (16 bytes)");
  LDM_CHECK_ARG(d->n > 0 && d->n <= d->dy_ld && d->c > 0 && d->c <= d->c_ld, "wgrad: n / c out of range");
  const int64_t M = (int64_t)d->batch * d->h * d->w;
  LDM_CHECK_ARG(M > 0 && M < (1ll << 31), "wgrad: M out of range");
  // tcgen05 kernel first (power-of-two maps); LDMB200_WGRAD_MMA=1 forces the warp-level kernel (A/B comparisons)
  static const bool force_mma = getenv("LDMB200_WGRAD_MMA") && atoi(getenv("LDMB200_WGRAD_MMA")) != 0;
  if (!force_mma) {
    const int rc = wgrad_tc_launch(d, (cudaStream_t)stream);
    if (rc <= 0) return rc;
  }
  WgParams p;
  p.dy = (const uint16_t*)d->dy;
  p.x = (const uint16_t*)d->x;
  p.dw = d->dw;
  p.n_stride = d->n_stride; p.c_stride = d->c_stride; p.tap_stride = d->tap_stride;
  p.dy_ld = d->dy_ld; p.n = d->n; p.c_ld = d->c_ld; p.c = d->c;
  p.batch = d->batch; p.h = d->h; p.w = d->w; p.M = (int)M;
  p.ks = d->ksize; p.stride = d->stride; p.taps = d->ksize * d->ksize;
  p.tiles_n = ceil_div(d->n, WG_BN);
  p.tiles_c = ceil_div(d->c, WG_BC);
  p.chunks = ceil_div((int)M, WG_KP);
  const int base_units = p.tiles_n * p.tiles_c * p.taps;
  int splits = ceil_div(sm_count() * 4, base_units);
  splits = std::max(1, std::min(splits, ceil_div(p.chunks, 8)));
  p.chunks_per_split = ceil_div(p.chunks, splits);
  p.splits = ceil_div(p.chunks, p.chunks_per_split);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM);
    if (e != cudaSuccess) {
      set_error("wgrad: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
      return LDMB200_ERR_CUDA;
    }
    attr_set = true;
  }
  const int64_t grid = (int64_t)base_units * p.splits;
  LDM_CHECK_ARG(grid < (1ll << 31), "wgrad: grid too large");
  wgrad_mma_kernel<<<(unsigned)grid, WG_THREADS, WG_SMEM, (cudaStream_t)stream>>>(p);
  LDM_CHECK_LAUNCH("wgrad");
  return 0;
}
