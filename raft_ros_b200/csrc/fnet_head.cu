// Producer-side fusion: the feature encoder's output layer -- a 1x1 convolution, `conv2` of
// BasicEncoder / SmallEncoder (reference core/extractor.py:144,184 and :217,256: 128 -> 256, small
// 96 -> 128) -- emitting its result directly as the K-major bf16 hi / lo operand maps that the
// tensor-core build reads through TMA.  The reference writes the feature maps as fp32 NCHW
// (core/raft.py:99-103, `.float()`), and the build's layout prepass (build_tc.cu: prep_kernel)
// re-reads them to transpose and split; here the fp32 (B, 256, h, w) maps never exist:
//
//     y[b, p, o] = bias[o] + sum_c W[o, c] * x[b, c, p]      ->  hi = bf16(y), lo = bf16(y - hi)
//
// It is a GEMM with M = pixels, N = Cout, K = Cin on tcgen05: converter warps read x (NCHW fp32,
// coalesced along the pixels), split it into bf16 hi + lo and write both as K-major SWIZZLE_128B
// operand rows into shared memory; the weights (hi + lo, converted once per model) stay resident in
// shared memory for the whole kernel; three MMAs per K step (hi*hi + lo*hi + hi*lo, fp32 accumulator in
// TMEM) give fp32-class products, the same error-compensated arithmetic as RC_PREC_BF16X3 of the build.
// The epilogue adds the bias, splits again and writes (B, h*w, Cout) bf16 rows, 512 B per pixel.
//
// One CTA per SM, persistent over tiles of 128 pixels, 13 warps:
//   warps 0-7   converters (thread = pixel x half of the input channels)
//   warps 8-11  epilogue (thread <-> pixel = TMEM lane)
//   warp 12     weights (one TMA load at start) and MMA issue
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace fh {

using namespace tc;

constexpr int kConvWarps = 8, kEpiWarps = 4, kMmaWarp = 12, kThreads = 13 * 32;
constexpr int TMPIX = 128;
constexpr int A_KBLOCK = TMPIX * 64 * 2;  // 16 KB: 128 pixels x 64 channels bf16
constexpr int kMaxKB = 2;                 // Cin <= 128
constexpr int kMaxCout = 256;

struct Params {
  const float *x;       // (Bt, Cin, HW) fp32
  const float *bias;    // (Cout)
  __nv_bfloat16 *hi;    // (Bt, HW, Cout)
  __nv_bfloat16 *lo;    // same, or null
  int Bt, Cin, Cout, HW, KB;
  int tiles_per_b, total_tiles;
};

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t *>(&v);
}
__device__ __forceinline__ void sts128u(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
// one whole 32 B sector per thread (16 B pieces make the L1 send every sector twice, half masked)
__device__ __forceinline__ void stg256u(void *p, const uint32_t (&a)[4], const uint32_t (&b)[4]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]),
               "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3])
               : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
conv1x1_prep_kernel(const __grid_constant__ CUtensorMap tmW, const Params P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  const int KB = P.KB, Cout = P.Cout;
  const uint32_t w_block = (uint32_t)Cout * 128u;           // one [Cout][64] weight block
  // layout: A hi [KB] | A lo [KB] | W hi [KB] | W lo [KB] | bias | barriers
  const uint32_t a_u32 = base_u32, w_u32 = base_u32 + 2 * kMaxKB * A_KBLOCK;
  float *bias_s = reinterpret_cast<float *>(smem + 2 * kMaxKB * A_KBLOCK + 2 * kMaxKB * kMaxCout * 128);
  uint64_t *bars = reinterpret_cast<uint64_t *>(bias_s + kMaxCout);
  uint64_t *w_full = bars, *a_full = bars + 1, *a_empty = bars + 2, *acc_full = bars + 3, *acc_empty = bars + 5;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 7);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();

  if (warp == kMmaWarp && lane == 0) {
    mbar_init(w_full, 1);
    mbar_init(a_full, kConvWarps);
    mbar_init(a_empty, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kEpiWarps);
    }
    fence_barrier_init();
    prefetch_tmap(&tmW);
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  // channels Cin .. 64*KB-1 of the x operand are padding: zero, never written again
  for (int i = threadIdx.x; i < 2 * kMaxKB * A_KBLOCK / 16; i += kThreads) sts128(a_u32 + i * 16, 0.f, 0.f, 0.f, 0.f);
  fence_proxy_async_smem();
  pdl_wait();
  for (int i = threadIdx.x; i < kMaxCout; i += kThreads) bias_s[i] = i < Cout ? __ldg(P.bias + i) : 0.0f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kConvWarps) {
    // ===================== converters: x (fp32, pixel-contiguous) -> bf16 hi / lo operand rows =====================
    const int prow = (warp & 3) * 32 + lane;   // pixel of the tile = operand row
    const int chalf = warp >> 2;               // this thread's half of the input channels
    const int cper = (P.Cin + 1) / 2, cbeg = ((chalf * cper) + 7) & ~7;  // multiples of 8 (Cin is a multiple of 16)
    const int cend = chalf ? P.Cin : (((cper) + 7) & ~7);
    const uint32_t rbase = (uint32_t)((prow >> 3) * 1024 + (prow & 7) * 128);
    const int rx = prow & 7;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x, ++it) {
      const int b = tile / P.tiles_per_b, pix = (tile - b * P.tiles_per_b) * TMPIX + prow;
      const bool pv = pix < P.HW;
      const float *xp = P.x + (size_t)b * P.Cin * P.HW + pix;
      // all of this thread's channels are requested before it waits for the operand buffer, so the loads of tile
      // t+1 are in flight while the MMAs of tile t run
      float v[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) v[i] = (pv && cbeg + i < cend) ? __ldg(xp + (size_t)(cbeg + i) * P.HW) : 0.0f;
      mbar_wait(a_empty, (it & 1) ^ 1);  // the MMAs of the previous tile have read the operand buffer
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        const int c = cbeg + 8 * g;
        if (c < cend) {
          uint32_t h[4], l[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float a0 = v[8 * g + 2 * j], a1 = v[8 * g + 2 * j + 1];
            const __nv_bfloat162 hh = __floats2bfloat162_rn(a0, a1);
            h[j] = *reinterpret_cast<const uint32_t *>(&hh);
            l[j] = pack_bf16(a0 - __bfloat162float(hh.x), a1 - __bfloat162float(hh.y));
          }
          const uint32_t off = (uint32_t)((c >> 6) * A_KBLOCK) + rbase + (uint32_t)(((((c & 63) >> 3) ^ rx)) << 4);
          sts128u(a_u32 + off, h[0], h[1], h[2], h[3]);
          sts128u(a_u32 + kMaxKB * A_KBLOCK + off, l[0], l[1], l[2], l[3]);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_full);
    }
  } else if (warp < kConvWarps + kEpiWarps) {
    // ===================== epilogue: + bias, split, (B, HW, Cout) bf16 rows =====================
    const int e = warp - kConvWarps;
    int acc = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
      const int b = tile / P.tiles_per_b, pix = (tile - b * P.tiles_per_b) * TMPIX + e * 32 + lane;
      const bool pv = pix < P.HW;
      mbar_wait(&acc_full[acc], phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + (uint32_t)(acc * 256);
      const size_t orow = ((size_t)b * P.HW + pix) * Cout;
      for (int c0 = 0; c0 < Cout; c0 += 32) {
        float v[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
        if (pv) {
#pragma unroll
          for (int g2 = 0; g2 < 2; ++g2) {  // 16 channels = one 32 B sector of the hi row and one of the lo row
            uint32_t h[2][4], l[2][4];
#pragma unroll
            for (int g = 0; g < 2; ++g)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int c = 16 * g2 + 8 * g + 2 * j;
                const float a0 = v[c] + bias_s[c0 + c], a1 = v[c + 1] + bias_s[c0 + c + 1];
                const __nv_bfloat162 hh = __floats2bfloat162_rn(a0, a1);
                h[g][j] = *reinterpret_cast<const uint32_t *>(&hh);
                l[g][j] = pack_bf16(a0 - __bfloat162float(hh.x), a1 - __bfloat162float(hh.y));
              }
            stg256u(P.hi + orow + c0 + 16 * g2, h[0], h[1]);
            if (P.lo) stg256u(P.lo + orow + c0 + 16 * g2, l[0], l[1]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
      acc ^= 1;
      if (acc == 0) phase ^= 1;
    }
  } else {
    // ===================== weights + MMA issue =====================
    if (lane == 0) {
      mbar_expect_tx(w_full, 2u * (uint32_t)KB * w_block);
      for (int part = 0; part < 2; ++part)
        for (int kb = 0; kb < KB; ++kb)
          tma_load_3d(smem + 2 * kMaxKB * A_KBLOCK + (part * kMaxKB + kb) * (kMaxCout * 128), &tmW, w_full, kb * 64, 0, part);
      mbar_wait(w_full, 0);
      tc_fence_after();
      // kind::f16: D = f32, A = B = bf16, K-major, M = 128, N = Cout
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(Cout >> 3) << 17) | ((uint32_t)(TMPIX >> 4) << 24);
      int acc = 0;
      uint32_t accphase = 0, it = 0;
      for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x, ++it) {
        mbar_wait(&acc_empty[acc], accphase ^ 1);
        mbar_wait(a_full, it & 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
        uint32_t first = 1;
        // x_hi * W_hi + x_lo * W_hi + x_hi * W_lo
#pragma unroll 1
        for (int pass = 0; pass < 3; ++pass) {
          const uint32_t a_part = (pass == 1) ? kMaxKB * A_KBLOCK : 0u;
          const uint32_t w_part = (pass == 2) ? kMaxKB * (kMaxCout * 128) : 0u;
          for (int kb = 0; kb < KB; ++kb) {
            const uint64_t adesc = make_sw128_desc(a_u32 + a_part + kb * A_KBLOCK);
            const uint64_t bdesc = make_sw128_desc(w_u32 + w_part + kb * (kMaxCout * 128));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, first ? 0u : 1u);
              first = 0;
            }
          }
        }
        umma_commit(a_empty);
        umma_commit(&acc_full[acc]);
        acc ^= 1;
        if (acc == 0) accphase ^= 1;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
}

constexpr int kSmemBytes = 1024 + 2 * kMaxKB * A_KBLOCK + 2 * kMaxKB * kMaxCout * 128 + kMaxCout * 4 + 128;
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

// weights (Cout, Cin) fp32 -> [2 parts: hi, lo][Cout][Kp] bf16, Kp = Cin rounded up to 64, zero padding
__global__ void conv1x1_prep_weight_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ wp, int Cout, int Cin, int Kp) {
  const int total = Cout * Kp;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int o = i / Kp, c = i - o * Kp;
    const float v = c < Cin ? __ldg(w + (size_t)o * Cin + c) : 0.0f;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    wp[i] = h;
    wp[total + i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

}  // namespace fh

size_t conv1x1_weight_bytes(int Cout, int Cin) { return (size_t)2 * Cout * ((Cin + 63) / 64 * 64) * 2; }

int launch_conv1x1_prep_weight(const float *w, void *wp, int Cout, int Cin, cudaStream_t st) {
  const int Kp = (Cin + 63) / 64 * 64;
  fh::conv1x1_prep_weight_kernel<<<(Cout * Kp + 255) / 256, 256, 0, st>>>(w, static_cast<__nv_bfloat16 *>(wp), Cout, Cin, Kp);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_conv1x1_prep(const float *x, const void *wp, const float *bias, void *hi, void *lo, int Bt, int Cin, int Cout, int HW,
                        cudaStream_t st) {
  using namespace fh;
  const int Kp = (Cin + 63) / 64 * 64;
  CUtensorMap tm;
  {
    const unsigned long long dims[3] = {(unsigned long long)Kp, (unsigned long long)Cout, 2};
    const unsigned long long strides[2] = {(unsigned long long)Kp * 2, (unsigned long long)Kp * 2 * Cout};
    const unsigned int box[3] = {64, (unsigned int)Cout, 1};
    if (int rc = encode_tmap_sw128(&tm, wp, 3, dims, strides, box, "rc_fnet_conv1x1_prep")) return rc;
  }
  Params P{};
  P.x = x;
  P.bias = bias;
  P.hi = static_cast<__nv_bfloat16 *>(hi);
  P.lo = static_cast<__nv_bfloat16 *>(lo);
  P.Bt = Bt; P.Cin = Cin; P.Cout = Cout; P.HW = HW; P.KB = Kp / 64;
  P.tiles_per_b = (HW + TMPIX - 1) / TMPIX;
  P.total_tiles = Bt * P.tiles_per_b;
  RC_CUDA(cudaFuncSetAttribute(conv1x1_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int dev = 0, sms = 148;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = P.total_tiles < sms ? P.total_tiles : sms;
  RC_CUDA(launch_pdl(conv1x1_prep_kernel, dim3((unsigned)grid), dim3(kThreads), (size_t)kSmemBytes, st, tm, P));
  return RC_OK;
}

}  // namespace rc
