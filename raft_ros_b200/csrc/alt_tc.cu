// On-the-fly correlation on tensor cores: AlternateCorrBlock.__call__ for all pyramid levels in
// one launch, without ever materialising the volume.  sm_100a only.
//
// Replaces the per-level loop of the reference's AlternateCorrBlock.__call__ (core/corr.py:74-91)
// and the alt_cuda_corr forward kernel under it (alt_cuda_corr/correlation_kernel.cu:18-119), whose
// accumulator lives in global memory and which touches 102 KB of fmap2 per pixel per level.
//
// Idea: neighbouring query pixels look at overlapping windows, so a TILE of 8x16 query pixels
// needs the dot products with one small bounding BOX of fmap2 pixels per level.  That is a dense
// 128 x (box) x C contraction: the tile of fmap1 (A, 128 rows) against 32x8 chunks of the box
// (B, 256 rows, a 4-D TMA box straight out of the NHWC feature map, zero-filled outside the map =
// the reference's zero padding) on tcgen05, fp32 accumulation in TMEM.  The epilogue thread that
// owns a query pixel picks its (2r+2)^2 patch out of the accumulator rows (tcgen05.ld of a chunk row,
// then every register goes to its patch slot under a predicate: the column offset is per pixel),
// forms the (2r+1)^2 bilinear taps and writes them in the final (B, L*(2r+1)^2, H, W) layout.  The redundancy (box
// vs. window) is 2-5x for coherent flow and still far cheaper than CUDA-core dot products when
// the flow is incoherent; correctness never depends on coherence (the box just grows).
//
//   plan kernel : per (tile, level) bounding box of floor(coords / 2^l) -> chunk grid, in scratch
//   main kernel : persistent, warp 4 = TMA producer, warp 5 = MMA issuer, warps 0-3 = epilogue
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace alttc {

using namespace tc;

constexpr int R = 4;                 // window radius this kernel is specialised for
constexpr int RD = 2 * R + 1, PN = RD + 1, NTAP = RD * RD, NPATCH = PN * PN;
constexpr int QH = 8, QW = 16;       // query tile (rows x cols) = 128 pixels = TMEM lanes
constexpr int CW = 32, CHH = 8;      // target chunk (cols x rows) = 256 = TMEM columns
constexpr int TK = 64, UK = 16;
constexpr int A_BYTES = 128 * TK * 2, B_BYTES = 256 * TK * 2;
constexpr int kStageBytes = A_BYTES + B_BYTES;   // one operand pair per stage
constexpr int kStages = 3;
constexpr int kEpiWarps = 4, kProducerWarp = 4, kMmaWarp = 5, kThreads = 192;
constexpr int kPatchStride = NPATCH + 1;         // 101: odd -> conflict-free across lanes
constexpr int kPatchBytes = 128 * kPatchStride * 4;
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kPatchBytes + 256;
constexpr uint32_t kIdescA = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
constexpr int kMaxLevels = 4;

struct Maps {
  CUtensorMap a[2];               // fmap1 hi, lo: (Cp, W1, H1, B), box (64, 16, 8, 1)
  CUtensorMap b[kMaxLevels][2];   // fmap2 level l hi, lo: (Cp, W2, H2, B), box (64, 32, 8, 1)
};

struct Params {
  int B, H1, W1, L, k_blocks, npass;
  int tiles_x, tiles_y;
  long long total_tiles;
  int h2[kMaxLevels], w2[kMaxLevels];
  const float *coords;   // (B, 2, H1, W1)
  float *out;            // (B, L*81, H1, W1)
  const int4 *plan;      // [tile][level] = {box x0, box y0, chunks in x, chunks in y}
  float scale;
};

// ------------------------------------------------------------------ plan
__global__ void __launch_bounds__(128)
alt_plan_kernel(const float *__restrict__ coords, int4 *__restrict__ plan, int B, int H1, int W1, int L, int tiles_x,
                int tiles_y, long long total_tiles, int h2_0, int w2_0) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long tile = (long long)blockIdx.x * 4 + warp;
  if (tile >= total_tiles) return;
  const int per_item = tiles_x * tiles_y;
  const int b = (int)(tile / per_item);
  const int rem = (int)(tile - (long long)b * per_item);
  const int ty0 = (rem / tiles_x) * QH, tx0 = (rem % tiles_x) * QW;
  const size_t HW = (size_t)H1 * W1;
  float cx[4], cy[4];
  bool ok[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = lane + 32 * i;
    const int y = ty0 + m / QW, x = tx0 + m % QW;
    ok[i] = y < H1 && x < W1;
    cx[i] = ok[i] ? __ldg(coords + ((size_t)b * 2 + 0) * HW + (size_t)y * W1 + x) : 0.0f;
    cy[i] = ok[i] ? __ldg(coords + ((size_t)b * 2 + 1) * HW + (size_t)y * W1 + x) : 0.0f;
  }
  for (int l = 0; l < L; ++l) {
    const float inv = 1.0f / (float)(1 << l);
    const int h2 = h2_0 >> l, w2 = w2_0 >> l;
    int xmin = 0x7fffffff, xmax = -0x7fffffff, ymin = 0x7fffffff, ymax = -0x7fffffff;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (!ok[i]) continue;
      float fr;
      // window [x0, x0+PN) x [y0, y0+PN), clipped to the map: only that part can be non-zero
      const int x0 = safe_floor_int(cx[i] * inv, &fr) - R, y0 = safe_floor_int(cy[i] * inv, &fr) - R;
      const int xa = max(x0, 0), xb = min(x0 + PN, w2), ya = max(y0, 0), yb = min(y0 + PN, h2);
      if (xa < xb && ya < yb) {
        xmin = min(xmin, xa); xmax = max(xmax, xb);
        ymin = min(ymin, ya); ymax = max(ymax, yb);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      xmin = min(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
      xmax = max(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
      ymin = min(ymin, __shfl_xor_sync(0xffffffffu, ymin, o));
      ymax = max(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
    }
    if (lane == 0) {
      int4 e = make_int4(0, 0, 0, 0);
      if (xmin < xmax && ymin < ymax) e = make_int4(xmin, ymin, (xmax - xmin + CW - 1) / CW, (ymax - ymin + CHH - 1) / CHH);
      plan[tile * L + l] = e;
    }
  }
}

// ------------------------------------------------------------------ main kernel
__global__ void __launch_bounds__(kThreads, 1)
alt_tc_kernel(const __grid_constant__ Maps M, const Params P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  const uint32_t patch_u32 = base_u32 + kStages * kStageBytes;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes + kPatchBytes);
  uint64_t *full = bars, *empty = bars + kStages, *tfull = bars + 2 * kStages, *tempty = tfull + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kProducerWarp && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int per_item = P.tiles_x * P.tiles_y;
  const int steps = P.k_blocks * P.npass;  // operand-pair steps per chunk

  if (warp == kProducerWarp) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        const int b = (int)(t / per_item);
        const int rem = (int)(t - (long long)b * per_item);
        const int ty0 = (rem / P.tiles_x) * QH, tx0 = (rem % P.tiles_x) * QW;
        for (int l = 0; l < P.L; ++l) {
          const int4 pl = __ldg(P.plan + t * P.L + l);
          for (int cyi = 0; cyi < pl.w; ++cyi)
            for (int cxi = 0; cxi < pl.z; ++cxi) {
              const int ox = pl.x + cxi * CW, oy = pl.y + cyi * CHH;
              for (int s = 0; s < steps; ++s) {
                const int kb = s / P.npass, pr = s - kb * P.npass;  // pr: 0 hi*hi, 1 hi*lo, 2 lo*hi
                mbar_wait(&empty[stage], phase ^ 1);
                uint8_t *sp = smem + stage * kStageBytes;
                mbar_expect_tx(&full[stage], kStageBytes);
                tma_load_4d(sp, &M.a[pr == 2 ? 1 : 0], &full[stage], kb * TK, tx0, ty0, b);
                tma_load_4d(sp + A_BYTES, &M.b[l][pr == 1 ? 1 : 0], &full[stage], kb * TK, ox, oy, b);
                if (++stage == kStages) { stage = 0; phase ^= 1; }
              }
            }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    if (lane == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (long long t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        for (int l = 0; l < P.L; ++l) {
          const int4 pl = __ldg(P.plan + t * P.L + l);
          const int chunks = pl.z * pl.w;
          for (int c = 0; c < chunks; ++c) {
            mbar_wait(&tempty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
            for (int s = 0; s < steps; ++s) {
              mbar_wait(&full[stage], phase);
              tc_fence_after();
              const uint32_t sa = base_u32 + stage * kStageBytes;
              const uint64_t ad = make_sw128_desc(sa), bd = make_sw128_desc(sa + A_BYTES);
#pragma unroll
              for (int k = 0; k < TK / UK; ++k) {
                const uint64_t koff = (uint64_t)((k * UK * 2) >> 4);
                umma_bf16(d_tmem, ad + koff, bd + koff, kIdescA, (s | k) != 0 ? 1u : 0u);
              }
              umma_commit(&empty[stage]);
              if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            umma_commit(&tfull[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
          }
        }
      }
    }
  } else {
    // ===================== epilogue: thread <-> query pixel =====================
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const uint32_t patch = patch_u32 + (uint32_t)m * (kPatchStride * 4);
    int acc = 0;
    uint32_t acc_phase = 0;
    const size_t HW = (size_t)P.H1 * P.W1;
    const int CH = P.L * NTAP;
    for (long long t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      const int b = (int)(t / per_item);
      const int rem = (int)(t - (long long)b * per_item);
      const int y = (rem / P.tiles_x) * QH + m / QW, x = (rem % P.tiles_x) * QW + m % QW;
      const bool pvalid = y < P.H1 && x < P.W1;
      float cx = 0.0f, cy = 0.0f;
      if (pvalid) {
        cx = __ldg(P.coords + ((size_t)b * 2 + 0) * HW + (size_t)y * P.W1 + x);
        cy = __ldg(P.coords + ((size_t)b * 2 + 1) * HW + (size_t)y * P.W1 + x);
      }
      for (int l = 0; l < P.L; ++l) {
        const float inv = 1.0f / (float)(1 << l);
        float dx, dy;
        const int x0 = safe_floor_int(cx * inv, &dx) - R, y0 = safe_floor_int(cy * inv, &dy) - R;
#pragma unroll 4
        for (int i = 0; i < NPATCH; ++i) sts32(patch + i * 4, 0.0f);
        const int4 pl = __ldg(P.plan + t * P.L + l);
        for (int cyi = 0; cyi < pl.w; ++cyi)
          for (int cxi = 0; cxi < pl.z; ++cxi) {
            const int ox = pl.x + cxi * CW, oy = pl.y + cyi * CHH;
            mbar_wait(&tfull[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
            const bool xhit = pvalid && x0 + PN > ox && x0 < ox + CW;
#pragma unroll 1
            for (int rr = 0; rr < CHH; ++rr) {
              const int py = oy + rr - y0;
              const bool need = xhit && (unsigned)py < (unsigned)PN;
              if (__any_sync(0xffffffffu, need)) {  // tcgen05.ld is warp-collective
                float v[32];
                tmem_ld32(taddr + rr * CW, v);
                tmem_ld_wait();
                if (need) {
                  // window column px of this pixel is accumulator column cbase + px: every register goes
                  // straight to its patch slot under a predicate (no bounce through a row buffer; the
                  // epilogue has issue slots to spare, shared-memory round trips it has not)
                  const int cbase = x0 - ox;
                  const uint32_t prow = patch + (uint32_t)((py * PN - cbase) * 4);
#pragma unroll
                  for (int c = 0; c < CW; ++c)
                    if ((unsigned)(c - cbase) < (unsigned)PN) sts32(prow + c * 4, v[c]);
                }
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
          }
        // bilinear taps from the patch, straight into the final layout
        if (pvalid) {
          const float w00 = (1.0f - dy) * (1.0f - dx) * P.scale, w01 = (1.0f - dy) * dx * P.scale;
          const float w10 = dy * (1.0f - dx) * P.scale, w11 = dy * dx * P.scale;
          float *o = P.out + ((size_t)b * CH + (size_t)l * NTAP) * HW + (size_t)y * P.W1 + x;
#pragma unroll
          for (int i = 0; i < RD; ++i) {      // x offset (slow index)
            float col0[PN], col1[PN];
#pragma unroll
            for (int j = 0; j < PN; ++j) {
              col0[j] = lds32(patch + (j * PN + i) * 4);
              col1[j] = lds32(patch + (j * PN + i + 1) * 4);
            }
#pragma unroll
            for (int j = 0; j < RD; ++j) {    // y offset (fast index)
              float s = col0[j] * w00;
              s = fmaf(col1[j], w01, s);
              s = fmaf(col0[j + 1], w10, s);
              s = fmaf(col1[j + 1], w11, s);
              st_stream(o + (size_t)(i * RD + j) * HW, s);
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------ layout prepass
// fmap (B, C, HW) fp32 NCHW -> (B, HW, Cp) bf16 K-major "hi" (+ "lo" residual), zero channel padding
__global__ void __launch_bounds__(256)
alt_prep_kernel(const float *__restrict__ f, __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo, int C,
                int Cp, int HW, int with_lo) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const float *src = f + (size_t)b * C * HW;
  const int p0 = blockIdx.x * 32, d0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int d = d0 + ty + 8 * i, p = p0 + tx;
    tile[ty + 8 * i][tx] = (d < C && p < HW) ? __ldg(src + (size_t)d * HW + p) : 0.0f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 8 * i, d = d0 + tx;
    if (p < HW && d < Cp) {
      const float x = tile[tx][ty + 8 * i];
      const __nv_bfloat16 xh = __float2bfloat16_rn(x);
      const size_t o = ((size_t)b * HW + p) * Cp + d;
      hi[o] = xh;
      if (with_lo) lo[o] = __float2bfloat16_rn(x - __bfloat162float(xh));
    }
  }
}

// ------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

static int map4(CUtensorMap *tm, const void *base, int Cp, int W, int H, int B, int bw, int bh) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return RC_ERR_CUDA;
  }
  const cuuint64_t dims[4] = {(cuuint64_t)Cp, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  const cuuint64_t strides[3] = {(cuuint64_t)Cp * 2, (cuuint64_t)W * Cp * 2, (cuuint64_t)H * W * Cp * 2};
  const cuuint32_t box[4] = {TK, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  const cuuint32_t ones[4] = {1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(base), dims, strides, box, ones,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return RC_ERR_CUDA;
  }
  return RC_OK;
}

static int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  return n;
}

}  // namespace alttc

int launch_alt_prep(const float *fmap, void *hi, void *lo, int B, int C, int HW, cudaStream_t st) {
  const int Cp = (C + 63) / 64 * 64;
  dim3 grid((unsigned)((HW + 31) / 32), (unsigned)(Cp / 32), (unsigned)B);
  alttc::alt_prep_kernel<<<grid, 256, 0, st>>>(fmap, static_cast<__nv_bfloat16 *>(hi), static_cast<__nv_bfloat16 *>(lo), C,
                                              Cp, HW, lo != nullptr ? 1 : 0);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

size_t alt_tc_plan_bytes(int B, int H1, int W1, int L) {
  const long long tiles = (long long)B * ((H1 + alttc::QH - 1) / alttc::QH) * ((W1 + alttc::QW - 1) / alttc::QW);
  return (size_t)tiles * L * sizeof(int4);
}

int launch_alt_tc(const void *f1_hi, const void *f1_lo, const void *const *f2_hi, const void *const *f2_lo,
                  const float *coords, float *out, int B, int H1, int W1, int Cp, int L, int radius, int precision,
                  float scale, void *plan_ws, size_t plan_bytes, cudaStream_t st) {
  using namespace alttc;
  if (radius != R || L < 1 || L > kMaxLevels) {
    set_error("rc_alt_corr_pyramid_fwd: the tensor-core path covers radius %d and 1..%d levels (got r=%d, L=%d)", R,
              kMaxLevels, radius, L);
    return RC_ERR_UNSUPPORTED;
  }
  if (Cp % TK != 0) {
    set_error("rc_alt_corr_pyramid_fwd: channel pitch %d must be a multiple of %d", Cp, TK);
    return RC_ERR_INVALID;
  }
  const size_t need = alt_tc_plan_bytes(B, H1, W1, L);
  if (plan_ws == nullptr || plan_bytes < need || (reinterpret_cast<uintptr_t>(plan_ws) & 15) != 0) {
    set_error("rc_alt_corr_pyramid_fwd: plan scratch missing, misaligned or too small (%zu < %zu bytes)", plan_bytes, need);
    return RC_ERR_WORKSPACE;
  }
  const bool x3 = precision == RC_PREC_BF16X3;
  Maps M;
  if (int rc = map4(&M.a[0], f1_hi, Cp, W1, H1, B, QW, QH)) return rc;
  if (int rc = map4(&M.a[1], x3 ? f1_lo : f1_hi, Cp, W1, H1, B, QW, QH)) return rc;
  Params P{};
  for (int l = 0; l < kMaxLevels; ++l) {
    const int ll = l < L ? l : L - 1;
    P.h2[l] = H1 >> ll;
    P.w2[l] = W1 >> ll;
    if (P.h2[l] < 1 || P.w2[l] < 1) {
      set_error("rc_alt_corr_pyramid_fwd: level %d of a %dx%d grid is empty", ll, H1, W1);
      return RC_ERR_INVALID;
    }
    if (int rc = map4(&M.b[l][0], f2_hi[ll], Cp, P.w2[l], P.h2[l], B, CW, CHH)) return rc;
    if (int rc = map4(&M.b[l][1], x3 ? f2_lo[ll] : f2_hi[ll], Cp, P.w2[l], P.h2[l], B, CW, CHH)) return rc;
  }
  P.B = B; P.H1 = H1; P.W1 = W1; P.L = L;
  P.k_blocks = Cp / TK;
  P.npass = x3 ? 3 : 1;
  P.tiles_x = (W1 + QW - 1) / QW;
  P.tiles_y = (H1 + QH - 1) / QH;
  P.total_tiles = (long long)B * P.tiles_x * P.tiles_y;
  P.coords = coords;
  P.out = out;
  P.plan = static_cast<const int4 *>(plan_ws);
  P.scale = scale;
  {
    const unsigned blocks = (unsigned)((P.total_tiles + 3) / 4);
    alt_plan_kernel<<<blocks, 128, 0, st>>>(coords, static_cast<int4 *>(plan_ws), B, H1, W1, L, P.tiles_x, P.tiles_y,
                                            P.total_tiles, H1, W1);
    RC_CUDA(cudaGetLastError());
  }
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
  RC_CUDA(cudaFuncSetAttribute(alt_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  const long long sms = sm_count();
  const unsigned grid = (unsigned)(P.total_tiles < sms ? P.total_tiles : sms);
  alt_tc_kernel<<<grid, kThreads, kSmemBytes, st>>>(M, P);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

}  // namespace rc
