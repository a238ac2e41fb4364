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
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace alttc {

using namespace tc;

constexpr int QH = 8, QW = 16;       // query tile (rows x cols) = 128 pixels = TMEM lanes
// target chunk = 256 fmap2 pixels = TMEM columns, in one of two shapes chosen per (tile, level) by the plan:
// 32 x 8 (wide boxes: level 0 of a 16-wide tile) or 16 x 16 (the small boxes of the coarse levels fit one)
constexpr int CW0 = 32, CH0 = 8, CW1 = 16, CH1 = 16;
constexpr int TK = 64, UK = 16;
constexpr int A_BYTES = 128 * TK * 2, B_BYTES = 256 * TK * 2;
// one operand pair per stage; a CTA of a pair (NCTA = 2) stages its own 128 query pixels and HALF of the chunk
template <int NCTA>
struct StageCfg {
  static constexpr int kStageBytes = A_BYTES + B_BYTES / NCTA;
  static constexpr int kStages = NCTA == 2 ? 4 : 3;
};
constexpr int kRingBytes = 3 * (A_BYTES + B_BYTES);  // 144 KB: 3 stages of 48 KB, or 4 of 32 KB
constexpr int kEpiWarps = 4, kProducerWarp = 4, kMmaWarp = 5, kThreads = 192;
constexpr int kPatchStrideMax = 101;             // (2*4+2)^2 + 1: odd -> conflict-free across lanes
constexpr int kPatchBytes = 128 * kPatchStrideMax * 4;
constexpr int kSmemBytes = 1024 + kRingBytes + kPatchBytes + 256;
constexpr uint32_t kIdescA2 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
constexpr uint32_t kIdescA = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
// fp16 (kind::f16) or e4m3 (kind::f8f6f4) operands: both formats are code 0 (RC_PREC_FP16_FP8, see build_tc.cu)
constexpr uint32_t kIdescF2 = (1u << 4) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
constexpr uint32_t kIdescF = (1u << 4) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
constexpr int kMaxLevels = 4;

struct Maps {
  CUtensorMap a[2];                  // fmap1 hi, lo: (Cp, W1, H1, B), box (64, 16, 8, 1)
  CUtensorMap b[kMaxLevels][2][2];   // fmap2 level l, chunk shape s (box (64, 32, 8, 1) / (64, 16, 16, 1)), hi / lo
};

struct Params {
  int B, H1, W1, L, k_blocks, npass;
  int tiles_x, tiles_y;
  long long total_tiles;
  int h2[kMaxLevels], w2[kMaxLevels];
  const float *coords;   // (B, 2, H1, W1)
  float *out;            // (B, L*81, H1, W1)
  const int4 *plan;      // [tile][level] = {box x0, box y0, chunks in x | shape << 16, chunks in y}
  float scale;
  int f8;                       // RC_PREC_FP16_FP8: pass 0 = fp16 hi * hi, pass 1 = the 8-bit pair maps (both cross terms)
  const unsigned *amax1, *amax2;  //   bits of max |x| of fmap1 / fmap2 per batch item (their power-of-two normalisation)
};

__device__ __forceinline__ float alt_norm_mul(unsigned amax_bits) {  // = build_tc.cu's norm_mul
  if (amax_bits == 0u) return 1.0f;
  int f = 261 - (int)(amax_bits >> 23);
  f = f < 1 ? 1 : (f > 254 ? 254 : f);
  return __uint_as_float((unsigned)f << 23);
}

// ------------------------------------------------------------------ plan
template <int R, int NCTA>
__global__ void __launch_bounds__(128)
alt_plan_kernel(const float *__restrict__ coords, int4 *__restrict__ plan, int B, int H1, int W1, int L, int tiles_x,
                int tiles_y, long long total_tiles, int h2_0, int w2_0, int allow_square) {
  constexpr int PN = 2 * R + 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long tile = (long long)blockIdx.x * 4 + warp;
  pdl_launch_dependents();
  pdl_wait();
  if (tile >= total_tiles) return;
  const int per_item = tiles_x * tiles_y;
  const int b = (int)(tile / per_item);
  const int rem = (int)(tile - (long long)b * per_item);
  constexpr int NI = 4 * NCTA;  // query pixels per lane: the tile is QH * NCTA rows x QW columns
  const int ty0 = (rem / tiles_x) * (QH * NCTA), tx0 = (rem % tiles_x) * QW;
  const size_t HW = (size_t)H1 * W1;
  float cx[NI], cy[NI];
  bool ok[NI];
#pragma unroll
  for (int i = 0; i < NI; ++i) {
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
    for (int i = 0; i < NI; ++i) {
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
      if (xmin < xmax && ymin < ymax) {
        const int bw = xmax - xmin, bh = ymax - ymin;
        const int nx0 = (bw + CW0 - 1) / CW0, ny0 = (bh + CH0 - 1) / CH0, nx1 = (bw + CW1 - 1) / CW1, ny1 = (bh + CH1 - 1) / CH1;
        // fewer chunks = fewer MMAs: what counts in bf16x3 (tensor-bound).  A single bf16 pass is bound by the
        // epilogue, which skips whole accumulator rows of a 32 x 8 chunk more often: it keeps that shape.
        e = (allow_square && nx1 * ny1 < nx0 * ny0) ? make_int4(xmin, ymin, nx1 | (1 << 16), ny1) : make_int4(xmin, ymin, nx0, ny0);
      }
      plan[tile * L + l] = e;
    }
  }
}

// ------------------------------------------------------------------ main kernel
// NCTA == 2: clusters of two CTAs share one 16 x 16 query tile (rows 0-7 / 8-15) and its boxes: a pair contracts its
// 256 query pixels with a chunk in one tcgen05.mma.cta_group::2 (each CTA stages its own A and half of the chunk),
// so a box -- which grows by the window, not by the tile -- is paid for by twice the pixels.
template <int R, int NCTA>
__global__ void __launch_bounds__(kThreads, 1)
alt_tc_kernel(const __grid_constant__ Maps M, const Params P) {
  constexpr int kStages = StageCfg<NCTA>::kStages, kStageBytes = StageCfg<NCTA>::kStageBytes;
  const uint32_t rank = (NCTA == 2) ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const unsigned unit0 = blockIdx.x / NCTA, unit_step = gridDim.x / NCTA;
  constexpr int RD = 2 * R + 1, PN = RD + 1, NTAP = RD * RD, NPATCH = PN * PN;
  constexpr int kPatchStride = NPATCH + 1;  // odd -> conflict-free across lanes
  static_assert(kPatchStride <= kPatchStrideMax, "patch buffer");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  const uint32_t patch_u32 = base_u32 + kRingBytes;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kRingBytes + kPatchBytes);
  uint64_t *full = bars, *empty = bars + kStages, *tfull = bars + 2 * kStages, *tempty = tfull + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();

  if (warp == kProducerWarp && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], NCTA);  // pair: the leader's barrier also takes the peer's arrival and bytes
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], kEpiWarps * NCTA);
    }
    fence_barrier_init();
  }
  if (warp == kMmaWarp) {
    if (NCTA == 2) tmem_alloc2(tmem_slot, 512);
    else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();  // both CTAs' barriers are initialised before any remote use
  tc_fence_after();
  pdl_wait();  // the plan (previous kernel), the coordinates and the feature maps are read from here on
  const uint32_t tmem_base = *tmem_slot;
  const int per_item = P.tiles_x * P.tiles_y;
  const int steps = P.k_blocks * P.npass;  // operand-pair steps per chunk
  const int a_lo_pass = P.f8 ? 1 : 2;      // the pass that takes fmap1's second map (bf16x3: lo * hi is the third)

  if (warp == kProducerWarp) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long t = unit0; t < P.total_tiles; t += unit_step) {
        const int b = (int)(t / per_item);
        const int rem = (int)(t - (long long)b * per_item);
        const int ty0 = (rem / P.tiles_x) * (QH * NCTA) + (int)rank * QH, tx0 = (rem % P.tiles_x) * QW;
        for (int l = 0; l < P.L; ++l) {
          const int4 pl = __ldg(P.plan + t * P.L + l);
          const int shape = pl.z >> 16, ncx = pl.z & 0xffff;
          const int cw = shape ? CW1 : CW0, chh = shape ? CH1 : CH0;
          for (int cyi = 0; cyi < pl.w; ++cyi)
            for (int cxi = 0; cxi < ncx; ++cxi) {
              const int ox = pl.x + cxi * cw, oy = pl.y + cyi * chh + (int)rank * (chh / NCTA);  // this CTA's rows of the chunk
              for (int s = 0; s < steps; ++s) {
                const int kb = s / P.npass, pr = s - kb * P.npass;  // pr: 0 hi*hi, 1 hi*lo, 2 lo*hi
                mbar_wait(&empty[stage], phase ^ 1);
                uint8_t *sp = smem + stage * kStageBytes;
                if (NCTA == 1) {
                  mbar_expect_tx(&full[stage], kStageBytes);
                  tma_load_4d(sp, &M.a[pr == a_lo_pass ? 1 : 0], &full[stage], kb * TK, tx0, ty0, b);
                  tma_load_4d(sp + A_BYTES, &M.b[l][shape][pr == 1 ? 1 : 0], &full[stage], kb * TK, ox, oy, b);
                } else {
                  if (leader) mbar_expect_tx(&full[stage], 2 * kStageBytes);  // both CTAs' bytes land on the leader's barrier
                  tma2_load_4d(sp, &M.a[pr == a_lo_pass ? 1 : 0], &full[stage], kb * TK, tx0, ty0, b);
                  tma2_load_4d(sp + A_BYTES, &M.b[l][shape][pr == 1 ? 1 : 0], &full[stage], kb * TK, ox, oy, b);
                  if (!leader) mbar_arrive_leader(&full[stage]);
                }
                if (++stage == kStages) { stage = 0; phase ^= 1; }
              }
            }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    if (lane == 0 && leader) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (long long t = unit0; t < P.total_tiles; t += unit_step) {
        for (int l = 0; l < P.L; ++l) {
          const int4 pl = __ldg(P.plan + t * P.L + l);
          const int chunks = (pl.z & 0xffff) * pl.w;
          for (int c = 0; c < chunks; ++c) {
            mbar_wait(&tempty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
            for (int s = 0; s < steps; ++s) {
              mbar_wait(&full[stage], phase);
              tc_fence_after();
              const uint32_t sa = base_u32 + stage * kStageBytes;
              const uint64_t ad = make_sw128_desc(sa), bd = make_sw128_desc(sa + A_BYTES);
              if (P.f8 && (s & 1)) {  // npass = 2: odd steps are the 8-bit pair maps, K = 32 per instruction
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  if (NCTA == 1) umma_f8(d_tmem, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), kIdescF, 1u);
                  else umma2_f8(d_tmem, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), kIdescF2, 1u);
                }
              } else {
                const uint32_t id1 = P.f8 ? kIdescF : kIdescA, id2 = P.f8 ? kIdescF2 : kIdescA2;
#pragma unroll
                for (int k = 0; k < TK / UK; ++k) {
                  const uint64_t koff = (uint64_t)((k * UK * 2) >> 4);
                  if (NCTA == 1) umma_bf16(d_tmem, ad + koff, bd + koff, id1, (s | k) != 0 ? 1u : 0u);
                  else umma2_bf16(d_tmem, ad + koff, bd + koff, id2, (s | k) != 0 ? 1u : 0u);
                }
              }
              if (NCTA == 1) umma_commit(&empty[stage]);
              else umma2_commit_both(&empty[stage]);
              if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            if (NCTA == 1) umma_commit(&tfull[acc]);
            else umma2_commit_both(&tfull[acc]);
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
    for (long long t = unit0; t < P.total_tiles; t += unit_step) {
      const int b = (int)(t / per_item);
      const int rem = (int)(t - (long long)b * per_item);
      const int y = (rem / P.tiles_x) * (QH * NCTA) + (int)rank * QH + m / QW, x = (rem % P.tiles_x) * QW + m % QW;
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
        // windows wholly inside the level are wholly written by the chunks of the plan's box: only the others need zeros
        if (!(x0 >= 0 && y0 >= 0 && x0 + PN <= P.w2[l] && y0 + PN <= P.h2[l])) {
#pragma unroll 4
          for (int i = 0; i < NPATCH; ++i) sts32(patch + i * 4, 0.0f);
        }
        const int4 pl = __ldg(P.plan + t * P.L + l);
        const int shape = pl.z >> 16, ncx = pl.z & 0xffff;
        const int cw = shape ? CW1 : CW0, chh = shape ? CH1 : CH0;
        for (int cyi = 0; cyi < pl.w; ++cyi)
          for (int cxi = 0; cxi < ncx; ++cxi) {
            const int ox = pl.x + cxi * cw, oy = pl.y + cyi * chh;
            mbar_wait(&tfull[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
            const bool xhit = pvalid && x0 + PN > ox && x0 < ox + cw;
            // window column px of this pixel is chunk column cbase + px: every register goes straight to
            // its patch slot under a predicate (no bounce through a row buffer; the epilogue has issue
            // slots to spare, shared-memory round trips it has not)
            const int cbase = x0 - ox;
            if (shape == 0) {
              // 32 x 8 chunk: one chunk row per load; rows that no window of the warp touches are skipped
#pragma unroll 1
              for (int rr = 0; rr < CH0; ++rr) {
                const int py = oy + rr - y0;
                const bool need = xhit && (unsigned)py < (unsigned)PN;
                if (__any_sync(0xffffffffu, need)) {  // tcgen05.ld is warp-collective
                  float v[32];
                  tmem_ld32(taddr + rr * CW0, v);
                  tmem_ld_wait();
                  if (need) {
                    const uint32_t prow = patch + (uint32_t)((py * PN - cbase) * 4);
#pragma unroll
                    for (int c = 0; c < CW0; ++c)
                      if ((unsigned)(c - cbase) < (unsigned)PN) sts32(prow + c * 4, v[c]);
                  }
                }
              }
            } else {
              // 16 x 16 chunk: two chunk rows per load
#pragma unroll 1
              for (int rr = 0; rr < CH1 / 2; ++rr) {
                const int py0 = oy + 2 * rr - y0;
                const bool need0 = xhit && (unsigned)py0 < (unsigned)PN;
                const bool need1 = xhit && (unsigned)(py0 + 1) < (unsigned)PN;
                if (__any_sync(0xffffffffu, need0 || need1)) {
                  float v[32];
                  tmem_ld32(taddr + rr * 32, v);
                  tmem_ld_wait();
                  if (need0) {
                    const uint32_t prow = patch + (uint32_t)((py0 * PN - cbase) * 4);
#pragma unroll
                    for (int c = 0; c < CW1; ++c)
                      if ((unsigned)(c - cbase) < (unsigned)PN) sts32(prow + c * 4, v[c]);
                  }
                  if (need1) {
                    const uint32_t prow = patch + (uint32_t)(((py0 + 1) * PN - cbase) * 4);
#pragma unroll
                    for (int c = 0; c < CW1; ++c)
                      if ((unsigned)(c - cbase) < (unsigned)PN) sts32(prow + c * 4, v[CW1 + c]);
                  }
                }
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (NCTA == 1) mbar_arrive(&tempty[acc]);
              else mbar_arrive_leader(&tempty[acc]);  // the pair's MMA thread lives in the leader CTA
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
          }
        // bilinear taps from the patch, straight into the final layout
        if (pvalid) {
          float sc = P.scale;
          if (P.f8) sc = sc * (1.0f / alt_norm_mul(__ldg(P.amax1 + b))) * (1.0f / alt_norm_mul(__ldg(P.amax2 + b)));
          const float w00 = (1.0f - dy) * (1.0f - dx) * sc, w01 = (1.0f - dy) * dx * sc;
          const float w10 = dy * (1.0f - dx) * sc, w11 = dy * dx * sc;
          float *o = P.out + ((size_t)b * CH + (size_t)l * NTAP) * HW + (size_t)y * P.W1 + x;
          float col0[PN], col1[PN];
#pragma unroll
          for (int j = 0; j < PN; ++j) col1[j] = lds32(patch + (j * PN) * 4);
#pragma unroll
          for (int i = 0; i < RD; ++i) {      // x offset (slow index); a patch column is read once and used for two offsets
#pragma unroll
            for (int j = 0; j < PN; ++j) {
              col0[j] = col1[j];
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
  if (NCTA == 2) cluster_sync_all();  // the peer may still be signalling our barriers
  tc_fence_after();
  if (warp == kMmaWarp) {
    if (NCTA == 2) tmem_dealloc2(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
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

// The same for pyramid level LVL >= 1 of the feature map WITHOUT materialising the pooled fp32 map: the
// reference pools iteratively (core/corr.py:68-72, F.avg_pool2d(.., 2, stride=2) of the previous level,
// floor mode), each level rounded to fp32; the value here is formed by the same nested 2x2 means in the
// same order (((a + b) + c) + d) * 0.25, so it is bit-identical to the chain.
template <int LVL>
__device__ __forceinline__ float pooled_at(const float *__restrict__ plane, int W, int y, int x) {
  if constexpr (LVL == 0) {
    return __ldg(plane + (size_t)y * W + x);
  } else {
    float s = pooled_at<LVL - 1>(plane, W, 2 * y, 2 * x) + pooled_at<LVL - 1>(plane, W, 2 * y, 2 * x + 1);
    s += pooled_at<LVL - 1>(plane, W, 2 * y + 1, 2 * x);
    s += pooled_at<LVL - 1>(plane, W, 2 * y + 1, 2 * x + 1);
    return s * 0.25f;
  }
}

template <int LVL>
__global__ void __launch_bounds__(256)
alt_prep_pool_kernel(const float *__restrict__ f, __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo, int C,
                     int Cp, int H, int W, int with_lo) {
  __shared__ float tile[32][33];
  const int Hl = H >> LVL, Wl = W >> LVL, HWl = Hl * Wl;
  const int b = blockIdx.z;
  const float *src = f + (size_t)b * C * H * W;
  const int p0 = blockIdx.x * 32, d0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int d = d0 + ty + 8 * i, p = p0 + tx;
    float v = 0.0f;
    if (d < C && p < HWl) v = pooled_at<LVL>(src + (size_t)d * H * W, W, p / Wl, p % Wl);
    tile[ty + 8 * i][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 8 * i, d = d0 + tx;
    if (p < HWl && d < Cp) {
      const float x = tile[tx][ty + 8 * i];
      const __nv_bfloat16 xh = __float2bfloat16_rn(x);
      const size_t o = ((size_t)b * HWl + p) * Cp + d;
      hi[o] = xh;
      if (with_lo) lo[o] = __float2bfloat16_rn(x - __bfloat162float(xh));
    }
  }
}

// RC_PREC_FP16_FP8 flavour of the two kernels above: the map is multiplied by the power of two of its batch item's
// largest magnitude (alt_norm_mul), hi = fp16, and the second map holds, per pixel and 64 channels, 128 bytes of e4m3:
// [16 lo | hi / 16] for fmap1 (side 0), [hi / 16 | 16 lo] for fmap2 (side 1), so that ONE 8-bit contraction over the 128
// gives both cross terms (build_tc.cu explains the arithmetic).  Pooled levels are means, so level 0's magnitude bounds them.
template <int LVL>
__global__ void __launch_bounds__(256)
alt_prep_f8_kernel(const float *__restrict__ f, __half *__restrict__ hi, uint8_t *__restrict__ pair, const unsigned *__restrict__ amax,
                   int side, int C, int Cp, int H, int W) {
  __shared__ float tile[32][33];
  const int Hl = H >> LVL, Wl = W >> LVL, HWl = Hl * Wl;
  const int b = blockIdx.z;
  const float *src = f + (size_t)b * C * H * W;
  const int p0 = blockIdx.x * 32, d0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float s = alt_norm_mul(__ldg(amax + b));
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int d = d0 + ty + 8 * i, p = p0 + tx;
    float v = 0.0f;
    if (d < C && p < HWl) v = pooled_at<LVL>(src + (size_t)d * H * W, W, p / Wl, p % Wl) * s;
    tile[ty + 8 * i][tx] = v;
  }
  __syncthreads();
  const int off_lo = side ? 64 : 0, off_hi = side ? 0 : 64;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 8 * i, d = d0 + tx;
    if (p < HWl && d < Cp) {
      const float x = tile[tx][ty + 8 * i];
      const __half xh = __float2half_rn(x);
      const float h = __half2float(xh);
      hi[((size_t)b * HWl + p) * Cp + d] = xh;
      uint8_t *row = pair + ((size_t)b * HWl + p) * Cp * 2 + (size_t)(d >> 6) * 128 + (d & 63);
      row[off_lo] = (uint8_t)__nv_cvt_float_to_fp8((x - h) * 16.0f, __NV_SATFINITE, __NV_E4M3);
      row[off_hi] = (uint8_t)__nv_cvt_float_to_fp8(h * 0.0625f, __NV_SATFINITE, __NV_E4M3);
    }
  }
}

// largest magnitude per batch item (amax zero-filled by the caller; non-negative floats order like their bit patterns)
__global__ void __launch_bounds__(256)
alt_absmax_item_kernel(const float *__restrict__ f, unsigned *__restrict__ amax, long long per_item) {
  const float *src = f + (size_t)blockIdx.y * per_item;
  float m = 0.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < per_item; i += (long long)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(__ldg(src + i)));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.0f) atomicMax(amax + blockIdx.y, __float_as_uint(m));
}

// ------------------------------------------------------------------ host
static int map4(CUtensorMap *tm, const void *base, int Cp, int W, int H, int B, int bw, int bh) {
  const unsigned long long dims[4] = {(unsigned long long)Cp, (unsigned long long)W, (unsigned long long)H, (unsigned long long)B};
  const unsigned long long strides[3] = {(unsigned long long)Cp * 2, (unsigned long long)W * Cp * 2, (unsigned long long)H * W * Cp * 2};
  const unsigned int box[4] = {TK, (unsigned int)bw, (unsigned int)bh, 1};
  return encode_tmap_sw128(tm, base, 4, dims, strides, box, "rc_alt_corr_pyramid");
}

static int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  return n;
}

// Everything about one AlternateCorrBlock that does not change between calls: the tensor maps of the
// prepared feature maps and the geometry.  A plain host-side blob the caller keeps (no allocation here).
struct Setup {
  Maps M;
  Params P;   // coords / out / plan / scale are filled per call
  int radius, ncta;
  unsigned magic;
};
constexpr unsigned kSetupMagic = 0x52434154u;

}  // namespace alttc

int launch_alt_prep(const float *fmap, void *hi, void *lo, int B, int C, int HW, cudaStream_t st) {
  const int Cp = (C + 63) / 64 * 64;
  dim3 grid((unsigned)((HW + 31) / 32), (unsigned)(Cp / 32), (unsigned)B);
  alttc::alt_prep_kernel<<<grid, 256, 0, st>>>(fmap, static_cast<__nv_bfloat16 *>(hi), static_cast<__nv_bfloat16 *>(lo), C,
                                              Cp, HW, lo != nullptr ? 1 : 0);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_alt_prep_pooled(const float *fmap, void *hi, void *lo, int B, int C, int H, int W, int level, cudaStream_t st) {
  using namespace alttc;
  if (level == 0) return launch_alt_prep(fmap, hi, lo, B, C, H * W, st);
  const int Cp = (C + 63) / 64 * 64, HWl = (H >> level) * (W >> level);
  if (HWl <= 0) {
    set_error("rc_alt_prep_fmap_level: level %d of a %dx%d map is empty", level, H, W);
    return RC_ERR_INVALID;
  }
  dim3 grid((unsigned)((HWl + 31) / 32), (unsigned)(Cp / 32), (unsigned)B);
  __nv_bfloat16 *h = static_cast<__nv_bfloat16 *>(hi), *l = static_cast<__nv_bfloat16 *>(lo);
  const int wl = lo != nullptr ? 1 : 0;
  switch (level) {
    case 1: alt_prep_pool_kernel<1><<<grid, 256, 0, st>>>(fmap, h, l, C, Cp, H, W, wl); break;
    case 2: alt_prep_pool_kernel<2><<<grid, 256, 0, st>>>(fmap, h, l, C, Cp, H, W, wl); break;
    case 3: alt_prep_pool_kernel<3><<<grid, 256, 0, st>>>(fmap, h, l, C, Cp, H, W, wl); break;
    default:
      set_error("rc_alt_prep_fmap_level: level %d unsupported (0..3)", level);
      return RC_ERR_UNSUPPORTED;
  }
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_alt_absmax(const float *fmap, unsigned *amax, int B, long long per_item, cudaStream_t st) {
  const unsigned bx = (unsigned)std::min<long long>(256, (per_item + 255) / 256);
  alttc::alt_absmax_item_kernel<<<dim3(bx, (unsigned)B), 256, 0, st>>>(fmap, amax, per_item);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_alt_prep_f8(const float *fmap, void *hi, void *pair, const unsigned *amax, int side, int B, int C, int H, int W,
                       int level, cudaStream_t st) {
  using namespace alttc;
  const int Cp = (C + 63) / 64 * 64, HWl = (H >> level) * (W >> level);
  if (HWl <= 0) {
    set_error("rc_alt_prep_fmap_level_f8: level %d of a %dx%d map is empty", level, H, W);
    return RC_ERR_INVALID;
  }
  dim3 grid((unsigned)((HWl + 31) / 32), (unsigned)(Cp / 32), (unsigned)B);
  __half *h = static_cast<__half *>(hi);
  uint8_t *p8 = static_cast<uint8_t *>(pair);
  switch (level) {
    case 0: alt_prep_f8_kernel<0><<<grid, 256, 0, st>>>(fmap, h, p8, amax, side, C, Cp, H, W); break;
    case 1: alt_prep_f8_kernel<1><<<grid, 256, 0, st>>>(fmap, h, p8, amax, side, C, Cp, H, W); break;
    case 2: alt_prep_f8_kernel<2><<<grid, 256, 0, st>>>(fmap, h, p8, amax, side, C, Cp, H, W); break;
    case 3: alt_prep_f8_kernel<3><<<grid, 256, 0, st>>>(fmap, h, p8, amax, side, C, Cp, H, W); break;
    default:
      set_error("rc_alt_prep_fmap_level_f8: level %d unsupported (0..3)", level);
      return RC_ERR_UNSUPPORTED;
  }
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

// One CTA per 8 x 16 query tile by default; RAFT_CORR_ALT_CTAS=2 selects CTA pairs (a 16 x 16 tile per pair).  Measured at
// 2160x3840: the pairs issue 37 % fewer MMAs (15 instead of 24 chunks per 256 query pixels) but every epilogue thread then
// scans the chunks of the larger box, and the epilogue becomes the bound: 0.434 vs 0.406 ms (bf16x3), 0.231 vs 0.219 (bf16).
static int alt_ctas() {
  static const int n = [] {
    const char *e = getenv("RAFT_CORR_ALT_CTAS");
    return (e && e[0] == '2') ? 2 : 1;
  }();
  return n;
}

size_t alt_tc_plan_bytes(int B, int H1, int W1, int L) {
  const long long tiles = (long long)B * ((H1 + alttc::QH - 1) / alttc::QH) * ((W1 + alttc::QW - 1) / alttc::QW);
  return (size_t)tiles * L * sizeof(int4);
}

size_t alt_tc_setup_bytes() { return sizeof(alttc::Setup); }

int alt_tc_setup(void *blob, const void *f1_hi, const void *f1_lo, const void *const *f2_hi, const void *const *f2_lo, int B,
                 int H1, int W1, int H2, int W2, int Cp, int L, int radius, int precision) {
  using namespace alttc;
  if ((radius != 3 && radius != 4) || L < 1 || L > kMaxLevels) {
    set_error("rc_alt_corr_pyramid: the tensor-core path covers radius 3 or 4 and 1..%d levels (got r=%d, L=%d)", kMaxLevels,
              radius, L);
    return RC_ERR_UNSUPPORTED;
  }
  if (Cp % TK != 0) {
    set_error("rc_alt_corr_pyramid: channel pitch %d must be a multiple of %d", Cp, TK);
    return RC_ERR_INVALID;
  }
  const bool f8 = precision == RC_PREC_FP16_FP8;
  const bool x3 = precision == RC_PREC_BF16X3 || f8;  // two maps per feature map
  const int ncta = alt_ctas();
  Setup S{};
  S.ncta = ncta;
  if (int rc = map4(&S.M.a[0], f1_hi, Cp, W1, H1, B, QW, QH)) return rc;
  if (int rc = map4(&S.M.a[1], x3 ? f1_lo : f1_hi, Cp, W1, H1, B, QW, QH)) return rc;
  Params &P = S.P;
  for (int l = 0; l < kMaxLevels; ++l) {
    const int ll = l < L ? l : L - 1;
    P.h2[l] = H2 >> ll;
    P.w2[l] = W2 >> ll;
    if (P.h2[l] < 1 || P.w2[l] < 1) {
      set_error("rc_alt_corr_pyramid: level %d of a %dx%d grid is empty", ll, H2, W2);
      return RC_ERR_INVALID;
    }
    for (int sh = 0; sh < 2; ++sh) {
      const int bw = sh ? CW1 : CW0, bh = (sh ? CH1 : CH0) / ncta;  // a CTA of a pair loads half of the chunk's rows
      if (int rc = map4(&S.M.b[l][sh][0], f2_hi[ll], Cp, P.w2[l], P.h2[l], B, bw, bh)) return rc;
      if (int rc = map4(&S.M.b[l][sh][1], x3 ? f2_lo[ll] : f2_hi[ll], Cp, P.w2[l], P.h2[l], B, bw, bh)) return rc;
    }
  }
  P.B = B; P.H1 = H1; P.W1 = W1; P.L = L;
  P.k_blocks = Cp / TK;
  P.npass = f8 ? 2 : (x3 ? 3 : 1);
  P.f8 = f8 ? 1 : 0;
  P.tiles_x = (W1 + QW - 1) / QW;
  P.tiles_y = (H1 + QH * ncta - 1) / (QH * ncta);
  P.total_tiles = (long long)B * P.tiles_x * P.tiles_y;
  S.radius = radius;
  S.magic = kSetupMagic;
  memcpy(blob, &S, sizeof(S));
  return RC_OK;
}

template <int R, int NCTA>
static int alt_tc_run_r(const alttc::Setup &S, cudaStream_t st) {
  using namespace alttc;
  const Params &P = S.P;
  {
    const unsigned blocks = (unsigned)((P.total_tiles + 3) / 4);
    RC_CUDA(launch_pdl(alt_plan_kernel<R, NCTA>, dim3(blocks), dim3(128), 0, st, P.coords, const_cast<int4 *>(P.plan), P.B, P.H1,
                       P.W1, P.L, P.tiles_x, P.tiles_y, P.total_tiles, P.h2[0], P.w2[0], P.npass >= 2 ? 1 : 0));
  }
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
  auto kern = alt_tc_kernel<R, NCTA>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  const long long units = sm_count() / NCTA;
  const unsigned grid = (unsigned)((P.total_tiles < units ? P.total_tiles : units) * NCTA);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = NCTA;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  RC_CUDA(cudaLaunchKernelEx(&cfg, kern, S.M, P));
  return RC_OK;
}

int alt_tc_run(const void *blob, const float *coords, float *out, float scale, void *plan_ws, size_t plan_bytes,
               cudaStream_t st) {
  using namespace alttc;
  alignas(64) Setup S;
  memcpy(&S, blob, sizeof(S));
  if (S.magic != kSetupMagic) {
    set_error("rc_alt_corr_pyramid_run: the setup blob was not initialised by rc_alt_corr_pyramid_setup");
    return RC_ERR_INVALID;
  }
  if (S.P.f8 && (S.P.amax1 == nullptr || S.P.amax2 == nullptr)) {
    set_error("rc_alt_corr_pyramid_run: RC_PREC_FP16_FP8 needs rc_alt_corr_pyramid_set_norm after the setup");
    return RC_ERR_INVALID;
  }
  const size_t need = alt_tc_plan_bytes(S.P.B, S.P.H1, S.P.W1, S.P.L);
  if (plan_ws == nullptr || plan_bytes < need || (reinterpret_cast<uintptr_t>(plan_ws) & 15) != 0) {
    set_error("rc_alt_corr_pyramid: plan scratch missing, misaligned or too small (%zu < %zu bytes)", plan_bytes, need);
    return RC_ERR_WORKSPACE;
  }
  S.P.coords = coords;
  S.P.out = out;
  S.P.plan = static_cast<const int4 *>(plan_ws);
  S.P.scale = scale;
  if (S.ncta == 2) return S.radius == 3 ? alt_tc_run_r<3, 2>(S, st) : alt_tc_run_r<4, 2>(S, st);
  return S.radius == 3 ? alt_tc_run_r<3, 1>(S, st) : alt_tc_run_r<4, 1>(S, st);
}

int alt_tc_set_norm(void *blob, const unsigned *amax1, const unsigned *amax2) {
  using namespace alttc;
  alignas(64) Setup S;
  memcpy(&S, blob, sizeof(S));
  if (S.magic != kSetupMagic) {
    set_error("rc_alt_corr_pyramid_set_norm: the setup blob was not initialised by rc_alt_corr_pyramid_setup");
    return RC_ERR_INVALID;
  }
  S.P.amax1 = amax1;
  S.P.amax2 = amax2;
  memcpy(blob, &S, sizeof(S));
  return RC_OK;
}

int launch_alt_tc(const void *f1_hi, const void *f1_lo, const void *const *f2_hi, const void *const *f2_lo,
                  const float *coords, float *out, int B, int H1, int W1, int Cp, int L, int radius, int precision,
                  float scale, void *plan_ws, size_t plan_bytes, cudaStream_t st) {
  alignas(64) unsigned char blob[sizeof(alttc::Setup)];
  if (int rc = alt_tc_setup(blob, f1_hi, f1_lo, f2_hi, f2_lo, B, H1, W1, H1, W1, Cp, L, radius, precision)) return rc;
  return alt_tc_run(blob, coords, out, scale, plan_ws, plan_bytes, st);
}

}  // namespace rc
