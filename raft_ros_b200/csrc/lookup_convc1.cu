// Pyramid lookup FUSED into its only consumer: the 1x1 convolution `convc1` + ReLU of the motion
// encoder (reference core/update.py:83,90 -- raft-small: :66,72), whose input is exactly the
// (B, L*(2r+1)^2, h, w) tensor that CorrBlock.__call__ (core/corr.py:29-50) returns:
//
//     out[b, o, p] = relu( bias[o] + sum_c W[o, c] * lookup(coords)[b, c, p] )        c = l*(2r+1)^2 + i*(2r+1) + j
//
// The 324 taps of a query pixel never reach global memory: the gather warps of the tiled lookup
// (lookup_tiled.cu, same window arithmetic) write them as 16-bit K-major SWIZZLE_128B operand rows
// into shared memory, and a tcgen05 GEMM (M = 128 query pixels = TMEM lanes, N = Cout <= 256,
// K = 96 per level, fp32 accumulator in TMEM) contracts them with the weights, level by level.
// Per launch at batch 8 @ 55x128 that removes the 73 MB fp32 output write of the lookup and the
// convolution's re-read of it; what is written instead is the (B, Cout, h, w) fp32 activation.
//
// Operand arithmetic: fp16 (what the reference's --mixed_precision autocast gives convc1) or bf16,
// single pass, fp32 accumulate, fp32 bias + ReLU + output.  The fp32 (non-autocast) convolution has
// no tensor-core equivalent in one pass; this entry point is a stated-bounds fast path next to
// rc_corr_lookup_tiled_fwd, not a replacement for it.
//
// One CTA per SM, persistent over tiles of 128 consecutive query pixels of one frame pair, 22 warps:
//   warps 0-15  gather: warp g owns pixels 8g .. 8g+7 of the tile, one work item per level (lane = pixel x
//               quarter, cp.async patch buffer, separable bilinear taps from registers -- the lookup kernel's
//               item), taps -> A[level parity] in shared memory.  Sixteen warps with one buffer each rather
//               than eight with two: the item is bound by instruction latency, which more warps hide.
//   warps 16-19 epilogue: thread <-> pixel, tcgen05.ld 32 output channels at a time, bias, ReLU,
//               one coalesced 128 B store per channel and warp
//   warp 20     TMA producer of the weight blocks ([Cout][64] 16-bit, SWIZZLE_128B, from L2)
//   warp 21     MMA issuer (tcgen05.mma kind::f16, 128 x Cout x 16), TMEM owner
// K order inside a level: the quarter-lane that forms a tap also fixes its K index (slot t of quarter q sits
// in chunk (SLOTS/8)*q + t/8, SLOTS = 24 for r = 4, 16 for r = 3; rc_corr_convc1_prep_weight lays the weights
// out the same way), so a tap's shared-memory address is the sum of two per-lane registers and two neighbouring
// taps leave as one packed 32-bit store.
// TMEM holds two accumulators (2 x 256 columns): the epilogue of tile t overlaps tile t+1.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace fc {

using namespace tc;

constexpr int kGatherWarps = 16, kEpiWarps = 4;
constexpr int kProducerWarp = 20, kMmaWarp = 21, kThreads = 22 * 32;
constexpr int TMPIX = 128;                         // query pixels per tile (TMEM lanes)
constexpr int kPix = 8;                            // query pixels per gather item
constexpr int PS = 16;                             // patch row stride (floats): the 16 fetched columns
constexpr int kPatch = 10 * PS + 12;               // per-pixel patch (floats); the pad spreads the banks
constexpr int kStageFloats = kPix * kPatch;
constexpr int kStages = 1;
constexpr int A_KBLOCK = TMPIX * 64 * 2;           // one 64-wide K block of the tap operand: 16 KB
constexpr int A_BYTES = 2 * A_KBLOCK;              // taps of one level: K padded to 128
constexpr int W_BLOCK = 256 * 64 * 2;              // one weight block: up to 256 rows x 64 K
constexpr int kAccCols = 256;
constexpr int kSmemBytes = 1024 + 2 * A_BYTES + 2 * W_BLOCK + kGatherWarps * kStages * kStageFloats * 4 + 1024 /*bias*/ + 256;
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

struct Params {
  ConstLevelPtrs lv;
  const float *coords;
  const float *bias;
  float *out;
  int B, h, w, L, Cout;
  int tiles_per_b, total_tiles;
};

struct Ctx {
  int lvl, b, p0, valid;
  int x0, y0;
  float fx, fy;
};

__device__ __forceinline__ void sts16(uint32_t a, unsigned short v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(v) : "memory");
}

// V16: fp16 volume (tiles of 2 x 16 halves); RR: window radius 3 or 4; BF16OPS: bf16 instead of fp16 operands
template <bool V16, int RR, bool BF16OPS>
__global__ void __launch_bounds__(kThreads, 1)
lookup_convc1_kernel(const __grid_constant__ CUtensorMap tmW, const Params P) {
  constexpr int R = RR, RD = 2 * R + 1, PN = RD + 1, NTAP = RD * RD;
  constexpr int KSTEPS = (NTAP + 15) / 16;   // 6 (r = 4: 81 -> 96), 4 (r = 3: 49 -> 64)
  constexpr int KBLK = (KSTEPS + 3) / 4;     // 64-wide K blocks per level: 2, 1
  constexpr int KP = KBLK * 64;              // K per level in the prepared weights
  static_assert(R == 3 || R == 4, "tap assignment below is written for radius 3 and 4");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  const uint32_t a_u32 = base_u32, w_u32 = base_u32 + 2 * A_BYTES;
  float *gather = reinterpret_cast<float *>(smem + 2 * A_BYTES + 2 * W_BLOCK);
  float *bias_s = gather + kGatherWarps * kStages * kStageFloats;
  uint64_t *bars = reinterpret_cast<uint64_t *>(bias_s + 256);
  uint64_t *a_full = bars, *a_empty = bars + 2, *w_full = bars + 4, *w_empty = bars + 6, *acc_full = bars + 8,
           *acc_empty = bars + 10;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N = P.h * P.w, L = P.L, Cout = P.Cout;
  pdl_launch_dependents();

  if (warp == kProducerWarp && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&a_full[i], kGatherWarps);
      mbar_init(&a_empty[i], 1);
      mbar_init(&w_full[i], 1);
      mbar_init(&w_empty[i], 1);
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kEpiWarps);
    }
    fence_barrier_init();
    prefetch_tmap(&tmW);
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  // the K padding of the tap operand (columns NTAP .. 127 of every row) is zero and is never written again
  for (int i = threadIdx.x; i < 2 * A_BYTES / 16; i += kThreads) sts128(a_u32 + i * 16, 0.f, 0.f, 0.f, 0.f);
  fence_proxy_async_smem();
  pdl_wait();  // nothing of an earlier kernel (weights, bias, coordinates, volume) is read before this
  for (int i = threadIdx.x; i < 256; i += kThreads) bias_s[i] = i < Cout ? __ldg(P.bias + i) : 0.0f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kGatherWarps) {
    // ===================== gather: taps of this warp's 8 pixels, level by level -> A =====================
    const int tu = lane >> 2, tc = lane & 3;
    float *patchbuf = gather + warp * kStageFloats;
    const uint32_t patch_u32 = smem_u32(patchbuf);
    const uint32_t dst_lane = (uint32_t)(tu * kPatch + tc * 4) * 4u;
    // operand row of this lane's pixel: K-major, 128 B rows, 16 B chunk c at position c ^ (row & 7).  This lane's
    // taps are slots 0 .. SLOTS-1 of quarter tc, i.e. K = SLOTS*tc + t: chunk (SLOTS/8)*tc + (t >> 3), a constant
    // offset (t & 7) inside it.
    constexpr int SLOTS = KSTEPS * 4;  // 24 (r = 4), 16 (r = 3)
    constexpr int NCH = SLOTS / 8;     // chunks per quarter: 3, 2
    const int row = warp * 8 + tu, rx = row & 7;
    uint32_t chunk_off[NCH];           // byte offset of this lane's chunks inside an A buffer
#pragma unroll
    for (int m = 0; m < NCH; ++m) {
      const int ch = NCH * tc + m;
      chunk_off[m] = (uint32_t)((ch >> 3) * A_KBLOCK + (row >> 3) * 1024 + rx * 128 + (((ch & 7) ^ rx) << 4));
    }
    // inside its chunk, quarter tc keeps slot t at element (t + 2 tc) & 7: the four quarters of a pixel then hit
    // four different 32-bit words with the same store instruction, and the 32 lanes 32 different banks
    uint32_t word_off[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) word_off[q] = (uint32_t)(((q + tc) & 3) << 2);
    auto pack2 = [](float a, float b) -> uint32_t {  // a -> low half (K even), b -> high half
      uint32_t r;
      if constexpr (BF16OPS) asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
      else asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
      return r;
    };
    auto copy16 = [&](uint32_t dst, const float *src, bool ok) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
    };
    // the lookup kernel's copy issue (lookup_tiled.cu): lane (u, k) fetches the k-th 16 B piece of the PN
    // window rows of pixel u, zero-filled outside the level
    auto issue = [&](const Ctx &c) {
      const int hl = P.h >> c.lvl, wl = P.w >> c.lvl;
      constexpr int CW = V16 ? 16 : 8, PC = V16 ? 8 : 4;
      const int hp = (hl + 1) & ~1, wp = (wl + CW - 1) & ~(CW - 1), tpr = wp / CW;
      const int wb = (wl + PC - 1) & ~(PC - 1);
      const int cg = (c.x0 & ~(PC - 1)) + PC * tc;
      const bool want = PC * tc < (c.x0 & (PC - 1)) + PN;
      const bool okx = want && (unsigned)cg < (unsigned)wb;
      constexpr int G = kTileGroup;
      const int ng = (N + G - 1) / G;
      const size_t grp = (size_t)c.b * ng + c.p0 / G;
      const int u = c.p0 % G + tu, sw = (u >> 1) & 3, half = (cg / PC) & 1;
      const size_t gwords = (size_t)(hp * wp) * G / (V16 ? 2 : 1);
      const float *ibase = pick_level(P.lv, c.lvl) + grp * gwords + u * 16;  // always inside the level
      const int trs = tpr * 16 * G;
      const int pos0 = ((half ^ sw) << 2), pos1 = (((2 | half) ^ sw) << 2);
      // two arithmetic sequences of offsets (r even / r odd), as in lookup_tiled.cu
      const int par = c.y0 & 1, t0 = (c.y0 >> 1) * trs + (cg / CW) * (16 * G);
      const int tA = t0 + (par ? pos1 : pos0), tB = t0 + (par ? trs + pos0 : pos1);
      constexpr uint32_t RB = V16 ? 48u : (uint32_t)PS * 4u;
      const uint32_t dst = patch_u32 + dst_lane;
#pragma unroll
      for (int r = 0; r < PN; ++r) {
        const bool ok = okx && (unsigned)(c.y0 + r) < (unsigned)hl;
        const int rel = ((r & 1) ? tB : tA) + (r >> 1) * trs;
        if (!V16 || tc < 3) copy16(dst + (uint32_t)r * RB, ibase + (ok ? rel : 0), ok);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };

    uint32_t cnt = 0;  // levels done: A buffer = cnt & 1
    float ncx = 0.f, ncy = 0.f;  // coordinates of the next tile's pixel, loaded a tile ahead
    auto load_coords = [&](int tile, float &x, float &y) {
      x = 0.f;
      y = 0.f;
      if (tile < P.total_tiles) {
        const int b = tile / P.tiles_per_b, p = (tile - b * P.tiles_per_b) * TMPIX + warp * 8 + tu;
        if (p < N) {
          x = __ldg(P.coords + ((size_t)b * 2 + 0) * N + p);
          y = __ldg(P.coords + ((size_t)b * 2 + 1) * N + p);
        }
      }
    };
    load_coords(blockIdx.x, ncx, ncy);
    for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
      const float cx = ncx, cy = ncy;
      load_coords(tile + gridDim.x, ncx, ncy);
      const int b = tile / P.tiles_per_b, p0 = (tile - b * P.tiles_per_b) * TMPIX + warp * 8;
      const bool pv = p0 + tu < N;
      for (int l = 0; l < L; ++l, ++cnt) {
        Ctx cur{l, b, p0, 1, 0, -(1 << 28), 0.f, 0.f};
        if (pv) {
          const float inv = 1.0f / (float)(1 << l);
          cur.x0 = safe_floor_int(cx * inv, &cur.fx) - R;
          cur.y0 = safe_floor_int(cy * inv, &cur.fy) - R;
        }
        issue(cur);
        const uint32_t buf = cnt & 1;
        mbar_wait(&a_empty[buf], ((cnt >> 1) & 1) ^ 1);  // the MMAs that read this buffer two levels ago are done
        const uint32_t a_buf = a_u32 + buf * A_BYTES;
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        const float dx = cur.fx, dy = cur.fy, ex = 1.0f - dx, ey = 1.0f - dy;
        const float *pbase = patchbuf + tu * kPatch;
        auto at = [&](int j, int i) -> float {
          if constexpr (V16) return __half2float(reinterpret_cast<const __half *>(pbase)[j * 24 + (cur.x0 & 7) + i]);
          else return pbase[j * PS + (cur.x0 & 3) + i];
        };
        float tap[SLOTS];  // this lane's taps in slot order; unused slots stay zero
#pragma unroll
        for (int t = 0; t < SLOTS; ++t) tap[t] = 0.0f;
        if (tc < RD / 2) {
          // slots 0 .. RD-1: window column 2tc, slots RD .. 2RD-1: column 2tc+1 (both with j = 0 .. RD-1)
          float h0[PN], h1[PN];
#pragma unroll
          for (int j = 0; j < PN; ++j) {
            const float a = at(j, 2 * tc), bb = at(j, 2 * tc + 1), d = at(j, 2 * tc + 2);
            h0[j] = fmaf(bb, dx, a * ex);
            h1[j] = fmaf(d, dx, bb * ex);
          }
#pragma unroll
          for (int j = 0; j < RD; ++j) {
            tap[j] = fmaf(h0[j + 1], dy, h0[j] * ey);
            tap[RD + j] = fmaf(h1[j + 1], dy, h1[j] * ey);
          }
        }
        if constexpr (R == 4) {
          // window column 8, rows shared out over the quarters: quarter tc takes j = tc, tc+4 (and 8 for tc = 0) as
          // slots 18, 19, 20
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const int j = tc + 4 * k;
            if (j < RD && (k < 2 || tc == 0)) {
              const float a = at(j, 8), bb = at(j, 9), c2 = at(j + 1, 8), d2 = at(j + 1, 9);
              const float top = fmaf(bb, dx, a * ex), bot = fmaf(d2, dx, c2 * ex);
              tap[2 * RD + k] = fmaf(bot, dy, top * ey);
            }
          }
        } else if (tc == RD / 2) {
          // r = 3: the fourth quarter owns the last column (6) on its own: slots 0 .. 6
          float hc[PN];
#pragma unroll
          for (int j = 0; j < PN; ++j) hc[j] = fmaf(at(j, RD), dx, at(j, RD - 1) * ex);
#pragma unroll
          for (int j = 0; j < RD; ++j) tap[j] = fmaf(hc[j + 1], dy, hc[j] * ey);
        }
        // pairs of slots leave as one 32-bit store; slots a quarter never fills are left alone (they are zero)
        constexpr int USED = (R == 4) ? 22 : 14;  // rounded up to even
#pragma unroll
        for (int t = 0; t < USED; t += 2)
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(a_buf + chunk_off[t >> 3] + word_off[(t & 7) >> 1]), "r"(pack2(tap[t], tap[t + 1]))
                       : "memory");
        // this warp's 8 rows of the level are in place: make them visible to the tensor-core (async) proxy
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full[buf]);
      }
    }
  } else if (warp < kGatherWarps + kEpiWarps) {
    // ===================== epilogue: bias + ReLU, thread <-> pixel =====================
    const int e = warp - kGatherWarps;  // = warp % 4: the TMEM lane quarter this warp may read
    int acc = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
      const int b = tile / P.tiles_per_b, p = (tile - b * P.tiles_per_b) * TMPIX + e * 32 + lane;
      mbar_wait(&acc_full[acc], phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + (uint32_t)(acc * kAccCols);
      float *o = P.out + (size_t)b * Cout * N + p;
      const bool pv = p < N;
      int c0 = 0;
      for (; c0 + 32 <= Cout; c0 += 32) {
        float v[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
        if (pv) {
#pragma unroll
          for (int i = 0; i < 32; ++i) st_stream(o + (size_t)(c0 + i) * N, fmaxf(v[i] + bias_s[c0 + i], 0.0f));
        }
      }
      if (c0 < Cout) {  // Cout is a multiple of 16
        float v[16];
        tmem_ld16(taddr + c0, v);
        tmem_ld_wait();
        if (pv) {
#pragma unroll
          for (int i = 0; i < 16; ++i) st_stream(o + (size_t)(c0 + i) * N, fmaxf(v[i] + bias_s[c0 + i], 0.0f));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
      acc ^= 1;
      if (acc == 0) phase ^= 1;
    }
  } else if (warp == kProducerWarp) {
    // ===================== weight blocks: L2 -> shared memory =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x)
        for (int l = 0; l < L; ++l)
          for (int kb = 0; kb < KBLK; ++kb) {
            mbar_wait(&w_empty[stage], phase ^ 1);
            mbar_expect_tx(&w_full[stage], (uint32_t)Cout * 128u);
            tma_load_3d(smem + 2 * A_BYTES + stage * W_BLOCK, &tmW, &w_full[stage], l * KP + kb * 64, 0, 0);
            if (++stage == 2) { stage = 0; phase ^= 1; }
          }
    }
  } else {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      // kind::f16: D = f32; A, B = f16 (format 0) or bf16 (format 1), both K-major; M = 128, N = Cout
      const uint32_t idesc = (1u << 4) | (BF16OPS ? ((1u << 7) | (1u << 10)) : 0u) | ((uint32_t)(Cout >> 3) << 17) |
                             ((uint32_t)(TMPIX >> 4) << 24);
      int ws = 0, acc = 0;
      uint32_t wphase = 0, accphase = 0, cnt = 0;
      for (int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
        mbar_wait(&acc_empty[acc], accphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * kAccCols);
        for (int l = 0; l < L; ++l, ++cnt) {
          const uint32_t buf = cnt & 1;
          mbar_wait(&a_full[buf], (cnt >> 1) & 1);
          tc_fence_after();
          for (int kb = 0; kb < KBLK; ++kb) {
            mbar_wait(&w_full[ws], wphase);
            tc_fence_after();
            const uint64_t adesc = make_sw128_desc(a_u32 + buf * A_BYTES + kb * A_KBLOCK);
            const uint64_t bdesc = make_sw128_desc(w_u32 + ws * W_BLOCK);
            const int nk = KSTEPS - 4 * kb < 4 ? KSTEPS - 4 * kb : 4;
            for (int k = 0; k < nk; ++k)
              umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (l | kb | k) != 0 ? 1u : 0u);
            umma_commit(&w_empty[ws]);
            if (++ws == 2) { ws = 0; wphase ^= 1; }
          }
          umma_commit(&a_empty[buf]);
        }
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

// weights (Cout, L*(2r+1)^2) fp32 -> (Cout, L*KP) 16-bit in the kernel's K order: inside a level, slot t of quarter q
// (K = SLOTS*q + 8*(t/8) + (t + 2q) % 8) holds the tap that quarter-lane q forms as its t-th (see the gather above);
// unused slots are zero.
//   r = 4 (SLOTS 24): q = 0..3: t < 9 -> tap (i = 2q, j = t); t < 18 -> (2q+1, t-9); t = 18, 19 -> (8, q), (8, q+4);
//                     q = 0 also t = 20 -> (8, 8)
//   r = 3 (SLOTS 16): q = 0..2: t < 7 -> (2q, t); t < 14 -> (2q+1, t-7); q = 3: t < 7 -> (6, t)
// tap (i, j) is input channel i*(2r+1) + j of the level (i = x offset, j = y offset).
__device__ __forceinline__ int slot_tap(int r, int q, int t) {
  const int rd = 2 * r + 1;
  if (r == 4) {
    if (t < 9) return (2 * q) * rd + t;
    if (t < 18) return (2 * q + 1) * rd + (t - 9);
    if (t == 18) return 8 * rd + q;
    if (t == 19) return 8 * rd + q + 4;
    if (t == 20 && q == 0) return 8 * rd + 8;
    return -1;
  }
  if (q < 3) {
    if (t < 7) return (2 * q) * rd + t;
    if (t < 14) return (2 * q + 1) * rd + (t - 7);
    return -1;
  }
  return t < 7 ? 6 * rd + t : -1;
}

template <bool BF16OPS>
__global__ void convc1_prep_weight_kernel(const float *__restrict__ w, unsigned short *__restrict__ wp, int Cout, int L,
                                          int r, int KP) {
  const int total = Cout * L * KP, ntap = (2 * r + 1) * (2 * r + 1), slots = (r == 4) ? 24 : 16;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int o = i / (L * KP), rem = i - o * (L * KP), l = rem / KP, k = rem - l * KP;
    const int q = k / slots, kk = k - q * slots;
    const int t = (kk & ~7) | ((kk - 2 * q) & 7);  // element (t + 2q) & 7 of its chunk holds slot t (bank spreading, see the gather)
    const int tap = q < 4 ? slot_tap(r, q, t) : -1;
    const float v = tap >= 0 ? __ldg(w + (size_t)o * (L * ntap) + l * ntap + tap) : 0.0f;
    unsigned short rr;
    if (BF16OPS) rr = __bfloat16_as_ushort(__float2bfloat16_rn(v));
    else asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(rr) : "f"(v));
    wp[i] = rr;
  }
}

static int kp_of(int r) { return ((((2 * r + 1) * (2 * r + 1) + 15) / 16 + 3) / 4) * 64; }

template <bool V16, int RR, bool BF16OPS>
static int launch_t(const CUtensorMap &tm, const Params &P, cudaStream_t st) {
  auto kern = lookup_convc1_kernel<V16, RR, BF16OPS>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int dev = 0, sms = 148;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = P.total_tiles < sms ? P.total_tiles : sms;
  RC_CUDA(launch_pdl(kern, dim3((unsigned)grid), dim3(kThreads), (size_t)kSmemBytes, st, tm, P));
  return RC_OK;
}

}  // namespace fc

size_t convc1_weight_bytes(int Cout, int L, int r) { return (size_t)Cout * L * fc::kp_of(r) * 2; }

int launch_convc1_prep_weight(const float *w, void *wp, int Cout, int L, int r, int bf16_ops, cudaStream_t st) {
  const int KP = fc::kp_of(r);
  const int total = Cout * L * KP;
  const int blocks = (total + 255) / 256;
  if (bf16_ops) fc::convc1_prep_weight_kernel<true><<<blocks, 256, 0, st>>>(w, static_cast<unsigned short *>(wp), Cout, L, r, KP);
  else fc::convc1_prep_weight_kernel<false><<<blocks, 256, 0, st>>>(w, static_cast<unsigned short *>(wp), Cout, L, r, KP);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_lookup_convc1(const ConstLevelPtrs &lv, bool vol16, const float *coords, const void *wp, const float *bias,
                         float *out, int B, int h, int w, int L, int r, int Cout, int bf16_ops, cudaStream_t st) {
  using namespace fc;
  const int KP = kp_of(r);
  CUtensorMap tm;
  {
    const unsigned long long dims[3] = {(unsigned long long)L * KP, (unsigned long long)Cout, 1};
    const unsigned long long strides[2] = {(unsigned long long)L * KP * 2, (unsigned long long)L * KP * 2 * Cout};
    const unsigned int box[3] = {64, (unsigned int)Cout, 1};
    if (int rc = encode_tmap_sw128(&tm, wp, 3, dims, strides, box, "rc_corr_lookup_convc1_fwd")) return rc;
  }
  Params P{};
  P.lv = lv;
  P.coords = coords;
  P.bias = bias;
  P.out = out;
  P.B = B; P.h = h; P.w = w; P.L = L; P.Cout = Cout;
  P.tiles_per_b = (h * w + TMPIX - 1) / TMPIX;
  P.total_tiles = B * P.tiles_per_b;
#define RC_FC(V, RR_)                                                                   \
  (bf16_ops ? launch_t<V, RR_, true>(tm, P, st) : launch_t<V, RR_, false>(tm, P, st))
  if (r == 4) return vol16 ? RC_FC(true, 4) : RC_FC(false, 4);
  if (r == 3) return vol16 ? RC_FC(true, 3) : RC_FC(false, 3);
#undef RC_FC
  set_error("rc_corr_lookup_convc1_fwd: radius %d unsupported (3 and 4 only)", r);
  return RC_ERR_UNSUPPORTED;
}

}  // namespace rc
