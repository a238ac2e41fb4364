// Stride-1 "same" convolutions of the per-iteration update (reference core/update.py: the motion encoder's convc2 /
// convf2 / conv, SepConvGRU's 1x5 and 5x1 gates, the flow and mask heads) as an implicit GEMM on tcgen05 whose filter
// taps cost no copies:
//
//     out[b, o, y, x] = act( bias[o] + sum_{c, ky, kx} W[o, c, ky, kx] * in[b, c, y + ky - KH/2, x + kx - KW/2] )
//
// A tile is 128 output pixels: one row segment (or 8 rows x 16 columns when the filter is one column wide).  Converter warps read the input once per 64-channel block
// (NCHW fp32, coalesced along x), round it to fp16 and lay the tile WITH ITS HALO -- KH rows of 128 + KW - 1 pixels --
// out in shared memory as K-major SWIZZLE_128B operand rows (one 128 B row per halo pixel).  Filter tap (ky, kx) then
// is the SAME buffer read from row ky * (128 + KW - 1) + kx on: the A descriptor of each tap's MMAs just starts that
// many rows (x 128 B) later -- the 128 B swizzle is a function of the shared-memory address alone, so a start that is
// not on a 1024 B boundary reads the rows exactly as they were written (base-offset field 0; measured: setting it to the
// start row's phase gives wrong results) -- and the KH * KW taps re-use one conversion.  Weights ([tap][Cout][64] fp16 blocks, prepared once per model) stream from L2 through a TMA ring;
// the accumulator (128 pixels x Cout <= 256, fp32) lives in TMEM, double-buffered; the epilogue (thread <-> pixel) adds
// the bias, applies the activation and writes NCHW fp32, one coalesced 128 B store per channel and warp.  The input
// may be the channel concatenation of two tensors (core/update.py:41,47,52,58,95,134 `torch.cat([...], dim=1)`), read in
// place.  Operands are fp16 (11-bit mantissa: the class of TF32, which is what cuDNN runs these convolutions in by
// torch's default, and of the reference's --mixed_precision autocast), one pass, fp32 accumulation.
//
// One CTA per SM, persistent, 15 warps: 0-7 converters, 8-11 epilogue, 12 weight TMA, 13 input TMA, 14 MMA issue.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace cv {

using namespace tc;

constexpr int kConvWarps = 8, kEpiWarps = 4, kProducerWarp = 12, kInputWarp = 13, kMmaWarp = 14, kThreads = 15 * 32;
constexpr int TMPIX = 128;
constexpr int kMaxHaloPix = 400;                    // 3 x 130 (3x3), 12 x 16 (5x1 on 8 x 16 tiles), 1 x 132 (1x5)
constexpr int kHaloBytes = kMaxHaloPix * 128;       // one 64-channel block of the halo tile: 50 KB
constexpr int kInStages = 12;                       // fp32 input staging: 8 channels of the halo tile per stage, filled by TMA (at most; P.in_stages)
constexpr int kMaxBoxPix = 448;                     // staged box: rows rounded up to 8 columns (3 x 136)
constexpr int kInStageBytes = 8 * kMaxBoxPix * 4;  //   [8 channels][rows][columns rounded up to 4] fp32 = 12.5 KB
constexpr int kMaxStages = 12;                      // weight ring: as many blocks of (filters of this CTA) x 64 K as the rest holds
constexpr int kTailBytes = 1024 /*bias*/ + 512 /*barriers*/;
constexpr int kSmemBytes = 227 * 1024;              // all of it: what the halo and staging buffers of this filter shape leave goes to the ring
constexpr int kDataBytes = kSmemBytes - 1024 /*alignment slack*/ - kTailBytes;
static_assert(2 * kHaloBytes + 4 * kInStageBytes + 2 * 32 * 1024 <= kDataBytes, "four input stages and two 32 KB weight stages must always fit");
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

struct Params {
  const float *x1, *x2;  // input = channels [0, C1) of x1 followed by [0, Cin - C1) of x2 (x2 may be null)
  const float *bias;
  float *out;
  long long x1_bs, x2_bs, out_bs;  // floats from one batch item to the next (a channel slice of a wider tensor has more than C*H*W)
  int epi, split;                  // RC_EPI_*: what the epilogue does with act(acc + bias)
  int cstore;                      // filters [0, cstore) are stored, the rest (zero padding of the filter count) dropped
  const float *aux1, *aux2;
  long long aux1_bs, aux2_bs;
  int B, C1, Cin, Cout, H, W, KH, KW, act;
  int TH, TW;  // output tile: TH rows x TW columns = 128 pixels (1 x 128, or 8 x 16 for filters one column wide)
  int ncb, tiles_x, tiles_y, total_tiles;
  int BH, BW;            // staged box of the input: TH + KH - 1 rows of (TW + KW - 1 rounded up to 4) columns -- or the whole
  int box_x0, box_y0;    //   image extent where that is smaller (a box may not exceed the tensor): then it starts at 0 (flags)
  int box_shift;         // columns between the box's 16 B-aligned start and the halo tile's first column
  int nparts, Cn;        // a work item = (pixel tile, one of nparts groups of Cn filters): small images still fill the SMs
  int stages, stage_bytes;
  int in_stages;
  int halo_bytes, in_stage_bytes, in_off, w_off;  // shared-memory layout of this launch: [2 halo][kInStages staging][ring][bias, barriers]
};

// The activation of a whole register chunk, branched ONCE per chunk: a per-element switch is if-converted by the compiler
// into evaluating every activation for every element (exp, reciprocal and tanh for a ReLU layer: measured 4.4 us of a
// 18 us kernel).  tanh(v) = 1 - 2 / (1 + e^(2v)): absolute error ~1e-7, exact limits at both ends.
template <int N>
__device__ __forceinline__ void activate_all(float (&v)[N], int act) {
  if (act == RC_ACT_RELU) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = fmaxf(v[i], 0.0f);
  } else if (act == RC_ACT_SIGMOID) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = __fdividef(1.0f, 1.0f + __expf(-v[i]));
  } else if (act == RC_ACT_TANH) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * v[i]));
  }
}

// TMA_IN: the input reaches the converters through a ring of TMA boxes (rows of 16 B multiples, i.e. W % 4 == 0, and a
// first tensor whose channel count is a multiple of 8); the box's zero fill outside the image IS the padding.  A box
// starts on a 16 B boundary of the image row (whole rows above the image are fine, a start at column x0 - 1 is not).  Otherwise
// the converters read global memory themselves (any width; slower: few bytes in flight per SM).
template <bool TMA_IN, bool EPI>
__global__ void __launch_bounds__(kThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX1,
               const __grid_constant__ CUtensorMap tmX2, const Params P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  const int kInOff = P.in_off, kWOff = P.w_off, kHaloB = P.halo_bytes, kInStageB = P.in_stage_bytes;
  const uint32_t halo_u32 = base_u32, w_u32 = base_u32 + kWOff;
  float *bias_s = reinterpret_cast<float *>(smem + kDataBytes);
  uint64_t *bars = reinterpret_cast<uint64_t *>(bias_s + 256);
  uint64_t *h_full = bars, *h_empty = bars + 2, *acc_full = bars + 4, *acc_empty = bars + 6, *w_full = bars + 8,
           *w_empty = bars + 8 + kMaxStages, *s_full = bars + 8 + 2 * kMaxStages, *s_empty = s_full + kInStages;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(s_empty + kInStages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int KH = P.KH, KW = P.KW, T = KH * KW, ph = KH >> 1, pw = KW >> 1;
  const int TH = P.TH, TW = P.TW;
  const int RW = TW + 2 * pw, HP = (TH + 2 * ph) * RW;  // halo tile: TH + KH - 1 rows of RW pixels
  const int Cout = P.Cout, Cn = P.Cn, H = P.H, W = P.W;
  const int items = P.total_tiles * P.nparts;
  const size_t HW = (size_t)H * W;
  pdl_launch_dependents();

  if (warp == kProducerWarp && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&h_full[i], kConvWarps);
      mbar_init(&h_empty[i], 1);
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kEpiWarps);
    }
    for (int i = 0; i < kMaxStages; ++i) {
      mbar_init(&w_full[i], 1);
      mbar_init(&w_empty[i], 1);
    }
    for (int i = 0; i < kInStages; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], kConvWarps);
    }
    fence_barrier_init();
    prefetch_tmap(&tmW);
    if (TMA_IN) {
      prefetch_tmap(&tmX1);
      prefetch_tmap(&tmX2);
    }
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  pdl_wait();
  for (int i = threadIdx.x; i < 256; i += kThreads) bias_s[i] = (i < Cout && P.bias) ? __ldg(P.bias + i) : 0.0f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kConvWarps) {
    // ===================== converters: input tile with halo -> fp16 operand rows, one 64-channel block at a time =====================
    const int tid = threadIdx.x;
    uint32_t n = 0;  // channel blocks done (over all tiles): buffer n & 1
    int ist = 0;     // input staging ring
    uint32_t iphase = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      const int tile = item / P.nparts;
      const int txi = tile % P.tiles_x, by = tile / P.tiles_x, tyi = by % P.tiles_y, b = by / P.tiles_y;
      const int x0 = txi * TW, y = tyi * TH;
      // TMA path: this thread's halo pixels tid and tid + 256 (HP <= 400): offset in a staged box (-1: outside the box, i.e.
      // zero), row offset and swizzle phase in the operand tile
      const int plane = P.BH * P.BW;
      int hp_src0 = -1, hp_src1 = -1;
      uint32_t hp_dst0 = 0, hp_dst1 = 0, hp_sw0 = 0, hp_sw1 = 0;
      bool hp_ok0 = false, hp_ok1 = false;
      if constexpr (TMA_IN) {
        // a box pinned to the image origin (image narrower / lower than the halo tile) is shifted against the tile
        const int dx = P.box_x0 ? pw - x0 : -P.box_shift, dy = P.box_y0 ? ph - y : 0;
        auto place = [&](int hi, int &src_off, uint32_t &dst, uint32_t &sw, bool &ok) {
          ok = hi < HP;
          const int r = hi / RW, xx = hi - r * RW;
          const int br = r - dy, bx = xx - dx;
          src_off = (ok && (unsigned)br < (unsigned)P.BH && (unsigned)bx < (unsigned)P.BW) ? br * P.BW + bx : -1;
          dst = (uint32_t)hi * 128u;
          sw = (uint32_t)(hi & 7);
        };
        place(tid, hp_src0, hp_dst0, hp_sw0, hp_ok0);
        place(tid + kConvWarps * 32, hp_src1, hp_dst1, hp_sw1, hp_ok1);
      }
      for (int cb = 0; cb < P.ncb; ++cb, ++n) {
        const uint32_t buf = n & 1;
        mbar_wait(&h_empty[buf], ((n >> 1) & 1) ^ 1);  // the MMAs that read this buffer two blocks ago are done
        const uint32_t hb = halo_u32 + buf * kHaloB;
        if constexpr (TMA_IN) {
          // eight channels at a time arrive as one fp32 box [8][BH][BW]; a thread turns a halo pixel's 8 values into one
          // 16 B chunk of its operand row (consecutive lanes: consecutive pixels, conflict-free reads).  A thread owns
          // the same (up to two) halo pixels for the whole tile: their offsets are formed once (hp_* below).
          for (int sub = 0; sub < 8; ++sub) {
            const bool live = cb * 64 + sub * 8 < P.Cin;  // channel padding of the last block: zeros, nothing staged
            const float *src = reinterpret_cast<const float *>(smem + kInOff + ist * kInStageB);
            if (live) mbar_wait(&s_full[ist], iphase);
            float v0[8], v1[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v0[j] = (live && hp_src0 >= 0) ? src[hp_src0 + j * plane] : 0.0f;
              v1[j] = (live && hp_src1 >= 0) ? src[hp_src1 + j * plane] : 0.0f;
            }
            uint32_t w0[4], w1[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w0[j]) : "f"(v0[2 * j + 1]), "f"(v0[2 * j]));
              asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w1[j]) : "f"(v1[2 * j + 1]), "f"(v1[2 * j]));
            }
            if (hp_ok0)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(hb + hp_dst0 + (uint32_t)((sub ^ hp_sw0) << 4)), "r"(w0[0]),
                           "r"(w0[1]), "r"(w0[2]), "r"(w0[3])
                           : "memory");
            if (hp_ok1)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(hb + hp_dst1 + (uint32_t)((sub ^ hp_sw1) << 4)), "r"(w1[0]),
                           "r"(w1[1]), "r"(w1[2]), "r"(w1[3])
                           : "memory");
            if (live) {
              __syncwarp();
              if (lane == 0) mbar_arrive(&s_empty[ist]);
              if (++ist == P.in_stages) { ist = 0; iphase ^= 1; }
            }
          }
        } else {
        // 8-channel chunk c8 of halo pixel hi (consecutive threads: consecutive pixels); two chunks per thread in flight
        auto load8 = [&](int it, float (&v)[8]) {
          const int c8 = it / HP, hi = it - c8 * HP;
          const int r = hi / RW, xx = hi - r * RW;
          const int iy = y - ph + r, ix = x0 - pw + xx;
          const bool inside = it < HP * 8 && (unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W;
          const int c0 = cb * 64 + c8 * 8;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int c = c0 + j;
            const float *src = c < P.C1 ? P.x1 + (size_t)b * P.x1_bs + (size_t)c * HW : P.x2 + (size_t)b * P.x2_bs + (size_t)(c - P.C1) * HW;
            v[j] = (inside && c < P.Cin) ? __ldg(src + (size_t)iy * W + ix) : 0.0f;
          }
        };
        auto store8 = [&](int it, const float (&v)[8]) {
          if (it >= HP * 8) return;
          const int c8 = it / HP, hi = it - c8 * HP;
          uint32_t w4[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w4[j]) : "f"(v[2 * j + 1]), "f"(v[2 * j]));
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(hb + (uint32_t)hi * 128u + (uint32_t)((c8 ^ (hi & 7)) << 4)),
                       "r"(w4[0]), "r"(w4[1]), "r"(w4[2]), "r"(w4[3])
                       : "memory");
        };
        constexpr int kStep = kConvWarps * 32;
        for (int it = tid; it < HP * 8; it += 2 * kStep) {
          float va[8], vb[8];
          load8(it, va);
          load8(it + kStep, vb);
          store8(it, va);
          store8(it + kStep, vb);
        }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&h_full[buf]);
      }
    }
  } else if (warp < kConvWarps + kEpiWarps) {
    // ===================== epilogue: bias + activation, thread <-> pixel =====================
    const int e = warp - kConvWarps;
    int acc = 0;
    uint32_t phase = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      const int tile = item / P.nparts, n0 = (item - tile * P.nparts) * Cn;  // this item's filters: [n0, n0 + Cn)
      const int txi = tile % P.tiles_x, by = tile / P.tiles_x, tyi = by % P.tiles_y, b = by / P.tiles_y;
      const int m = e * 32 + lane;  // output pixel of the tile, row-major
      const int y = tyi * TH + m / TW, x = txi * TW + m % TW;
      const bool pv = x < W && y < H;
      const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + (uint32_t)(acc * 256);
      const size_t pix = (size_t)y * W + x;
      float *o = P.out + (size_t)b * P.out_bs + (size_t)n0 * HW + pix;
      if constexpr (EPI) {
        // RC_EPI_MUL_FROM: channels from `split` on are multiplied by aux1[b, c - split] (the GRU's r * h)
        // RC_EPI_GRU:      out = (1 - z) * h + z * act(.), z = aux1[b, c], h = aux2[b, c]            (core/update.py:46,53)
        // The operands of 16 filters are fetched one chunk ahead -- the first before the accumulator is even complete --
        // so that their L2 round trips overlap the MMAs and the previous chunk's stores instead of adding up.
        const bool gru = P.epi == RC_EPI_GRU;
        const float *a1 = P.aux1 + (size_t)b * P.aux1_bs + pix;
        const float *a2 = gru ? P.aux2 + (size_t)b * P.aux2_bs + pix : a1;
        auto fetch = [&](int c0, float (&u)[16], float (&w)[16]) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int c = n0 + c0 + i;
            if (gru) {
              u[i] = __ldg(a1 + (size_t)c * HW);
              w[i] = __ldg(a2 + (size_t)c * HW);
            } else {
              u[i] = c >= P.split ? __ldg(a1 + (size_t)(c - P.split) * HW) : 1.0f;
              w[i] = 0.0f;
            }
          }
        };
        float u[16], w[16];
        if (pv) fetch(0, u, w);
        mbar_wait(&acc_full[acc], phase);
        tc_fence_after();
        for (int c0 = 0; c0 < Cn; c0 += 16) {
          float v[16], un[16], wn[16];
          tmem_ld16(taddr + c0, v);
          if (pv && c0 + 16 < Cn) fetch(c0 + 16, un, wn);
          tmem_ld_wait();
          float bv[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) bv[i] = bias_s[n0 + c0 + i];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] += bv[i];
          activate_all(v, P.act);
          if (gru) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = (1.0f - u[i]) * w[i] + u[i] * v[i];
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] *= u[i];
          }
          if (pv) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (n0 + c0 + i < P.cstore) st_stream(o + (size_t)(c0 + i) * HW, v[i]);
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            u[i] = un[i];
            w[i] = wn[i];
          }
        }
      } else {
        mbar_wait(&acc_full[acc], phase);
        tc_fence_after();
        const float *bs = bias_s + n0;
        int c0 = 0;
        for (; c0 + 32 <= Cn; c0 += 32) {
          float v[32];
          tmem_ld32(taddr + c0, v);
          float bv[32];  // (before the stores: their "memory" clobber would otherwise pin every bias load to its own store)
#pragma unroll
          for (int i = 0; i < 32; ++i) bv[i] = bs[c0 + i];
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] += bv[i];
          activate_all(v, P.act);
          if (pv) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (n0 + c0 + i < P.cstore) st_stream(o + (size_t)(c0 + i) * HW, v[i]);
          }
        }
        if (c0 < Cn) {  // Cn is a multiple of 16
          float v[16];
          tmem_ld16(taddr + c0, v);
          float bv[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) bv[i] = bs[c0 + i];
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] += bv[i];
          activate_all(v, P.act);
          if (pv) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (n0 + c0 + i < P.cstore) st_stream(o + (size_t)(c0 + i) * HW, v[i]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
      acc ^= 1;
      if (acc == 0) phase ^= 1;
    }
  } else if (warp == kProducerWarp) {
    // ===================== weight blocks [Cout][64] of (channel block, tap): L2 -> shared memory =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int n0 = (item % P.nparts) * Cn;
        for (int cb = 0; cb < P.ncb; ++cb)
          for (int t = 0; t < T; ++t) {
            mbar_wait(&w_empty[stage], phase ^ 1);
            mbar_expect_tx(&w_full[stage], (uint32_t)Cn * 128u);
            tma_load_3d(smem + kWOff + stage * P.stage_bytes, &tmW, &w_full[stage], 0, n0, cb * T + t);
            if (++stage == P.stages) { stage = 0; phase ^= 1; }
          }
      }
    }
  } else if (warp == kInputWarp) {
    // ===================== input boxes: 8 channels x halo tile, fp32, zero-filled outside the image =====================
    if (TMA_IN && lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t bytes = (uint32_t)(8 * P.BH * P.BW * 4);
      for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int tile = item / P.nparts;
        const int txi = tile % P.tiles_x, by = tile / P.tiles_x, tyi = by % P.tiles_y, b = by / P.tiles_y;
        const int xs = P.box_x0 ? 0 : txi * TW - pw - P.box_shift, ys = P.box_y0 ? 0 : tyi * TH - ph;
        for (int c = 0; c < P.Cin; c += 8) {
          mbar_wait(&s_empty[stage], phase ^ 1);
          mbar_expect_tx(&s_full[stage], bytes);
          if (c < P.C1) tma_load_4d(smem + kInOff + stage * kInStageB, &tmX1, &s_full[stage], xs, ys, c, b);
          else tma_load_4d(smem + kInOff + stage * kInStageB, &tmX2, &s_full[stage], xs, ys, c - P.C1, b);
          if (++stage == P.in_stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      // kind::f16: D = f32, A = B = f16, K-major, M = 128, N = Cout
      const uint32_t idesc = (1u << 4) | ((uint32_t)(Cn >> 3) << 17) | ((uint32_t)(TMPIX >> 4) << 24);
      int ws = 0, acc = 0;
      uint32_t wphase = 0, accphase = 0, n = 0;
      for (int item = blockIdx.x; item < items; item += gridDim.x) {
        mbar_wait(&acc_empty[acc], accphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
        uint32_t first = 1;
        for (int cb = 0; cb < P.ncb; ++cb, ++n) {
          const uint32_t buf = n & 1;
          mbar_wait(&h_full[buf], (n >> 1) & 1);
          tc_fence_after();
          const uint32_t hb = halo_u32 + buf * kHaloB;
          for (int t = 0; t < T; ++t) {
            const int ky = t / KW, kx = t - ky * KW;
            mbar_wait(&w_full[ws], wphase);
            tc_fence_after();
            // tap (ky, kx): the tile's 128 output pixels (row-major) see halo pixels ky * RW + kx .. + 127 -- consecutive
            // because the tile is a single row (RW = TW + KW - 1 only matters within it) or the filter a single column (RW = TW)
            const uint64_t adesc = make_sw128_desc(hb + (uint32_t)(ky * RW + kx) * 128u);
            const uint64_t bdesc = make_sw128_desc(w_u32 + ws * P.stage_bytes);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, first ? 0u : 1u);
              first = 0;
            }
            umma_commit(&w_empty[ws]);
            if (++ws == P.stages) { ws = 0; wphase ^= 1; }
          }
          umma_commit(&h_empty[buf]);
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

// weights (Cout, Cin, KH, KW) fp32 -> [channel block][tap][Cout][64] fp16, channels beyond Cin zero
__global__ void conv_prep_weight_kernel(const float *__restrict__ w, unsigned short *__restrict__ wp, int Cout, int Cin, int T, int ncb) {
  const int total = ncb * T * Cout * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c64 = i & 63, o = (i >> 6) % Cout, bt = (i >> 6) / Cout, t = bt % T, cb = bt / T;
    const int c = cb * 64 + c64;
    const float v = c < Cin ? __ldg(w + ((size_t)o * Cin + c) * T + t) : 0.0f;
    unsigned short r;
    asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(r) : "f"(v));
    wp[i] = r;
  }
}

}  // namespace cv

size_t conv_weight_bytes(int Cout, int Cin, int KH, int KW) { return (size_t)((Cin + 63) / 64) * KH * KW * Cout * 64 * 2; }

int launch_conv_prep_weight(const float *w, void *wp, int Cout, int Cin, int KH, int KW, cudaStream_t st) {
  const int ncb = (Cin + 63) / 64, T = KH * KW;
  const int total = ncb * T * Cout * 64;
  cv::conv_prep_weight_kernel<<<(total + 255) / 256, 256, 0, st>>>(w, static_cast<unsigned short *>(wp), Cout, Cin, T, ncb);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_conv_tc(const float *x1, int C1, const float *x2, const void *wp, const float *bias, float *out, int B, int Cin, int Cout,
                   int H, int W, int KH, int KW, int act, cudaStream_t st, const ConvExtra *ex) {
  using namespace cv;
  const int ncb = (Cin + 63) / 64, T = KH * KW;
  int dev = 0, sms = 148;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  Params P{};
  P.x1 = x1; P.x2 = x2; P.bias = bias; P.out = out;
  const long long hw = (long long)H * W;
  P.x1_bs = ex && ex->x1_bs ? ex->x1_bs : hw * C1;
  P.x2_bs = ex && ex->x2_bs ? ex->x2_bs : hw * (Cin - C1);
  P.out_bs = ex && ex->out_bs ? ex->out_bs : hw * Cout;
  P.cstore = ex && ex->store_channels > 0 ? ex->store_channels : Cout;
  if (ex) {
    P.epi = ex->epi; P.split = ex->split;
    P.aux1 = ex->aux1; P.aux2 = ex->aux2;
    P.aux1_bs = ex->aux1_bs; P.aux2_bs = ex->aux2_bs;
  }
  P.B = B; P.C1 = C1; P.Cin = Cin; P.Cout = Cout; P.H = H; P.W = W; P.KH = KH; P.KW = KW; P.act = act;
  P.ncb = ncb;
  if (KW == 1 && KH > 1) { P.TH = 8; P.TW = 16; }  // vertical filters: a square-ish tile keeps the halo at 1.5x instead of KH x
  else { P.TH = 1; P.TW = TMPIX; }
  P.tiles_x = (W + P.TW - 1) / P.TW;
  P.tiles_y = (H + P.TH - 1) / P.TH;
  P.total_tiles = B * P.tiles_y * P.tiles_x;
  // few tiles (one frame pair): split the filters over up to 4 CTAs per tile so that the whole GPU works on it
  P.nparts = 1;
  for (int p = 2; p <= 4; ++p)
    if (P.total_tiles * p <= sms && Cout % (16 * p) == 0) P.nparts = p;
  P.Cn = Cout / P.nparts;
  P.stage_bytes = (P.Cn * 128 + 1023) / 1024 * 1024;
  P.BH = P.TH + KH - 1;
  // a box must START on a 16 B boundary of the image row (a start at column x0 - 1 was measured to fault at run time;
  // whole rows below the image are fine): it starts up to 3 columns left of the halo tile and is that much wider
  P.box_shift = (((KW >> 1) + 3) & ~3) - (KW >> 1);
  P.BW = (P.box_shift + P.TW + KW - 1 + 3) & ~3;
  if (P.BW > W) { P.BW = W; P.box_x0 = 1; P.box_shift = 0; }  // (W % 4 == 0 on the TMA path)
  if (P.BH > H) { P.BH = H; P.box_y0 = 1; }
  {
    const int halo_pix = (P.TH + KH - 1) * (P.TW + KW - 1);
    P.halo_bytes = (halo_pix * 128 + 1023) / 1024 * 1024;
    P.in_stage_bytes = (8 * P.BH * P.BW * 4 + 127) / 128 * 128;
    P.in_off = 2 * P.halo_bytes;
    // the weights get two stages or 96 KB (measured: 48 -> 96 KB is 57 -> 51 us on convc2 at batch 8 and nothing elsewhere),
    // the input staging what is left (4 .. kInStages boxes in flight), the weights then the remainder
    const int w_min = 2 * P.stage_bytes > 96 * 1024 ? 2 * P.stage_bytes : 96 * 1024;
    int ins = (kDataBytes - P.in_off - w_min - 1024) / P.in_stage_bytes;
    P.in_stages = ins < 4 ? 4 : (ins > kInStages ? kInStages : ins);
    P.w_off = (P.in_off + P.in_stages * P.in_stage_bytes + 1023) / 1024 * 1024;
    const int ring = kDataBytes - P.w_off;
    P.stages = ring / P.stage_bytes < kMaxStages ? ring / P.stage_bytes : kMaxStages;
  }
  CUtensorMap tm;
  {
    const unsigned long long dims[3] = {64, (unsigned long long)Cout, (unsigned long long)ncb * T};
    const unsigned long long strides[2] = {128, 128ull * Cout};
    const unsigned int box[3] = {64, (unsigned int)P.Cn, 1};
    if (int rc = encode_tmap_sw128(&tm, wp, 3, dims, strides, box, "rc_conv2d_fwd")) return rc;
  }
  const int items = P.total_tiles * P.nparts;
  const int grid = items < sms ? items : sms;
  const int C2 = Cin - C1;
  const bool tma_in = W % 4 == 0 && (reinterpret_cast<uintptr_t>(x1) & 15) == 0 && P.BH * P.BW <= kMaxBoxPix && P.x1_bs % 4 == 0 &&
                      C1 >= 8 && (C2 == 0 || (C2 >= 8 && C1 % 8 == 0 && (reinterpret_cast<uintptr_t>(x2) & 15) == 0 && P.x2_bs % 4 == 0));
  if (tma_in) {
    CUtensorMap tx1, tx2;
    const unsigned int box[4] = {(unsigned int)P.BW, (unsigned int)P.BH, 8, 1};
    {
      const unsigned long long dims[4] = {(unsigned long long)W, (unsigned long long)H, (unsigned long long)C1, (unsigned long long)B};
      const unsigned long long strides[3] = {4ull * W, 4ull * W * H, 4ull * (unsigned long long)P.x1_bs};
      if (int rc = encode_tmap_f32(&tx1, x1, 4, dims, strides, box, "rc_conv2d_fwd")) return rc;
    }
    if (C2 > 0) {
      const unsigned long long dims[4] = {(unsigned long long)W, (unsigned long long)H, (unsigned long long)C2, (unsigned long long)B};
      const unsigned long long strides[3] = {4ull * W, 4ull * W * H, 4ull * (unsigned long long)P.x2_bs};
      if (int rc = encode_tmap_f32(&tx2, x2, 4, dims, strides, box, "rc_conv2d_fwd")) return rc;
    } else {
      tx2 = tx1;
    }
    auto kern = P.epi != RC_EPI_NONE ? conv_tc_kernel<true, true> : conv_tc_kernel<true, false>;
    RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    RC_CUDA(launch_pdl(kern, dim3((unsigned)grid), dim3(kThreads), (size_t)kSmemBytes, st, tm, tx1, tx2, P));
  } else {
    auto kern = P.epi != RC_EPI_NONE ? conv_tc_kernel<false, true> : conv_tc_kernel<false, false>;
    RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    RC_CUDA(launch_pdl(kern, dim3((unsigned)grid), dim3(kThreads), (size_t)kSmemBytes, st, tm, tm, tm, P));
  }
  return RC_OK;
}

}  // namespace rc
