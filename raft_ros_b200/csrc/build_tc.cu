// Tensor-core build: all-pairs volume on tcgen05 (UMMA) with the 1/sqrt(D) scale and the whole
// average-pool pyramid fused into the epilogue.  sm_100a only.
//
// Replaces CorrBlock.corr + the pyramid loop of CorrBlock.__init__ (reference
// core/corr.py:52-60 and :21-27): cuBLAS SGEMM + divide + 3x avg_pool2d, which write the
// 198 MB (440x1024) volume, re-read and re-write it for the scale and re-read each level for
// pooling.  Here every pyramid element is written to HBM exactly once and never re-read.
//
// Data flow
//   prep kernel : fmap (B,D,h,w) fp32 NCHW -> (B,h,w,Dp) bf16 K-major "hi" (+ "lo" residual
//                 for RC_PREC_BF16X3: x ~= hi + lo, 16 mantissa bits), Dp = D rounded up to 64.
//   main kernel : persistent, one CTA per SM, 10 warps:
//                   warp 8  TMA producer  (A: 128 query pixels x 64 ch; B: an 8x32 SPATIAL patch
//                           of fmap2 x 64 ch via a 4-D box, so the 256 columns of a tile are a
//                           spatial patch and every 2x2/4x4/8x8 pooling window is tile-local)
//                   warp 9  MMA issuer    (tcgen05.mma 128x256x16, bf16 -> fp32 in TMEM; BF16X3
//                           issues hi*hi + hi*lo + lo*hi into the same accumulator)
//                   warps 0-7 epilogue    (warp = TMEM lane quarter x upper/lower half of the patch;
//                           tcgen05.ld one spatial row pair at a time; thread <-> query pixel so
//                           all pooling is in registers.  Row layout: rows are transposed through a
//                           swizzled per-warp smem buffer and written with st.global.v4, 512 B per
//                           instruction.  Tiled layout (what CorrBlock uses): two whole tiles per
//                           pixel are staged in global order and leave as one 4 KB bulk copy; the
//                           V16 flavour packs them to fp16 first)
//                 TMEM holds two 128x256 fp32 accumulators (all 512 columns): the MMAs of tile
//                 i+1 overlap the epilogue of tile i.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {

namespace tc {

constexpr int TM = 128;        // query pixels per tile (TMEM lanes)
constexpr int PH = 8;          // patch rows
constexpr int PW = 32;         // patch cols
constexpr int TN = PH * PW;    // 256 target pixels per tile (TMEM columns)
constexpr int TK = 64;         // channels per stage: 128 B rows, SWIZZLE_128B
constexpr int UK = 16;         // K of one tcgen05.mma (bf16)
constexpr int A_BYTES = TM * TK * 2;
constexpr int B_BYTES = TN * TK * 2;
constexpr int kEpiWarps = 8;           // warps 0..7
constexpr int kProducerWarp = 8;
constexpr int kMmaWarp = 9;
constexpr int kThreads = 320;
constexpr int kStgBytesPerWarp = 4096;  // one 32x32 fp32 transposition buffer per epilogue warp
constexpr int kTmemCols = 512;

struct Params {
  int B, N, h, w, L;
  int k_blocks;                 // Dp / 64
  int m_tiles, nt_x, nt_y;      // tiles along query pixels / patch grid
  long long total_tiles;
  float scale;
  const float *inv_sa, *inv_sb; // NPASS == 2: 1 / (power-of-two normalisation) of every pixel of fmap1 / fmap2, [B][N] each
  int vec_mask;                 // bit l: rows of level l are 16 B aligned -> vector stores
  int tpr[4], ng;               // TILED output only: tiles per tile-row of level l; groups of query pixels per item
  long long gstride[4];         //   floats per group of level l (= kTileGroup * hp * wp)
  LevelPtrs lv;
};

// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major, M=128, N=256
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
// the CTA pair's instruction: M=256 (128 rows in each CTA), N=256 (128 B rows held by each CTA)
constexpr uint32_t kIdesc2 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)((2 * TM) >> 4) << 24);

// the same shape with fp16 operands (kind::f16) or e4m3 operands (kind::f8f6f4): both formats are code 0
constexpr uint32_t kIdescF = (1u << 4) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
constexpr uint32_t kIdescF2 = (1u << 4) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)((2 * TM) >> 4) << 24);

// RC_PREC_FP16_FP8 (NPASS == 2).  Every PIXEL of a feature map (its D-vector) is scaled by a power of two s so that its largest magnitude lands
// in [128, 256) and split x*s = hi + lo, hi = fp16(x*s) (lo is then exact in fp32, |lo| <= 2^-4).  Of
//   (a*sa)(b*sb) = ah*bh + (ah*bl + al*bh) + al*bl
// the first term runs in fp16; the bracket is ONE contraction over 2x the channels of 8-bit operands
//   A8 = [ e4m3(16 al) | e4m3(ah / 16) ],  B8 = [ e4m3(bh / 16) | e4m3(16 bl) ]      (128 B per pixel and 64 channels)
// -- the 16s cancel in the products and put both factors in e4m3's normal range (16|lo| <= 1, |hi|/16 < 16; e4m3
// keeps 2^-6 .. 448) -- at the 8-bit rate, i.e. for the tensor time of one more fp16 term: 2 units instead of
// BF16X3's 3.  Rounding the cross terms' factors to 4 significant bits is a ~4 % error on terms that are 2^-12 of
// the result; al*bl (2^-22) is dropped as in BF16X3.  The epilogue multiplies element (p, q) by scale / (sa[p] * sb[q]):
// per-pixel factors cost it one shuffle and one multiply per element, and cost the prepass nothing (a pixel's 256
// channels are in one block) -- a per-map factor needed a separate pass over the feature maps (20 us at batch 8).
__device__ __forceinline__ float norm_mul(unsigned amax_bits) {
  // 2^(7 - e) for max|x| in [2^e, 2^(e+1)); 1 for an all-zero map; clamped to the finite powers of two
  if (amax_bits == 0u) return 1.0f;
  int f = 261 - (int)(amax_bits >> 23);
  f = f < 1 ? 1 : (f > 254 ? 254 : f);
  return __uint_as_float((unsigned)f << 23);
}

template <int NPASS, int NCTA>
struct Cfg {
  static constexpr int kOps = (NPASS >= 2) ? 2 : 1;          // hi (+ lo, or the 8-bit pair)
  static constexpr int kBBytes = B_BYTES / NCTA;             // a CTA of a pair stages half of the patch
  static constexpr int kStageBytes = kOps * (A_BYTES + kBBytes);
  static constexpr int kStages = 192 * 1024 / kStageBytes;   // 4 / 2 (single CTA), 6 / 3 (pair)
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * kStageBytes + kEpiWarps * kStgBytesPerWarp +
                                    256 /*barriers*/;
};

// TILED: every row p of a level is cut into [hp/2][wp/8] tiles of 2 rows x 8 floats (hp = ceil2,
// wp = ceil8 of the level's size, zeros in the padding that a lookup can touch) and the tiles of
// G = 32 consecutive query pixels are interleaved: [group][tile][pixel u][4 chunks of 16 B], chunk
// k = 2*(row & 1) + (column half) of pixel u at position k ^ ((u >> 1) & 3) (include/raftcorr.h) --
// the layout lookup_tiled.cu reads.  It costs the epilogue nothing: a warp (32 query pixels = one
// group) holds, per row pair of the 8x32 patch, four whole tiles per pixel; two tiles of the group
// are 4 KB of contiguous global memory, staged in that order (conflict-free thanks to the chunk
// swizzle) and moved by one bulk copy; every tile is 64 B aligned whatever h and w are.
// V16 (TILED only): the volume is stored as fp16 (saturating), tiles of 2 rows x 16 halves = the
// same 64 B / 4 chunks per pixel, so the layout code is shared: the warp's row pair is two tiles.
template <int NPASS, int NCTA, bool TILED, bool V16 = false>
__global__ void __launch_bounds__(kThreads, 1)
build_tc_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
                const Params P) {
  using C = Cfg<NPASS, NCTA>;
  static_assert(!V16 || TILED, "the 16-bit volume exists in the tiled layout only");
  // NCTA == 2: the kernel runs as clusters of two CTAs on one TPC.  The pair computes a 256 x 256
  // tile with tcgen05.mma.cta_group::2: each CTA stages its own 128 query pixels (A) and HALF of
  // the patch (B, 4 of the 8 rows), which halves the B-operand shared-memory traffic per SM;
  // each CTA's TMEM still receives its 128 rows x all 256 columns, so the epilogue is unchanged.
  const uint32_t rank = (NCTA == 2) ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const unsigned unit0 = blockIdx.x / NCTA, unit_step = gridDim.x / NCTA;
  extern __shared__ uint8_t smem_raw[];
  // 1024 B alignment (SWIZZLE_128B atoms) by offsetting inside the shared window
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  uint8_t *stage_base = smem;
  const uint32_t stg_u32 = base_u32 + C::kStages * C::kStageBytes;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::kStages * C::kStageBytes + kEpiWarps * kStgBytesPerWarp);
  uint64_t *full = bars;                       // [kStages]
  uint64_t *empty = bars + C::kStages;         // [kStages]
  uint64_t *tfull = bars + 2 * C::kStages;     // [2]
  uint64_t *tempty = tfull + 2;                // [2]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();  // persistent and wholly resident: the next kernel's CTAs follow ours as they leave

  if (warp == kProducerWarp && lane == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full[s], NCTA);  // pair: the leader's barrier also takes the peer's arrival and bytes
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], kEpiWarps * NCTA);
    }
    fence_barrier_init();
    prefetch_tmap(&tmA_hi);
    prefetch_tmap(&tmB_hi);
    if (NPASS >= 2) {
      prefetch_tmap(&tmA_lo);
      prefetch_tmap(&tmB_lo);
    }
  }
  if (warp == kMmaWarp) {
    if (NCTA == 2) tmem_alloc2(tmem_slot, kTmemCols);
    else tmem_alloc(tmem_slot, kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();  // both CTAs' barriers are initialised before any remote use
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const long long tiles_per_item = (long long)P.m_tiles * P.nt_x * P.nt_y;
  const int nt = P.nt_x * P.nt_y;

  if (warp == kProducerWarp) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      // Everything above (barriers, TMEM, tensor-map prefetch) ran while the layout prepass was still
      // draining; its output is read from here on.  The MMA and epilogue roles only ever act on what
      // this thread has loaded, so they are ordered behind this wait as well.
      pdl_wait();
      for (long long t = unit0; t < P.total_tiles; t += unit_step) {
        const int b = (int)(t / tiles_per_item);
        const int rem = (int)(t - (long long)b * tiles_per_item);
        const int mu = rem / nt, n = rem - mu * nt;
        const int m = mu * NCTA + (int)rank;
        const int ny = n / P.nt_x, nx = n - ny * P.nt_x;
        const int by = ny * PH + (int)rank * (PH / NCTA);  // this CTA's rows of the patch
        for (int kb = 0; kb < P.k_blocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t *sp = stage_base + stage * C::kStageBytes;
          if (NCTA == 1) {
            mbar_expect_tx(&full[stage], C::kStageBytes);
            tma_load_3d(sp, &tmA_hi, &full[stage], kb * TK, m * TM, b);
            tma_load_4d(sp + C::kOps * A_BYTES, &tmB_hi, &full[stage], kb * TK, nx * PW, by, b);
            if (NPASS >= 2) {
              tma_load_3d(sp + A_BYTES, &tmA_lo, &full[stage], kb * TK, m * TM, b);
              tma_load_4d(sp + C::kOps * A_BYTES + C::kBBytes, &tmB_lo, &full[stage], kb * TK, nx * PW, by, b);
            }
          } else {
            // both CTAs' bytes are counted on the LEADER's barrier
            if (leader) mbar_expect_tx(&full[stage], 2 * C::kStageBytes);
            tma2_load_3d(sp, &tmA_hi, &full[stage], kb * TK, m * TM, b);
            tma2_load_4d(sp + C::kOps * A_BYTES, &tmB_hi, &full[stage], kb * TK, nx * PW, by, b);
            if (NPASS >= 2) {
              tma2_load_3d(sp + A_BYTES, &tmA_lo, &full[stage], kb * TK, m * TM, b);
              tma2_load_4d(sp + C::kOps * A_BYTES + C::kBBytes, &tmB_lo, &full[stage], kb * TK, nx * PW, by, b);
            }
            if (!leader) mbar_arrive_leader(&full[stage]);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer (leader CTA only in pair mode) =====================
    if (lane == 0 && leader) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (long long t = unit0; t < P.total_tiles; t += unit_step) {
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TN);
        for (int kb = 0; kb < P.k_blocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = base_u32 + stage * C::kStageBytes;
          const uint64_t a_hi = make_sw128_desc(sa);
          const uint64_t b_hi = make_sw128_desc(sa + C::kOps * A_BYTES);
          const uint64_t a_lo = make_sw128_desc(sa + A_BYTES);
          const uint64_t b_lo = make_sw128_desc(sa + C::kOps * A_BYTES + C::kBBytes);
          if constexpr (NPASS == 2) {
            // fp16 products, then the two cross terms as one e4m3 contraction (the "lo" maps hold the 8-bit pair)
#pragma unroll
            for (int k = 0; k < TK / UK; ++k) {
              const uint64_t koff = (uint64_t)((k * UK * 2) >> 4);
              const uint32_t accum = (kb | k) != 0 ? 1u : 0u;
              if (NCTA == 1) umma_bf16(d_tmem, a_hi + koff, b_hi + koff, kIdescF, accum);
              else umma2_bf16(d_tmem, a_hi + koff, b_hi + koff, kIdescF2, accum);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {  // 128 B rows = 4 x K 32 of 8-bit operands
              const uint64_t koff = (uint64_t)(k * 2);
              if (NCTA == 1) umma_f8(d_tmem, a_lo + koff, b_lo + koff, kIdescF, 1u);
              else umma2_f8(d_tmem, a_lo + koff, b_lo + koff, kIdescF2, 1u);
            }
          } else
#pragma unroll
          for (int k = 0; k < TK / UK; ++k) {
            const uint64_t koff = (uint64_t)((k * UK * 2) >> 4);  // 32 B per K step inside the 128 B row
            const uint32_t accum = (kb | k) != 0 ? 1u : 0u;
            if (NCTA == 1) {
              umma_bf16(d_tmem, a_hi + koff, b_hi + koff, kIdesc, accum);
              if (NPASS == 3) {
                umma_bf16(d_tmem, a_hi + koff, b_lo + koff, kIdesc, 1u);
                umma_bf16(d_tmem, a_lo + koff, b_hi + koff, kIdesc, 1u);
              }
            } else {
              umma2_bf16(d_tmem, a_hi + koff, b_hi + koff, kIdesc2, accum);
              if (NPASS == 3) {
                umma2_bf16(d_tmem, a_hi + koff, b_lo + koff, kIdesc2, 1u);
                umma2_bf16(d_tmem, a_lo + koff, b_hi + koff, kIdesc2, 1u);
              }
            }
          }
          // smem slot reusable (in both CTAs) once these MMAs have read it
          if (NCTA == 1) umma_commit(&empty[stage]);
          else umma2_commit_both(&empty[stage]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        // accumulator complete -> epilogue (of both CTAs)
        if (NCTA == 1) umma_commit(&tfull[acc]);
        else umma2_commit_both(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue: 8 warps =====================
    // warp e: TMEM lane quarter q = e & 3 (32 query pixels, thread <-> pixel) and row half
    // hf = e >> 2 (spatial rows 0-3 / 4-7 of the 8x32 patch).  Pooling up to level 2 is in
    // registers; the two halves meet once per tile for the single level-3 row.
    const int q = warp & 3, hf = warp >> 2;
    const uint32_t buf = stg_u32 + warp * kStgBytesPerWarp;
    const uint32_t buf_peer = stg_u32 + (warp | 4) * kStgBytesPerWarp;  // hf=1 warp of this quarter
    const int bar_full = 1 + q, bar_free = 5 + q;                       // named barriers of the pair
    int acc = 0;
    uint32_t acc_phase = 0;
    const int h = P.h, w = P.w, N = P.N, L = P.L;
    const int h1 = h >> 1, w1 = w >> 1, h2 = h >> 2, w2 = w >> 2, h3 = h >> 3, w3 = w >> 3;
    // row layout: floats per volume row; tiled layout: floats per group of G = 32 query pixels
    const size_t pitch0 = TILED ? (size_t)P.gstride[0] : (size_t)h * w, pitch1 = TILED ? (size_t)P.gstride[1] : (size_t)h1 * w1,
                 pitch2 = TILED ? (size_t)P.gstride[2] : (size_t)h2 * w2, pitch3 = TILED ? (size_t)P.gstride[3] : (size_t)h3 * w3;
    const bool vec0 = P.vec_mask & 1, vec1 = P.vec_mask & 2, vec2 = P.vec_mask & 4, vec3 = P.vec_mask & 8;
    bool peer_pending = false;  // hf=1: our staging buffer still holds the level-2 row for the peer
    if (NPASS == 2) pdl_wait();  // the normalisation exponents below come from the kernels before this one
    for (long long t = unit0; t < P.total_tiles; t += unit_step) {
      const int b = (int)(t / tiles_per_item);
      float out_scale = P.scale;
      const int rem = (int)(t - (long long)b * tiles_per_item);
      const int mu = rem / nt, n = rem - mu * nt;
      const int m = mu * NCTA + (int)rank;
      const int ny = n / P.nt_x, nx = n - ny * P.nt_x;
      const int prow0 = m * TM + q * 32;                 // first query pixel of this warp
      const int rows_valid = min(32, N - prow0);         // may be <= 0
      const size_t grow0 = (size_t)b * N + prow0;        // first row of the volume for this warp
      const int x0 = nx * PW;
      // row layout: first element of this warp's patch columns in row grow0 of each level;
      // tiled layout: the patch's first tile (tile-row 0) of the warp's group of 32 query pixels
      constexpr int G = kTileGroup, TS = 16 * G;  // query pixels per group, floats per tile of a group
      const size_t grp0 = (size_t)b * P.ng + prow0 / G;  // prow0 is a multiple of 32 = G
      // (V16: the same expressions in 32-bit words; a tile is 16 columns wide, so half as many tiles per patch)
      float *const g0 = V16     ? P.lv.p[0] + grp0 * pitch0 + nx * (2 * TS)
                        : TILED ? P.lv.p[0] + grp0 * pitch0 + nx * (4 * TS)
                                : P.lv.p[0] + grow0 * pitch0 + x0;
      float *const g1 = V16     ? P.lv.p[1] + grp0 * pitch1 + nx * TS
                        : TILED ? P.lv.p[1] + grp0 * pitch1 + nx * (2 * TS)
                                : P.lv.p[1] + grow0 * pitch1 + (x0 >> 1);
      float *const g2 = V16     ? P.lv.p[2] + grp0 * pitch2 + (nx >> 1) * TS + lane * 16
                        : TILED ? P.lv.p[2] + grp0 * pitch2 + nx * TS
                                : P.lv.p[2] + grow0 * pitch2 + (x0 >> 2);
      float *const g3 = V16     ? P.lv.p[3] + grp0 * pitch3 + (nx >> 2) * TS + lane * 16
                        : TILED ? P.lv.p[3] + grp0 * pitch3 + (nx >> 1) * TS + lane * 16
                                : P.lv.p[3] + (grow0 + lane) * pitch3 + (x0 >> 3);  // direct: thread <-> row
      // TILED: the staging buffer may still be being read by this warp's last bulk store
      auto buf_ready = [&]() {
        if (lane == 0) bulk_wait_read0();
        __syncwarp();
      };
      // two whole tiles per query pixel -> global: one 4 KB bulk copy when the group and both tiles
      // are complete (the warp moves on while the copy engine drains it), per-lane stores at edges
      auto put_tiles = [&](const float *ch, float *gdst, int tiles_valid) {
        buf_ready();
        if (rows_valid >= 32 && tiles_valid >= 2) bulk_tiles32(buf, ch, lane, gdst);
        else store_tiles32<G>(buf, ch, lane, gdst, rows_valid, tiles_valid);
      };
      // V16: pack two neighbouring columns into one word of fp16 (saturating: no infinities)
      auto pack = [](float lo_col, float hi_col) {
        uint32_t r;
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_col), "f"(lo_col));
        return __uint_as_float(r);
      };
      // NPASS == 2: undo the power-of-two normalisation of this thread's query pixel (row) and of the target pixels
      // (columns; lane i holds column x0 + i of this warp's four patch rows, handed round by shuffles below).  Exact.
      // Fetched before the accumulator is complete, so the round trip overlaps the MMAs.
      float sb[4] = {1.0f, 1.0f, 1.0f, 1.0f};
      if constexpr (NPASS == 2) {
        const float ra = lane < rows_valid ? __ldg(P.inv_sa + (size_t)b * N + prow0 + lane) : 0.0f;
        out_scale = P.scale * ra;
#pragma unroll
        for (int r4 = 0; r4 < 4; ++r4) {
          const int yy = ny * PH + hf * 4 + r4;
          sb[r4] = (yy < h && x0 + lane < w) ? __ldg(P.inv_sb + (size_t)b * N + (size_t)yy * w + x0 + lane) : 0.0f;
        }
      }
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * TN + hf * (TN / 2));
      if (L > 3 && hf == 1 && peer_pending) {
        // the peer must have read last tile's level-2 row before this buffer is reused
        asm volatile("bar.sync %0, 64;" ::"r"(bar_free) : "memory");
        peer_pending = false;
      }
      float l1prev[16], l2[8];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float v0[32], v1[32];
        tmem_ld32(taddr + (2 * j) * PW, v0);
        tmem_ld32(taddr + (2 * j + 1) * PW, v1);
        tmem_ld_wait();
        if (j == 1) {
          // all TMEM reads of this tile are done: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (NCTA == 1) mbar_arrive(&tempty[acc]);
            else mbar_arrive_leader(&tempty[acc]);  // the pair's MMA thread lives in the leader CTA
          }
        }
        if constexpr (NPASS == 2) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            v0[i] *= out_scale * __shfl_sync(0xffffffffu, sb[2 * j], i);
            v1[i] *= out_scale * __shfl_sync(0xffffffffu, sb[2 * j + 1], i);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            v0[i] *= out_scale;
            v1[i] *= out_scale;
          }
        }
        const int y0 = ny * PH + hf * 4 + 2 * j;
        if constexpr (V16) {
          // rows y0, y0+1 x 32 columns = two tiles of 2 x 16 halves
          if (y0 < h) {
            float ch[32];
#pragma unroll
            for (int t2 = 0; t2 < 2; ++t2)
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                ch[16 * t2 + i] = pack(v0[16 * t2 + 2 * i], v0[16 * t2 + 2 * i + 1]);
                ch[16 * t2 + 8 + i] = pack(v1[16 * t2 + 2 * i], v1[16 * t2 + 2 * i + 1]);
              }
            put_tiles(ch, g0 + (size_t)(y0 >> 1) * (P.tpr[0] * TS), P.tpr[0] - nx * 2);
          }
        } else if constexpr (TILED) {
          // rows y0, y0+1 = one tile-row; columns beyond w (and row h of an odd h) are zero accumulators
          if (y0 < h) {
            float *const gt = g0 + (size_t)(y0 >> 1) * (P.tpr[0] * TS);
            const int tiles = P.tpr[0] - nx * 4;  // tiles of this tile-row from the patch's first one on
#pragma unroll
            for (int c2 = 0; c2 < 2; ++c2) {
              float ch[32];
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                ch[i] = v0[16 * c2 + i];
                ch[8 + i] = v1[16 * c2 + i];
                ch[16 + i] = v0[16 * c2 + 8 + i];
                ch[24 + i] = v1[16 * c2 + 8 + i];
              }
              put_tiles(ch, gt + c2 * (2 * TS), tiles - 2 * c2);
            }
          }
        } else {
          if (y0 < h) store_rows<32>(buf, v0, lane, g0 + (size_t)y0 * w, pitch0, rows_valid, w - x0, vec0);
          if (y0 + 1 < h) store_rows<32>(buf, v1, lane, g0 + (size_t)(y0 + 1) * w, pitch0, rows_valid, w - x0, vec0);
        }
        if (L > 1) {
          float l1[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float s = v0[2 * i] + v0[2 * i + 1];
            s += v1[2 * i];
            s += v1[2 * i + 1];
            l1[i] = s * 0.25f;
          }
          const int y1 = ny * (PH / 2) + hf * 2 + j;
          if constexpr (TILED) {
            // pooled cells that hang over the level's edge (odd sizes) are padding: they must read as zero
            if ((nx + 1) * 16 > w1 || y1 >= h1) {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (nx * 16 + i >= w1 || y1 >= h1) l1[i] = 0.0f;
            }
            if (V16 && j == 1 && y1 - 1 < h1) {
              // two rows x 16 columns = ONE tile, 64 B per pixel: a 2 KB bulk copy per warp (per-lane
              // 16 B stores at the swizzled positions for an incomplete group)
              if (nx < P.tpr[1]) {
                float *gt = g1 + (size_t)(y1 >> 1) * (P.tpr[1] * TS);
                const int sw = tile_sw(lane);
                float4 c4[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float *src = (k >> 1) ? l1 : l1prev;
                  const int c0 = (k & 1) * 8;
                  c4[k] = make_float4(pack(src[c0], src[c0 + 1]), pack(src[c0 + 2], src[c0 + 3]),
                                      pack(src[c0 + 4], src[c0 + 5]), pack(src[c0 + 6], src[c0 + 7]));
                }
                if (rows_valid >= 32) {
                  buf_ready();
#pragma unroll
                  for (int k = 0; k < 4; ++k) sts128(buf + lane * 64 + ((k ^ sw) << 4), c4[k].x, c4[k].y, c4[k].z, c4[k].w);
                  fence_proxy_async_smem();
                  __syncwarp();
                  if (lane == 0) {
                    bulk_store(gt, buf, 2048);
                    bulk_commit();
                  }
                } else if (lane < rows_valid) {
#pragma unroll
                  for (int k = 0; k < 4; ++k) stg128(gt + lane * 16 + ((k ^ sw) << 2), c4[k]);
                }
              }
            } else if (!V16 && j == 1 && y1 - 1 < h1) {
              float ch[32];
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                ch[i] = l1prev[i];
                ch[8 + i] = l1[i];
                ch[16 + i] = l1prev[8 + i];
                ch[24 + i] = l1[8 + i];
              }
              put_tiles(ch, g1 + (size_t)(y1 >> 1) * (P.tpr[1] * TS), P.tpr[1] - nx * 2);
            }
          } else {
            if (y1 < h1) store_rows<16>(buf, l1, lane, g1 + (size_t)y1 * w1, pitch1, rows_valid, w1 - (x0 >> 1), vec1);
          }
          if (L > 2 || TILED) {
            if (j == 0) {
#pragma unroll
              for (int i = 0; i < 16; ++i) l1prev[i] = l1[i];
            } else if (L > 2) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                float s = l1prev[2 * i] + l1prev[2 * i + 1];
                s += l1[2 * i];
                s += l1[2 * i + 1];
                l2[i] = s * 0.25f;
              }
              const int y2 = ny * (PH / 4) + hf;
              if constexpr (TILED) {
                if ((nx + 1) * 8 > w2 || y2 >= h2) {
#pragma unroll
                  for (int i = 0; i < 8; ++i)
                    if (nx * 8 + i >= w2 || y2 >= h2) l2[i] = 0.0f;
                }
                if constexpr (V16) {
                  // one patch = half a level-2 tile (8 of its 16 columns); this warp holds row hf of it:
                  // chunk 2 hf + (nx & 1), 16 B per pixel
                  if (2 * ny < h2 && (nx >> 1) < P.tpr[2] && lane < rows_valid)
                    stg128(g2 + (size_t)ny * (P.tpr[2] * TS) + ((((hf << 1) | (nx & 1)) ^ tile_sw(lane)) << 2),
                           make_float4(pack(l2[0], l2[1]), pack(l2[2], l2[3]), pack(l2[4], l2[5]), pack(l2[6], l2[7])));
                } else if (2 * ny < h2) {
                  // one patch = one level-2 tile; this warp holds its row hf (8 floats = one 32 B sector)
                  buf_ready();
                  store_halftile8<G>(buf, l2, lane, g2 + (size_t)ny * (P.tpr[2] * TS), hf, rows_valid, nx < P.tpr[2]);
                }
              } else {
                if (y2 < h2) store_rows<8>(buf, l2, lane, g2 + (size_t)y2 * w2, pitch2, rows_valid, w2 - (x0 >> 2), vec2);
              }
            }
          }
        }
      }
      if (L > 3) {
        if (hf == 1) {
          // publish this half's level-2 row (8 floats per query pixel) through our staging buffer
          if constexpr (TILED) buf_ready();
          sts128(buf + lane * 32, l2[0], l2[1], l2[2], l2[3]);
          sts128(buf + lane * 32 + 16, l2[4], l2[5], l2[6], l2[7]);
          __threadfence_block();
          asm volatile("bar.arrive %0, 64;" ::"r"(bar_full) : "memory");
          peer_pending = true;
        } else {
          asm volatile("bar.sync %0, 64;" ::"r"(bar_full) : "memory");
          const float4 pa = lds128(buf_peer + lane * 32), pb = lds128(buf_peer + lane * 32 + 16);
          asm volatile("bar.arrive %0, 64;" ::"r"(bar_free) : "memory");
          const float lo[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
          float l3[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float s = l2[2 * i] + l2[2 * i + 1];
            s += lo[2 * i];
            s += lo[2 * i + 1];
            l3[i] = s * 0.25f;
          }
          const int cv3 = w3 - (x0 >> 3);
          if constexpr (TILED) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (i >= cv3 || ny >= h3) l3[i] = 0.0f;
            if constexpr (V16) {
              // an eighth of a level-3 tile: row ny & 1, columns 4 (nx & 3) .. +3 = half of chunk
              // 2 (ny & 1) + ((nx >> 1) & 1).  The last patch column also zeroes the half chunk
              // that no patch covers (a lookup may fetch it).
              if (ny < ((h3 + 1) & ~1) && (nx >> 2) < P.tpr[3] && lane < rows_valid) {
                float *d = g3 + (size_t)(ny >> 1) * (P.tpr[3] * TS) +
                           (((((ny & 1) << 1) | ((nx >> 1) & 1)) ^ tile_sw(lane)) << 2) + (nx & 1) * 2;
                *reinterpret_cast<float2 *>(d) = make_float2(pack(l3[0], l3[1]), pack(l3[2], l3[3]));
                if ((nx & 1) == 0 && nx + 1 >= P.nt_x) *reinterpret_cast<float2 *>(d + 2) = make_float2(0.f, 0.f);
              }
            } else
            // a quarter of a level-3 tile: row ny & 1, columns 4 (nx & 1) .. +3 (16 B, always aligned)
            if (ny < ((h3 + 1) & ~1) && (nx >> 1) < P.tpr[3] && lane < rows_valid)
              stg128(g3 + (size_t)(ny >> 1) * (P.tpr[3] * TS) + (((((ny & 1) << 1) | (nx & 1)) ^ tile_sw(lane)) << 2),
                     make_float4(l3[0], l3[1], l3[2], l3[3]));
          } else if (ny < h3 && lane < rows_valid) {
            float *d = g3 + (size_t)ny * w3;
            if (vec3 && cv3 >= 4) {
              stg128(d, make_float4(l3[0], l3[1], l3[2], l3[3]));
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                if (i < cv3) st_stream(d + i, l3[i]);
            }
          }
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (L > 3 && hf == 1 && peer_pending) asm volatile("bar.sync %0, 64;" ::"r"(bar_free) : "memory");
    if (TILED && lane == 0) bulk_wait_all();  // our shared memory must outlive the last bulk store
  }

  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();  // the peer may still be reading our smem / signalling our barriers
  tc_fence_after();
  if (warp == kMmaWarp) {
    if (NCTA == 2) tmem_dealloc2(tmem_base, kTmemCols);
    else tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------ prep: NCHW fp32 -> NHWC bf16 hi/lo
__global__ void __launch_bounds__(256)
prep_kernel(const float *__restrict__ f1, const float *__restrict__ f2, __nv_bfloat16 *__restrict__ a_hi,
            __nv_bfloat16 *__restrict__ a_lo, __nv_bfloat16 *__restrict__ b_hi, __nv_bfloat16 *__restrict__ b_lo,
            int D, int Dp, int N, int with_lo) {
  // one block: 64 channels x 32 pixels.  Loads are coalesced along the pixels (128 B per warp),
  // stores along the channels as bf16 pairs (128 B per warp = one whole row of the K-major tile).
  __shared__ float tile[64][33];
  const int which = blockIdx.z & 1, b = blockIdx.z >> 1;
  const float *src = (which ? f2 : f1) + (size_t)b * D * N;
  __nv_bfloat16 *hi = (which ? b_hi : a_hi) + (size_t)b * N * Dp;
  __nv_bfloat16 *lo = (which ? b_lo : a_lo) + (size_t)b * N * Dp;
  const int p0 = blockIdx.x * 32, d0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  pdl_launch_dependents();
  pdl_wait();
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int d = d0 + ty + 8 * i, p = p0 + tx;
    tile[ty + 8 * i][tx] = (d < D && p < N) ? __ldg(src + (size_t)d * N + p) : 0.0f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int pl = ty + 8 * i, p = p0 + pl, d = d0 + 2 * tx;
    if (p < N) {  // Dp is a multiple of 64: d + 1 < Dp
      const float x0 = tile[2 * tx][pl], x1 = tile[2 * tx + 1][pl];
      const __nv_bfloat162 h2 = __floats2bfloat162_rn(x0, x1);
      *reinterpret_cast<__nv_bfloat162 *>(hi + (size_t)p * Dp + d) = h2;
      if (with_lo)
        *reinterpret_cast<__nv_bfloat162 *>(lo + (size_t)p * Dp + d) =
            __floats2bfloat162_rn(x0 - __bfloat162float(h2.x), x1 - __bfloat162float(h2.y));
    }
  }
}

// ------------------------------------------------------------------ prep for RC_PREC_FP16_FP8
// NCHW fp32 -> K-major fp16 "hi" map and the 8-bit pair map (see norm_mul above), each pixel normalised by its own power
// of two; 1 / that power goes to inv_s[b][p] for the build's epilogue.  One block = 32 pixels x all channels: first the
// pixels' largest magnitudes (coalesced along the pixels), then the 64-channel blocks as in prep_kernel (the block's
// 32 KB of input are re-read from L1 / L2).
__global__ void __launch_bounds__(256)
prep_f16f8_kernel(const float *__restrict__ f1, const float *__restrict__ f2, __half *__restrict__ a_hi, uint8_t *__restrict__ a_8,
                  __half *__restrict__ b_hi, uint8_t *__restrict__ b_8, float *__restrict__ inv_sa, float *__restrict__ inv_sb,
                  int D, int Dp, int N) {
  __shared__ float tile[64][33];
  __shared__ float pmax[8][33];
  __shared__ float pscale[32];
  const int which = blockIdx.y & 1, b = blockIdx.y >> 1;
  const float *src = (which ? f2 : f1) + (size_t)b * D * N;
  __half *hi = (which ? b_hi : a_hi) + (size_t)b * N * Dp;
  uint8_t *p8 = (which ? b_8 : a_8) + (size_t)b * N * Dp * 2;
  float *inv_s = (which ? inv_sb : inv_sa) + (size_t)b * N;
  const int p0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  pdl_launch_dependents();
  pdl_wait();
  {
    float m = 0.0f;
    const int p = p0 + tx;
    if (p < N) {
#pragma unroll 8
      for (int d = ty; d < D; d += 8) m = fmaxf(m, fabsf(__ldg(src + (size_t)d * N + p)));
    }
    pmax[ty][tx] = m;
    __syncthreads();
    if (ty == 0) {
#pragma unroll
      for (int i = 1; i < 8; ++i) m = fmaxf(m, pmax[i][tx]);
      const float s = norm_mul(__float_as_uint(m));
      pscale[tx] = s;
      if (p < N) inv_s[p] = 1.0f / s;  // exact: a power of two
    }
    __syncthreads();
  }
  // the lo halves of the 128 B row pair with the other map's hi halves: A = [lo | hi], B = [hi | lo]
  const int off_lo = which ? 64 : 0, off_hi = which ? 0 : 64;
  for (int kb = 0; kb < Dp / 64; ++kb) {
    const int d0 = kb * 64;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int d = d0 + ty + 8 * i, p = p0 + tx;
      tile[ty + 8 * i][tx] = (d < D && p < N) ? __ldg(src + (size_t)d * N + p) * pscale[tx] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int pl = ty + 8 * i, p = p0 + pl, d = d0 + 2 * tx;
      if (p < N) {
        const float x0 = tile[2 * tx][pl], x1 = tile[2 * tx + 1][pl];
        const __half2 h2 = __floats2half2_rn(x0, x1);
        *reinterpret_cast<__half2 *>(hi + (size_t)p * Dp + d) = h2;
        const float h0 = __low2float(h2), h1 = __high2float(h2);
        uint8_t *row = p8 + (size_t)p * Dp * 2 + (size_t)kb * 128 + 2 * tx;
        *reinterpret_cast<uint16_t *>(row + off_lo) =
            __nv_cvt_float2_to_fp8x2(make_float2((x0 - h0) * 16.0f, (x1 - h1) * 16.0f), __NV_SATFINITE, __NV_E4M3);
        *reinterpret_cast<uint16_t *>(row + off_hi) =
            __nv_cvt_float2_to_fp8x2(make_float2(h0 * 0.0625f, h1 * 0.0625f), __NV_SATFINITE, __NV_E4M3);
      }
    }
    __syncthreads();
  }
}

// The same for Dp <= 256 (every RAFT model) in ONE pass over the input: the block's 32 pixels x Dp channels stay in shared
// memory between finding the pixels' magnitudes and converting them (32 loads in flight per thread, no re-read).
__global__ void __launch_bounds__(256)
prep_f16f8_small_kernel(const float *__restrict__ f1, const float *__restrict__ f2, __half *__restrict__ a_hi,
                        uint8_t *__restrict__ a_8, __half *__restrict__ b_hi, uint8_t *__restrict__ b_8, float *__restrict__ inv_sa,
                        float *__restrict__ inv_sb, int D, int Dp, int N) {
  __shared__ float tile[256][33];
  __shared__ float pmax[8][33];
  __shared__ float pscale[32];
  const int which = blockIdx.y & 1, b = blockIdx.y >> 1;
  const float *src = (which ? f2 : f1) + (size_t)b * D * N;
  __half *hi = (which ? b_hi : a_hi) + (size_t)b * N * Dp;
  uint8_t *p8 = (which ? b_8 : a_8) + (size_t)b * N * Dp * 2;
  float *inv_s = (which ? inv_sb : inv_sa) + (size_t)b * N;
  const int p0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  pdl_launch_dependents();
  pdl_wait();
  const int p = p0 + tx;
  float m = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const int d = ty + 8 * i;
    float v = 0.0f;
    if (d < D && p < N) v = __ldg(src + (size_t)d * N + p);
    if (d < Dp) tile[d][tx] = v;
    m = fmaxf(m, fabsf(v));
  }
  pmax[ty][tx] = m;
  __syncthreads();
  if (ty == 0) {
#pragma unroll
    for (int i = 1; i < 8; ++i) m = fmaxf(m, pmax[i][tx]);
    const float s = norm_mul(__float_as_uint(m));
    pscale[tx] = s;
    if (p < N) inv_s[p] = 1.0f / s;  // exact: a power of two
  }
  __syncthreads();
  const int off_lo = which ? 64 : 0, off_hi = which ? 0 : 64;
  for (int kb = 0; kb < Dp / 64; ++kb) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int pl = ty + 8 * i, pp = p0 + pl, d = kb * 64 + 2 * tx;
      if (pp < N) {
        const float sc = pscale[pl];
        const float x0 = tile[d][pl] * sc, x1 = tile[d + 1][pl] * sc;
        const __half2 h2 = __floats2half2_rn(x0, x1);
        *reinterpret_cast<__half2 *>(hi + (size_t)pp * Dp + d) = h2;
        const float h0 = __low2float(h2), h1 = __high2float(h2);
        uint8_t *row = p8 + (size_t)pp * Dp * 2 + (size_t)kb * 128 + 2 * tx;
        *reinterpret_cast<uint16_t *>(row + off_lo) =
            __nv_cvt_float2_to_fp8x2(make_float2((x0 - h0) * 16.0f, (x1 - h1) * 16.0f), __NV_SATFINITE, __NV_E4M3);
        *reinterpret_cast<uint16_t *>(row + off_hi) =
            __nv_cvt_float2_to_fp8x2(make_float2(h0 * 0.0625f, h1 * 0.0625f), __NV_SATFINITE, __NV_E4M3);
      }
    }
  }
}

// ------------------------------------------------------------------ host side
static int make_map(CUtensorMap *tm, void *base, int rank, const cuuint64_t *dims, const cuuint64_t *strides,
                    const cuuint32_t *box) {
  return encode_tmap_sw128(tm, base, rank, reinterpret_cast<const unsigned long long *>(dims),
                           reinterpret_cast<const unsigned long long *>(strides), box, "rc_corr_build_pyramid");
}

static int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  return n;
}

template <int NPASS, int NCTA, bool TILED, bool V16 = false>
static int launch(const CUtensorMap &ah, const CUtensorMap &al, const CUtensorMap &bh, const CUtensorMap &bl,
                  const Params &P, cudaStream_t st) {
  using C = Cfg<NPASS, NCTA>;
  static_assert(C::kSmemBytes <= 227 * 1024, "shared memory budget");
  auto kern = build_tc_kernel<NPASS, NCTA, TILED, V16>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  const long long units = sm_count() / NCTA;  // one CTA (or CTA pair) per SM (per TPC)
  const unsigned grid = (unsigned)((P.total_tiles < units ? P.total_tiles : units) * NCTA);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = NCTA;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // set up while the prepass drains
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  RC_CUDA(cudaLaunchKernelEx(&cfg, kern, ah, al, bh, bl, P));
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

// CTA pairs (cta_group::2) unless RAFT_CORR_TC_CTAS=1 asks for the single-CTA kernel
static int ctas_per_tile() {
  static const int n = [] {
    const char *e = getenv("RAFT_CORR_TC_CTAS");
    return (e && e[0] == '1') ? 1 : 2;
  }();
  return n;
}

}  // namespace tc

static inline int round_up64(int d) { return (d + 63) / 64 * 64; }

size_t build_tc_workspace_bytes(int B, int D, int h, int w, int precision) {
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0) return 0;
  const size_t per_map = (size_t)B * h * w * round_up64(D) * 2;  // bf16
  const size_t aligned = (per_map + 255) / 256 * 256;
  if (precision == RC_PREC_FP16_FP8)  // + 1 / (normalisation) of every pixel of both maps
    return aligned * 4 + ((size_t)2 * B * h * w * 4 + 255) / 256 * 256;
  return aligned * 2 * (precision == RC_PREC_BF16X3 ? 2 : 1);
}

static int build_from_operands(__nv_bfloat16 *a_hi, __nv_bfloat16 *a_lo, __nv_bfloat16 *b_hi, __nv_bfloat16 *b_lo, int B,
                               int Dp, int h, int w, const LevelPtrs &lv, int L, int precision, float scale, cudaStream_t st,
                               bool tiled, const float *inv_s = nullptr);

static int launch_build_tc_impl(const float *f1, const float *f2, int B, int D, int h, int w, const LevelPtrs &lv,
                                int L, int precision, float scale, void *ws, size_t ws_bytes, cudaStream_t st,
                                bool tiled) {
  using namespace tc;
  const size_t need = build_tc_workspace_bytes(B, D, h, w, precision);
  if (ws == nullptr || ws_bytes < need) {
    set_error("rc_corr_build_pyramid: workspace too small (%zu < %zu bytes)", ws_bytes, need);
    return RC_ERR_WORKSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(ws) & 255) != 0) {
    set_error("rc_corr_build_pyramid: workspace must be 256-byte aligned");
    return RC_ERR_INVALID;
  }
  const int N = h * w, Dp = round_up64(D);
  const bool f8 = precision == RC_PREC_FP16_FP8;
  const bool x3 = precision == RC_PREC_BF16X3 || f8;  // two operand maps per feature map
  const bool v16 = precision == RC_PREC_BF16_VOL16;
  if (v16 && !tiled) {
    set_error("RC_PREC_BF16_VOL16 exists in the tiled layout only (rc_corr_build_pyramid_tiled)");
    return RC_ERR_UNSUPPORTED;
  }
  const int ncta = tc::ctas_per_tile();
  const size_t per_map = ((size_t)B * N * Dp * 2 + 255) / 256 * 256;
  uint8_t *base = static_cast<uint8_t *>(ws);
  __nv_bfloat16 *a_hi = reinterpret_cast<__nv_bfloat16 *>(base);
  __nv_bfloat16 *b_hi = reinterpret_cast<__nv_bfloat16 *>(base + per_map);
  __nv_bfloat16 *a_lo = x3 ? reinterpret_cast<__nv_bfloat16 *>(base + 2 * per_map) : a_hi;
  __nv_bfloat16 *b_lo = x3 ? reinterpret_cast<__nv_bfloat16 *>(base + 3 * per_map) : b_hi;

  const dim3 grid((unsigned)((N + 31) / 32), (unsigned)(Dp / 64), (unsigned)(2 * B));
  if (f8) {
    float *inv_s = reinterpret_cast<float *>(base + 4 * per_map);  // [2][B][N]
    RC_CUDA(launch_pdl(Dp <= 256 ? prep_f16f8_small_kernel : prep_f16f8_kernel, dim3((unsigned)((N + 31) / 32), (unsigned)(2 * B)), dim3(256), 0, st, f1, f2,
                       reinterpret_cast<__half *>(a_hi), reinterpret_cast<uint8_t *>(a_lo), reinterpret_cast<__half *>(b_hi),
                       reinterpret_cast<uint8_t *>(b_lo), inv_s, inv_s + (size_t)B * N, D, Dp, N));
    return build_from_operands(a_hi, a_lo, b_hi, b_lo, B, Dp, h, w, lv, L, precision, scale, st, tiled, inv_s);
  }
  RC_CUDA(launch_pdl(prep_kernel, grid, dim3(256), 0, st, f1, f2, a_hi, a_lo, b_hi, b_lo, D, Dp, N, x3 ? 1 : 0));
  return build_from_operands(a_hi, a_lo, b_hi, b_lo, B, Dp, h, w, lv, L, precision, scale, st, tiled);
}

// The tensor-core build proper, on operands already in the K-major bf16 hi (/ lo) layout (B, h*w, Dp): what the
// layout prepass above writes, and what rc_fnet_conv1x1_prep (fnet_head.cu) emits straight from the encoder.
static int build_from_operands(__nv_bfloat16 *a_hi, __nv_bfloat16 *a_lo, __nv_bfloat16 *b_hi, __nv_bfloat16 *b_lo, int B,
                               int Dp, int h, int w, const LevelPtrs &lv, int L, int precision, float scale, cudaStream_t st,
                               bool tiled, const float *inv_s) {
  using namespace tc;
  const int N = h * w;
  const bool f8 = precision == RC_PREC_FP16_FP8;
  const bool x3 = precision == RC_PREC_BF16X3;
  const bool v16 = precision == RC_PREC_BF16_VOL16;
  const int ncta = tc::ctas_per_tile();

  CUtensorMap tmAh, tmAl, tmBh, tmBl;
  {
    const cuuint64_t dims[3] = {(cuuint64_t)Dp, (cuuint64_t)N, (cuuint64_t)B};
    const cuuint64_t strides[2] = {(cuuint64_t)Dp * 2, (cuuint64_t)N * Dp * 2};
    const cuuint32_t box[3] = {TK, TM, 1};
    if (int rc = make_map(&tmAh, a_hi, 3, dims, strides, box)) return rc;
    if (int rc = make_map(&tmAl, a_lo, 3, dims, strides, box)) return rc;
  }
  {
    const cuuint64_t dims[4] = {(cuuint64_t)Dp, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
    const cuuint64_t strides[3] = {(cuuint64_t)Dp * 2, (cuuint64_t)w * Dp * 2, (cuuint64_t)N * Dp * 2};
    const cuuint32_t box[4] = {TK, PW, (cuuint32_t)(PH / ncta), 1};  // a CTA of a pair loads half the patch
    if (int rc = make_map(&tmBh, b_hi, 4, dims, strides, box)) return rc;
    if (int rc = make_map(&tmBl, b_lo, 4, dims, strides, box)) return rc;
  }

  Params P{};
  P.B = B; P.N = N; P.h = h; P.w = w; P.L = L;
  P.k_blocks = Dp / TK;
  P.m_tiles = (N + TM * ncta - 1) / (TM * ncta);  // tiles (or tile pairs) along the query pixels
  P.nt_x = (w + PW - 1) / PW;
  P.nt_y = (h + PH - 1) / PH;
  P.total_tiles = (long long)B * P.m_tiles * P.nt_x * P.nt_y;
  P.scale = scale;
  P.inv_sa = inv_s;
  P.inv_sb = inv_s ? inv_s + (size_t)B * N : nullptr;
  P.lv = lv;
  P.vec_mask = 0;
  for (int l = 0; l < 4 && l < L; ++l) {
    const int wl = w >> l;
    const uintptr_t a = reinterpret_cast<uintptr_t>(lv.p[l]);
    if (wl % 4 == 0 && (a & 15) == 0) P.vec_mask |= 1 << l;
  }
  for (int l = L; l < 4; ++l) P.lv.p[l] = lv.p[0];  // never dereferenced; keeps the table non-null
  for (int l = 0; l < 4; ++l) {
    // tile = 2 rows x 8 floats, or 2 rows x 16 halves; gstride in 32-bit words
    const int cw = v16 ? 16 : 8;
    const int hp = ((h >> l) + 1) & ~1, wp = ((w >> l) + cw - 1) & ~(cw - 1);
    P.tpr[l] = wp / cw;
    P.gstride[l] = (long long)hp * wp * kTileGroup / (v16 ? 2 : 1);
  }

  P.ng = (N + kTileGroup - 1) / kTileGroup;

  int rc;
  if (tiled) {
    if (L > 4) {
      set_error("rc_corr_build_pyramid_tiled: at most 4 levels (got %d)", L);
      return RC_ERR_UNSUPPORTED;
    }
    for (int l = 0; l < L; ++l)
      if ((reinterpret_cast<uintptr_t>(lv.p[l]) & 63) != 0) {
        set_error("rc_corr_build_pyramid_tiled: level %d must be 64-byte aligned", l);
        return RC_ERR_INVALID;
      }
    if (v16)
      rc = ncta == 2 ? tc::launch<1, 2, true, true>(tmAh, tmAl, tmBh, tmBl, P, st)
                     : tc::launch<1, 1, true, true>(tmAh, tmAl, tmBh, tmBl, P, st);
    else if (f8)
      rc = ncta == 2 ? tc::launch<2, 2, true>(tmAh, tmAl, tmBh, tmBl, P, st) : tc::launch<2, 1, true>(tmAh, tmAl, tmBh, tmBl, P, st);
    else if (ncta == 2)
      rc = x3 ? tc::launch<3, 2, true>(tmAh, tmAl, tmBh, tmBl, P, st) : tc::launch<1, 2, true>(tmAh, tmAl, tmBh, tmBl, P, st);
    else
      rc = x3 ? tc::launch<3, 1, true>(tmAh, tmAl, tmBh, tmBl, P, st) : tc::launch<1, 1, true>(tmAh, tmAl, tmBh, tmBl, P, st);
    return rc;
  }
  if (f8)
    rc = ncta == 2 ? tc::launch<2, 2, false>(tmAh, tmAl, tmBh, tmBl, P, st) : tc::launch<2, 1, false>(tmAh, tmAl, tmBh, tmBl, P, st);
  else if (ncta == 2)
    rc = x3 ? tc::launch<3, 2, false>(tmAh, tmAl, tmBh, tmBl, P, st) : tc::launch<1, 2, false>(tmAh, tmAl, tmBh, tmBl, P, st);
  else
    rc = x3 ? tc::launch<3, 1, false>(tmAh, tmAl, tmBh, tmBl, P, st) : tc::launch<1, 1, false>(tmAh, tmAl, tmBh, tmBl, P, st);
  if (rc) return rc;
  if (L > 4) return launch_pool_tail(lv, B, h, w, 4, L, st);
  return RC_OK;
}

int launch_build_tc_tiled_prepared(const void *a_hi, const void *a_lo, const void *b_hi, const void *b_lo, int B, int Dp, int h,
                                   int w, const LevelPtrs &lv, int L, int precision, float scale, cudaStream_t st) {
  auto p = [](const void *q) { return static_cast<__nv_bfloat16 *>(const_cast<void *>(q)); };
  const bool x3 = precision == RC_PREC_BF16X3;
  return build_from_operands(p(a_hi), x3 ? p(a_lo) : p(a_hi), p(b_hi), x3 ? p(b_lo) : p(b_hi), B, Dp, h, w, lv, L, precision, scale,
                             st, true);
}

int launch_build_tc(const float *f1, const float *f2, int B, int D, int h, int w, const LevelPtrs &lv, int L,
                    int precision, float scale, void *ws, size_t ws_bytes, cudaStream_t st) {
  return launch_build_tc_impl(f1, f2, B, D, h, w, lv, L, precision, scale, ws, ws_bytes, st, false);
}

int launch_build_tc_tiled(const float *f1, const float *f2, int B, int D, int h, int w, const LevelPtrs &lv, int L,
                          int precision, float scale, void *ws, size_t ws_bytes, cudaStream_t st) {
  return launch_build_tc_impl(f1, f2, B, D, h, w, lv, L, precision, scale, ws, ws_bytes, st, true);
}

// ------------------------------------------------------------------ tiled <-> row layout of one level
// Row layout: (B*N, hl, wl).  Tiled layout: [B][ceil(N/G)][hp/2][wp/8][G pixels][4 chunks of 16 B] -- the
// tiles of G = 32 consecutive query pixels are interleaved (a group's copies of one tile are 2 KB of
// contiguous bytes) and pixel u keeps chunk k = 2*(row & 1) + (column half) at position k ^ ((u >> 1) & 3).  The loop walks the TILED array (the padded, larger one): converting to it
// writes every padding element (zero), converting from it skips them.
template <bool TO_TILED>
__global__ void __launch_bounds__(256)
relayout_kernel(const float *__restrict__ src, float *__restrict__ dst, int B, int N, int hl, int wl) {
  const int hp = (hl + 1) & ~1, wp = (wl + 7) & ~7, tpr = wp >> 3;
  constexpr int G = kTileGroup;
  const int ng = (N + G - 1) / G;
  const long long gstride = (long long)hp * wp * G, total = (long long)B * ng * gstride;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long grp = i / gstride;
    const int e = (int)(i - grp * gstride);
    const int tile = e / (16 * G), u = (e >> 4) % G;
    const int k = ((e >> 2) & 3) ^ ((u >> 1) & 3);  // chunk stored at this position
    const int ty = tile / tpr, tx = tile - ty * tpr;
    const int y = ty * 2 + (k >> 1), x = tx * 8 + (k & 1) * 4 + (e & 3);
    const int b = (int)(grp / ng), p = (int)(grp - (long long)b * ng) * G + u;
    const bool inside = y < hl && x < wl && p < N;
    const long long r = ((long long)b * N + p) * hl * wl + (long long)y * wl + x;
    if (TO_TILED) dst[i] = inside ? __ldg(src + r) : 0.0f;
    else if (inside) dst[r] = __ldg(src + i);
  }
}

int launch_relayout(const float *src, float *dst, int B, int N, int hl, int wl, bool to_tiled, cudaStream_t st) {
  const long long total = (long long)B * ((N + kTileGroup - 1) / kTileGroup) * kTileGroup * ((hl + 1) & ~1) * ((wl + 7) & ~7);
  if (total <= 0) return RC_OK;
  long long blocks = (total + 255) / 256;
  if (blocks > 148LL * 32) blocks = 148LL * 32;
  if (to_tiled) relayout_kernel<true><<<(unsigned)blocks, 256, 0, st>>>(src, dst, B, N, hl, wl);
  else relayout_kernel<false><<<(unsigned)blocks, 256, 0, st>>>(src, dst, B, N, hl, wl);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

// The same for the 16-bit volume: tiled fp16 (tiles of 2 rows x 16 halves) <-> row-layout fp32.
template <bool TO_TILED>
__global__ void __launch_bounds__(256)
relayout16_kernel(const void *__restrict__ src_, void *__restrict__ dst_, int B, int N, int hl, int wl) {
  const int hp = (hl + 1) & ~1, wp = (wl + 15) & ~15, tpr = wp >> 4;
  constexpr int G = kTileGroup;
  const int ng = (N + G - 1) / G;
  const long long gstride = (long long)hp * wp * G, total = (long long)B * ng * gstride;  // in halves
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long grp = i / gstride;
    const int e = (int)(i - grp * gstride);
    const int tile = e / (32 * G), u = (e >> 5) % G;   // 32 halves per pixel per tile
    const int k = ((e >> 3) & 3) ^ ((u >> 1) & 3);    // chunk stored at this position
    const int ty = tile / tpr, tx = tile - ty * tpr;
    const int y = ty * 2 + (k >> 1), x = tx * 16 + (k & 1) * 8 + (e & 7);
    const int b = (int)(grp / ng), p = (int)(grp - (long long)b * ng) * G + u;
    const bool inside = y < hl && x < wl && p < N;
    const long long r = ((long long)b * N + p) * hl * wl + (long long)y * wl + x;
    if (TO_TILED) {
      const float v = inside ? __ldg(static_cast<const float *>(src_) + r) : 0.0f;
      static_cast<__half *>(dst_)[i] = __float2half_rn(fminf(fmaxf(v, -65504.0f), 65504.0f));
    } else if (inside) {
      static_cast<float *>(dst_)[r] = __half2float(static_cast<const __half *>(src_)[i]);
    }
  }
}

int launch_relayout16(const void *src, void *dst, int B, int N, int hl, int wl, bool to_tiled, cudaStream_t st) {
  const long long total = (long long)B * ((N + kTileGroup - 1) / kTileGroup) * kTileGroup * ((hl + 1) & ~1) * ((wl + 15) & ~15);
  if (total <= 0) return RC_OK;
  long long blocks = (total + 255) / 256;
  if (blocks > 148LL * 32) blocks = 148LL * 32;
  if (to_tiled) relayout16_kernel<true><<<(unsigned)blocks, 256, 0, st>>>(src, dst, B, N, hl, wl);
  else relayout16_kernel<false><<<(unsigned)blocks, 256, 0, st>>>(src, dst, B, N, hl, wl);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

// =====================================================================================
// Allocation-free variant of the tensor-core build (no counterpart in the reference).
//
// Some hosts of this library must not allocate device memory while they run: a real-time
// inference loop such as the ROS node's worker thread cannot afford cudaMalloc's unbounded
// latency and implicit device-wide synchronisation, and a process whose device memory has been
// handed over in full to another pool (a framework's caching allocator, a pre-partitioned
// embedded budget) would simply see cudaMalloc fail.  For them the scratch that
// rc_corr_build_pyramid normally receives from the caller lives in the fixed array below, which
// is part of the library image: every GPU that loads libraftcorr.so holds these
// RC_STATIC_SCRATCH_BYTES until the process exits, whether or not this entry point is used.
//
// Rules of use (also in include/raftcorr.h):
//  * sizes whose scratch need exceeds RC_STATIC_SCRATCH_BYTES are rejected, nothing is launched;
//  * the first call on each device resolves the device address of the array (one
//    cudaGetSymbolAddress with that device current); make that first call before any stream
//    capture is begun on that device;
//  * all calls on one device share the one array and it keeps whatever the previous call left.
//    Nothing in it has to start at zero: the layout prepass rewrites every byte the main kernel
//    reads (channel padding included) before it is read.  Two calls that overlap on the GPU
//    (different streams, no dependency) race on it: keep them ordered.
// =====================================================================================
static __device__ __align__(256) unsigned char g_static_scratch[RC_STATIC_SCRATCH_BYTES];

namespace {
std::mutex g_scratch_mu;
void *g_scratch_addr[64] = {};  // per device ordinal; filled lazily, never freed
}  // namespace

// In host code `g_static_scratch` is only a handle for the symbol, not a device address: the
// address a kernel can use has to be asked for, per device.
static int static_scratch_ptr(void **out) {
  int dev = -1;
  RC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("rc_corr_build_pyramid_noalloc: device ordinal %d not supported", dev);
    return RC_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
    if (p == nullptr) {
      set_error("rc_corr_build_pyramid_noalloc: static scratch has no device address");
      return RC_ERR_CUDA;
    }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return RC_OK;
}

int launch_build_tc_noalloc(const float *f1, const float *f2, int B, int D, int h, int w, const LevelPtrs &lv,
                            int L, int precision, float scale, cudaStream_t st) {
  const size_t need = build_tc_workspace_bytes(B, D, h, w, precision);
  if (need == 0 || need > (size_t)RC_STATIC_SCRATCH_BYTES) {
    set_error("rc_corr_build_pyramid_noalloc: this problem needs %zu bytes of scratch but the built-in scratch "
              "holds %zu; use rc_corr_build_pyramid with a caller-provided workspace",
              need, (size_t)RC_STATIC_SCRATCH_BYTES);
    return RC_ERR_WORKSPACE;
  }
  void *ws = nullptr;
  if (int rc = static_scratch_ptr(&ws)) return rc;
  return launch_build_tc(f1, f2, B, D, h, w, lv, L, precision, scale, ws, (size_t)RC_STATIC_SCRATCH_BYTES, st);
}

}  // namespace rc
