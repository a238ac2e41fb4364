// Pyramid lookup on the TILED volume layout (r = 3 or 4): warp-autonomous, persistent, cp.async-fed.
//
// Same result as lookup_fwd_kernel<R> (reference core/corr.py:29-50) on the TILED layout of
// include/raftcorr.h: every row p of a level is cut into tiles of 2 rows x 8 floats (64 B, the
// plane padded with zeros to an even height and a multiple-of-8 width), the tiles of 32 consecutive
// query pixels are interleaved, and each pixel's four 16 B chunks sit at swizzled positions:
//     level l : [B][ceil(N/32)][hp/2][wp/8][32 pixels][4 chunks][4 floats]
// A 10x10 window then lies in 5-6 x 2-3 tiles = whole DRAM granules instead of ten 40 B segments
// that straddle them, and -- because neighbouring query pixels look at neighbouring places -- the
// tiles the 8 pixels of a work item need sit next to each other (DRAM reads per launch: 235 MB in
// the row layout, 132 MB here).  The shape is what the build epilogue produces for free
// (build_tc.cu).  The V16 flavour reads an fp16 volume (tiles of 2 x 16 halves).
//
// Work item = 8 consecutive query pixels of one level, owned by ONE warp (no block-level sync, no
// shuffles: lane = pixel x quarter in every phase):
//   fetch   the 4 lanes of pixel u load its coordinates (one iteration ahead of their use) and
//           derive the window origin / fractions;
//   issue   lane (u, k) copies the k-th 16 B piece of the 10 rows of pixel u's patch global ->
//           shared with cp.async (zero-filled outside the map) into a two-stage ring: the copies
//           of item i+1 are in flight while item i is consumed;
//   taps    lane (u, c) owns window columns 2c, 2c+1 (and a share of column 8) of pixel u:
//           separable bilinear interpolation from registers, 21 taps per lane;
//   store   straight from registers: the 8 pixels of a warp are contiguous in the output, so every
//           store instruction writes whole 32 B sectors.
#include <cuda_fp16.h>

#include "common.cuh"

namespace rc {

#ifndef RC_TL_STAGES
#define RC_TL_STAGES 2
#endif
#ifndef RC_TL_PS
#define RC_TL_PS 20
#endif
#ifndef RC_TL_EXP
#define RC_TL_EXP 0  // profiling experiments only: 1 = no output stores, 2 = no volume reads
#endif
#if RC_TL_EXP == 1
#define RC_TL_ST(p, v) do { const float v_ = (v); if (v_ == 1.2345e30f) st_stream((p), v_); } while (0)
#else
#define RC_TL_ST(p, v) st_stream((p), (v))
#endif
constexpr int kTlWarps = 4, kTlThreads = kTlWarps * 32, kTlPix = 8, kTlStages = RC_TL_STAGES;
constexpr int kTlPS = RC_TL_PS;                 // patch row stride (floats): 16 fetched (+ pad), 16 B aligned
constexpr int kTlPatch = 10 * kTlPS + 4;        // per-pixel patch (floats), padded to spread banks
constexpr int kTlStageFloats = kTlPix * kTlPatch;

__device__ __forceinline__ void cp_async16_zfill(float *dst_smem, const float *src, bool valid) {
  const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(dst_smem));
  const int n = valid ? 16 : 0;  // src-size 0: nothing is read, the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// One work item as seen by lane (u, k), u = lane >> 2 the pixel, k = lane & 3: the item's place in
// the problem (warp-uniform) and pixel u's window.  The four lanes of a pixel hold the same values
// (they load the same coordinates), so neither the copies nor the taps need a shuffle.
struct TlRaw {
  int lvl, b, p0, valid;
  float cx, cy;
};
struct TlCtx {
  int lvl, b, p0, valid;
  int x0, y0;  // window origin (y0 far outside for a pixel beyond the image: every row zero-filled)
  float fx, fy;
};

// V16: the volume is fp16, tiles of 2 rows x 16 halves (same 64 B, same 4 chunks of 16 B = 8 halves):
// a window row is 2-3 pieces of 8 halves from a multiple-of-8 column, kept as halves in shared memory
// (48 B per row) and converted when the taps read them; interpolation and output stay fp32.
// RR = window radius: 4 (raft-things) or 3 (raft-small, the ROS node's `is_small`).
template <bool V16, int RR>
__global__ void __launch_bounds__(kTlThreads)
lookup_tiled_fwd_kernel(ConstLevelPtrs lv, const float *__restrict__ coords, float *__restrict__ out, int B, int h, int w,
                        int L, int groups_per_item, int total_items) {
  constexpr int R = RR, RD = 2 * R + 1, PN = RD + 1, NTAP = RD * RD, PS = kTlPS;
  static_assert(R == 3 || R == 4, "tap assignment below is written for radius 3 and 4");
  extern __shared__ __align__(16) float smem[];
  const int N = h * w;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *ring = smem + warp * (kTlStages * kTlStageFloats);
  const int tu = lane >> 2, tc = lane & 3;  // pixel 0..7 of the item; 16 B piece of a patch row / column pair

  // ---- item decoder: items are numbered level-major (lvl, b, group); a warp visits it0, it0 + stride, ...
  // The decoder runs ahead of the pipeline and advances by `stride` with carries instead of divisions.
  const int stride = gridDim.x * kTlWarps;
  const int it0 = blockIdx.x * kTlWarps + warp;
  const int per_level = B * groups_per_item;
  int dl = it0 / per_level, db = (it0 - dl * per_level) / groups_per_item, dg = it0 - dl * per_level - db * groups_per_item;
  const int s_l = stride / per_level, s_b = (stride - s_l * per_level) / groups_per_item,
            s_g = stride - s_l * per_level - s_b * groups_per_item;
  // Programmatic dependent launch (the launcher sets cudaLaunchAttributeProgrammaticStreamSerialization):
  // this grid is persistent and wholly resident, so the next kernel in the stream may be set up at once
  // -- its CTAs take the places of ours as they exit -- while everything above ran before the previous
  // kernel in the stream had finished.  Nothing produced by earlier kernels (coordinates, volume) is
  // read, and nothing is written, before the wait.
  pdl_launch_dependents();
  pdl_wait();
  auto next_raw = [&]() {
    TlRaw r{dl, db, dg * kTlPix, dl < L ? 1 : 0, 0.f, 0.f};
    if (r.valid) {
      const int p = r.p0 + tu;
      if (p < N) {
        // the 4 lanes of a pixel read the same word: one 32 B sector per plane for the warp
        r.cx = __ldg(coords + ((size_t)r.b * 2 + 0) * N + p);
        r.cy = __ldg(coords + ((size_t)r.b * 2 + 1) * N + p);
      }
    }
    dg += s_g;
    if (dg >= groups_per_item) { dg -= groups_per_item; ++db; }
    db += s_b;
    if (db >= B) { db -= B; ++dl; }
    dl += s_l;
    return r;
  };
  auto resolve = [&](const TlRaw &r) {
    TlCtx c{r.lvl, r.b, r.p0, r.valid, 0, -(1 << 28), 0.f, 0.f};
    if (r.valid && r.p0 + tu < N) {
      const float inv = 1.0f / (float)(1 << r.lvl);
      c.x0 = safe_floor_int(r.cx * inv, &c.fx) - R;
      c.y0 = safe_floor_int(r.cy * inv, &c.fy) - R;
    }
    return c;
  };

  const unsigned ring_u32 = static_cast<unsigned>(__cvta_generic_to_shared(ring));
  const unsigned dst_lane = (unsigned)(tu * kTlPatch + tc * 4) * 4u;
  auto copy16 = [&](unsigned dst, const float *src, bool ok) {
#if RC_TL_EXP == 2
    ok = false;
#endif
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
  };
  // Copies of one item: lane (u, k) fetches columns xg+4k .. +3 of the 10 window rows of pixel u,
  // xg = x0 rounded down to a multiple of 4 (the 4th piece only matters at offset 3).  src-size 0
  // zero-fills: rows outside the level, pieces outside it (the build zeroes columns wl ..
  // ceil4(wl)-1, so whole pieces are fetched up to that bound), pixels beyond the image.
  auto issue = [&](const TlCtx &c, int stage) {
    if (c.valid) {
      const int hl = h >> c.lvl, wl = w >> c.lvl;
      constexpr int CW = V16 ? 16 : 8;       // columns per tile
      constexpr int PC = V16 ? 8 : 4;        // columns per 16 B piece
      const int hp = (hl + 1) & ~1, wp = (wl + CW - 1) & ~(CW - 1), tpr = wp / CW;  // tiles per tile-row
      const int wb = (wl + PC - 1) & ~(PC - 1);
      const int cg = (c.x0 & ~(PC - 1)) + PC * tc;
      // pieces a window of PN columns needs from a multiple-of-PC start (r = 4: 3, or 4 at offset 3 in
      // fp32; 2, or 3 at offset 7 in fp16)
      const bool want = PC * tc < (c.x0 & (PC - 1)) + PN;  // piece k holds window columns iff PC*k < offset + PN
      const bool okx = want && (unsigned)cg < (unsigned)wb;
      // group of G = 32 query pixels: [tile][pixel u][4 chunks], chunk k = 2*(row & 1) + (column half) stored
      // at position k ^ ((u >> 1) & 3) (the build's bank-conflict-free staging order, build_tc.cu).
      // Offsets are in 32-bit words for both element types (a pixel's tile is 16 words).
      constexpr int G = kTileGroup;
      const int ng = (N + G - 1) / G;
      const size_t grp = (size_t)c.b * ng + c.p0 / G;
      const int u = c.p0 % G + tu, sw = (u >> 1) & 3, half = (cg / PC) & 1;
      const size_t gwords = (size_t)(hp * wp) * G / (V16 ? 2 : 1);
      // this pixel's slot in its group: always inside the level, so it doubles as the (unread) source of zero fills
      const float *ibase = pick_level(lv, c.lvl) + grp * gwords + u * 16;
      const int trs = tpr * 16 * G;  // words per tile-row of a group
      const int pos0 = ((half ^ sw) << 2), pos1 = (((2 | half) ^ sw) << 2);  // chunk position for even / odd rows
      // window row r is level row y0 + r: rows alternate between the two chunk positions and advance one tile-row
      // every second row, so the offsets are two arithmetic sequences (A: r even, B: r odd)
      const int par = c.y0 & 1, t0 = (c.y0 >> 1) * trs + (cg / CW) * (16 * G);
      const int tA = t0 + (par ? pos1 : pos0), tB = t0 + (par ? trs + pos0 : pos1);
      constexpr unsigned RB = V16 ? 48u : (unsigned)PS * 4u;  // bytes per patch row in shared memory
      const unsigned dst = ring_u32 + (unsigned)(stage * kTlStageFloats) * 4u + dst_lane;
#pragma unroll
      for (int r = 0; r < PN; ++r) {
        const bool ok = okx && (unsigned)(c.y0 + r) < (unsigned)hl;
        const int rel = ((r & 1) ? tB : tA) + (r >> 1) * trs;
        // (fp16 rows hold three pieces: the 4th lane of a pixel has nothing to write)
        if (!V16 || tc < 3) copy16(dst + (unsigned)r * RB, ibase + (ok ? rel : 0), ok);
      }
    }
    cp_async_commit();  // always commit (possibly empty) so that the wait counts stay uniform
  };

  // ring of kTlStages patch buffers: while item i is consumed, items i+1 .. i+D are in flight
  constexpr int D = kTlStages - 1;
  TlCtx q_[D + 1];  // q_[0] = item being consumed, q_[d] = item d steps ahead (coordinates resolved)
#pragma unroll
  for (int d = 0; d <= D; ++d) q_[d] = resolve(next_raw());
  TlRaw raw = next_raw();
#pragma unroll
  for (int d = 0; d < D; ++d) issue(q_[d], d);
  int stage = 0;
  const int CH = L * NTAP;

  for (int it = it0; it < total_items; it += stride) {
    const TlCtx nn = resolve(raw);  // item D+1 ahead: its coordinates were loaded an iteration ago
    raw = next_raw();
    issue(q_[D], (stage + D) % kTlStages);
    cp_async_wait<D>();
    __syncwarp();
    const TlCtx cur = q_[0];

    // ---- taps: lane (tu, tc) owns window columns 2tc, 2tc+1 and a share of column 8 of pixel tu
    const float dx = cur.fx, dy = cur.fy;
    const float ex = 1.0f - dx, ey = 1.0f - dy;
    const float *pbase = ring + stage * kTlStageFloats + tu * kTlPatch;
    // element (row j, window column i) of the patch, as fp32
    auto at = [&](int j, int i) -> float {
      if constexpr (V16) return __half2float(reinterpret_cast<const __half *>(pbase)[j * 24 + (cur.x0 & 7) + i]);
      else return pbase[j * PS + (cur.x0 & 3) + i];
    };
    const int p = cur.p0 + tu;
    float *o = out + ((size_t)cur.b * CH + (size_t)cur.lvl * NTAP) * N + p;
    const bool pv = p < N;
    if (tc < RD / 2) {
      // columns 2tc, 2tc+1: horizontal interpolation of the PN rows, then vertical pairs
      float h0[PN], h1[PN];
#pragma unroll
      for (int j = 0; j < PN; ++j) {
        const float a = at(j, 2 * tc), b = at(j, 2 * tc + 1), d = at(j, 2 * tc + 2);
        h0[j] = fmaf(b, dx, a * ex);
        h1[j] = fmaf(d, dx, b * ex);
      }
      if (pv) {
        float *o0 = o + (size_t)(2 * tc * RD) * N;
#pragma unroll
        for (int j = 0; j < RD; ++j) {
          RC_TL_ST(o0 + (size_t)j * N, fmaf(h0[j + 1], dy, h0[j] * ey));
          RC_TL_ST(o0 + (size_t)(RD + j) * N, fmaf(h1[j + 1], dy, h1[j] * ey));
        }
      }
    }
    if constexpr (R == 4) {
      // column 8, rows shared out over the 4 lanes of the pixel: lane tc takes j = tc, tc+4 (and 8 for tc = 0)
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int j = tc + 4 * k;
        if (j < RD && (k < 2 || tc == 0)) {
          const float a = at(j, 8), b = at(j, 9), c2 = at(j + 1, 8), d2 = at(j + 1, 9);
          const float top = fmaf(b, dx, a * ex), bot = fmaf(d2, dx, c2 * ex);
          if (pv) RC_TL_ST(o + (size_t)(8 * RD + j) * N, fmaf(bot, dy, top * ey));
        }
      }
    } else if (tc == RD / 2) {
      // r = 3: three lanes own the column pairs 0-5, the fourth the last column (6) on its own
      float hc[PN];
#pragma unroll
      for (int j = 0; j < PN; ++j) hc[j] = fmaf(at(j, RD), dx, at(j, RD - 1) * ex);
      if (pv) {
#pragma unroll
        for (int j = 0; j < RD; ++j) RC_TL_ST(o + (size_t)((RD - 1) * RD + j) * N, fmaf(hc[j + 1], dy, hc[j] * ey));
      }
    }
    __syncwarp();
    stage = (stage + 1) % kTlStages;
#pragma unroll
    for (int d = 0; d < D; ++d) q_[d] = q_[d + 1];
    q_[D] = nn;
  }
  cp_async_wait<0>();
}

template <bool V16, int RR>
static int launch_tiled_r(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h, int w, int L,
                          cudaStream_t st, const char *fn) {
  const int N = h * w;
  const int groups = (N + kTlPix - 1) / kTlPix;
  const long long total = (long long)L * B * groups;
  if (total + 4LL * 148 * 32 * kTlWarps > 0x7fffffffLL) {
    set_error("%s: problem too large", fn);
    return RC_ERR_UNSUPPORTED;
  }
  const size_t smem = sizeof(float) * (size_t)(kTlWarps * kTlStages * kTlStageFloats);
  // resident CTAs per device: asked once per device (this runs every RAFT iteration)
  static int resident[64] = {};  // (one table per instantiation)
  int dev = 0;
  RC_CUDA(cudaGetDevice(&dev));
  int cached = (dev >= 0 && dev < 64) ? __atomic_load_n(&resident[dev], __ATOMIC_ACQUIRE) : 0;
  if (cached == 0) {
    int sms = 0, ctas_per_sm = 0;
    RC_CUDA(cudaFuncSetAttribute(lookup_tiled_fwd_kernel<V16, RR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    RC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, lookup_tiled_fwd_kernel<V16, RR>, kTlThreads, smem));
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    cached = sms * ctas_per_sm;
    if (dev >= 0 && dev < 64) __atomic_store_n(&resident[dev], cached, __ATOMIC_RELEASE);
  }
  long long grid = cached;
  const long long need = (total + kTlWarps - 1) / kTlWarps;
  if (grid > need) grid = need;
  RC_CUDA(launch_pdl(lookup_tiled_fwd_kernel<V16, RR>, dim3((unsigned)grid), dim3(kTlThreads), smem, st, lv, coords, out, B, h,
                     w, L, groups, (int)total));
  return RC_OK;
}

template <bool V16>
static int launch_tiled(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h, int w, int L, int r,
                        cudaStream_t st, const char *fn) {
  if (r == 4) return launch_tiled_r<V16, 4>(lv, coords, out, B, h, w, L, st, fn);
  if (r == 3) return launch_tiled_r<V16, 3>(lv, coords, out, B, h, w, L, st, fn);
  set_error("%s: radius %d unsupported on the tiled layout (3 and 4 only)", fn, r);
  return RC_ERR_UNSUPPORTED;
}

int launch_lookup_tiled_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h, int w, int L, int r,
                            cudaStream_t st) {
  return launch_tiled<false>(lv, coords, out, B, h, w, L, r, st, "rc_corr_lookup_tiled_fwd");
}

int launch_lookup_tiled16_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h, int w, int L, int r,
                              cudaStream_t st) {
  return launch_tiled<true>(lv, coords, out, B, h, w, L, r, st, "rc_corr_lookup_tiled16_fwd");
}

}  // namespace rc
