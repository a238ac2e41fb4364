// Pyramid lookup (forward) and its deterministic backward.
//
// Forward replaces CorrBlock.__call__ (reference core/corr.py:29-50) and the
// bilinear_sampler/grid_sample chain under it (core/utils/utils.py:57-71): ~80 torch launches
// and a materialised (B*h*w, 9, 9, 2) grid per level become ONE launch.
//
// HBM-bound gather.  Per query pixel p and level l the (2r+1)^2 bilinear taps only touch the
// (2r+2)^2 integer patch of row p of level l, so the kernel reads that patch once (100 values
// for r=4, not 81*4), forms the taps from shared memory and writes them in the final
// (B, L*(2r+1)^2, h, w) layout with 128-byte coalesced stores (lanes <-> adjacent pixels).
//
// Work split: one CTA = 32 consecutive query pixels of one batch item x one level.
//   phase 1 (warp per pixel, lanes <-> patch elements): issue all patch loads of 4 pixels
//            (16 independent loads per lane in flight), park them in a per-warp patch buffer,
//            lanes <-> taps compute the bilinear sums into a [tap][33] tile (conflict-free).
//   phase 2 (lanes <-> pixels): one coalesced 128 B store per tap channel.
#include "common.cuh"

namespace rc {

#ifndef RC_LOOKUP_THREADS
#define RC_LOOKUP_THREADS 128
#endif
#ifndef RC_LOOKUP_UNROLL
#define RC_LOOKUP_UNROLL 4
#endif
constexpr int kLookupThreads = RC_LOOKUP_THREADS;
constexpr int kLookupWarps = kLookupThreads / 32;
constexpr int kTilePix = 32;
constexpr int kPixPerWarp = kTilePix / kLookupWarps;  // 8
constexpr int kUnroll = RC_LOOKUP_UNROLL < kPixPerWarp ? RC_LOOKUP_UNROLL : kPixPerWarp;  // pixels in flight per warp

template <int R>
__global__ void __launch_bounds__(kLookupThreads)
lookup_fwd_kernel(ConstLevelPtrs lv, const float *__restrict__ coords, float *__restrict__ out,
                  int B, int h, int w, int L, int tiles_per_item) {
  constexpr int RD = 2 * R + 1;
  constexpr int PN = RD + 1;
  constexpr int NPATCH = PN * PN;
  constexpr int NTAP = RD * RD;
  constexpr int NE = (NPATCH + 31) / 32;
  constexpr int NT = (NTAP + 31) / 32;
  constexpr int TS = kTilePix + 1;  // padded tile stride

  extern __shared__ float smem[];
  float *tile = smem;                                  // [NTAP][TS]
  float *patch_all = smem + NTAP * TS;                 // [warps][kUnroll][NPATCH]

  const int lvl = blockIdx.y;
  const int b = blockIdx.x / tiles_per_item;
  const int p0 = (blockIdx.x - b * tiles_per_item) * kTilePix;
  const int N = h * w;
  const int hl = h >> lvl, wl = w >> lvl;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float inv = 1.0f / (float)(1 << lvl);  // exact: same value as coords / 2**lvl
  const float *__restrict__ vol = pick_level(lv, lvl);
  float *patch = patch_all + warp * (kUnroll * NPATCH);

  // per-lane constants of the patch element(s) and tap(s) this lane owns, hoisted out of the
  // pixel loop: the kernel is issue-bound on integer address math otherwise
  int e_py[NE], e_px[NE], e_rel[NE];
#pragma unroll
  for (int k = 0; k < NE; ++k) {
    const int e = lane + 32 * k;
    e_py[k] = e / PN;
    e_px[k] = e - e_py[k] * PN;
    e_rel[k] = e_py[k] * wl + e_px[k];
  }
  int t_src[NT], t_dst[NT];
#pragma unroll
  for (int k = 0; k < NT; ++k) {
    const int t = lane + 32 * k;
    const int i = t / RD, j = t - i * RD;  // i: x offset, j: y offset
    t_src[k] = j * PN + i;
    t_dst[k] = t * TS;
  }
  const int row_elems = hl * wl;

  // Per-pixel scalars (window origin, fractions, interior flag) are computed ONCE, by lane u for
  // pixel u of this warp, from one coalesced coordinate load, and broadcast with shuffles; doing
  // them redundantly in all 32 lanes was a quarter of the kernel's (issue-bound) instructions.
  const int pw0 = p0 + warp * kPixPerWarp;
  int my_x0 = 0, my_y0 = -(1 << 28), my_in = 0;  // invalid pixel: every element out of range -> zeros
  float my_fx = 0.f, my_fy = 0.f;
  if (lane < kPixPerWarp && pw0 + lane < N) {
    const float cx = __ldg(coords + ((size_t)b * 2 + 0) * N + pw0 + lane) * inv;
    const float cy = __ldg(coords + ((size_t)b * 2 + 1) * N + pw0 + lane) * inv;
    my_x0 = safe_floor_int(cx, &my_fx) - R;
    my_y0 = safe_floor_int(cy, &my_fy) - R;
    my_in = (my_y0 >= 0 && my_x0 >= 0 && my_y0 + PN <= hl && my_x0 + PN <= wl) ? 1 : 0;
  }
  const float *__restrict__ base = vol + ((size_t)b * N + pw0) * (size_t)row_elems;  // pixel u: + u * row_elems

#pragma unroll 1
  for (int g = 0; g < kPixPerWarp; g += kUnroll) {
    float v[kUnroll][NE];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int x0 = __shfl_sync(0xffffffffu, my_x0, g + u);
      const int y0 = __shfl_sync(0xffffffffu, my_y0, g + u);
      const int interior = __shfl_sync(0xffffffffu, my_in, g + u);
      const float *row = base + (g + u) * row_elems + (y0 * wl + x0);
      // whole (2r+2)^2 patch inside the level: no per-element bounds tests (warp-uniform branch)
      if (interior) {
#pragma unroll
        for (int k = 0; k < NE; ++k)
          v[u][k] = ((NPATCH % 32 == 0) || lane + 32 * k < NPATCH) ? ld_stream(row + e_rel[k]) : 0.0f;
      } else {
#pragma unroll
        for (int k = 0; k < NE; ++k) {
          const bool ok = (lane + 32 * k < NPATCH) && (unsigned)(y0 + e_py[k]) < (unsigned)hl &&
                          (unsigned)(x0 + e_px[k]) < (unsigned)wl;
          v[u][k] = ok ? ld_stream(row + e_rel[k]) : 0.0f;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
#pragma unroll
      for (int k = 0; k < NE; ++k)
        if ((NPATCH % 32 == 0) || lane + 32 * k < NPATCH) patch[u * NPATCH + lane + 32 * k] = v[u][k];
    __syncwarp();
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int pix = warp * kPixPerWarp + g + u;
      const float dx = __shfl_sync(0xffffffffu, my_fx, g + u), dy = __shfl_sync(0xffffffffu, my_fy, g + u);
      const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx;
      const float w10 = dy * (1.0f - dx), w11 = dy * dx;
      const float *pp = patch + u * NPATCH;
#pragma unroll
      for (int k = 0; k < NT; ++k) {
        if ((NTAP % 32 == 0) || lane + 32 * k < NTAP) {
          const float *q = pp + t_src[k];
          float s = q[0] * w00;
          s = fmaf(q[1], w01, s);
          s = fmaf(q[PN], w10, s);
          s = fmaf(q[PN + 1], w11, s);
          tile[t_dst[k] + pix] = s;
        }
      }
    }
    __syncwarp();
  }
  __syncthreads();

  const int CH = L * NTAP;
  const int p = p0 + lane;
  if (p < N) {
    float *o = out + ((size_t)b * CH + (size_t)lvl * NTAP + warp) * N + p;
    const float *src = tile + warp * TS + lane;
    const size_t ostep = (size_t)kLookupWarps * N;
#pragma unroll 7
    for (int t = warp; t < NTAP; t += kLookupWarps) {
      st_stream(o, *src);
      o += ostep;
      src += kLookupWarps * TS;
    }
  }
}

// ------------------------------------------------------------------ backward
// Deterministic scatter-add: warp per (pixel, level); the (2r+2)^2 patch gradient is formed
// in shared memory from the (2r+1)^2 output cotangents in a fixed order and added to row p
// of the grad pyramid with plain read-modify-write (no other query touches that row).
template <int R>
__global__ void __launch_bounds__(kLookupThreads)
lookup_bwd_kernel(LevelPtrs glv, const float *__restrict__ coords, const float *__restrict__ gout,
                  int B, int h, int w, int L, int tiles_per_item) {
  constexpr int RD = 2 * R + 1;
  constexpr int PN = RD + 1;
  constexpr int NPATCH = PN * PN;
  constexpr int NTAP = RD * RD;
  constexpr int NE = (NPATCH + 31) / 32;
  constexpr int TS = kTilePix + 1;

  extern __shared__ float smem[];
  float *tile = smem;  // [NTAP][TS] cotangents of this level, transposed in via coalesced loads

  const int lvl = blockIdx.y;
  const int b = blockIdx.x / tiles_per_item;
  const int p0 = (blockIdx.x - b * tiles_per_item) * kTilePix;
  const int N = h * w;
  const int hl = h >> lvl, wl = w >> lvl;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float inv = 1.0f / (float)(1 << lvl);
  const int CH = L * NTAP;

  {
    const int p = p0 + lane;
    const float *gi = gout + ((size_t)b * CH + (size_t)lvl * NTAP) * N + p;
    for (int t = warp; t < NTAP; t += kLookupWarps)
      tile[t * TS + lane] = (p < N) ? ld_stream(gi + (size_t)t * N) : 0.0f;
  }
  __syncthreads();

  float *__restrict__ gvol = pick_level(glv, lvl);
  // The read-modify-write of a patch is latency-bound (one dependent DRAM round trip per pixel), so
  // PB pixels are handled together: all their loads are issued before the first add / store.  Rows
  // of different query pixels never alias, which is what makes the reordering legal.
  constexpr int PB = 2;
  static_assert(kPixPerWarp % PB == 0, "pixel batch");
#pragma unroll 1
  for (int u0 = 0; u0 < kPixPerWarp; u0 += PB) {
    float *dst[PB][NE];
    float add[PB][NE];
#pragma unroll
    for (int v = 0; v < PB; ++v) {
      const int pix = warp * kPixPerWarp + u0 + v;
      const int p = p0 + pix;
      const bool pvalid = p < N;  // warp-uniform
      float dx = 0.f, dy = 0.f;
      int x0 = 0, y0 = 0;
      if (pvalid) {
        const float cx = __ldg(coords + ((size_t)b * 2 + 0) * N + p) * inv;
        const float cy = __ldg(coords + ((size_t)b * 2 + 1) * N + p) * inv;
        x0 = safe_floor_int(cx, &dx) - R;
        y0 = safe_floor_int(cy, &dy) - R;
      }
      const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx;
      const float w10 = dy * (1.0f - dx), w11 = dy * dx;
      float *row = gvol + ((size_t)b * N + (pvalid ? p : 0)) * ((size_t)hl * wl);
#pragma unroll
      for (int k = 0; k < NE; ++k) {
        const int e = lane + 32 * k;
        const int py = e / PN, px = e - py * PN;
        const int gy = y0 + py, gx = x0 + px;
        dst[v][k] = nullptr;
        add[v][k] = 0.0f;
        if (pvalid && e < NPATCH && gy >= 0 && gy < hl && gx >= 0 && gx < wl) {
          // patch cell (py,px) receives from taps (i,j) in {px-1,px} x {py-1,py}
          float s = 0.0f;
          if (py < RD && px < RD) s = tile[(px * RD + py) * TS + pix] * w00;
          if (py < RD && px > 0) s = fmaf(tile[((px - 1) * RD + py) * TS + pix], w01, s);
          if (py > 0 && px < RD) s = fmaf(tile[(px * RD + py - 1) * TS + pix], w10, s);
          if (py > 0 && px > 0) s = fmaf(tile[((px - 1) * RD + py - 1) * TS + pix], w11, s);
          dst[v][k] = row + (size_t)gy * wl + gx;
          add[v][k] = s;
        }
      }
    }
    float old[PB][NE];
#pragma unroll
    for (int v = 0; v < PB; ++v)
#pragma unroll
      for (int k = 0; k < NE; ++k) old[v][k] = dst[v][k] ? *dst[v][k] : 0.0f;
#pragma unroll
    for (int v = 0; v < PB; ++v)
#pragma unroll
      for (int k = 0; k < NE; ++k)
        if (dst[v][k]) *dst[v][k] = old[v][k] + add[v][k];
  }
}

template <int R>
static int launch_fwd_r(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h,
                        int w, int L, cudaStream_t st) {
  constexpr int RD = 2 * R + 1, PN = RD + 1;
  const int N = h * w;
  const int tiles = (N + kTilePix - 1) / kTilePix;
  const size_t smem = sizeof(float) * (RD * RD * (kTilePix + 1) + kLookupWarps * kUnroll * PN * PN);
  if (smem > 48 * 1024)
    RC_CUDA(cudaFuncSetAttribute(lookup_fwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)(B * tiles), (unsigned)L);
  lookup_fwd_kernel<R><<<grid, kLookupThreads, smem, st>>>(lv, coords, out, B, h, w, L, tiles);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

template <int R>
static int launch_bwd_r(const float *gout, const float *coords, const LevelPtrs &glv, int B, int h,
                        int w, int L, cudaStream_t st) {
  constexpr int RD = 2 * R + 1;
  const int N = h * w;
  const int tiles = (N + kTilePix - 1) / kTilePix;
  const size_t smem = sizeof(float) * (RD * RD * (kTilePix + 1));
  if (smem > 48 * 1024)
    RC_CUDA(cudaFuncSetAttribute(lookup_bwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)(B * tiles), (unsigned)L);
  lookup_bwd_kernel<R><<<grid, kLookupThreads, smem, st>>>(glv, coords, gout, B, h, w, L, tiles);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

#define RC_RADIUS_SWITCH(r, CALL)          \
  switch (r) {                             \
    case 1: return CALL(1);                \
    case 2: return CALL(2);                \
    case 3: return CALL(3);                \
    case 4: return CALL(4);                \
    case 5: return CALL(5);                \
    case 6: return CALL(6);                \
    case 7: return CALL(7);                \
    case 8: return CALL(8);                \
    default:                               \
      set_error("radius %d unsupported (1..%d)", r, RC_MAX_RADIUS); \
      return RC_ERR_UNSUPPORTED;           \
  }

int launch_lookup_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h,
                      int w, int L, int r, cudaStream_t st) {
#define CALL(R) launch_fwd_r<R>(lv, coords, out, B, h, w, L, st)
  RC_RADIUS_SWITCH(r, CALL)
#undef CALL
}

int launch_lookup_bwd(const float *gout, const float *coords, const LevelPtrs &glv, int B, int h,
                      int w, int L, int r, cudaStream_t st) {
#define CALL(R) launch_bwd_r<R>(gout, coords, glv, B, h, w, L, st)
  RC_RADIUS_SWITCH(r, CALL)
#undef CALL
}

}  // namespace rc
