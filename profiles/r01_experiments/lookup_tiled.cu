// Pyramid lookup on the TILED volume layout (r = 4): persistent, cp.async-pipelined gather.
//
// Same result as lookup_fwd_kernel<4> (reference core/corr.py:29-50), but every row p of a level
// is stored as 4x4-float tiles (64 B each, tiles of a tile-row contiguous, the plane padded to
// multiples of 4 with zeros) instead of h_l rows of w_l floats.  A 10x10 window then lies in
// 3-4 x 3-4 tiles, i.e. in a few contiguous 192-256 B runs, instead of ten 40 B segments that each
// straddle DRAM fill granules: the lookup's DRAM read traffic (1.9x its algorithmic bytes in the
// row layout) drops to ~1.15x, and the patch arrives as ~40 aligned 16 B copies instead of 100
// scalar loads.
//   level l of row p : [hp/4][wp/4][4][4] floats, hp = ceil4(h>>l), wp = ceil4(w>>l)
//
// The row-layout kernel is latency-bound: each CTA pays coords -> patch -> taps in series and the
// time is (waves) x (that latency).  Here CTAs are persistent and every warp keeps a ring of
// patch buffers filled by cp.async (global -> shared, no registers): while it forms the taps of
// one group of 4 pixels, the copies of the next two groups and the coordinates of the next item
// are in flight.
#include "common.cuh"

namespace rc {

#ifndef RC_TL_GROUP
#define RC_TL_GROUP 4
#endif
#ifndef RC_TL_STAGES
#define RC_TL_STAGES 3
#endif
constexpr int kTlThreads = 128, kTlWarps = 4, kTlTile = 32, kTlPixPerWarp = 8, kTlGroup = RC_TL_GROUP, kTlStages = RC_TL_STAGES;
constexpr int kTlGroupsPerItem = kTlPixPerWarp / kTlGroup;
constexpr int kTlPS = 20, kTlPatch = 10 * kTlPS;   // patch: 10 rows x 16 fetched floats (+pad), floats
constexpr int kTlStageFloats = kTlGroup * kTlPatch;

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

__global__ void __launch_bounds__(kTlThreads)
lookup_tiled_fwd_kernel(ConstLevelPtrs lv, const float *__restrict__ coords, float *__restrict__ out, int B, int h, int w,
                        int L, int tiles_per_item, int total_items) {
  constexpr int R = 4, RD = 9, NTAP = 81, NT = 3, TS = kTlTile + 1, PS = kTlPS;
  extern __shared__ float smem[];
  float *tile = smem;                       // [NTAP][TS]
  float *ring_all = smem + NTAP * TS + 3;   // 16 B aligned (81*33 + 3 = 2676 floats)

  const int N = h * w;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *ring = ring_all + warp * (kTlStages * kTlStageFloats);
  const int lr = lane >> 2, lx = lane & 3;  // patch row 0..7 and x-tile 0..3 of this lane
  const int pl = lr * PS + lx * 4;          // where this lane's 16 bytes land in a patch
  int t_src[NT], t_dst[NT];
#pragma unroll
  for (int k = 0; k < NT; ++k) {
    const int t = lane + 32 * k;
    const int i = t / RD, j = t - i * RD;  // i: x offset, j: y offset
    t_src[k] = j * PS + i;
    t_dst[k] = t * TS;
  }

  // One decoded work item (32 pixels of one level) plus the per-pixel scalars of this warp's 8
  // pixels, held by lane u (< 8) for pixel u.  Decoded once, when its coordinates are fetched.
  struct Ctx {
    int lvl, b, p0, valid;
    int x0, y0;
    float fx, fy;
  };
  auto fetch = [&](int it) {
    Ctx c{0, 0, 0, 0, 0, -(1 << 28), 0.f, 0.f};  // invalid pixel: rows far above the map -> zero-filled patch
    if (it < total_items) {
      const int per_level = B * tiles_per_item;  // level-major order: same-level work is contiguous
      c.lvl = it / per_level;
      const int rem = it - c.lvl * per_level;
      c.b = rem / tiles_per_item;
      c.p0 = (rem - c.b * tiles_per_item) * kTlTile;
      c.valid = 1;
      const int p = c.p0 + warp * kTlPixPerWarp + lane;
      if (lane < kTlPixPerWarp && p < N) {
        const float inv = 1.0f / (float)(1 << c.lvl);
        const float cx = __ldg(coords + ((size_t)c.b * 2 + 0) * N + p) * inv;
        const float cy = __ldg(coords + ((size_t)c.b * 2 + 1) * N + p) * inv;
        c.x0 = safe_floor_int(cx, &c.fx) - R;
        c.y0 = safe_floor_int(cy, &c.fy) - R;
      }
    }
    return c;
  };
  // issue the copies of pixels [g, g + kTlGroup) of item c into ring stage `stage`
  auto issue = [&](const Ctx &c, int g, int stage) {
    if (c.valid) {
      const int hl = h >> c.lvl, wl = w >> c.lvl;
      const int hp = (hl + 3) & ~3, wp = (wl + 3) & ~3, twn = wp >> 2, tws = twn * 16;
      const int row_elems = hp * wp;
      const float *level0 = pick_level(lv, c.lvl);  // also the (never read) source of zero-fill copies
      const float *base = level0 + ((size_t)c.b * N + c.p0 + warp * kTlPixPerWarp) * (size_t)row_elems;
      float *dst = ring + stage * kTlStageFloats + pl;
#pragma unroll
      for (int u = 0; u < kTlGroup; ++u) {
        const int x0 = __shfl_sync(0xffffffffu, c.x0, g + u);
        const int y0 = __shfl_sync(0xffffffffu, c.y0, g + u);
        const int tx = (x0 >> 2) + lx;  // arithmetic shift = floor division
        const bool xok = (unsigned)tx < (unsigned)twn && (lx < 3 || (x0 & 3) == 3);
        const float *rowp = base + (g + u) * row_elems + tx * 16;
        const int ya = y0 + lr, yb = ya + 8;
        const bool oka = xok && (unsigned)ya < (unsigned)hl;
        const bool okb = xok && (unsigned)yb < (unsigned)hl;
        cp_async16_zfill(dst + u * kTlPatch, oka ? rowp + (ya >> 2) * tws + (ya & 3) * 4 : level0, oka);
        if (lr < 2) cp_async16_zfill(dst + u * kTlPatch + 8 * PS, okb ? rowp + (yb >> 2) * tws + (yb & 3) * 4 : level0, okb);
      }
    }
    cp_async_commit();  // always commit (possibly empty) so that the wait counts stay uniform
  };

  // ---- software pipeline over the sequence of pixel groups of this CTA's items: while group n is
  // consumed, groups n+1 .. n+D are in flight (D = stages - 1 <= groups per item)
  constexpr int D = kTlStages - 1, GPI = kTlGroupsPerItem;
  static_assert(D >= 1 && D <= GPI, "prefetch distance must not exceed the groups of one item");
  const int stride = gridDim.x;
  int it_cur = blockIdx.x;
  Ctx cur = fetch(it_cur);
  Ctx nxt = fetch(it_cur + stride);
#pragma unroll
  for (int d = 0; d < D; ++d) issue(cur, d * kTlGroup, d);
  int stage = 0;

  for (; it_cur < total_items; it_cur += stride) {
    const Ctx nn = fetch(it_cur + 2 * stride);  // coordinates two items ahead: a whole item of slack
#pragma unroll
    for (int gi = 0; gi < GPI; ++gi) {
      const int pre_stage = (stage + D) % kTlStages;
      if (gi + D < GPI) issue(cur, (gi + D) * kTlGroup, pre_stage);
      else issue(nxt, (gi + D - GPI) * kTlGroup, pre_stage);
      cp_async_wait<D>();
      __syncwarp();
      const float *pbase = ring + stage * kTlStageFloats;
#pragma unroll
      for (int u = 0; u < kTlGroup; ++u) {
        const int pix = warp * kTlPixPerWarp + gi * kTlGroup + u;
        const float dx = __shfl_sync(0xffffffffu, cur.fx, gi * kTlGroup + u);
        const float dy = __shfl_sync(0xffffffffu, cur.fy, gi * kTlGroup + u);
        const int off = __shfl_sync(0xffffffffu, cur.x0, gi * kTlGroup + u) & 3;
        const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx;
        const float w10 = dy * (1.0f - dx), w11 = dy * dx;
        const float *pp = pbase + u * kTlPatch + off;
#pragma unroll
        for (int k = 0; k < NT; ++k) {
          if (lane + 32 * k < NTAP) {
            const float *q = pp + t_src[k];
            float s = q[0] * w00;
            s = fmaf(q[1], w01, s);
            s = fmaf(q[PS], w10, s);
            s = fmaf(q[PS + 1], w11, s);
            tile[t_dst[k] + pix] = s;
          }
        }
      }
      __syncwarp();
      stage = (stage + 1) % kTlStages;
    }
    __syncthreads();
    {
      const int CH = L * NTAP;
      const int p = cur.p0 + lane;
      if (p < N) {
        float *o = out + ((size_t)cur.b * CH + (size_t)cur.lvl * NTAP + warp) * N + p;
        const float *src = tile + warp * TS + lane;
        const size_t ostep = (size_t)kTlWarps * N;
#pragma unroll 7
        for (int t = warp; t < NTAP; t += kTlWarps) {
          st_stream(o, *src);
          o += ostep;
          src += kTlWarps * TS;
        }
      }
    }
    __syncthreads();
    cur = nxt;
    nxt = nn;
  }
  cp_async_wait<0>();
}

int launch_lookup_tiled_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h, int w, int L, int r,
                            cudaStream_t st) {
  if (r != 4) {
    set_error("rc_corr_lookup_tiled_fwd: radius %d unsupported on the tiled layout (4 only)", r);
    return RC_ERR_UNSUPPORTED;
  }
  const int N = h * w;
  const int tiles = (N + kTlTile - 1) / kTlTile;
  const long long total = (long long)L * B * tiles;
  if (total > 0x7fffffffLL) {
    set_error("rc_corr_lookup_tiled_fwd: problem too large");
    return RC_ERR_UNSUPPORTED;
  }
  const size_t smem = sizeof(float) * (81 * (kTlTile + 1) + 3 + kTlWarps * kTlStages * kTlStageFloats);
  static int ctas_per_sm = 0, sms = 0;
  if (ctas_per_sm == 0) {
    RC_CUDA(cudaFuncSetAttribute(lookup_tiled_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, n = 0;
    RC_CUDA(cudaGetDevice(&dev));
    RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    RC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, lookup_tiled_fwd_kernel, kTlThreads, smem));
    ctas_per_sm = n > 0 ? n : 1;
  }
  long long grid = (long long)sms * ctas_per_sm;
  if (grid > total) grid = total;
  lookup_tiled_fwd_kernel<<<(unsigned)grid, kTlThreads, smem, st>>>(lv, coords, out, B, h, w, L, tiles, (int)total);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

}  // namespace rc
