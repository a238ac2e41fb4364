// On-the-fly windowed correlation (no materialised volume): alt_cuda_corr.forward/backward.
//
// Replaces the reference's only native code, alt_cuda_corr/correlation_kernel.cu:
// corr_forward_kernel (:18-119) and corr_backward_kernel (:122-256).  Results match it; the
// structure does not: the reference keeps its accumulator in global memory (4 read-modify-
// writes per grid point per 32-channel chunk into a pre-zeroed output) and relies on implicit
// warp lock-step.  Here every (2r+2)^2 dot product is reduced in registers, the (2r+1)^2
// bilinear taps are formed in shared memory and each output element is written exactly once,
// coalesced, so the output needs no zero fill.
//
// Forward work split: one CTA = 32 consecutive query pixels of one (b, n); a warp owns a pixel
// at a time, lanes <-> channels (float4, coalesced 512 B rows of fmap2), butterfly reduce.
#include "common.cuh"

namespace rc {

constexpr int kAltThreads = 128;
constexpr int kAltWarps = kAltThreads / 32;
constexpr int kAltTile = 32;
constexpr int kAltPixPerWarp = kAltTile / kAltWarps;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int R>
__global__ void __launch_bounds__(kAltThreads)
alt_fwd_kernel(const float *__restrict__ f1, const float *__restrict__ f2,
               const float *__restrict__ coords, float *__restrict__ corr, int Nn, int H1, int W1,
               int H2, int W2, int C, int tiles_per_item) {
  constexpr int RD = 2 * R + 1, PN = RD + 1, NPATCH = PN * PN, NTAP = RD * RD;
  constexpr int NT = (NTAP + 31) / 32;
  constexpr int TS = kAltTile + 1;
  extern __shared__ float smem[];
  float *tile = smem;                       // [NTAP][TS]
  float *dots_all = smem + NTAP * TS;       // [warps][NPATCH]

  const int HW1 = H1 * W1;
  const int bn = blockIdx.x / tiles_per_item;  // b*Nn + n
  const int b = bn / Nn;
  const int p0 = (blockIdx.x - bn * tiles_per_item) * kAltTile;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *dots = dots_all + warp * NPATCH;
  const int C4 = C >> 2;

#pragma unroll 1
  for (int u = 0; u < kAltPixPerWarp; ++u) {
    const int pix = warp * kAltPixPerWarp + u;
    const int p = p0 + pix;
    if (p >= HW1) break;  // warp-uniform
    const float2 cxy = __ldg(reinterpret_cast<const float2 *>(coords) + (size_t)bn * HW1 + p);
    float dx, dy;
    const int x0 = safe_floor_int(cxy.x, &dx) - R;
    const int y0 = safe_floor_int(cxy.y, &dy) - R;
    const float4 *a4 = reinterpret_cast<const float4 *>(f1 + ((size_t)b * HW1 + p) * C);
    const float4 *b4 = reinterpret_cast<const float4 *>(f2 + (size_t)b * H2 * W2 * C);
#pragma unroll 1
    for (int e = 0; e < NPATCH; ++e) {
      const int py = e / PN, px = e - py * PN;
      const int gy = y0 + py, gx = x0 + px;
      float s = 0.0f;
      if (gy >= 0 && gy < H2 && gx >= 0 && gx < W2) {  // warp-uniform
        const float4 *q4 = b4 + ((size_t)gy * W2 + gx) * C4;
        for (int c = lane; c < C4; c += 32) {
          const float4 av = __ldg(a4 + c), bv = __ldg(q4 + c);
          s = fmaf(av.x, bv.x, s);
          s = fmaf(av.y, bv.y, s);
          s = fmaf(av.z, bv.z, s);
          s = fmaf(av.w, bv.w, s);
        }
        s = warp_sum(s);
      }
      if (lane == 0) dots[e] = s;
    }
    __syncwarp();
    const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx;
    const float w10 = dy * (1.0f - dx), w11 = dy * dx;
#pragma unroll
    for (int k = 0; k < NT; ++k) {
      const int t = lane + 32 * k;
      if (t < NTAP) {
        const int i = t / RD, j = t - i * RD;  // t = oy + RD*ox  ->  i = ox, j = oy
        const float *q = dots + j * PN + i;
        float s = q[0] * w00;
        s = fmaf(q[1], w01, s);
        s = fmaf(q[PN], w10, s);
        s = fmaf(q[PN + 1], w11, s);
        tile[t * TS + pix] = s;
      }
    }
    __syncwarp();
  }
  __syncthreads();
  const int p = p0 + lane;
  if (p < HW1) {
    float *o = corr + (size_t)bn * NTAP * HW1 + p;
    for (int t = warp; t < NTAP; t += kAltWarps) o[(size_t)t * HW1] = tile[t * TS + lane];
  }
}

// ---- deterministic fmap2_grad: fixed-point scale and helper kernels
// maxima[0] = max |corr_grad|, maxima[1] = max |fmap1| (as float bit patterns, non-negative: integer max = float max).
// One contribution is g * f1 with |g| <= max|corr_grad| (a convex combination of taps), at most `count` of them meet
// in one accumulator: scale = 2^floor(61 - log2(count * bound)) keeps every partial sum below 2^62.
__device__ __forceinline__ double det_scale(const float *maxima, double count) {
  const double bound = (double)maxima[0] * (double)maxima[1] * count;
  if (!(bound > 0.0) || !isfinite(bound)) return 1.0;
  int e;
  frexp(bound, &e);  // bound < 2^e
  return ldexp(1.0, 61 - e);
}

__global__ void alt_absmax_kernel(const float *__restrict__ a, long long na, const float *__restrict__ b, long long nb,
                                  unsigned int *__restrict__ maxima) {
  float ma = 0.0f, mb = 0.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < na; i += (long long)gridDim.x * blockDim.x)
    ma = fmaxf(ma, fabsf(__ldg(a + i)));
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += (long long)gridDim.x * blockDim.x)
    mb = fmaxf(mb, fabsf(__ldg(b + i)));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ma = fmaxf(ma, __shfl_xor_sync(0xffffffffu, ma, o));
    mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(maxima + 0, __float_as_uint(ma));  // non-negative floats order like their bit patterns
    atomicMax(maxima + 1, __float_as_uint(mb));
  }
}

__global__ void alt_det_finish_kernel(const long long *__restrict__ acc, float *__restrict__ g2, long long n,
                                      const float *__restrict__ maxima, double count) {
  const double inv = 1.0 / det_scale(maxima, count);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    g2[i] = (float)((double)acc[i] * inv);
}

// Backward: a warp owns one query pixel (all N coordinate sets), lanes <-> channels.
// fmap1_grad is accumulated in registers and written once; fmap2_grad is a scatter into a
// caller-zeroed buffer (float atomics, as in the reference :237).
// DET: fmap2_grad is accumulated as 64-bit FIXED-POINT integers (integer addition is associative, so the
// result does not depend on the order in which the atomics land: bit-reproducible) into `acc`, scaled by a
// power of two chosen from the data so that no sum can overflow; alt_det_finish_kernel converts back.
template <int R, bool DET>
__global__ void __launch_bounds__(kAltThreads)
alt_bwd_kernel(const float *__restrict__ f1, const float *__restrict__ f2,
               const float *__restrict__ coords, const float *__restrict__ cg,
               float *__restrict__ g1, float *__restrict__ g2, int Nn, int H1, int W1, int H2,
               int W2, int C, int B, unsigned long long *__restrict__ acc64 = nullptr,
               const float *__restrict__ maxima = nullptr) {
  constexpr int RD = 2 * R + 1, PN = RD + 1, NPATCH = PN * PN, NTAP = RD * RD;
  constexpr int NE = (NPATCH + 31) / 32;
  __shared__ float gs_all[kAltWarps][NPATCH];
  const int HW1 = H1 * W1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long gp = (long long)blockIdx.x * kAltWarps + warp;  // b*HW1 + p
  if (gp >= (long long)B * HW1) return;
  const int b = (int)(gp / HW1), p = (int)(gp - (long long)b * HW1);
  float *gs = gs_all[warp];
  const int C4 = C >> 2;
  const float4 *a4 = reinterpret_cast<const float4 *>(f1 + ((size_t)b * HW1 + p) * C);
  const float4 *b4 = reinterpret_cast<const float4 *>(f2 + (size_t)b * H2 * W2 * C);
  float *g2b = g2 + (size_t)b * H2 * W2 * C;
  unsigned long long *accb = DET ? acc64 + (size_t)b * H2 * W2 * C : nullptr;
  const double fx_scale = DET ? det_scale(maxima, (double)Nn * HW1) : 0.0;

  for (int c0 = 0; c0 < C4; c0 += 32) {
    const int c = c0 + lane;
    const bool cok = c < C4;
    const float4 av = cok ? __ldg(a4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int n = 0; n < Nn; ++n) {
      const size_t bn = (size_t)b * Nn + n;
      const float2 cxy = __ldg(reinterpret_cast<const float2 *>(coords) + bn * HW1 + p);
      float dx, dy;
      const int x0 = safe_floor_int(cxy.x, &dx) - R;
      const int y0 = safe_floor_int(cxy.y, &dy) - R;
      const float *gptr = cg + bn * NTAP * HW1 + p;
      __syncwarp();
#pragma unroll
      for (int k = 0; k < NE; ++k) {
        const int e = lane + 32 * k;
        if (e < NPATCH) {
          const int iy = e / PN, ix = e - iy * PN;
          float g = 0.0f;
          if (iy > 0 && ix > 0) g += __ldg(gptr + (size_t)((iy - 1) + RD * (ix - 1)) * HW1) * dy * dx;
          if (iy > 0 && ix < RD) g += __ldg(gptr + (size_t)((iy - 1) + RD * ix) * HW1) * dy * (1.0f - dx);
          if (iy < RD && ix > 0) g += __ldg(gptr + (size_t)(iy + RD * (ix - 1)) * HW1) * (1.0f - dy) * dx;
          if (iy < RD && ix < RD) g += __ldg(gptr + (size_t)(iy + RD * ix) * HW1) * (1.0f - dy) * (1.0f - dx);
          gs[e] = g;
        }
      }
      __syncwarp();
#pragma unroll 1
      for (int e = 0; e < NPATCH; ++e) {
        const int iy = e / PN, ix = e - iy * PN;
        const int gy = y0 + iy, gx = x0 + ix;
        if (gy < 0 || gy >= H2 || gx < 0 || gx >= W2) continue;  // warp-uniform
        const float g = gs[e];
        if (cok) {
          const size_t off = ((size_t)gy * W2 + gx) * C4 + c;
          const float4 bv = __ldg(b4 + off);
          acc.x = fmaf(g, bv.x, acc.x);
          acc.y = fmaf(g, bv.y, acc.y);
          acc.z = fmaf(g, bv.z, acc.z);
          acc.w = fmaf(g, bv.w, acc.w);
          if constexpr (DET) {
            unsigned long long *dst = accb + off * 4;
            atomicAdd(dst + 0, (unsigned long long)__double2ll_rn((double)(g * av.x) * fx_scale));
            atomicAdd(dst + 1, (unsigned long long)__double2ll_rn((double)(g * av.y) * fx_scale));
            atomicAdd(dst + 2, (unsigned long long)__double2ll_rn((double)(g * av.z) * fx_scale));
            atomicAdd(dst + 3, (unsigned long long)__double2ll_rn((double)(g * av.w) * fx_scale));
          } else {
            float *dst = g2b + off * 4;
            atomicAdd(dst + 0, g * av.x);
            atomicAdd(dst + 1, g * av.y);
            atomicAdd(dst + 2, g * av.z);
            atomicAdd(dst + 3, g * av.w);
          }
        }
      }
    }
    if (cok) reinterpret_cast<float4 *>(g1 + ((size_t)b * HW1 + p) * C)[c] = acc;
  }
}

template <int R>
static int alt_fwd_r(const float *f1, const float *f2, const float *coords, float *corr, int B,
                     int Nn, int H1, int W1, int H2, int W2, int C, cudaStream_t st) {
  constexpr int RD = 2 * R + 1, PN = RD + 1;
  const int HW1 = H1 * W1;
  const int tiles = (HW1 + kAltTile - 1) / kAltTile;
  const size_t smem = sizeof(float) * (RD * RD * (kAltTile + 1) + kAltWarps * PN * PN);
  if (smem > 48 * 1024)
    RC_CUDA(cudaFuncSetAttribute(alt_fwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long blocks = (long long)B * Nn * tiles;
  if (blocks > 2147483647LL) { set_error("alt_corr_fwd: grid too large"); return RC_ERR_UNSUPPORTED; }
  alt_fwd_kernel<R><<<(unsigned)blocks, kAltThreads, smem, st>>>(f1, f2, coords, corr, Nn, H1, W1, H2, W2, C, tiles);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

template <int R>
static int alt_bwd_r(const float *f1, const float *f2, const float *coords, const float *cg,
                     float *g1, float *g2, int B, int Nn, int H1, int W1, int H2, int W2, int C,
                     cudaStream_t st) {
  const long long pixels = (long long)B * H1 * W1;
  const long long blocks = (pixels + kAltWarps - 1) / kAltWarps;
  if (blocks > 2147483647LL) { set_error("alt_corr_bwd: grid too large"); return RC_ERR_UNSUPPORTED; }
  alt_bwd_kernel<R, false><<<(unsigned)blocks, kAltThreads, 0, st>>>(f1, f2, coords, cg, g1, g2, Nn, H1, W1, H2, W2, C, B);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

template <int R>
static int alt_bwd_det_r(const float *f1, const float *f2, const float *coords, const float *cg, float *g1, float *g2, int B,
                         int Nn, int H1, int W1, int H2, int W2, int C, void *ws, cudaStream_t st) {
  constexpr int RD = 2 * R + 1;
  const long long pixels = (long long)B * H1 * W1;
  const long long blocks = (pixels + kAltWarps - 1) / kAltWarps;
  if (blocks > 2147483647LL) { set_error("alt_corr_bwd: grid too large"); return RC_ERR_UNSUPPORTED; }
  const long long n2 = (long long)B * H2 * W2 * C;
  unsigned int *maxima = static_cast<unsigned int *>(ws);               // 2 words (in a 256 B header)
  unsigned long long *acc = reinterpret_cast<unsigned long long *>(static_cast<unsigned char *>(ws) + 256);
  RC_CUDA(cudaMemsetAsync(ws, 0, 256 + (size_t)n2 * 8, st));
  alt_absmax_kernel<<<148 * 8, 256, 0, st>>>(cg, (long long)B * Nn * RD * RD * H1 * W1, f1, pixels * C, maxima);
  RC_CUDA(cudaGetLastError());
  alt_bwd_kernel<R, true><<<(unsigned)blocks, kAltThreads, 0, st>>>(f1, f2, coords, cg, g1, g2, Nn, H1, W1, H2, W2, C, B, acc,
                                                                  reinterpret_cast<const float *>(maxima));
  RC_CUDA(cudaGetLastError());
  alt_det_finish_kernel<<<148 * 8, 256, 0, st>>>(reinterpret_cast<const long long *>(acc), g2, n2,
                                                reinterpret_cast<const float *>(maxima), (double)Nn * H1 * W1);
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

int launch_alt_fwd(const float *f1, const float *f2, const float *coords, float *corr, int B,
                   int N, int H1, int W1, int H2, int W2, int C, int r, cudaStream_t st) {
#define CALL(R) alt_fwd_r<R>(f1, f2, coords, corr, B, N, H1, W1, H2, W2, C, st)
  RC_RADIUS_SWITCH(r, CALL)
#undef CALL
}

int launch_alt_bwd(const float *f1, const float *f2, const float *coords, const float *cg,
                   float *g1, float *g2, int B, int N, int H1, int W1, int H2, int W2, int C,
                   int r, cudaStream_t st) {
#define CALL(R) alt_bwd_r<R>(f1, f2, coords, cg, g1, g2, B, N, H1, W1, H2, W2, C, st)
  RC_RADIUS_SWITCH(r, CALL)
#undef CALL
}

}  // namespace rc

namespace rc {

size_t alt_bwd_det_workspace_bytes(int B, int H2, int W2, int C) { return 256 + (size_t)B * H2 * W2 * C * 8; }

int launch_alt_bwd_det(const float *f1, const float *f2, const float *coords, const float *cg, float *g1, float *g2, int B, int N,
                       int H1, int W1, int H2, int W2, int C, int r, void *ws, cudaStream_t st) {
#define CALL(R) alt_bwd_det_r<R>(f1, f2, coords, cg, g1, g2, B, N, H1, W1, H2, W2, C, ws, st)
  RC_RADIUS_SWITCH(r, CALL)
#undef CALL
}

}  // namespace rc
