// Generic pixel-coordinate bilinear sampler: the reference's bilinear_sampler
// (core/utils/utils.py:57-71 = F.grid_sample(img, grid, align_corners=True), bilinear,
// zeros padding) without the normalise / un-normalise round trip.
//   img (M, C, H, W), coords (M, Ho, Wo, 2) [x, y] in pixels -> out (M, C, Ho, Wo)
//   optional mask (M, Ho, Wo, 1): the reference's rule, evaluated as the reference evaluates it
//   (utils.py:61-62,67-69): xg = 2*x/(W-1) - 1 in fp32 (one rounding per operation, IEEE division),
//   mask = -1 < xg < 1 and -1 < yg < 1.  It differs from "0 < x < W-1" only within an ulp of the two
//   borders, where the normalised coordinate rounds onto +-1.
#include "common.cuh"

namespace rc {

__global__ void bilinear_fwd_kernel(const float *__restrict__ img, const float *__restrict__ coords,
                                    float *__restrict__ out, float *__restrict__ mask, long long sites,
                                    int C, int H, int W, int HoWo) {
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < sites;
       s += (long long)gridDim.x * blockDim.x) {
    const long long m = s / HoWo;
    const int o = (int)(s - m * HoWo);
    const float2 c = __ldg(reinterpret_cast<const float2 *>(coords) + s);
    float dx, dy;
    const int x0 = safe_floor_int(c.x, &dx), y0 = safe_floor_int(c.y, &dy);
    const bool xa = x0 >= 0 && x0 < W, xb = x0 + 1 >= 0 && x0 + 1 < W;
    const bool ya = y0 >= 0 && y0 < H, yb = y0 + 1 >= 0 && y0 + 1 < H;
    const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx;
    const float w10 = dy * (1.0f - dx), w11 = dy * dx;
    if (mask) {
      const float xg = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, c.x), (float)(W - 1)), 1.0f);
      const float yg = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, c.y), (float)(H - 1)), 1.0f);
      mask[s] = (xg > -1.0f && yg > -1.0f && xg < 1.0f && yg < 1.0f) ? 1.0f : 0.0f;
    }
    for (int ch = 0; ch < C; ++ch) {
      const float *im = img + ((size_t)m * C + ch) * ((size_t)H * W);
      float v = 0.0f;
      if (ya && xa) v = __ldg(im + (size_t)y0 * W + x0) * w00;
      if (ya && xb) v = fmaf(__ldg(im + (size_t)y0 * W + x0 + 1), w01, v);
      if (yb && xa) v = fmaf(__ldg(im + (size_t)(y0 + 1) * W + x0), w10, v);
      if (yb && xb) v = fmaf(__ldg(im + (size_t)(y0 + 1) * W + x0 + 1), w11, v);
      out[((size_t)m * C + ch) * HoWo + o] = v;
    }
  }
}

// d(out)/d(img): scatter with float atomics into a caller-zeroed gimg (as ATen's
// grid_sampler backward does); coords receive no gradient here.
__global__ void bilinear_bwd_kernel(const float *__restrict__ gout, const float *__restrict__ coords,
                                    float *__restrict__ gimg, long long sites, int C, int H, int W,
                                    int HoWo) {
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < sites;
       s += (long long)gridDim.x * blockDim.x) {
    const long long m = s / HoWo;
    const int o = (int)(s - m * HoWo);
    const float2 c = __ldg(reinterpret_cast<const float2 *>(coords) + s);
    float dx, dy;
    const int x0 = safe_floor_int(c.x, &dx), y0 = safe_floor_int(c.y, &dy);
    const bool xa = x0 >= 0 && x0 < W, xb = x0 + 1 >= 0 && x0 + 1 < W;
    const bool ya = y0 >= 0 && y0 < H, yb = y0 + 1 >= 0 && y0 + 1 < H;
    const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx;
    const float w10 = dy * (1.0f - dx), w11 = dy * dx;
    for (int ch = 0; ch < C; ++ch) {
      const float g = __ldg(gout + ((size_t)m * C + ch) * HoWo + o);
      float *im = gimg + ((size_t)m * C + ch) * ((size_t)H * W);
      if (ya && xa) atomicAdd(im + (size_t)y0 * W + x0, g * w00);
      if (ya && xb) atomicAdd(im + (size_t)y0 * W + x0 + 1, g * w01);
      if (yb && xa) atomicAdd(im + (size_t)(y0 + 1) * W + x0, g * w10);
      if (yb && xb) atomicAdd(im + (size_t)(y0 + 1) * W + x0 + 1, g * w11);
    }
  }
}

static int grid_for(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = 148LL * 32;
  return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

int launch_bilinear_fwd(const float *img, const float *coords, float *out, float *mask, int M, int C,
                        int H, int W, int Ho, int Wo, cudaStream_t st) {
  const long long sites = (long long)M * Ho * Wo;
  bilinear_fwd_kernel<<<grid_for(sites), 256, 0, st>>>(img, coords, out, mask, sites, C, H, W, Ho * Wo);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_bilinear_bwd(const float *gout, const float *coords, float *gimg, int M, int C, int H,
                        int W, int Ho, int Wo, cudaStream_t st) {
  const long long sites = (long long)M * Ho * Wo;
  bilinear_bwd_kernel<<<grid_for(sites), 256, 0, st>>>(gout, coords, gimg, sites, C, H, W, Ho * Wo);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

}  // namespace rc
