// The two small pieces of the per-iteration update that are neither correlation nor a tensor-core convolution
// (reference core/update.py:86,92 and core/raft.py:123-132), each one launch instead of a cuDNN convolution wrapped in
// two layout conversions, or three element-wise kernels on strided views:
//   small_conv_relu_kernel : convf1, the K x K (7 x 7) convolution of the 2-channel flow field + bias + ReLU
//   flow_update_kernel     : coords1 += delta_flow; flow = coords1 - coords0 (written where the next iteration's
//                            motion encoder and GRU read it)
#include "common.cuh"

namespace rc {
namespace glue {

constexpr int kMaxTaps = 4 * 7 * 7;  // Cin <= 4, K <= 7
constexpr int OG = 16;               // output channels per thread

// out[b, o, y, x] = relu(bias[o] + sum_{c, ky, kx} w[o, c, ky, kx] * in[b, c, y + ky - K/2, x + kx - K/2])
// A block owns 128 pixels of one row and OG filters: the input rows with their halo (zero outside the image) and the
// filters ([tap][OG], read as broadcast float4) sit in shared memory; a thread owns one pixel.
constexpr int TWS = 128;
__global__ void __launch_bounds__(TWS)
small_conv_relu_kernel(const float *__restrict__ in, const float *__restrict__ w, const float *__restrict__ bias,
                       float *__restrict__ out, int Cin, int Cout, int H, int W, int K, int relu, int tiles_x) {
  __shared__ __align__(16) float ws[kMaxTaps * OG];
  __shared__ float bs[OG];
  __shared__ float tile[4 * 7 * (TWS + 6)];
  const int taps = Cin * K * K, og = blockIdx.y, b = blockIdx.z, r = K >> 1, RWS = TWS + K - 1;
  const int y = blockIdx.x / tiles_x, x0 = (blockIdx.x - y * tiles_x) * TWS;
  pdl_launch_dependents();
  for (int i = threadIdx.x; i < taps * OG; i += TWS) {  // coalesced read of OG filters, transposed into [tap][OG]
    const int o = i / taps, t = i - o * taps;
    ws[t * OG + o] = og * OG + o < Cout ? __ldg(w + (size_t)(og * OG + o) * taps + t) : 0.0f;
  }
  if (threadIdx.x < OG) bs[threadIdx.x] = (bias && og * OG + threadIdx.x < Cout) ? __ldg(bias + og * OG + threadIdx.x) : 0.0f;
  pdl_wait();
  const float *src = in + (size_t)b * Cin * H * W;
  for (int i = threadIdx.x; i < Cin * K * RWS; i += TWS) {
    const int row = i / RWS, col = i - row * RWS, c = row / K, ky = row - c * K;
    const int iy = y + ky - r, ix = x0 + col - r;
    tile[i] = ((unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W) ? __ldg(src + ((size_t)c * H + iy) * W + ix) : 0.0f;
  }
  __syncthreads();
  const int x = x0 + threadIdx.x;
  float acc[OG];
#pragma unroll
  for (int i = 0; i < OG; ++i) acc[i] = bs[i];
  int t = 0;
  for (int row = 0; row < Cin * K; ++row) {
    const float *trow = tile + row * RWS + threadIdx.x;
    for (int kx = 0; kx < K; ++kx, ++t) {
      const float v = trow[kx];
      const float4 *w4 = reinterpret_cast<const float4 *>(ws + t * OG);
#pragma unroll
      for (int q = 0; q < OG / 4; ++q) {
        const float4 f = w4[q];
        acc[4 * q + 0] = fmaf(f.x, v, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(f.y, v, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(f.z, v, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(f.w, v, acc[4 * q + 3]);
      }
    }
  }
  if (x >= W) return;
  float *o = out + ((size_t)b * Cout + og * OG) * H * W + (size_t)y * W + x;
#pragma unroll
  for (int i = 0; i < OG; ++i)
    if (og * OG + i < Cout) o[(size_t)i * H * W] = relu ? fmaxf(acc[i], 0.0f) : acc[i];
}

// coords1 (B, 2, H, W) += delta (B, 2, H, W with its own batch stride); flow = coords1 - coords0 -> flow_out and, if
// given, flow_out2 (each with its own batch stride): core/raft.py:123,131-132 for the next iteration
__global__ void __launch_bounds__(256)
flow_update_kernel(float *__restrict__ coords1, const float *__restrict__ delta, long long delta_bs,
                   const float *__restrict__ coords0, float *__restrict__ flow_out, long long flow_bs,
                   float *__restrict__ flow_out2, long long flow2_bs, int B, int HW2) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * HW2) return;
  const int b = (int)(i / HW2), e = (int)(i - (long long)b * HW2);
  const float c1 = coords1[i] + __ldg(delta + (size_t)b * delta_bs + e);
  coords1[i] = c1;
  const float f = c1 - __ldg(coords0 + i);
  flow_out[(size_t)b * flow_bs + e] = f;
  if (flow_out2) flow_out2[(size_t)b * flow2_bs + e] = f;
}

}  // namespace glue

int launch_small_conv_relu(const float *in, const float *w, const float *bias, float *out, int B, int Cin, int Cout, int H,
                           int W, int K, int relu, cudaStream_t st) {
  using namespace glue;
  const int tiles_x = (W + TWS - 1) / TWS;
  const dim3 grid((unsigned)(H * tiles_x), (unsigned)((Cout + OG - 1) / OG), (unsigned)B);
  RC_CUDA(launch_pdl(small_conv_relu_kernel, grid, dim3(TWS), 0, st, in, w, bias, out, Cin, Cout, H, W, K, relu, tiles_x));
  return RC_OK;
}

int launch_flow_update(float *coords1, const float *delta, long long delta_bs, const float *coords0, float *flow_out,
                       long long flow_bs, float *flow_out2, long long flow2_bs, int B, int H, int W, cudaStream_t st) {
  const long long n = (long long)B * 2 * H * W;
  RC_CUDA(launch_pdl(glue::flow_update_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, coords1, delta, delta_bs,
                     coords0, flow_out, flow_bs, flow_out2, flow2_bs, B, 2 * H * W));
  return RC_OK;
}

}  // namespace rc
