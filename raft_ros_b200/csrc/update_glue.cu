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
constexpr int OG = 8;                // filters per thread
constexpr int FG = 32;               // filters per block: one warp per OG of them
constexpr int PX = 4;                // consecutive pixels per thread
constexpr int TWS = 32 * PX;         // pixels of a row per block
constexpr int RWP = TWS + 8;         // pitch of a staged input row (halo of up to 3 + 3, rounded to a float4)

// out[b, o, y, x] = relu(bias[o] + sum_{c, ky, kx} w[o, c, ky, kx] * in[b, c, y + ky - K/2, x + kx - K/2])
// A block owns 128 pixels of one row and FG filters: the input rows with their halo (zero outside the image) and the
// filters ([tap][FG]) sit in shared memory.  A thread owns PX consecutive pixels x OG filters (32 accumulators): per
// input row it reads 12 inputs (three conflict-free float4) and, per tap, its 8 filters as two broadcast float4 --
// 32 FMAs per 2 shared-memory loads.
__global__ void __launch_bounds__(FG / OG * 32)
small_conv_relu_kernel(const float *__restrict__ in, const float *__restrict__ w, const float *__restrict__ bias,
                       float *__restrict__ out, int Cin, int Cout, int H, int W, int K, int relu, int tiles_x) {
  __shared__ __align__(16) float ws[kMaxTaps * FG];
  __shared__ float bs[FG];
  __shared__ __align__(16) float tile[4 * 7 * RWP];
  const int taps = Cin * K * K, fg = blockIdx.y, b = blockIdx.z, r = K >> 1;
  const int y = blockIdx.x / tiles_x, x0 = (blockIdx.x - y * tiles_x) * TWS;
  const int lane = threadIdx.x & 31, og = threadIdx.x >> 5;
  pdl_launch_dependents();
  // (both staging loops are unrolled so that eight loads are in flight per thread: with one at a time their latency,
  // ~50 round trips per block, was the whole kernel: 20 us for 88 MFLOP)
#pragma unroll 8
  for (int i = threadIdx.x; i < taps * FG; i += FG / OG * 32) {  // coalesced read of FG filters, transposed into [tap][FG]
    const int o = i / taps, t = i - o * taps;
    ws[t * FG + o] = fg * FG + o < Cout ? __ldg(w + (size_t)(fg * FG + o) * taps + t) : 0.0f;
  }
  if (threadIdx.x < FG) bs[threadIdx.x] = (bias && fg * FG + threadIdx.x < Cout) ? __ldg(bias + fg * FG + threadIdx.x) : 0.0f;
  pdl_wait();
  const float *src = in + (size_t)b * Cin * H * W;
#pragma unroll 8
  for (int i = threadIdx.x; i < Cin * K * RWP; i += FG / OG * 32) {  // staged column j <-> image column x0 + j - r
    const int row = i / RWP, col = i - row * RWP, c = row / K, ky = row - c * K;
    const int iy = y + ky - r, ix = x0 + col - r;
    tile[i] = ((unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W) ? __ldg(src + ((size_t)c * H + iy) * W + ix) : 0.0f;
  }
  __syncthreads();
  float acc[PX][OG];
#pragma unroll
  for (int p = 0; p < PX; ++p)
#pragma unroll
    for (int i = 0; i < OG; ++i) acc[p][i] = bs[og * OG + i];
  for (int row = 0; row < Cin * K; ++row) {
    float v[12];
    const float4 *t4 = reinterpret_cast<const float4 *>(tile + row * RWP + PX * lane);
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const float4 f = t4[q];
      v[4 * q] = f.x; v[4 * q + 1] = f.y; v[4 * q + 2] = f.z; v[4 * q + 3] = f.w;
    }
    const float *wrow = ws + (size_t)row * K * FG + og * OG;
#pragma unroll
    for (int kx = 0; kx < 7; ++kx) {
      if (kx < K) {
        const float4 *w4 = reinterpret_cast<const float4 *>(wrow + kx * FG);
#pragma unroll
        for (int q = 0; q < OG / 4; ++q) {
          const float4 f = w4[q];
#pragma unroll
          for (int p = 0; p < PX; ++p) {
            acc[p][4 * q + 0] = fmaf(f.x, v[p + kx], acc[p][4 * q + 0]);
            acc[p][4 * q + 1] = fmaf(f.y, v[p + kx], acc[p][4 * q + 1]);
            acc[p][4 * q + 2] = fmaf(f.z, v[p + kx], acc[p][4 * q + 2]);
            acc[p][4 * q + 3] = fmaf(f.w, v[p + kx], acc[p][4 * q + 3]);
          }
        }
      }
    }
  }
  const int x = x0 + PX * lane;
  if (x >= W) return;
  float *o = out + ((size_t)b * Cout + fg * FG + og * OG) * H * W + (size_t)y * W + x;
  const bool vec = x + PX <= W && (reinterpret_cast<uintptr_t>(o) & 15) == 0 && (((size_t)H * W) & 3) == 0;
#pragma unroll
  for (int i = 0; i < OG; ++i) {
    if (fg * FG + og * OG + i >= Cout) break;
    float q[PX];
#pragma unroll
    for (int p = 0; p < PX; ++p) q[p] = relu ? fmaxf(acc[p][i], 0.0f) : acc[p][i];
    float *oc = o + (size_t)i * H * W;
    if (vec) {
      *reinterpret_cast<float4 *>(oc) = make_float4(q[0], q[1], q[2], q[3]);
    } else {
#pragma unroll
      for (int p = 0; p < PX; ++p)
        if (x + p < W) oc[p] = q[p];
    }
  }
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
  const dim3 grid((unsigned)(H * tiles_x), (unsigned)((Cout + FG - 1) / FG), (unsigned)B);
  RC_CUDA(launch_pdl(small_conv_relu_kernel, grid, dim3(FG / OG * 32), 0, st, in, w, bias, out, Cin, Cout, H, W, K, relu, tiles_x));
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
