// Convex upsampling of the flow field, RAFT.upsample_flow (reference core/raft.py:72-83), as one kernel:
//     out[n, c, 8y+i, 8x+j] = sum_k softmax_k(mask[n, k*64 + i*8 + j, y, x]) * 8 * flow[n, c, y + k/3 - 1, x + k%3 - 1]
// (zero outside the map: F.unfold(..., padding=1)).  The reference runs view / softmax / unfold / mul / sum /
// permute / reshape with three full-size temporaries; here the 576 mask values of a coarse pixel are read once
// (coalesced: lanes <-> 32 neighbouring pixels of a row), the softmax and the 9-term sums are formed in registers
// and the 8x8 block leaves through a small shared-memory transposition as whole 1 KB rows.  It is one piece of the
// per-iteration update that follows the lookup; inference only (no gradient).
#include "common.cuh"

namespace rc {

constexpr int kUpWarps = 4;

__global__ void __launch_bounds__(kUpWarps * 32)
convex_upsample_kernel(const float *__restrict__ flow, const float *__restrict__ mask, float *__restrict__ out, int N, int H,
                       int W, int xblocks, long long items) {
  __shared__ __align__(16) float stage[kUpWarps][2][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t HW = (size_t)H * W;
  pdl_launch_dependents();
  pdl_wait();
  for (long long it = (long long)blockIdx.x * kUpWarps + warp; it < items; it += (long long)gridDim.x * kUpWarps) {
    // one work item = one of the 8 output rows of a block of 32 coarse pixels (enough warps to fill the machine
    // at batch 1 too)
    const int i = (int)(it & 7);
    const long long it8 = it >> 3;
    const int xb = (int)(it8 % xblocks);
    const long long r = it8 / xblocks;
    const int y = (int)(r % H), n = (int)(r / H);
    const int x = xb * 32 + lane;
    const bool pv = x < W;
    // the 3x3 neighbourhood of 8 * flow, zero outside the map
    float f0[9], f1[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
      const bool ok = pv && yy >= 0 && yy < H && xx >= 0 && xx < W;
      f0[k] = ok ? 8.0f * __ldg(flow + ((size_t)n * 2 + 0) * HW + (size_t)yy * W + xx) : 0.0f;
      f1[k] = ok ? 8.0f * __ldg(flow + ((size_t)n * 2 + 1) * HW + (size_t)yy * W + xx) : 0.0f;
    }
    const float *mp = mask + (size_t)n * 576 * HW + (size_t)y * W + x;
    float *o0 = out + ((size_t)n * 2 + 0) * 64 * HW + (size_t)(8 * y) * (8 * W) + (size_t)xb * 256;
    float *o1 = o0 + 64 * HW;
    const int cols = min(256, 8 * (W - xb * 32));  // valid output columns of this block of a row
    {
      float a0[8], a1[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float m[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) m[k] = pv ? ld_stream(mp + (size_t)(k * 64 + i * 8 + j) * HW) : 0.0f;
        float mx = m[0];
#pragma unroll
        for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k]);
        float s = 0.0f, u0 = 0.0f, u1 = 0.0f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          const float e = __expf(m[k] - mx);  // ex2.approx: 2^-22 relative, far inside the 2e-6 bound; the kernel is issue-bound on this
          s += e;
          u0 = fmaf(e, f0[k], u0);
          u1 = fmaf(e, f1[k], u1);
        }
        const float inv = 1.0f / s;
        a0[j] = u0 * inv;
        a1[j] = u1 * inv;
      }
      // lane x holds columns 8x .. 8x+7 of output row 8y+i: through shared memory so that a store instruction
      // writes 512 contiguous bytes
      __syncwarp();
      float4 *s0 = reinterpret_cast<float4 *>(&stage[warp][0][lane * 8]), *s1 = reinterpret_cast<float4 *>(&stage[warp][1][lane * 8]);
      s0[0] = make_float4(a0[0], a0[1], a0[2], a0[3]);
      s0[1] = make_float4(a0[4], a0[5], a0[6], a0[7]);
      s1[0] = make_float4(a1[0], a1[1], a1[2], a1[3]);
      s1[1] = make_float4(a1[4], a1[5], a1[6], a1[7]);
      __syncwarp();
      const size_t rowoff = (size_t)i * (8 * W);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c4 = (h * 32 + lane) * 4;
        if (c4 < cols) {  // cols is a multiple of 8
          const float4 v0 = *reinterpret_cast<const float4 *>(&stage[warp][0][c4]);
          const float4 v1 = *reinterpret_cast<const float4 *>(&stage[warp][1][c4]);
          *reinterpret_cast<float4 *>(o0 + rowoff + c4) = v0;
          *reinterpret_cast<float4 *>(o1 + rowoff + c4) = v1;
        }
      }
    }
  }
}

int launch_convex_upsample(const float *flow, const float *mask, float *out, int N, int H, int W, cudaStream_t st) {
  const int xblocks = (W + 31) / 32;
  const long long items = (long long)N * H * xblocks * 8;
  long long blocks = (items + kUpWarps - 1) / kUpWarps;
  if (blocks > 148LL * 32) blocks = 148LL * 32;
  if (blocks < 1) blocks = 1;
  RC_CUDA(launch_pdl(convex_upsample_kernel, dim3((unsigned)blocks), dim3(kUpWarps * 32), 0, st, flow, mask, out, N, H, W, xblocks,
                     items));
  return RC_OK;
}

}  // namespace rc
