// RC_PREC_FP32 build: all-pairs volume + scale + pyramid in one pass on CUDA cores.
//
// Replaces CorrBlock.corr (reference core/corr.py:52-60) and the avg_pool2d pyramid of
// CorrBlock.__init__ (core/corr.py:21-27).  This is the exact-arithmetic path (fp32 FMA, same
// class as cuBLAS SGEMM); the tensor-core path is build_tc.cu.  It shares the tiling idea
// that makes the fusion possible: the N tile of the contraction is a SPATIAL 8x16 patch of
// fmap2, so every 2x2 / 4x4 / 8x8 pooling window lies inside one CTA tile and each pyramid
// level is written to HBM exactly once and never re-read.
//
// Tile: 64 query pixels (rows of the volume) x 128 target pixels (8 rows x 16 cols of fmap2),
// K chunks of 16, 256 threads, 4x8 accumulators per thread.
#include "common.cuh"

namespace rc {

constexpr int BM = 64;    // query pixels per CTA
constexpr int PH = 8;     // patch rows
constexpr int PW = 16;    // patch cols
constexpr int BN = PH * PW;
constexpr int BK = 16;
constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
build_simt_kernel(const float *__restrict__ f1, const float *__restrict__ f2, LevelPtrs lv, int D,
                  int h, int w, int L, float scale, int m_tiles, int patches_x) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  __shared__ float Cs[BM][BN + 1];

  const int N = h * w;
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * BM;
  const int py0 = (blockIdx.y / patches_x) * PH;
  const int px0 = (blockIdx.y % patches_x) * PW;
  const int tid = threadIdx.x;
  const int ty = tid >> 4;  // 0..15 -> rows ty*4..+3
  const int tx = tid & 15;  // 0..15 -> cols tx*8..+7

  const float *A = f1 + (size_t)b * D * N;
  const float *Bm = f2 + (size_t)b * D * N;

  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;

  for (int k0 = 0; k0 < D; k0 += BK) {
    // A chunk: BK x BM (contiguous along p)
    for (int e = tid; e < BK * BM; e += kThreads) {
      const int kk = e / BM, m = e - kk * BM;
      const int p = p0 + m, k = k0 + kk;
      As[kk][m] = (p < N && k < D) ? __ldg(A + (size_t)k * N + p) : 0.0f;
    }
    // B chunk: BK x (8x16 patch)
    for (int e = tid; e < BK * BN; e += kThreads) {
      const int kk = e / BN, n = e - kk * BN;
      const int y = py0 + n / PW, x = px0 + (n % PW), k = k0 + kk;
      Bs[kk][n] = (y < h && x < w && k < D) ? __ldg(Bm + (size_t)k * N + (size_t)y * w + x) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], bb[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 8; ++j) bb[j] = Bs[kk][tx * 8 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) Cs[ty * 4 + i][tx * 8 + j] = acc[i][j] * scale;
  __syncthreads();

  // level 0: rows of 16 contiguous floats
  for (int e = tid; e < BM * BN; e += kThreads) {
    const int m = e / BN, n = e - m * BN;
    const int p = p0 + m, y = py0 + n / PW, x = px0 + (n % PW);
    if (p < N && y < h && x < w) lv.p[0][((size_t)b * N + p) * N + (size_t)y * w + x] = Cs[m][n];
  }
  // coarser levels, pooled hierarchically in place (level l from level l-1, like the reference)
  int ph = PH, pw = PW;
  for (int l = 1; l < L; ++l) {
    const int hh = h >> l, ww = w >> l;
    const int nh = ph >> 1, nw = pw >> 1;
    if (nh == 0 || nw == 0) break;  // deeper levels are handled by the tail kernel
    float r[(BM * (PH / 2) * (PW / 2) + kThreads - 1) / kThreads];
    int cnt = 0;
    for (int e = tid; e < BM * nh * nw; e += kThreads, ++cnt) {
      const int m = e / (nh * nw), n = e - m * (nh * nw);
      const int yy = n / nw, xx = n - yy * nw;
      const float *s = &Cs[m][(2 * yy) * PW + 2 * xx];
      float sum = s[0] + s[1];
      sum += s[PW];
      sum += s[PW + 1];
      r[cnt] = sum * 0.25f;
    }
    __syncthreads();
    cnt = 0;
    for (int e = tid; e < BM * nh * nw; e += kThreads, ++cnt) {
      const int m = e / (nh * nw), n = e - m * (nh * nw);
      const int yy = n / nw, xx = n - yy * nw;
      Cs[m][yy * PW + xx] = r[cnt];
      const int p = p0 + m, y = (py0 >> l) + yy, x = (px0 >> l) + xx;
      if (p < N && y < hh && x < ww) pick_level(lv, l)[((size_t)b * N + p) * ((size_t)hh * ww) + (size_t)y * ww + x] = r[cnt];
    }
    __syncthreads();
    ph = nh;
    pw = nw;
  }
}

// Levels deeper than 3 (patch exhausted) are pooled from the previous level by a plain kernel.
__global__ void pool_tail_kernel(const float *__restrict__ src, float *__restrict__ dst, size_t rows,
                                 int hs, int ws, int hd, int wd) {
  const size_t total = rows * hd * wd;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t rp = i / ((size_t)hd * wd);
    const int n = (int)(i - rp * hd * wd);
    const int y = n / wd, x = n - y * wd;
    const float *s = src + rp * hs * ws + (size_t)(2 * y) * ws + 2 * x;
    float sum = s[0] + s[1];
    sum += s[ws];
    sum += s[ws + 1];
    dst[i] = sum * 0.25f;
  }
}

int launch_pool_tail(const LevelPtrs &lv, int B, int h, int w, int first, int L, cudaStream_t st) {
  const size_t rows = (size_t)B * h * w;
  for (int l = first; l < L; ++l) {
    const int hs = h >> (l - 1), ws = w >> (l - 1), hd = h >> l, wd = w >> l;
    const size_t total = rows * hd * wd;
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    pool_tail_kernel<<<blocks, 256, 0, st>>>(lv.p[l - 1], lv.p[l], rows, hs, ws, hd, wd);
    RC_CUDA(cudaGetLastError());
  }
  return RC_OK;
}

int launch_build_simt(const float *f1, const float *f2, int B, int D, int h, int w,
                      const LevelPtrs &lv, int L, float scale, cudaStream_t st) {
  const int N = h * w;
  const int m_tiles = (N + BM - 1) / BM;
  const int patches_x = (w + PW - 1) / PW, patches_y = (h + PH - 1) / PH;
  dim3 grid((unsigned)m_tiles, (unsigned)(patches_x * patches_y), (unsigned)B);
  build_simt_kernel<<<grid, kThreads, 0, st>>>(f1, f2, lv, D, h, w, L, scale, m_tiles, patches_x);
  RC_CUDA(cudaGetLastError());
  if (L > 4) return launch_pool_tail(lv, B, h, w, 4, L, st);
  return RC_OK;
}

}  // namespace rc
