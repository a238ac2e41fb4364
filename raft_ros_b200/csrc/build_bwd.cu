// Gradient of the build (volume + pyramid) w.r.t. the two feature maps.
//
// Restates what autograd does through the reference's CorrBlock.__init__/corr
// (core/corr.py:24-27 avg_pool2d backward, :60 the 1/sqrt(D) scale, :58 matmul backward):
//   dC[p,q]     = scale * ( g0[p,q] + g1[p,q>>1]/4 + g2[p,q>>2]/16 + ... )   (floor-mode pooling:
//                 cells outside the pooled region receive nothing from coarser levels)
//   dfmap1[d,p] = sum_q dC[p,q] * fmap2[d,q]
//   dfmap2[d,q] = sum_p dC[p,q] * fmap1[d,p]
// Step 1 is one fused pass (grad pyramid read once, dC written once into the workspace);
// steps 2-3 are fp32 contractions on CUDA cores (fp32 like the reference's SGEMMs).
#include "common.cuh"

namespace rc {

__global__ void fold_grad_pyramid_kernel(ConstLevelPtrs glv, float *__restrict__ dC, size_t rows,
                                         int h, int w, int L, float scale) {
  const int N = h * w;
  const size_t total = rows * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t rp = i / N;
    const int q = (int)(i - rp * N);
    int y = q / w, x = q - y * w;
    float g = __ldg(glv.p[0] + i);
    float wgt = 1.0f;
    for (int l = 1; l < L; ++l) {
      const int hl = h >> l, wl = w >> l;
      y >>= 1;
      x >>= 1;
      wgt *= 0.25f;
      if (y >= hl || x >= wl) break;
      g = fmaf(wgt, __ldg(pick_level(glv, l) + rp * ((size_t)hl * wl) + (size_t)y * wl + x), g);
    }
    dC[i] = g * scale;
  }
}

// out[m][n] = sum_k A[m][k] * Bop[k][n], A row-major (lda), out row-major (ldc).
// B_KMAJOR: Bop[k][n] = Bm[n*ldb + k]  (rows of Bm are K-contiguous), else Bm[k*ldb + n].
template <bool B_KMAJOR>
__global__ void __launch_bounds__(256)
sgemm_kernel(const float *__restrict__ A, const float *__restrict__ Bm, float *__restrict__ Cm, int M,
             int Nn, int K, int lda, int ldb, int ldc, size_t strideA, size_t strideB, size_t strideC) {
  constexpr int TM = 64, TN = 64, TK = 16;
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  A += blockIdx.z * strideA;
  Bm += blockIdx.z * strideB;
  Cm += blockIdx.z * strideC;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += TK) {
    for (int e = tid; e < TM * TK; e += 256) {
      const int m = e / TK, kk = e - m * TK;  // K fastest: A rows are K-contiguous
      As[kk][m] = (m0 + m < M && k0 + kk < K) ? __ldg(A + (size_t)(m0 + m) * lda + k0 + kk) : 0.0f;
    }
    for (int e = tid; e < TN * TK; e += 256) {
      if (B_KMAJOR) {
        const int n = e / TK, kk = e - n * TK;
        Bs[kk][n] = (n0 + n < Nn && k0 + kk < K) ? __ldg(Bm + (size_t)(n0 + n) * ldb + k0 + kk) : 0.0f;
      } else {
        const int kk = e / TN, n = e - kk * TN;
        Bs[kk][n] = (n0 + n < Nn && k0 + kk < K) ? __ldg(Bm + (size_t)(k0 + kk) * ldb + n0 + n) : 0.0f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < Nn) Cm[(size_t)m * ldc + n] = acc[i][j];
    }
}

size_t build_bwd_workspace_bytes(int B, int D, int h, int w) {
  (void)D;
  const size_t N = (size_t)h * w;
  return sizeof(float) * (size_t)B * N * N;
}

int launch_build_bwd(const ConstLevelPtrs &glv, const float *f1, const float *f2, float *df1,
                     float *df2, int B, int D, int h, int w, int L, float scale, void *ws,
                     size_t ws_bytes, cudaStream_t st) {
  const int N = h * w;
  if (ws == nullptr || ws_bytes < build_bwd_workspace_bytes(B, D, h, w)) {
    set_error("rc_corr_build_bwd: workspace too small (%zu < %zu bytes)", ws_bytes,
              build_bwd_workspace_bytes(B, D, h, w));
    return RC_ERR_WORKSPACE;
  }
  float *dC = static_cast<float *>(ws);
  const size_t rows = (size_t)B * N;
  const size_t total = rows * N;
  const int fb = (int)((total + 255) / 256 < (size_t)148 * 32 ? (total + 255) / 256 : (size_t)148 * 32);
  fold_grad_pyramid_kernel<<<fb, 256, 0, st>>>(glv, dC, rows, h, w, L, scale);
  RC_CUDA(cudaGetLastError());
  // dfmap1[b] (D x N) = fmap2[b] (D x Nq) . dC[b]^T   -> A = fmap2, Bop[k=q][n=p] = dC[p][q] (K-major)
  {
    dim3 grid((unsigned)((N + 63) / 64), (unsigned)((D + 63) / 64), (unsigned)B);
    sgemm_kernel<true><<<grid, 256, 0, st>>>(f2, dC, df1, D, N, N, N, N, N, (size_t)D * N, (size_t)N * N, (size_t)D * N);
    RC_CUDA(cudaGetLastError());
  }
  // dfmap2[b] (D x N) = fmap1[b] (D x Np) . dC[b]    -> A = fmap1, Bop[k=p][n=q] = dC[p][q]
  {
    dim3 grid((unsigned)((N + 63) / 64), (unsigned)((D + 63) / 64), (unsigned)B);
    sgemm_kernel<false><<<grid, 256, 0, st>>>(f1, dC, df2, D, N, N, N, N, N, (size_t)D * N, (size_t)N * N, (size_t)D * N);
    RC_CUDA(cudaGetLastError());
  }
  return RC_OK;
}

}  // namespace rc
