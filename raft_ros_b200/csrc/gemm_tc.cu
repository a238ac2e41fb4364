// Tensor-core backward of the build: dfmap1 = dC . fmap2^T, dfmap2 = dC^T . fmap1 (matmul's
// backward at reference core/corr.py:58), with dC folded from the gradient pyramid
// (avg_pool2d backward, core/corr.py:26, and the 1/sqrt(D) of :60).  sm_100a only.
//
// Both contractions have K = N (the number of pixels) and an M of only D = 256, so they are two
// batched "K-major x K-major" GEMMs:
//     dfmap1[b][d][p] = sum_q fmap2[b][d][q] * dC [b][p][q]        A = fmap2 (D x N),  B = dC   (N x N)
//     dfmap2[b][d][q] = sum_p fmap1[b][d][p] * dCt[b][q][p]        A = fmap1 (D x N),  B = dC^T (N x N)
// run by one persistent tcgen05 kernel (128 x 256 tiles, bf16 hi/lo split operands, three MMAs per
// K step into one fp32 TMEM accumulator = fp32-class accuracy, like RC_PREC_BF16X3 forward).
// The prepass reads the gradient pyramid ONCE and writes dC and dC^T as bf16 hi/lo (the transpose
// goes through a shared-memory tile), so neither GEMM needs an MN-major operand.
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace tcg {

using namespace tc;

constexpr int GM = 128, GN = 256, GK = 64, UK = 16;
constexpr int GA_BYTES = GM * GK * 2, GB_BYTES = GN * GK * 2;
constexpr int kStageBytes = 2 * (GA_BYTES + GB_BYTES);  // hi + lo
constexpr int kStages = 2;
constexpr int kEpiWarps = 8, kProducerWarp = 8, kMmaWarp = 9, kThreads = 320;
constexpr int kStgBytesPerWarp = 4096;
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kEpiWarps * kStgBytesPerWarp + 256;
constexpr uint32_t kIdescG = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(GN >> 3) << 17) | ((uint32_t)(GM >> 4) << 24);

// Up to two GEMMs of identical shape run as ONE persistent launch (tile list = GEMM 0's tiles, then
// GEMM 1's): the two contractions of the build's backward have 216 tiles each at 50x90, batch 6 --
// 1.46 waves of 148 CTAs apiece, 2.92 together.
struct GemmMaps {
  CUtensorMap a_hi[2], a_lo[2], b_hi[2], b_lo[2];
};
struct GemmParams {
  int M, N, K, batch;
  int k_blocks, m_tiles, n_tiles;
  long long tiles_per_gemm, total_tiles;
  float *C[2];       // [batch][M][ldc]
  long long ldc, stride_c;
  int vec;           // rows of C are 16 B aligned
};

__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ GemmMaps TM, const GemmParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_u32 = smem_u32(smem_raw);
  const uint32_t base_u32 = (raw_u32 + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (base_u32 - raw_u32);
  const uint32_t stg_u32 = base_u32 + kStages * kStageBytes;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes + kEpiWarps * kStgBytesPerWarp);
  uint64_t *full = bars, *empty = bars + kStages, *tfull = bars + 2 * kStages, *tempty = tfull + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kProducerWarp && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], kEpiWarps);
    }
    fence_barrier_init();
    prefetch_tmap(&TM.a_hi[0]);
    prefetch_tmap(&TM.a_lo[0]);
    prefetch_tmap(&TM.b_hi[0]);
    prefetch_tmap(&TM.b_lo[0]);
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int per_item = P.m_tiles * P.n_tiles;

  if (warp == kProducerWarp) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        const int gi = t >= P.tiles_per_gemm ? 1 : 0;
        const long long tt = t - gi * P.tiles_per_gemm;
        const int b = (int)(tt / per_item);
        const int rem = (int)(tt - (long long)b * per_item);
        const int mt = rem / P.n_tiles, nt = rem - mt * P.n_tiles;
        for (int kb = 0; kb < P.k_blocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t *sp = smem + stage * kStageBytes;
          mbar_expect_tx(&full[stage], kStageBytes);
          tma_load_3d(sp, &TM.a_hi[gi], &full[stage], kb * GK, mt * GM, b);
          tma_load_3d(sp + GA_BYTES, &TM.a_lo[gi], &full[stage], kb * GK, mt * GM, b);
          tma_load_3d(sp + 2 * GA_BYTES, &TM.b_hi[gi], &full[stage], kb * GK, nt * GN, b);
          tma_load_3d(sp + 2 * GA_BYTES + GB_BYTES, &TM.b_lo[gi], &full[stage], kb * GK, nt * GN, b);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    if (lane == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (long long t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * GN);
        for (int kb = 0; kb < P.k_blocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = base_u32 + stage * kStageBytes;
          const uint64_t a_hi = make_sw128_desc(sa), a_lo = make_sw128_desc(sa + GA_BYTES);
          const uint64_t b_hi = make_sw128_desc(sa + 2 * GA_BYTES), b_lo = make_sw128_desc(sa + 2 * GA_BYTES + GB_BYTES);
#pragma unroll
          for (int k = 0; k < GK / UK; ++k) {
            const uint64_t koff = (uint64_t)((k * UK * 2) >> 4);
            umma_bf16(d_tmem, a_hi + koff, b_hi + koff, kIdescG, (kb | k) != 0 ? 1u : 0u);
            umma_bf16(d_tmem, a_hi + koff, b_lo + koff, kIdescG, 1u);
            umma_bf16(d_tmem, a_lo + koff, b_hi + koff, kIdescG, 1u);
          }
          umma_commit(&empty[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // epilogue: warp = TMEM lane quarter (32 rows of C) x column half (128 columns)
    const int q = warp & 3, hf = warp >> 2;
    const uint32_t buf = stg_u32 + warp * kStgBytesPerWarp;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      const int gi = t >= P.tiles_per_gemm ? 1 : 0;
      const long long tt = t - gi * P.tiles_per_gemm;
      const int b = (int)(tt / per_item);
      const int rem = (int)(tt - (long long)b * per_item);
      const int mt = rem / P.n_tiles, nt = rem - mt * P.n_tiles;
      const int row0 = mt * GM + q * 32;
      const int rows_valid = min(32, P.M - row0);
      const int col0 = nt * GN + hf * (GN / 2);
      float *g = (gi ? P.C[1] : P.C[0]) + (size_t)b * P.stride_c + (size_t)row0 * P.ldc + col0;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * GN + hf * (GN / 2));
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld32(taddr + c * 32, v);
        tmem_ld_wait();
        if (c == 3) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[acc]);
        }
        store_rows<32>(buf, v, lane, g + c * 32, (size_t)P.ldc, rows_valid, P.N - (col0 + c * 32), P.vec != 0);
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- prepasses
// fmap (rows x cols fp32, contiguous) -> bf16 hi / lo with a padded pitch
__global__ void split_rows_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ hi,
                                  __nv_bfloat16 *__restrict__ lo, long long rows, int cols, int pitch) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    const float x = __ldg(src + i);
    const __nv_bfloat16 xh = __float2bfloat16_rn(x);
    hi[r * pitch + c] = xh;
    lo[r * pitch + c] = __float2bfloat16_rn(x - __bfloat162float(xh));
  }
}

// gradient pyramid -> dC (p-major) and dC^T (q-major), both as bf16 hi/lo with pitch `pitch`.
// One CTA = a 64 x 64 tile of dC; a thread owns 4 consecutive columns (one 16 B read of level 0,
// 8 B bf16x4 writes), the transpose goes through shared memory.
struct __align__(8) bf16x4 {
  __nv_bfloat16 v[4];
};

__device__ __forceinline__ void split4(const float (&g)[4], bf16x4 &hi, bf16x4 &lo) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    hi.v[j] = __float2bfloat16_rn(g[j]);
    lo.v[j] = __float2bfloat16_rn(g[j] - __bfloat162float(hi.v[j]));
  }
}

// LV = number of pyramid levels (compile time: the level loop is unrolled and every per-column
// quantity -- the cell of each coarser level a column falls into, whether it is inside the floor-mode
// pooled area, whether its value differs from the previous column's -- is computed once per thread,
// outside the row loop; the first version recomputed them per element and was issue-bound).
template <int LV>
__global__ void __launch_bounds__(256)
fold_split_kernel(ConstLevelPtrs glv, __nv_bfloat16 *__restrict__ c_hi, __nv_bfloat16 *__restrict__ c_lo,
                  __nv_bfloat16 *__restrict__ t_hi, __nv_bfloat16 *__restrict__ t_lo, int h, int w, int pitch,
                  float scale, int vec_in) {
  __shared__ float tile[64][65];
  const int N = h * w;
  const int b = blockIdx.z;
  const int p0 = blockIdx.y * 64, q0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const size_t item = (size_t)b * N;
  const int q = q0 + tx * 4;
  // per-column geometry of this thread's 4 columns: the same for every row it handles
  int idx[LV > 1 ? LV - 1 : 1][4];  // cell of level l+1 under column j, or -1
  {
    int yy[4], xx[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      yy[j] = (q + j) / w;
      xx[j] = (q + j) - yy[j] * w;
    }
#pragma unroll
    for (int l = 1; l < LV; ++l) {
      const int hl = h >> l, wl = w >> l;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int y = yy[j] >> l, x = xx[j] >> l;
        idx[l - 1][j] = (q + j < N && y < hl && x < wl) ? y * wl + x : -1;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 16 * i;
    float g[4] = {0.f, 0.f, 0.f, 0.f};
    if (p < N && q < N) {
      const size_t rp = item + p;
      const float *g0 = glv.p[0] + rp * N + q;
      if (vec_in && q + 3 < N) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(g0));
        g[0] = v.x; g[1] = v.y; g[2] = v.z; g[3] = v.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (q + j < N) g[j] = __ldg(g0 + j);
      }
      float wgt = 1.0f;
#pragma unroll
      for (int l = 1; l < LV; ++l) {
        wgt *= 0.25f;
        const float *gl = glv.p[l] + rp * (size_t)((h >> l) * (w >> l));
        float v = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int id = idx[l - 1][j];
          if (id >= 0) {
            if (j == 0 || id != idx[l - 1][j - 1]) v = __ldg(gl + id);  // neighbours mostly share a cell
            g[j] = fmaf(wgt, v, g[j]);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] *= scale;
      bf16x4 hi, lo;
      split4(g, hi, lo);
      const size_t o = (item + p) * (size_t)pitch + q;  // 8 B aligned: pitch and q0 are multiples of 64
      *reinterpret_cast<bf16x4 *>(c_hi + o) = hi;
      *reinterpret_cast<bf16x4 *>(c_lo + o) = lo;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) tile[ty + 16 * i][tx * 4 + j] = g[j];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int qq = q0 + ty + 16 * i, pp = p0 + tx * 4;
    if (qq < N && pp < N) {
      float g[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] = tile[tx * 4 + j][ty + 16 * i];
      bf16x4 hi, lo;
      split4(g, hi, lo);
      const size_t o = (item + qq) * (size_t)pitch + pp;
      *reinterpret_cast<bf16x4 *>(t_hi + o) = hi;
      *reinterpret_cast<bf16x4 *>(t_lo + o) = lo;
    }
  }
}

// ---------------------------------------------------------------- host
static int map3(CUtensorMap *tm, void *base, cuuint64_t k, cuuint64_t rows, cuuint64_t batch, cuuint64_t pitch_elems,
                cuuint64_t batch_stride_elems, cuuint32_t box_rows) {
  const unsigned long long dims[3] = {k, rows, batch};
  const unsigned long long strides[2] = {pitch_elems * 2, batch_stride_elems * 2};
  const unsigned int box[3] = {GK, box_rows, 1};
  return encode_tmap_sw128(tm, base, 3, dims, strides, box, "rc_corr_build_bwd");
}

static int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  return n;
}

static inline size_t al256(size_t x) { return (x + 255) / 256 * 256; }
static inline int pitch_of(int n) { return (n + 63) / 64 * 64; }

}  // namespace tcg

size_t build_bwd_tc_workspace_bytes(int B, int D, int h, int w) {
  const size_t N = (size_t)h * w, pitch = (size_t)tcg::pitch_of((int)N);
  const size_t big = tcg::al256((size_t)B * N * pitch * 2);    // one of dC hi/lo, dC^T hi/lo
  const size_t small = tcg::al256((size_t)B * D * pitch * 2);  // one of fmap1/2 hi/lo
  return 4 * big + 4 * small;
}

// out_g[b][M=D][N] = sum_k A_g[b][D][k] * Bm_g[b][N][k] for g = 0, 1, one launch
struct GemmOperands {
  __nv_bfloat16 *a_hi, *a_lo, *b_hi, *b_lo;
  float *out;
};
static int run_gemms(const GemmOperands (&op)[2], int B, int D, int N, int pitch, cudaStream_t st) {
  using namespace tcg;
  GemmMaps M;
  for (int g = 0; g < 2; ++g) {
    if (int rc = map3(&M.a_hi[g], op[g].a_hi, N, D, B, pitch, (cuuint64_t)D * pitch, GM)) return rc;
    if (int rc = map3(&M.a_lo[g], op[g].a_lo, N, D, B, pitch, (cuuint64_t)D * pitch, GM)) return rc;
    if (int rc = map3(&M.b_hi[g], op[g].b_hi, N, N, B, pitch, (cuuint64_t)N * pitch, GN)) return rc;
    if (int rc = map3(&M.b_lo[g], op[g].b_lo, N, N, B, pitch, (cuuint64_t)N * pitch, GN)) return rc;
  }
  GemmParams P{};
  P.M = D; P.N = N; P.K = N; P.batch = B;
  P.k_blocks = (N + GK - 1) / GK;
  P.m_tiles = (D + GM - 1) / GM;
  P.n_tiles = (N + GN - 1) / GN;
  P.tiles_per_gemm = (long long)B * P.m_tiles * P.n_tiles;
  P.total_tiles = 2 * P.tiles_per_gemm;
  P.C[0] = op[0].out;
  P.C[1] = op[1].out;
  P.ldc = N;
  P.stride_c = (long long)D * N;
  P.vec = (N % 4 == 0 && (reinterpret_cast<uintptr_t>(op[0].out) & 15) == 0 && (reinterpret_cast<uintptr_t>(op[1].out) & 15) == 0) ? 1 : 0;
  RC_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  const long long sms = sm_count();
  const unsigned grid = (unsigned)(P.total_tiles < sms ? P.total_tiles : sms);
  gemm_tc_kernel<<<grid, kThreads, kSmemBytes, st>>>(M, P);
  RC_CUDA(cudaGetLastError());
  return RC_OK;
}

int launch_build_bwd_tc(const ConstLevelPtrs &glv, const float *f1, const float *f2, float *df1, float *df2, int B,
                        int D, int h, int w, int L, float scale, void *ws, size_t ws_bytes, cudaStream_t st) {
  using namespace tcg;
  const size_t need = build_bwd_tc_workspace_bytes(B, D, h, w);
  if (ws == nullptr || ws_bytes < need) {
    set_error("rc_corr_build_bwd: workspace too small (%zu < %zu bytes)", ws_bytes, need);
    return RC_ERR_WORKSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(ws) & 255) != 0) {
    set_error("rc_corr_build_bwd: workspace must be 256-byte aligned");
    return RC_ERR_INVALID;
  }
  const int N = h * w, pitch = pitch_of(N);
  const size_t big = al256((size_t)B * N * pitch * 2), small = al256((size_t)B * D * pitch * 2);
  uint8_t *base = static_cast<uint8_t *>(ws);
  auto *c_hi = reinterpret_cast<__nv_bfloat16 *>(base);
  auto *c_lo = reinterpret_cast<__nv_bfloat16 *>(base + big);
  auto *t_hi = reinterpret_cast<__nv_bfloat16 *>(base + 2 * big);
  auto *t_lo = reinterpret_cast<__nv_bfloat16 *>(base + 3 * big);
  auto *f1_hi = reinterpret_cast<__nv_bfloat16 *>(base + 4 * big);
  auto *f1_lo = reinterpret_cast<__nv_bfloat16 *>(base + 4 * big + small);
  auto *f2_hi = reinterpret_cast<__nv_bfloat16 *>(base + 4 * big + 2 * small);
  auto *f2_lo = reinterpret_cast<__nv_bfloat16 *>(base + 4 * big + 3 * small);
  {
    dim3 grid((unsigned)((N + 63) / 64), (unsigned)((N + 63) / 64), (unsigned)B);
    const int vec_in = (N % 4 == 0 && (reinterpret_cast<uintptr_t>(glv.p[0]) & 15) == 0) ? 1 : 0;
    switch (L) {
      case 1: fold_split_kernel<1><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      case 2: fold_split_kernel<2><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      case 3: fold_split_kernel<3><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      case 4: fold_split_kernel<4><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      case 5: fold_split_kernel<5><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      case 6: fold_split_kernel<6><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      case 7: fold_split_kernel<7><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
      default: fold_split_kernel<8><<<grid, 256, 0, st>>>(glv, c_hi, c_lo, t_hi, t_lo, h, w, pitch, scale, vec_in); break;
    }
    RC_CUDA(cudaGetLastError());
    const long long rows = (long long)B * D;
    const int blocks = 148 * 8;
    split_rows_kernel<<<blocks, 256, 0, st>>>(f1, f1_hi, f1_lo, rows, N, pitch);
    RC_CUDA(cudaGetLastError());
    split_rows_kernel<<<blocks, 256, 0, st>>>(f2, f2_hi, f2_lo, rows, N, pitch);
    RC_CUDA(cudaGetLastError());
  }
  const GemmOperands ops[2] = {{f2_hi, f2_lo, c_hi, c_lo, df1}, {f1_hi, f1_lo, t_hi, t_lo, df2}};
  return run_gemms(ops, B, D, N, pitch, st);
}

}  // namespace rc
