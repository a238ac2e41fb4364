/*
 * corr_oracle.c -- CPU restatement of RAFT's correlation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker for the CUDA kernels in
 * raft_ros_b200/csrc/.  Nothing in the product path may call it: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * Every function restates, in plain scalar C, what one piece of the reference
 * computes (citations are relative to the reference checkout):
 *
 *   oracle_corr_build        core/corr.py:52-60 (CorrBlock.corr) + :21-27 (pyramid)
 *   oracle_corr_lookup       core/corr.py:29-50 (CorrBlock.__call__) +
 *                            core/utils/utils.py:57-71 (bilinear_sampler -> grid_sample,
 *                            bilinear, zeros padding, align_corners=True)
 *   oracle_corr_lookup_bwd   autograd of the above w.r.t. the pyramid (coords are
 *                            detached at core/raft.py:123)
 *   oracle_corr_build_bwd    autograd of core/corr.py:24-27,52-60 (avg_pool2d, /sqrt(D), matmul)
 *   oracle_alt_fwd           alt_cuda_corr/correlation_kernel.cu:18-119
 *   oracle_alt_bwd           alt_cuda_corr/correlation_kernel.cu:122-256
 *
 * Parity pinning: the reference ships no golden vectors (no tests at all), so this
 * oracle is pinned by executing the reference's own Python (imported read-only from
 * the reference checkout) on seeded inputs; tests/golden/make_golden.py is the
 * generating script and tests/golden/ holds its outputs.  The arithmetic under the reference
 * is un-vendored PyTorch/ATen (torch.matmul, F.avg_pool2d, F.grid_sample); the
 * reference pins no version ("tested with PyTorch 1.6", README.md:45); fixtures were
 * generated with torch 2.11.0.
 *
 * Accumulation is in double so that the oracle is the most accurate party of any
 * comparison; results are rounded to float once.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_MAX_LEVELS 8

int oracle_version(void) { return 1; }

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* level sizes follow F.avg_pool2d(x, 2, stride=2) in floor mode: h_{l+1} = h_l / 2 */
static void level_dims(int h, int w, int l, int *hl, int *wl) {
  *hl = h >> l;
  *wl = w >> l;
}

/* ------------------------------------------------------------------------- */
/* core/corr.py:52-60 then :21-27                                             */
/* f1, f2: (B, D, h, w) fp32 NCHW.  levels[l]: (B*h*w, h_l, w_l) fp32.        */
/* ------------------------------------------------------------------------- */
int oracle_corr_build(const float *f1, const float *f2, int B, int D, int h, int w,
                      int L, float **levels) {
  if (L < 1 || L > ORACLE_MAX_LEVELS) return 1;
  const int N = h * w;
  /* the reference divides by torch.sqrt(torch.tensor(dim).float()) : a float sqrt */
  const float sd = sqrtf((float)D);
  float *l0 = levels[0];
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; ++b) {
    for (int p = 0; p < N; ++p) {
      const float *a = f1 + (size_t)b * D * N + p;
      const float *bb = f2 + (size_t)b * D * N;
      float *row = l0 + ((size_t)b * N + p) * N;
      for (int q0 = 0; q0 < N; q0 += 64) {
        double acc[64];
        const int qn = (N - q0) < 64 ? (N - q0) : 64;
        for (int j = 0; j < qn; ++j) acc[j] = 0.0;
        for (int d = 0; d < D; ++d) {
          const double av = a[(size_t)d * N];
          const float *br = bb + (size_t)d * N + q0;
          for (int j = 0; j < qn; ++j) acc[j] += av * (double)br[j];
        }
        for (int j = 0; j < qn; ++j) row[q0 + j] = (float)acc[j] / sd;
      }
    }
  }
  /* pyramid: each level is the floor-mode 2x2 mean of the previous one */
  for (int l = 1; l < L; ++l) {
    int hp, wp, hc, wc;
    level_dims(h, w, l - 1, &hp, &wp);
    level_dims(h, w, l, &hc, &wc);
    if (hc < 1 || wc < 1) return 2;
    const float *src = levels[l - 1];
    float *dst = levels[l];
    const size_t rows = (size_t)B * N;
#pragma omp parallel for schedule(static)
    for (size_t rp = 0; rp < rows; ++rp) {
      const float *s = src + rp * hp * wp;
      float *d = dst + rp * hc * wc;
      for (int y = 0; y < hc; ++y)
        for (int x = 0; x < wc; ++x) {
          const float *s0 = s + (size_t)(2 * y) * wp + 2 * x;
          float sum = s0[0] + s0[1];
          sum += s0[wp];
          sum += s0[wp + 1];
          d[y * wc + x] = sum * 0.25f;
        }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------- */
/* bilinear tap with zeros outside, pixel coordinates                         */
/* ------------------------------------------------------------------------- */
static inline float fetch0(const float *img, int H, int W, int y, int x) {
  return (y >= 0 && y < H && x >= 0 && x < W) ? img[(size_t)y * W + x] : 0.0f;
}

/*
 * roundtrip != 0 reproduces the float round trip of the reference's sampler:
 * core/utils/utils.py:61-62 normalises to [-1,1] (2*x/(W-1) - 1) and ATen's
 * grid_sampler un-normalises with ((g+1)/2)*(W-1) (align_corners=True), both in fp32.
 * roundtrip == 0 samples directly in pixel space (what a fused kernel does).
 */
static inline float sample_bilinear(const float *img, int H, int W, float x, float y,
                                    int roundtrip) {
  if (roundtrip) {
    float gx = 2.0f * x / (float)(W - 1) - 1.0f;
    float gy = 2.0f * y / (float)(H - 1) - 1.0f;
    x = ((gx + 1.0f) / 2.0f) * (float)(W - 1);
    y = ((gy + 1.0f) / 2.0f) * (float)(H - 1);
  }
  const float xf = floorf(x), yf = floorf(y);
  const int x0 = (int)xf, y0 = (int)yf;
  const float dx = x - xf, dy = y - yf;
  const float v00 = fetch0(img, H, W, y0, x0);
  const float v01 = fetch0(img, H, W, y0, x0 + 1);
  const float v10 = fetch0(img, H, W, y0 + 1, x0);
  const float v11 = fetch0(img, H, W, y0 + 1, x0 + 1);
  const double r = (double)v00 * (1.0 - dy) * (1.0 - dx) + (double)v01 * (1.0 - dy) * dx +
                   (double)v10 * dy * (1.0 - dx) + (double)v11 * dy * dx;
  return (float)r;
}

/* ------------------------------------------------------------------------- */
/* core/corr.py:29-50.  coords (B,2,h,w) [x,y]; out (B, L*(2r+1)^2, h, w).    */
/* channel = l*(2r+1)^2 + i*(2r+1) + j with i the x offset and j the y offset */
/* (delta = stack(meshgrid(dy,dx)) added to [x,y] centroids, core/corr.py:37-43) */
/* ------------------------------------------------------------------------- */
int oracle_corr_lookup(float *const *levels, const float *coords, int B, int h, int w,
                       int L, int r, int roundtrip, float *out) {
  if (L < 1 || L > ORACLE_MAX_LEVELS) return 1;
  const int N = h * w, rd = 2 * r + 1, CH = L * rd * rd;
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; ++b) {
    for (int p = 0; p < N; ++p) {
      const float cx = coords[((size_t)b * 2 + 0) * N + p];
      const float cy = coords[((size_t)b * 2 + 1) * N + p];
      for (int l = 0; l < L; ++l) {
        int hl, wl;
        level_dims(h, w, l, &hl, &wl);
        const float *img = levels[l] + ((size_t)b * N + p) * hl * wl;
        const float sc = (float)(1 << l);
        const float x = cx / sc, y = cy / sc;
        for (int i = 0; i < rd; ++i)
          for (int j = 0; j < rd; ++j) {
            const float v = sample_bilinear(img, hl, wl, x + (float)(i - r), y + (float)(j - r),
                                            roundtrip);
            out[((size_t)b * CH + (size_t)l * rd * rd + i * rd + j) * N + p] = v;
          }
      }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------- */
/* d(out)/d(pyramid): scatter-add of the bilinear weights (pixel space).      */
/* grad_levels must be zero-initialised by the caller or accumulate=1 adds.   */
/* ------------------------------------------------------------------------- */
int oracle_corr_lookup_bwd(const float *grad_out, const float *coords, int B, int h, int w,
                           int L, int r, float **grad_levels) {
  if (L < 1 || L > ORACLE_MAX_LEVELS) return 1;
  const int N = h * w, rd = 2 * r + 1, CH = L * rd * rd;
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; ++b) {
    for (int p = 0; p < N; ++p) {
      const float cx = coords[((size_t)b * 2 + 0) * N + p];
      const float cy = coords[((size_t)b * 2 + 1) * N + p];
      for (int l = 0; l < L; ++l) {
        int hl, wl;
        level_dims(h, w, l, &hl, &wl);
        float *g = grad_levels[l] + ((size_t)b * N + p) * hl * wl;
        const float sc = (float)(1 << l);
        const float x = cx / sc, y = cy / sc;
        const float xf = floorf(x), yf = floorf(y);
        const double dx = x - xf, dy = y - yf;
        /* accumulate the (rd+1)^2 integer patch in double, then add once */
        double patch[32 * 32];
        const int pn = rd + 1;
        if (pn > 32) continue;
        for (int k = 0; k < pn * pn; ++k) patch[k] = 0.0;
        for (int i = 0; i < rd; ++i)
          for (int j = 0; j < rd; ++j) {
            const double go =
                grad_out[((size_t)b * CH + (size_t)l * rd * rd + i * rd + j) * N + p];
            /* tap (i,j) touches patch columns i,i+1 (x) and rows j,j+1 (y) */
            patch[j * pn + i] += go * (1.0 - dy) * (1.0 - dx);
            patch[j * pn + i + 1] += go * (1.0 - dy) * dx;
            patch[(j + 1) * pn + i] += go * dy * (1.0 - dx);
            patch[(j + 1) * pn + i + 1] += go * dy * dx;
          }
        const int x0 = (int)xf - r, y0 = (int)yf - r;
        for (int py = 0; py < pn; ++py)
          for (int px = 0; px < pn; ++px) {
            const int yy = y0 + py, xx = x0 + px;
            if (yy >= 0 && yy < hl && xx >= 0 && xx < wl)
              g[(size_t)yy * wl + xx] += (float)patch[py * pn + px];
          }
      }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------- */
/* autograd of the build: grad pyramid -> df1, df2 (B,D,h,w).                 */
/* dL0_total = g0 + up(g1)/4 + up(g2)/16 + ...  (floor-mode pooling: cells    */
/* outside 2*h_{l+1} x 2*w_{l+1} receive nothing from coarser levels)         */
/* df1[d,p] = sum_q dC[p,q] f2[d,q] / sqrt(D);  df2[d,q] = sum_p dC[p,q] f1[d,p] / sqrt(D) */
/* ------------------------------------------------------------------------- */
int oracle_corr_build_bwd(float *const *grad_levels, const float *f1, const float *f2, int B,
                          int D, int h, int w, int L, float *df1, float *df2) {
  if (L < 1 || L > ORACLE_MAX_LEVELS) return 1;
  const int N = h * w;
  const double inv_sd = 1.0 / (double)sqrtf((float)D);
  double *dC = (double *)malloc(sizeof(double) * (size_t)N * N);
  double *acc2 = (double *)malloc(sizeof(double) * (size_t)D * N);
  if (!dC || !acc2) { free(dC); free(acc2); return 3; }
  for (int b = 0; b < B; ++b) {
#pragma omp parallel for schedule(static)
    for (int p = 0; p < N; ++p) {
      double *row = dC + (size_t)p * N;
      for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
          double g = grad_levels[0][((size_t)b * N + p) * N + (size_t)y * w + x];
          int yy = y, xx = x;
          double wgt = 1.0;
          for (int l = 1; l < L; ++l) {
            int hl, wl;
            level_dims(h, w, l, &hl, &wl);
            yy >>= 1; xx >>= 1; wgt *= 0.25;
            if (yy >= hl || xx >= wl) break; /* dropped by floor mode from here on */
            g += wgt * grad_levels[l][((size_t)b * N + p) * hl * wl + (size_t)yy * wl + xx];
          }
          row[(size_t)y * w + x] = g * inv_sd;
        }
    }
    const float *a = f1 + (size_t)b * D * N;
    const float *c = f2 + (size_t)b * D * N;
#pragma omp parallel for schedule(static)
    for (int d = 0; d < D; ++d) {
      for (int p = 0; p < N; ++p) {
        const double *row = dC + (size_t)p * N;
        double s = 0.0;
        for (int q = 0; q < N; ++q) s += row[q] * (double)c[(size_t)d * N + q];
        df1[((size_t)b * D + d) * N + p] = (float)s;
      }
      double *o = acc2 + (size_t)d * N;
      for (int q = 0; q < N; ++q) o[q] = 0.0;
      for (int p = 0; p < N; ++p) {
        const double *row = dC + (size_t)p * N;
        const double av = a[(size_t)d * N + p];
        for (int q = 0; q < N; ++q) o[q] += row[q] * av;
      }
      for (int q = 0; q < N; ++q) df2[((size_t)b * D + d) * N + q] = (float)o[q];
    }
  }
  free(dC);
  free(acc2);
  return 0;
}

/* ------------------------------------------------------------------------- */
/* alt_cuda_corr/correlation_kernel.cu:18-119                                 */
/* f1 (B,H1,W1,C), f2 (B,H2,W2,C) NHWC; coords (B,Nc,H1,W1,2) [x,y];          */
/* corr (B,Nc,rd*rd,H1,W1), tap index = oy + rd*ox, value = sum_c f1 . bilinear(f2)  */
/* at (x - r + ox, y - r + oy), zeros outside f2; NOT scaled by 1/sqrt(C).    */
/* ------------------------------------------------------------------------- */
int oracle_alt_fwd(const float *f1, const float *f2, const float *coords, int B, int Nc, int H1,
                   int W1, int H2, int W2, int C, int r, float *corr) {
  const int rd = 2 * r + 1, pn = rd + 1;
  if (pn > 32) return 1;
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; ++b)
    for (int p = 0; p < H1 * W1; ++p) {
      const float *a = f1 + ((size_t)b * H1 * W1 + p) * C;
      for (int n = 0; n < Nc; ++n) {
        const float *cp = coords + ((((size_t)b * Nc + n) * H1 * W1) + p) * 2;
        const float x = cp[0], y = cp[1];
        const float xf = floorf(x), yf = floorf(y);
        const double dx = x - xf, dy = y - yf;
        const int x0 = (int)xf - r, y0 = (int)yf - r;
        double dots[32 * 32];
        for (int iy = 0; iy < pn; ++iy)
          for (int ix = 0; ix < pn; ++ix) {
            const int yy = y0 + iy, xx = x0 + ix;
            double s = 0.0;
            if (yy >= 0 && yy < H2 && xx >= 0 && xx < W2) {
              const float *bb = f2 + (((size_t)b * H2 + yy) * W2 + xx) * C;
              for (int c = 0; c < C; ++c) s += (double)a[c] * (double)bb[c];
            }
            dots[iy * pn + ix] = s;
          }
        for (int ox = 0; ox < rd; ++ox)
          for (int oy = 0; oy < rd; ++oy) {
            const double v = dots[oy * pn + ox] * (1.0 - dy) * (1.0 - dx) +
                             dots[oy * pn + ox + 1] * (1.0 - dy) * dx +
                             dots[(oy + 1) * pn + ox] * dy * (1.0 - dx) +
                             dots[(oy + 1) * pn + ox + 1] * dy * dx;
            corr[((((size_t)b * Nc + n) * rd * rd) + (oy + rd * ox)) * H1 * W1 + p] = (float)v;
          }
      }
    }
  return 0;
}

/* ------------------------------------------------------------------------- */
/* alt_cuda_corr/correlation_kernel.cu:122-256.  coords_grad is identically   */
/* zero in the reference (:307, never written) and is not produced here.      */
/* Accumulation into f2_grad is done per batch item in a fixed order          */
/* (deterministic; the reference uses atomicAdd, :237).                       */
/* ------------------------------------------------------------------------- */
int oracle_alt_bwd(const float *f1, const float *f2, const float *coords, const float *corr_grad,
                   int B, int Nc, int H1, int W1, int H2, int W2, int C, int r, float *f1_grad,
                   float *f2_grad) {
  const int rd = 2 * r + 1, pn = rd + 1;
  if (pn > 32) return 1;
#pragma omp parallel for schedule(static)
  for (int b = 0; b < B; ++b) {
    double *g1 = (double *)calloc((size_t)H1 * W1 * C, sizeof(double));
    double *g2 = (double *)calloc((size_t)H2 * W2 * C, sizeof(double));
    for (int p = 0; p < H1 * W1; ++p) {
      const float *a = f1 + ((size_t)b * H1 * W1 + p) * C;
      for (int n = 0; n < Nc; ++n) {
        const float *cp = coords + ((((size_t)b * Nc + n) * H1 * W1) + p) * 2;
        const float x = cp[0], y = cp[1];
        const float xf = floorf(x), yf = floorf(y);
        const double dx = x - xf, dy = y - yf;
        const int x0 = (int)xf - r, y0 = (int)yf - r;
        const float *gp = corr_grad + (((size_t)b * Nc + n) * rd * rd) * H1 * W1 + p;
        for (int iy = 0; iy < pn; ++iy)
          for (int ix = 0; ix < pn; ++ix) {
            const int yy = y0 + iy, xx = x0 + ix;
            if (!(yy >= 0 && yy < H2 && xx >= 0 && xx < W2)) continue;
            double g = 0.0;
            if (iy > 0 && ix > 0) g += (double)gp[(size_t)((iy - 1) + rd * (ix - 1)) * H1 * W1] * dy * dx;
            if (iy > 0 && ix < rd) g += (double)gp[(size_t)((iy - 1) + rd * ix) * H1 * W1] * dy * (1.0 - dx);
            if (iy < rd && ix > 0) g += (double)gp[(size_t)(iy + rd * (ix - 1)) * H1 * W1] * (1.0 - dy) * dx;
            if (iy < rd && ix < rd) g += (double)gp[(size_t)(iy + rd * ix) * H1 * W1] * (1.0 - dy) * (1.0 - dx);
            const float *bb = f2 + (((size_t)b * H2 + yy) * W2 + xx) * C;
            double *o1 = g1 + (size_t)p * C;
            double *o2 = g2 + ((size_t)yy * W2 + xx) * C;
            for (int c = 0; c < C; ++c) {
              o1[c] += g * (double)bb[c];
              o2[c] += g * (double)a[c];
            }
          }
      }
    }
    for (size_t k = 0; k < (size_t)H1 * W1 * C; ++k) f1_grad[(size_t)b * H1 * W1 * C + k] = (float)g1[k];
    for (size_t k = 0; k < (size_t)H2 * W2 * C; ++k) f2_grad[(size_t)b * H2 * W2 * C + k] = (float)g2[k];
    free(g1);
    free(g2);
  }
  return 0;
}
