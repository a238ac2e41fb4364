// extern "C" surface of libraftcorr.so (declared in include/raftcorr.h).
// Argument validation + dispatch only; kernels live in the other translation units.
#include <cstdarg>
#include <cstdio>
#include <cuda.h>

#include <cmath>
#include <mutex>
#include <cstdlib>
#include <initializer_list>

#include "common.cuh"

namespace rc {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return RC_ERR_CUDA;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_tmap(::CUtensorMap_st *tm, const void *base, int rank, const unsigned long long *dims,
                       const unsigned long long *strides_bytes, const unsigned int *box, const char *who, bool f32_plain);

int encode_tmap_sw128(::CUtensorMap_st *tm, const void *base, int rank, const unsigned long long *dims,
                      const unsigned long long *strides_bytes, const unsigned int *box, const char *who) {
  return encode_tmap(tm, base, rank, dims, strides_bytes, box, who, false);
}

int encode_tmap_f32(::CUtensorMap_st *tm, const void *base, int rank, const unsigned long long *dims,
                    const unsigned long long *strides_bytes, const unsigned int *box, const char *who) {
  return encode_tmap(tm, base, rank, dims, strides_bytes, box, who, true);
}

static int encode_tmap(::CUtensorMap_st *tm, const void *base, int rank, const unsigned long long *dims,
                       const unsigned long long *strides_bytes, const unsigned int *box, const char *who, bool f32_plain) {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  if (!fn) {
    set_error("%s: cuTensorMapEncodeTiled is not available from the driver", who);
    return RC_ERR_CUDA;
  }
  cuuint64_t d[5], st[4];
  cuuint32_t b[5], ones[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) {
    d[i] = dims[i];
    b[i] = box[i];
    if (i + 1 < rank) st[i] = strides_bytes[i];
  }
  CUresult r = fn(tm, f32_plain ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                  const_cast<void *>(base), d, st, b, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  f32_plain ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("%s: cuTensorMapEncodeTiled failed with CUresult %d", who, (int)r);
    return RC_ERR_CUDA;
  }
  return RC_OK;
}

bool pdl_enabled() {
  static const bool on = [] {
    const char *e = getenv("RAFT_CORR_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

static int check_geometry(const char *fn, int B, int h, int w, int L) {
  RC_CHECK_ARG(B > 0 && h > 0 && w > 0, "%s: B, h, w must be positive (got %d, %d, %d)", fn, B, h, w);
  RC_CHECK_ARG(L >= 1 && L <= RC_MAX_LEVELS, "%s: num_levels %d outside 1..%d", fn, L, RC_MAX_LEVELS);
  RC_CHECK_ARG((h >> (L - 1)) >= 1 && (w >> (L - 1)) >= 1,
               "%s: level %d of a %dx%d grid is empty", fn, L - 1, h, w);
  RC_CHECK_ARG((long long)h * w <= 0x7fffffffLL / 4, "%s: feature grid too large", fn);
  return RC_OK;
}

}  // namespace rc

using namespace rc;

extern "C" {

int rc_version(void) { return 200; }

const char *rc_last_error(void) { return g_err; }

int rc_set_l2_fetch_granularity(int bytes) {
  RC_CHECK_ARG(bytes == 32 || bytes == 64 || bytes == 128, "rc_set_l2_fetch_granularity: bytes must be 32, 64 or 128");
  RC_CUDA(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes));
  return RC_OK;
}

int rc_get_l2_fetch_granularity(int *bytes) {
  RC_CHECK_ARG(bytes != nullptr, "rc_get_l2_fetch_granularity: null output");
  size_t v = 0;
  RC_CUDA(cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity));
  *bytes = (int)v;
  return RC_OK;
}

int rc_level_dims(int h, int w, int level, int *hl, int *wl) {
  RC_CHECK_ARG(hl && wl, "rc_level_dims: null output");
  RC_CHECK_ARG(level >= 0 && level < RC_MAX_LEVELS && h > 0 && w > 0, "rc_level_dims: bad arguments");
  *hl = h >> level;
  *wl = w >> level;
  RC_CHECK_ARG(*hl >= 1 && *wl >= 1, "rc_level_dims: level %d of %dx%d is empty", level, h, w);
  return RC_OK;
}

size_t rc_corr_build_workspace_bytes(int B, int D, int h, int w, int precision) {
  if (precision == RC_PREC_FP32) return 0;
  return build_tc_workspace_bytes(B, D, h, w, precision);
}

int rc_corr_build_pyramid(const float *fmap1, const float *fmap2, int B, int D, int h, int w,
                          float *const *level_ptrs, int num_levels, int precision, float scale,
                          void *workspace, size_t workspace_bytes, void *stream) {
  if (int rc = check_geometry("rc_corr_build_pyramid", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(fmap1 && fmap2 && level_ptrs, "rc_corr_build_pyramid: null pointer");
  RC_CHECK_ARG(D > 0, "rc_corr_build_pyramid: D must be positive");
  LevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_build_pyramid: level %d pointer is null", l);
    lv.p[l] = level_ptrs[l];
  }
  if (scale == 0.0f) scale = 1.0f / sqrtf((float)D);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (precision) {
    case RC_PREC_FP32:
      return launch_build_simt(fmap1, fmap2, B, D, h, w, lv, num_levels, scale, st);
    case RC_PREC_BF16X3:
    case RC_PREC_BF16:
    case RC_PREC_FP16_FP8:
      return launch_build_tc(fmap1, fmap2, B, D, h, w, lv, num_levels, precision, scale, workspace,
                             workspace_bytes, st);
    default:
      set_error("rc_corr_build_pyramid: unknown precision %d", precision);
      return RC_ERR_INVALID;
  }
}

// Allocation-free build: same contract as rc_corr_build_pyramid, scratch inside the library image
// (see the block comment in build_tc.cu and include/raftcorr.h).  Calls no allocating routine.
int rc_corr_build_pyramid_noalloc(const float *fmap1, const float *fmap2, int B, int D, int h, int w,
                                  float *const *level_ptrs, int num_levels, int precision, float scale,
                                  void *stream) {
  if (int rc = check_geometry("rc_corr_build_pyramid_noalloc", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(fmap1 && fmap2 && level_ptrs, "rc_corr_build_pyramid_noalloc: null pointer");
  RC_CHECK_ARG(D > 0, "rc_corr_build_pyramid_noalloc: D must be positive");
  RC_CHECK_ARG(precision == RC_PREC_BF16X3 || precision == RC_PREC_BF16,
               "rc_corr_build_pyramid_noalloc: only RC_PREC_BF16X3 / RC_PREC_BF16 use scratch; "
               "RC_PREC_FP32 goes through rc_corr_build_pyramid");
  LevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_build_pyramid_noalloc: level %d pointer is null", l);
    lv.p[l] = level_ptrs[l];
  }
  if (scale == 0.0f) scale = 1.0f / sqrtf((float)D);
  return launch_build_tc_noalloc(fmap1, fmap2, B, D, h, w, lv, num_levels, precision, scale,
                                 static_cast<cudaStream_t>(stream));
}

int rc_corr_lookup_fwd(const float *const *level_ptrs, const float *coords, float *out, int B,
                       int h, int w, int num_levels, int radius, void *stream) {
  if (int rc = check_geometry("rc_corr_lookup_fwd", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(level_ptrs && coords && out, "rc_corr_lookup_fwd: null pointer");
  RC_CHECK_ARG(radius >= 1, "rc_corr_lookup_fwd: radius must be >= 1");
  ConstLevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_lookup_fwd: level %d pointer is null", l);
    lv.p[l] = level_ptrs[l];
  }
  return launch_lookup_fwd(lv, coords, out, B, h, w, num_levels, radius, static_cast<cudaStream_t>(stream));
}

int rc_corr_tile_group(void) { return RC_TILE_GROUP; }

int rc_corr_tiled_level_elems(int B, int h, int w, int level, size_t *elems) {
  RC_CHECK_ARG(elems, "rc_corr_tiled_level_elems: null output");
  RC_CHECK_ARG(B > 0, "rc_corr_tiled_level_elems: B must be positive");
  int hl = 0, wl = 0;
  if (int rc = rc_level_dims(h, w, level, &hl, &wl)) return rc;
  const size_t groups = (size_t)B * (((size_t)h * w + RC_TILE_GROUP - 1) / RC_TILE_GROUP);
  *elems = groups * RC_TILE_GROUP * (size_t)((hl + 1) & ~1) * (size_t)((wl + 7) & ~7);
  return RC_OK;
}

int rc_corr_build_pyramid_tiled(const float *fmap1, const float *fmap2, int B, int D, int h, int w,
                                float *const *level_ptrs, int num_levels, int precision, float scale,
                                void *workspace, size_t workspace_bytes, void *stream) {
  if (int rc = check_geometry("rc_corr_build_pyramid_tiled", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(fmap1 && fmap2 && level_ptrs, "rc_corr_build_pyramid_tiled: null pointer");
  RC_CHECK_ARG(D > 0, "rc_corr_build_pyramid_tiled: D must be positive");
  if (precision != RC_PREC_BF16X3 && precision != RC_PREC_BF16 && precision != RC_PREC_BF16_VOL16 && precision != RC_PREC_FP16_FP8) {
    set_error("rc_corr_build_pyramid_tiled: only the tensor-core precisions write the tiled layout (got %d)",
              precision);
    return RC_ERR_UNSUPPORTED;
  }
  LevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_build_pyramid_tiled: level %d pointer is null", l);
    lv.p[l] = level_ptrs[l];
  }
  if (scale == 0.0f) scale = 1.0f / sqrtf((float)D);
  return launch_build_tc_tiled(fmap1, fmap2, B, D, h, w, lv, num_levels, precision, scale, workspace,
                               workspace_bytes, static_cast<cudaStream_t>(stream));
}

int rc_corr_build_pyramid_tiled_prepared(const void *fmap1_hi, const void *fmap1_lo, const void *fmap2_hi,
                                         const void *fmap2_lo, int B, int D, int h, int w, float *const *level_ptrs,
                                         int num_levels, int precision, float scale, void *stream) {
  const char *fn = "rc_corr_build_pyramid_tiled_prepared";
  if (int rc = check_geometry(fn, B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(fmap1_hi && fmap2_hi && level_ptrs, "%s: null pointer", fn);
  RC_CHECK_ARG(D > 0 && D % 64 == 0, "%s: the prepared maps have a channel pitch of D = %d, which must be a multiple of 64", fn, D);
  if (precision != RC_PREC_BF16X3 && precision != RC_PREC_BF16 && precision != RC_PREC_BF16_VOL16) {
    set_error("%s: only the tensor-core precisions write the tiled layout (got %d)", fn, precision);
    return RC_ERR_UNSUPPORTED;
  }
  RC_CHECK_ARG(precision != RC_PREC_BF16X3 || (fmap1_lo && fmap2_lo), "%s: RC_PREC_BF16X3 needs the lo parts", fn);
  for (const void *q : {fmap1_hi, fmap1_lo, fmap2_hi, fmap2_lo})
    RC_CHECK_ARG((reinterpret_cast<uintptr_t>(q) & 127) == 0, "%s: operand maps must be 128-byte aligned", fn);
  LevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "%s: level %d pointer is null", fn, l);
    lv.p[l] = level_ptrs[l];
  }
  if (scale == 0.0f) scale = 1.0f / sqrtf((float)D);
  return launch_build_tc_tiled_prepared(fmap1_hi, fmap1_lo, fmap2_hi, fmap2_lo, B, D, h, w, lv, num_levels, precision, scale,
                                        static_cast<cudaStream_t>(stream));
}

size_t rc_fnet_conv1x1_weight_bytes(int Cout, int Cin) {
  if (Cout <= 0 || Cin <= 0) return 0;
  return conv1x1_weight_bytes(Cout, Cin);
}

static int check_conv1x1(const char *fn, int Cout, int Cin) {
  RC_CHECK_ARG(Cout > 0 && Cout <= 256 && Cout % 64 == 0, "%s: Cout %d must be a multiple of 64, at most 256", fn, Cout);
  RC_CHECK_ARG(Cin > 0 && Cin <= 128 && Cin % 16 == 0, "%s: Cin %d must be a multiple of 16, at most 128", fn, Cin);
  return RC_OK;
}

int rc_fnet_conv1x1_prep_weight(const float *weight, void *wprep, int Cout, int Cin, void *stream) {
  RC_CHECK_ARG(weight && wprep, "rc_fnet_conv1x1_prep_weight: null pointer");
  if (int rc = check_conv1x1("rc_fnet_conv1x1_prep_weight", Cout, Cin)) return rc;
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0, "rc_fnet_conv1x1_prep_weight: wprep must be 128-byte aligned");
  return launch_conv1x1_prep_weight(weight, wprep, Cout, Cin, static_cast<cudaStream_t>(stream));
}

int rc_fnet_conv1x1_prep(const float *x, const void *wprep, const float *bias, void *hi, void *lo, int Bt, int Cin, int Cout,
                         int H, int W, void *stream) {
  RC_CHECK_ARG(x && wprep && bias && hi, "rc_fnet_conv1x1_prep: null pointer");
  RC_CHECK_ARG(Bt > 0 && H > 0 && W > 0 && (long long)Bt * ((long long)H * W / 128 + 1) < 0x7fffffffLL, "rc_fnet_conv1x1_prep: bad sizes");
  if (int rc = check_conv1x1("rc_fnet_conv1x1_prep", Cout, Cin)) return rc;
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0 && (reinterpret_cast<uintptr_t>(hi) & 127) == 0 &&
                   (reinterpret_cast<uintptr_t>(lo) & 127) == 0,
               "rc_fnet_conv1x1_prep: wprep, hi and lo must be 128-byte aligned");
  return launch_conv1x1_prep(x, wprep, bias, hi, lo, Bt, Cin, Cout, H * W, static_cast<cudaStream_t>(stream));
}

int rc_corr_tiled16_level_elems(int B, int h, int w, int level, size_t *elems) {
  RC_CHECK_ARG(elems, "rc_corr_tiled16_level_elems: null output");
  RC_CHECK_ARG(B > 0, "rc_corr_tiled16_level_elems: B must be positive");
  int hl = 0, wl = 0;
  if (int rc = rc_level_dims(h, w, level, &hl, &wl)) return rc;
  const size_t groups = (size_t)B * (((size_t)h * w + RC_TILE_GROUP - 1) / RC_TILE_GROUP);
  *elems = groups * RC_TILE_GROUP * (size_t)((hl + 1) & ~1) * (size_t)((wl + 15) & ~15);
  return RC_OK;
}

int rc_corr_lookup_tiled16_fwd(const void *const *level_ptrs, const float *coords, float *out, int B,
                               int h, int w, int num_levels, int radius, void *stream) {
  if (int rc = check_geometry("rc_corr_lookup_tiled16_fwd", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(level_ptrs && coords && out, "rc_corr_lookup_tiled16_fwd: null pointer");
  ConstLevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_lookup_tiled16_fwd: level %d pointer is null", l);
    RC_CHECK_ARG((reinterpret_cast<uintptr_t>(level_ptrs[l]) & 15) == 0, "rc_corr_lookup_tiled16_fwd: level %d must be 16-byte aligned", l);
    lv.p[l] = static_cast<const float *>(level_ptrs[l]);
  }
  return launch_lookup_tiled16_fwd(lv, coords, out, B, h, w, num_levels, radius, static_cast<cudaStream_t>(stream));
}

size_t rc_corr_convc1_weight_bytes(int Cout, int num_levels, int radius) {
  if (Cout <= 0 || num_levels <= 0 || num_levels > 4 || (radius != 3 && radius != 4)) return 0;
  return convc1_weight_bytes(Cout, num_levels, radius);
}

int rc_corr_convc1_prep_weight(const float *weight, void *wprep, int Cout, int num_levels, int radius, int bf16_ops,
                               void *stream) {
  RC_CHECK_ARG(weight && wprep, "rc_corr_convc1_prep_weight: null pointer");
  RC_CHECK_ARG(Cout > 0 && Cout <= 256 && Cout % 16 == 0, "rc_corr_convc1_prep_weight: Cout %d must be a multiple of 16, at most 256", Cout);
  RC_CHECK_ARG(num_levels >= 1 && num_levels <= 4, "rc_corr_convc1_prep_weight: num_levels %d outside 1..4", num_levels);
  RC_CHECK_ARG(radius == 3 || radius == 4, "rc_corr_convc1_prep_weight: radius %d unsupported (3 and 4 only)", radius);
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0, "rc_corr_convc1_prep_weight: wprep must be 128-byte aligned");
  return launch_convc1_prep_weight(weight, wprep, Cout, num_levels, radius, bf16_ops != 0, static_cast<cudaStream_t>(stream));
}

int rc_corr_lookup_convc1_fwd(const void *const *level_ptrs, int vol_fp16, const float *coords, const void *wprep,
                              const float *bias, float *out, int B, int h, int w, int num_levels, int radius, int Cout,
                              int bf16_ops, void *stream) {
  if (int rc = check_geometry("rc_corr_lookup_convc1_fwd", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(level_ptrs && coords && wprep && bias && out, "rc_corr_lookup_convc1_fwd: null pointer");
  RC_CHECK_ARG(Cout > 0 && Cout <= 256 && Cout % 16 == 0, "rc_corr_lookup_convc1_fwd: Cout %d must be a multiple of 16, at most 256", Cout);
  if (num_levels > 4 || (radius != 3 && radius != 4)) {
    set_error("rc_corr_lookup_convc1_fwd: radius 3 or 4 and at most 4 levels (got radius %d, %d levels)", radius, num_levels);
    return RC_ERR_UNSUPPORTED;
  }
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0, "rc_corr_lookup_convc1_fwd: wprep must be 128-byte aligned");
  RC_CHECK_ARG((long long)B * ((h * w + 127) / 128) < 0x7fffffffLL, "rc_corr_lookup_convc1_fwd: problem too large");
  ConstLevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_lookup_convc1_fwd: level %d pointer is null", l);
    RC_CHECK_ARG((reinterpret_cast<uintptr_t>(level_ptrs[l]) & 15) == 0, "rc_corr_lookup_convc1_fwd: level %d must be 16-byte aligned", l);
    lv.p[l] = static_cast<const float *>(level_ptrs[l]);
  }
  return launch_lookup_convc1(lv, vol_fp16 != 0, coords, wprep, bias, out, B, h, w, num_levels, radius, Cout, bf16_ops != 0,
                              static_cast<cudaStream_t>(stream));
}

int rc_corr_relayout16_level(const void *src, void *dst, int B, int N, int hl, int wl, int to_tiled,
                             void *stream) {
  RC_CHECK_ARG(src && dst, "rc_corr_relayout16_level: null pointer");
  RC_CHECK_ARG(B > 0 && N > 0 && hl > 0 && wl > 0, "rc_corr_relayout16_level: bad shape");
  return launch_relayout16(src, dst, B, N, hl, wl, to_tiled != 0, static_cast<cudaStream_t>(stream));
}

int rc_corr_relayout_level(const float *src, float *dst, int B, int N, int hl, int wl, int to_tiled,
                           void *stream) {
  RC_CHECK_ARG(src && dst, "rc_corr_relayout_level: null pointer");
  RC_CHECK_ARG(B > 0 && N > 0 && hl > 0 && wl > 0, "rc_corr_relayout_level: bad shape");
  return launch_relayout(src, dst, B, N, hl, wl, to_tiled != 0, static_cast<cudaStream_t>(stream));
}

int rc_corr_lookup_tiled_fwd(const float *const *level_ptrs, const float *coords, float *out, int B,
                             int h, int w, int num_levels, int radius, void *stream) {
  if (int rc = check_geometry("rc_corr_lookup_tiled_fwd", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(level_ptrs && coords && out, "rc_corr_lookup_tiled_fwd: null pointer");
  ConstLevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(level_ptrs[l] != nullptr, "rc_corr_lookup_tiled_fwd: level %d pointer is null", l);
    RC_CHECK_ARG((reinterpret_cast<uintptr_t>(level_ptrs[l]) & 15) == 0, "rc_corr_lookup_tiled_fwd: level %d must be 16-byte aligned", l);
    lv.p[l] = level_ptrs[l];
  }
  return launch_lookup_tiled_fwd(lv, coords, out, B, h, w, num_levels, radius, static_cast<cudaStream_t>(stream));
}

int rc_corr_lookup_bwd(const float *grad_out, const float *coords, float *const *grad_level_ptrs,
                       int B, int h, int w, int num_levels, int radius, void *stream) {
  if (int rc = check_geometry("rc_corr_lookup_bwd", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(grad_level_ptrs && coords && grad_out, "rc_corr_lookup_bwd: null pointer");
  RC_CHECK_ARG(radius >= 1, "rc_corr_lookup_bwd: radius must be >= 1");
  LevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(grad_level_ptrs[l] != nullptr, "rc_corr_lookup_bwd: level %d pointer is null", l);
    lv.p[l] = grad_level_ptrs[l];
  }
  return launch_lookup_bwd(grad_out, coords, lv, B, h, w, num_levels, radius, static_cast<cudaStream_t>(stream));
}

size_t rc_corr_build_bwd_workspace_bytes(int B, int D, int h, int w, int precision) {
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0) return 0;
  if (precision == RC_PREC_FP32) return build_bwd_workspace_bytes(B, D, h, w);
  return build_bwd_tc_workspace_bytes(B, D, h, w);
}

int rc_corr_build_bwd(const float *const *grad_level_ptrs, const float *fmap1, const float *fmap2,
                      float *dfmap1, float *dfmap2, int B, int D, int h, int w, int num_levels,
                      int precision, float scale, void *workspace, size_t workspace_bytes, void *stream) {
  if (int rc = check_geometry("rc_corr_build_bwd", B, h, w, num_levels)) return rc;
  RC_CHECK_ARG(grad_level_ptrs && fmap1 && fmap2 && dfmap1 && dfmap2, "rc_corr_build_bwd: null pointer");
  RC_CHECK_ARG(D > 0, "rc_corr_build_bwd: D must be positive");
  ConstLevelPtrs lv{};
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(grad_level_ptrs[l] != nullptr, "rc_corr_build_bwd: level %d pointer is null", l);
    lv.p[l] = grad_level_ptrs[l];
  }
  if (scale == 0.0f) scale = 1.0f / sqrtf((float)D);
  RC_CHECK_ARG(precision == RC_PREC_FP32 || precision == RC_PREC_BF16X3 || precision == RC_PREC_BF16,
               "rc_corr_build_bwd: unknown precision %d", precision);
  if (precision != RC_PREC_FP32)
    return launch_build_bwd_tc(lv, fmap1, fmap2, dfmap1, dfmap2, B, D, h, w, num_levels, scale, workspace,
                               workspace_bytes, static_cast<cudaStream_t>(stream));
  return launch_build_bwd(lv, fmap1, fmap2, dfmap1, dfmap2, B, D, h, w, num_levels, scale, workspace,
                          workspace_bytes, static_cast<cudaStream_t>(stream));
}

static int check_alt(const char *fn, int B, int N, int H1, int W1, int H2, int W2, int C, int radius) {
  RC_CHECK_ARG(B > 0 && N > 0 && H1 > 0 && W1 > 0 && H2 > 0 && W2 > 0 && C > 0,
               "%s: all sizes must be positive", fn);
  RC_CHECK_ARG(radius >= 1, "%s: radius must be >= 1", fn);
  if (C % 4 != 0) {
    set_error("%s: C=%d must be a multiple of 4 (the reference needs a multiple of 32)", fn, C);
    return RC_ERR_UNSUPPORTED;
  }
  return RC_OK;
}

int rc_alt_corr_fwd(const float *fmap1, const float *fmap2, const float *coords, float *corr,
                    int B, int N, int H1, int W1, int H2, int W2, int C, int radius, void *stream) {
  if (int rc = check_alt("rc_alt_corr_fwd", B, N, H1, W1, H2, W2, C, radius)) return rc;
  RC_CHECK_ARG(fmap1 && fmap2 && coords && corr, "rc_alt_corr_fwd: null pointer");
  return launch_alt_fwd(fmap1, fmap2, coords, corr, B, N, H1, W1, H2, W2, C, radius,
                        static_cast<cudaStream_t>(stream));
}

int rc_alt_prep_fmap(const float *fmap, void *hi, void *lo, int B, int C, int H, int W, void *stream) {
  RC_CHECK_ARG(fmap && hi, "rc_alt_prep_fmap: null pointer");
  RC_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "rc_alt_prep_fmap: all sizes must be positive");
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(hi) & 15) == 0 && (reinterpret_cast<uintptr_t>(lo) & 15) == 0,
               "rc_alt_prep_fmap: outputs must be 16-byte aligned");
  return launch_alt_prep(fmap, hi, lo, B, C, H * W, static_cast<cudaStream_t>(stream));
}

int rc_alt_prep_fmap_level(const float *fmap, void *hi, void *lo, int B, int C, int H, int W, int level, void *stream) {
  RC_CHECK_ARG(fmap && hi, "rc_alt_prep_fmap_level: null pointer");
  RC_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "rc_alt_prep_fmap_level: all sizes must be positive");
  RC_CHECK_ARG(level >= 0 && level <= 3, "rc_alt_prep_fmap_level: level %d outside 0..3", level);
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(hi) & 15) == 0 && (reinterpret_cast<uintptr_t>(lo) & 15) == 0,
               "rc_alt_prep_fmap_level: outputs must be 16-byte aligned");
  return launch_alt_prep_pooled(fmap, hi, lo, B, C, H, W, level, static_cast<cudaStream_t>(stream));
}

int rc_alt_absmax(const float *fmap, unsigned int *amax, int B, int C, int H, int W, void *stream) {
  RC_CHECK_ARG(fmap && amax, "rc_alt_absmax: null pointer");
  RC_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "rc_alt_absmax: all sizes must be positive");
  return launch_alt_absmax(fmap, amax, B, (long long)C * H * W, static_cast<cudaStream_t>(stream));
}

int rc_alt_prep_fmap_level_f8(const float *fmap, void *hi, void *pair, const unsigned int *amax, int side, int B, int C, int H,
                              int W, int level, void *stream) {
  RC_CHECK_ARG(fmap && hi && pair && amax, "rc_alt_prep_fmap_level_f8: null pointer");
  RC_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "rc_alt_prep_fmap_level_f8: all sizes must be positive");
  RC_CHECK_ARG(level >= 0 && level <= 3 && (side == 0 || side == 1), "rc_alt_prep_fmap_level_f8: level 0..3, side 0 (fmap1) or 1 (fmap2)");
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(hi) & 15) == 0 && (reinterpret_cast<uintptr_t>(pair) & 15) == 0,
               "rc_alt_prep_fmap_level_f8: outputs must be 16-byte aligned");
  return launch_alt_prep_f8(fmap, hi, pair, amax, side, B, C, H, W, level, static_cast<cudaStream_t>(stream));
}

int rc_alt_corr_pyramid_set_norm(void *setup, const unsigned int *amax1, const unsigned int *amax2) {
  RC_CHECK_ARG(setup && amax1 && amax2, "rc_alt_corr_pyramid_set_norm: null pointer");
  return alt_tc_set_norm(setup, amax1, amax2);
}

size_t rc_alt_corr_pyramid_setup_bytes(void) { return alt_tc_setup_bytes(); }

int rc_alt_corr_pyramid_setup(void *setup, const void *fmap1_hi, const void *fmap1_lo, const void *const *fmap2_hi,
                              const void *const *fmap2_lo, int B, int H1, int W1, int H2, int W2, int C, int num_levels,
                              int radius, int precision) {
  if (int rc = check_geometry("rc_alt_corr_pyramid_setup", B, H2, W2, num_levels)) return rc;
  RC_CHECK_ARG(setup && fmap1_hi && fmap2_hi, "rc_alt_corr_pyramid_setup: null pointer");
  RC_CHECK_ARG(H1 > 0 && W1 > 0 && C > 0, "rc_alt_corr_pyramid_setup: sizes must be positive");
  RC_CHECK_ARG(precision == RC_PREC_BF16X3 || precision == RC_PREC_BF16 || precision == RC_PREC_FP16_FP8,
               "rc_alt_corr_pyramid_setup: precision must be RC_PREC_BF16X3, RC_PREC_BF16 or RC_PREC_FP16_FP8");
  RC_CHECK_ARG(precision == RC_PREC_BF16 || (fmap1_lo && fmap2_lo), "rc_alt_corr_pyramid_setup: BF16X3 needs the lo parts");
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(fmap2_hi[l] != nullptr, "rc_alt_corr_pyramid_setup: level %d pointer is null", l);
    RC_CHECK_ARG(precision == RC_PREC_BF16 || fmap2_lo[l] != nullptr, "rc_alt_corr_pyramid_setup: level %d lo pointer is null", l);
  }
  return alt_tc_setup(setup, fmap1_hi, fmap1_lo, fmap2_hi, fmap2_lo, B, H1, W1, H2, W2, (C + 63) / 64 * 64, num_levels, radius,
                      precision);
}

int rc_alt_corr_pyramid_run(const void *setup, const float *coords, float *out, float scale, void *plan, size_t plan_bytes,
                            void *stream) {
  RC_CHECK_ARG(setup && coords && out, "rc_alt_corr_pyramid_run: null pointer");
  RC_CHECK_ARG(scale != 0.0f, "rc_alt_corr_pyramid_run: scale must be given (1/sqrt(C) for core/corr.py:91)");
  return alt_tc_run(setup, coords, out, scale, plan, plan_bytes, static_cast<cudaStream_t>(stream));
}

size_t rc_alt_corr_pyramid_plan_bytes(int B, int H, int W, int num_levels) {
  if (B <= 0 || H <= 0 || W <= 0 || num_levels <= 0) return 0;
  return alt_tc_plan_bytes(B, H, W, num_levels);
}

int rc_alt_corr_pyramid_fwd(const void *fmap1_hi, const void *fmap1_lo, const void *const *fmap2_hi,
                            const void *const *fmap2_lo, const float *coords, float *out, int B, int H, int W,
                            int C, int num_levels, int radius, int precision, float scale, void *plan,
                            size_t plan_bytes, void *stream) {
  if (int rc = check_geometry("rc_alt_corr_pyramid_fwd", B, H, W, num_levels)) return rc;
  RC_CHECK_ARG(fmap1_hi && fmap2_hi && coords && out, "rc_alt_corr_pyramid_fwd: null pointer");
  RC_CHECK_ARG(precision == RC_PREC_BF16X3 || precision == RC_PREC_BF16,
               "rc_alt_corr_pyramid_fwd: precision must be RC_PREC_BF16X3 or RC_PREC_BF16");
  RC_CHECK_ARG(precision == RC_PREC_BF16 || (fmap1_lo && fmap2_lo), "rc_alt_corr_pyramid_fwd: BF16X3 needs the lo parts");
  RC_CHECK_ARG(C > 0, "rc_alt_corr_pyramid_fwd: C must be positive");
  for (int l = 0; l < num_levels; ++l) {
    RC_CHECK_ARG(fmap2_hi[l] != nullptr, "rc_alt_corr_pyramid_fwd: level %d pointer is null", l);
    RC_CHECK_ARG(precision == RC_PREC_BF16 || fmap2_lo[l] != nullptr, "rc_alt_corr_pyramid_fwd: level %d lo pointer is null", l);
  }
  if (scale == 0.0f) scale = 1.0f / sqrtf((float)C);
  return launch_alt_tc(fmap1_hi, fmap1_lo, fmap2_hi, fmap2_lo, coords, out, B, H, W, (C + 63) / 64 * 64, num_levels,
                       radius, precision, scale, plan, plan_bytes, static_cast<cudaStream_t>(stream));
}

int rc_alt_corr_bwd(const float *fmap1, const float *fmap2, const float *coords,
                    const float *corr_grad, float *fmap1_grad, float *fmap2_grad, int B, int N,
                    int H1, int W1, int H2, int W2, int C, int radius, void *stream) {
  if (int rc = check_alt("rc_alt_corr_bwd", B, N, H1, W1, H2, W2, C, radius)) return rc;
  RC_CHECK_ARG(fmap1 && fmap2 && coords && corr_grad && fmap1_grad && fmap2_grad,
               "rc_alt_corr_bwd: null pointer");
  return launch_alt_bwd(fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, B, N, H1, W1, H2,
                        W2, C, radius, static_cast<cudaStream_t>(stream));
}

static int check_conv(const char *fn, int Cout, int Cin, int KH, int KW) {
  RC_CHECK_ARG(Cout > 0 && Cout <= 256 && Cout % 16 == 0, "%s: Cout %d must be a multiple of 16, at most 256", fn, Cout);
  RC_CHECK_ARG(Cin > 0 && Cin <= 4096, "%s: Cin %d out of range", fn, Cin);
  RC_CHECK_ARG(KH > 0 && KW > 0 && (KH & 1) && (KW & 1) && (KW == 1 ? (8 + KH - 1) * 16 : KH * (128 + KW - 1)) <= 400,
               "%s: %dx%d filters unsupported (odd sizes whose halo tile fits 400 pixels: 1x1, 3x3, 1x5, 3x5, KHx1 up to 17x1)", fn, KH, KW);
  return RC_OK;
}

size_t rc_conv_weight_bytes(int Cout, int Cin, int KH, int KW) {
  if (Cout <= 0 || Cin <= 0 || KH <= 0 || KW <= 0) return 0;
  return conv_weight_bytes(Cout, Cin, KH, KW);
}

int rc_conv_prep_weight(const float *weight, void *wprep, int Cout, int Cin, int KH, int KW, void *stream) {
  RC_CHECK_ARG(weight && wprep, "rc_conv_prep_weight: null pointer");
  if (int rc = check_conv("rc_conv_prep_weight", Cout, Cin, KH, KW)) return rc;
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0, "rc_conv_prep_weight: wprep must be 128-byte aligned");
  return launch_conv_prep_weight(weight, wprep, Cout, Cin, KH, KW, static_cast<cudaStream_t>(stream));
}

int rc_conv2d_fwd(const float *x1, int C1, const float *x2, const void *wprep, const float *bias, float *out, int B, int Cin,
                  int Cout, int H, int W, int KH, int KW, int act, void *stream) {
  RC_CHECK_ARG(x1 && wprep && out, "rc_conv2d_fwd: null pointer");
  RC_CHECK_ARG(B > 0 && H > 0 && W > 0 && (long long)B * H * ((W + 127) / 128) < 0x7fffffffLL, "rc_conv2d_fwd: bad sizes");
  if (int rc = check_conv("rc_conv2d_fwd", Cout, Cin, KH, KW)) return rc;
  RC_CHECK_ARG(C1 > 0 && C1 <= Cin && (C1 == Cin || x2 != nullptr), "rc_conv2d_fwd: C1 = %d of Cin = %d needs a second input", C1, Cin);
  RC_CHECK_ARG(act >= RC_ACT_NONE && act <= RC_ACT_TANH, "rc_conv2d_fwd: unknown activation %d", act);
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0, "rc_conv2d_fwd: wprep must be 128-byte aligned");
  return launch_conv_tc(x1, C1, x2, wprep, bias, out, B, Cin, Cout, H, W, KH, KW, act, static_cast<cudaStream_t>(stream));
}

int rc_conv2d_ex_fwd(const float *x1, int C1, long long x1_bs, const float *x2, long long x2_bs, const void *wprep,
                     const float *bias, float *out, long long out_bs, int store_channels, int B, int Cin, int Cout, int H,
                     int W, int KH, int KW, int act, int epilogue, int split, const float *aux1, long long aux1_bs,
                     const float *aux2, long long aux2_bs, void *stream) {
  const char *fn = "rc_conv2d_ex_fwd";
  RC_CHECK_ARG(store_channels >= 0 && store_channels <= Cout, "%s: store_channels %d outside 0..Cout", fn, store_channels);
  RC_CHECK_ARG(x1 && wprep && out, "%s: null pointer", fn);
  RC_CHECK_ARG(B > 0 && H > 0 && W > 0 && (long long)B * H * ((W + 127) / 128) < 0x7fffffffLL, "%s: bad sizes", fn);
  if (int rc = check_conv(fn, Cout, Cin, KH, KW)) return rc;
  RC_CHECK_ARG(C1 > 0 && C1 <= Cin && (C1 == Cin || x2 != nullptr), "%s: C1 = %d of Cin = %d needs a second input", fn, C1, Cin);
  RC_CHECK_ARG(act >= RC_ACT_NONE && act <= RC_ACT_TANH, "%s: unknown activation %d", fn, act);
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(wprep) & 127) == 0, "%s: wprep must be 128-byte aligned", fn);
  const long long hw = (long long)H * W;
  RC_CHECK_ARG((x1_bs == 0 || x1_bs >= hw * C1) && (x2_bs == 0 || x2_bs >= hw * (Cin - C1)) && (out_bs == 0 || out_bs >= hw * (store_channels ? store_channels : Cout)),
               "%s: a batch stride is smaller than the tensor it strides", fn);
  RC_CHECK_ARG(epilogue >= RC_EPI_NONE && epilogue <= RC_EPI_GRU, "%s: unknown epilogue %d", fn, epilogue);
  ConvExtra ex{};
  ex.x1_bs = x1_bs; ex.x2_bs = x2_bs; ex.out_bs = out_bs;
  ex.store_channels = store_channels;
  ex.epi = epilogue;
  if (epilogue == RC_EPI_MUL_FROM) {
    RC_CHECK_ARG(aux1 && split >= 0 && split <= Cout, "%s: RC_EPI_MUL_FROM needs aux1 and 0 <= split <= Cout", fn);
    ex.split = split; ex.aux1 = aux1; ex.aux1_bs = aux1_bs ? aux1_bs : hw * (Cout - split);
  } else if (epilogue == RC_EPI_GRU) {
    RC_CHECK_ARG(aux1 && aux2, "%s: RC_EPI_GRU needs aux1 (z) and aux2 (h)", fn);
    ex.aux1 = aux1; ex.aux2 = aux2;
    ex.aux1_bs = aux1_bs ? aux1_bs : hw * Cout; ex.aux2_bs = aux2_bs ? aux2_bs : hw * Cout;
  }
  return launch_conv_tc(x1, C1, x2, wprep, bias, out, B, Cin, Cout, H, W, KH, KW, act, static_cast<cudaStream_t>(stream), &ex);
}

int rc_small_conv_relu_fwd(const float *in, const float *weight, const float *bias, float *out, int B, int Cin, int Cout, int H,
                           int W, int K, int relu, void *stream) {
  RC_CHECK_ARG(in && weight && out, "rc_small_conv_relu_fwd: null pointer");
  RC_CHECK_ARG(B > 0 && B <= 65535 && Cout > 0 && H > 0 && W > 0, "rc_small_conv_relu_fwd: bad sizes");
  RC_CHECK_ARG(Cin >= 1 && Cin <= 4 && K >= 1 && K <= 7 && (K & 1), "rc_small_conv_relu_fwd: Cin 1..4 and K odd, at most 7 (got %d, %d)", Cin, K);
  return launch_small_conv_relu(in, weight, bias, out, B, Cin, Cout, H, W, K, relu, static_cast<cudaStream_t>(stream));
}

int rc_flow_update(float *coords1, const float *delta, long long delta_bs, const float *coords0, float *flow_out,
                   long long flow_bs, float *flow_out2, long long flow2_bs, int B, int H, int W, void *stream) {
  RC_CHECK_ARG(coords1 && delta && coords0 && flow_out, "rc_flow_update: null pointer");
  RC_CHECK_ARG(B > 0 && H > 0 && W > 0, "rc_flow_update: bad sizes");
  const long long hw2 = 2LL * H * W;
  return launch_flow_update(coords1, delta, delta_bs ? delta_bs : hw2, coords0, flow_out, flow_bs ? flow_bs : hw2, flow_out2,
                            flow2_bs ? flow2_bs : hw2, B, H, W, static_cast<cudaStream_t>(stream));
}

int rc_convex_upsample_fwd(const float *flow, const float *mask, float *out, int N, int H, int W, void *stream) {
  RC_CHECK_ARG(flow && mask && out, "rc_convex_upsample_fwd: null pointer");
  RC_CHECK_ARG(N > 0 && H > 0 && W > 0 && (long long)N * H * W < 0x7fffffffLL / 64, "rc_convex_upsample_fwd: bad sizes");
  RC_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "rc_convex_upsample_fwd: out must be 16-byte aligned");
  return launch_convex_upsample(flow, mask, out, N, H, W, static_cast<cudaStream_t>(stream));
}

size_t rc_alt_corr_bwd_det_workspace_bytes(int B, int H2, int W2, int C) {
  if (B <= 0 || H2 <= 0 || W2 <= 0 || C <= 0) return 0;
  return alt_bwd_det_workspace_bytes(B, H2, W2, C);
}

int rc_alt_corr_bwd_det(const float *fmap1, const float *fmap2, const float *coords, const float *corr_grad,
                        float *fmap1_grad, float *fmap2_grad, int B, int N, int H1, int W1, int H2, int W2, int C, int radius,
                        void *workspace, size_t workspace_bytes, void *stream) {
  if (int rc = check_alt("rc_alt_corr_bwd_det", B, N, H1, W1, H2, W2, C, radius)) return rc;
  RC_CHECK_ARG(fmap1 && fmap2 && coords && corr_grad && fmap1_grad && fmap2_grad, "rc_alt_corr_bwd_det: null pointer");
  const size_t need = alt_bwd_det_workspace_bytes(B, H2, W2, C);
  if (workspace == nullptr || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0) {
    set_error("rc_alt_corr_bwd_det: workspace missing, not 256-byte aligned or too small (%zu < %zu bytes)", workspace_bytes, need);
    return RC_ERR_WORKSPACE;
  }
  return launch_alt_bwd_det(fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, B, N, H1, W1, H2, W2, C, radius, workspace,
                            static_cast<cudaStream_t>(stream));
}

int rc_bilinear_sampler_fwd(const float *img, const float *coords, float *out, float *mask, int M,
                            int C, int H, int W, int Ho, int Wo, void *stream) {
  RC_CHECK_ARG(img && coords && out, "rc_bilinear_sampler_fwd: null pointer");
  RC_CHECK_ARG(M > 0 && C > 0 && H > 0 && W > 0 && Ho > 0 && Wo > 0,
               "rc_bilinear_sampler_fwd: all sizes must be positive");
  return launch_bilinear_fwd(img, coords, out, mask, M, C, H, W, Ho, Wo, static_cast<cudaStream_t>(stream));
}

int rc_bilinear_sampler_bwd(const float *grad_out, const float *coords, float *grad_img, int M, int C,
                            int H, int W, int Ho, int Wo, void *stream) {
  RC_CHECK_ARG(grad_out && coords && grad_img, "rc_bilinear_sampler_bwd: null pointer");
  RC_CHECK_ARG(M > 0 && C > 0 && H > 0 && W > 0 && Ho > 0 && Wo > 0,
               "rc_bilinear_sampler_bwd: all sizes must be positive");
  return launch_bilinear_bwd(grad_out, coords, grad_img, M, C, H, W, Ho, Wo, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
