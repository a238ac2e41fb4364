// Shared host/device helpers for libraftcorr (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/raftcorr.h"

struct CUtensorMap_st;  // <cuda.h>'s tensor-map descriptor (CUtensorMap), forward-declared: not every file includes cuda.h

namespace rc {

// thread-local last-error string (set by the host entry points)
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

#define RC_CHECK_ARG(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      rc::set_error(__VA_ARGS__);        \
      return RC_ERR_INVALID;             \
    }                                    \
  } while (0)

#define RC_CUDA(call)                                        \
  do {                                                       \
    cudaError_t e__ = (call);                                \
    if (e__ != cudaSuccess) return rc::cuda_fail(e__, #call); \
  } while (0)

// Programmatic dependent launch between this library's back-to-back kernels (on unless
// RAFT_CORR_PDL=0; read once per process)
bool pdl_enabled();

#ifdef __CUDACC__
// <<<grid, block, smem, st>>> with cudaLaunchAttributeProgrammaticStreamSerialization: the kernel may
// be set up while its predecessor in the stream drains; it calls pdl_wait() before touching memory.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// cuTensorMapEncodeTiled fetched through the runtime (the library links no driver stub: it loads on a machine without a
// driver); bf16 / 16-bit elements, SWIZZLE_128B, no interleave, zero fill outside the tensor.  `rank` entries of dims / box,
// rank-1 of strides (bytes).  One definition for every tensor-core kernel of the library (api.cu).
int encode_tmap_sw128(::CUtensorMap_st *tm, const void *base, int rank, const unsigned long long *dims,
                      const unsigned long long *strides_bytes, const unsigned int *box, const char *who);
// fp32 elements, no swizzle (zero fill outside the tensor, negative start coordinates included)
int encode_tmap_f32(::CUtensorMap_st *tm, const void *base, int rank, const unsigned long long *dims,
                      const unsigned long long *strides_bytes, const unsigned int *box, const char *who);

struct LevelPtrs {
  float *p[RC_MAX_LEVELS];
};
struct ConstLevelPtrs {
  const float *p[RC_MAX_LEVELS];
};

// TILED layout: query pixels whose tiles are interleaved (include/raftcorr.h) = one warp of the
// build epilogue
constexpr int kTileGroup = RC_TILE_GROUP;
static_assert(kTileGroup == 32, "RC_TILE_GROUP is the build epilogue's warp: 32");

// host launchers implemented in the .cu files (all return RC_* codes)
int launch_build_simt(const float *f1, const float *f2, int B, int D, int h, int w,
                      const LevelPtrs &lv, int L, float scale, cudaStream_t st);
int launch_build_tc(const float *f1, const float *f2, int B, int D, int h, int w,
                    const LevelPtrs &lv, int L, int precision, float scale, void *ws,
                    size_t ws_bytes, cudaStream_t st);
size_t build_tc_workspace_bytes(int B, int D, int h, int w, int precision);
int launch_build_tc_noalloc(const float *f1, const float *f2, int B, int D, int h, int w,
                            const LevelPtrs &lv, int L, int precision, float scale, cudaStream_t st);
int launch_build_tc_tiled(const float *f1, const float *f2, int B, int D, int h, int w,
                          const LevelPtrs &lv, int L, int precision, float scale, void *ws,
                          size_t ws_bytes, cudaStream_t st);
int launch_build_tc_tiled_prepared(const void *a_hi, const void *a_lo, const void *b_hi, const void *b_lo, int B, int Dp, int h,
                                   int w, const LevelPtrs &lv, int L, int precision, float scale, cudaStream_t st);
size_t conv1x1_weight_bytes(int Cout, int Cin);
int launch_conv1x1_prep_weight(const float *w, void *wp, int Cout, int Cin, cudaStream_t st);
int launch_conv1x1_prep(const float *x, const void *wp, const float *bias, void *hi, void *lo, int Bt, int Cin, int Cout, int HW,
                        cudaStream_t st);
int launch_relayout(const float *src, float *dst, int B, int N, int hl, int wl, bool to_tiled,
                    cudaStream_t st);
int launch_relayout16(const void *src, void *dst, int B, int N, int hl, int wl, bool to_tiled,
                      cudaStream_t st);
int launch_lookup_tiled16_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h,
                              int w, int L, int r, cudaStream_t st);
int launch_lookup_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h,
                      int w, int L, int r, cudaStream_t st);
int launch_lookup_tiled_fwd(const ConstLevelPtrs &lv, const float *coords, float *out, int B, int h,
                            int w, int L, int r, cudaStream_t st);
size_t convc1_weight_bytes(int Cout, int L, int r);
int launch_convc1_prep_weight(const float *w, void *wp, int Cout, int L, int r, int bf16_ops, cudaStream_t st);
int launch_lookup_convc1(const ConstLevelPtrs &lv, bool vol16, const float *coords, const void *wp, const float *bias,
                         float *out, int B, int h, int w, int L, int r, int Cout, int bf16_ops, cudaStream_t st);
int launch_lookup_bwd(const float *gout, const float *coords, const LevelPtrs &glv, int B, int h,
                      int w, int L, int r, cudaStream_t st);
size_t build_bwd_workspace_bytes(int B, int D, int h, int w);
int launch_build_bwd(const ConstLevelPtrs &glv, const float *f1, const float *f2, float *df1,
                     float *df2, int B, int D, int h, int w, int L, float scale, void *ws,
                     size_t ws_bytes, cudaStream_t st);
size_t build_bwd_tc_workspace_bytes(int B, int D, int h, int w);
int launch_build_bwd_tc(const ConstLevelPtrs &glv, const float *f1, const float *f2, float *df1,
                        float *df2, int B, int D, int h, int w, int L, float scale, void *ws,
                        size_t ws_bytes, cudaStream_t st);
int launch_alt_fwd(const float *f1, const float *f2, const float *coords, float *corr, int B,
                   int N, int H1, int W1, int H2, int W2, int C, int r, cudaStream_t st);
int launch_alt_bwd(const float *f1, const float *f2, const float *coords, const float *cg,
                   float *g1, float *g2, int B, int N, int H1, int W1, int H2, int W2, int C,
                   int r, cudaStream_t st);

size_t alt_bwd_det_workspace_bytes(int B, int H2, int W2, int C);
int launch_alt_bwd_det(const float *f1, const float *f2, const float *coords, const float *cg, float *g1, float *g2, int B, int N,
                       int H1, int W1, int H2, int W2, int C, int r, void *ws, cudaStream_t st);
int launch_alt_prep(const float *fmap, void *hi, void *lo, int B, int C, int HW, cudaStream_t st);
int launch_alt_prep_pooled(const float *fmap, void *hi, void *lo, int B, int C, int H, int W, int level, cudaStream_t st);
int launch_alt_absmax(const float *fmap, unsigned *amax, int B, long long per_item, cudaStream_t st);
int launch_alt_prep_f8(const float *fmap, void *hi, void *pair, const unsigned *amax, int side, int B, int C, int H, int W,
                       int level, cudaStream_t st);
int alt_tc_set_norm(void *blob, const unsigned *amax1, const unsigned *amax2);
size_t alt_tc_plan_bytes(int B, int H1, int W1, int L);
size_t alt_tc_setup_bytes();
int alt_tc_setup(void *blob, const void *f1_hi, const void *f1_lo, const void *const *f2_hi, const void *const *f2_lo, int B,
                 int H1, int W1, int H2, int W2, int Cp, int L, int radius, int precision);
int alt_tc_run(const void *blob, const float *coords, float *out, float scale, void *plan_ws, size_t plan_bytes,
               cudaStream_t st);
int launch_alt_tc(const void *f1_hi, const void *f1_lo, const void *const *f2_hi, const void *const *f2_lo,
                  const float *coords, float *out, int B, int H1, int W1, int Cp, int L, int radius,
                  int precision, float scale, void *plan_ws, size_t plan_bytes, cudaStream_t st);
int launch_pool_tail(const LevelPtrs &lv, int B, int h, int w, int first, int L, cudaStream_t st);
size_t conv_weight_bytes(int Cout, int Cin, int KH, int KW);
int launch_conv_prep_weight(const float *w, void *wp, int Cout, int Cin, int KH, int KW, cudaStream_t st);
// what rc_conv2d_ex_fwd adds to rc_conv2d_fwd: batch strides (channel slices of wider tensors) and the fused GRU epilogues
struct ConvExtra {
  long long x1_bs, x2_bs, out_bs;  // floats between batch items; 0 = dense
  int store_channels;              // only the first so many filters are stored (0 = all): zero-padded filters write nothing
  int epi, split;
  const float *aux1, *aux2;
  long long aux1_bs, aux2_bs;
};
int launch_conv_tc(const float *x1, int C1, const float *x2, const void *wp, const float *bias, float *out, int B, int Cin, int Cout,
                   int H, int W, int KH, int KW, int act, cudaStream_t st, const ConvExtra *ex = nullptr);
int launch_small_conv_relu(const float *in, const float *w, const float *bias, float *out, int B, int Cin, int Cout, int H,
                           int W, int K, int relu, cudaStream_t st);
int launch_flow_update(float *coords1, const float *delta, long long delta_bs, const float *coords0, float *flow_out,
                       long long flow_bs, float *flow_out2, long long flow2_bs, int B, int H, int W, cudaStream_t st);
int launch_convex_upsample(const float *flow, const float *mask, float *out, int N, int H, int W, cudaStream_t st);
int launch_bilinear_fwd(const float *img, const float *coords, float *out, float *mask, int M, int C,
                        int H, int W, int Ho, int Wo, cudaStream_t st);
int launch_bilinear_bwd(const float *gout, const float *coords, float *gimg, int M, int C, int H,
                        int W, int Ho, int Wo, cudaStream_t st);

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__
// kernel-parameter pointer tables are selected with an unrolled compare chain: a dynamic index
// would make the compiler copy the whole table to local memory
template <typename P>
__device__ __forceinline__ auto pick_level(const P &t, int lvl) {
  auto r = t.p[0];
#pragma unroll
  for (int l = 1; l < RC_MAX_LEVELS; ++l)
    if (lvl == l) r = t.p[l];
  return r;
}

// floor() of a coordinate as an int that is safe for any float (NaN/Inf/huge -> far outside the
// grid with a zero fraction); branch-free
__device__ __forceinline__ int safe_floor_int(float v, float *frac) {
  const float f = floorf(v);
  const bool sane = fabsf(v) < 16777216.0f;  // false for NaN / Inf / huge
  *frac = sane ? v - f : 0.0f;
  // fmaxf/fminf drop a NaN operand; everything beyond +-2^24 is outside any feature grid anyway
  return __float2int_rz(fminf(fmaxf(f, -16777216.0f), 16777216.0f));
}

// Programmatic dependent launch (no-ops in a kernel launched without the attribute).
//   pdl_launch_dependents: the next kernel of the stream may be set up as soon as every CTA of this
//     grid has got here (its CTAs still only become resident as ours leave).
//   pdl_wait: returns once every earlier kernel of the stream has completed and its writes are visible.
//     Nothing that an earlier kernel may have produced is read, and no global memory is written,
//     before it.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ float ld_stream(const float *p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream(float *p, float v) {
  asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
#endif

}  // namespace rc
