/*
 * raftcorr.h -- C ABI of libraftcorr.so: RAFT's correlation hot path on B200 (sm_100a).
 *
 * This is the drop-in boundary.  Every entry point takes plain device pointers, sizes and a
 * CUDA stream (as void*, i.e. a cudaStream_t); there are no torch types in any signature and
 * the library links nothing but the CUDA runtime.  Outputs are CALLER-allocated.  The
 * reference's extension instead allocates its outputs with torch::zeros
 * (alt_cuda_corr/correlation_kernel.cu:272-273,304-307); the Python host layer
 * (raft_ros_b200/) does that allocation with torch and then calls in here.
 *
 * Conventions
 *   - return value: RC_OK (0) or an RC_ERR_* code; rc_last_error() returns a thread-local
 *     message for the last failing call on this thread.  The reference signals errors by
 *     TORCH_CHECK -> RuntimeError (alt_cuda_corr/correlation.cpp:19-21); the Python layer
 *     turns a non-zero code into RuntimeError with the same wording where one exists.
 *   - all kernels are launched on `stream` of the CURRENT device; the library never
 *     synchronises the device and never allocates device memory.  (The reference launches
 *     on the legacy default stream, correlation_kernel.cu:278,313.)
 *   - all tensors are contiguous fp32 unless stated.  Level l of a pyramid built from an
 *     h x w feature grid is (B*h*w, h>>l, w>>l): floor-mode 2x2 average pooling, exactly
 *     F.avg_pool2d(x, 2, stride=2) at core/corr.py:26.
 *   - window channel order of a lookup: channel = l*(2r+1)^2 + i*(2r+1) + j where i is the
 *     x offset and j the y offset (core/corr.py:37-43; the same `iy + rd*ix` order as
 *     alt_cuda_corr/correlation_kernel.cu:92-95).
 */
#ifndef RAFTCORR_H_
#define RAFTCORR_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RC_OK 0
#define RC_ERR_INVALID 1     /* bad argument (null pointer, non-positive size, ...) */
#define RC_ERR_CUDA 2        /* a CUDA runtime / driver call failed */
#define RC_ERR_UNSUPPORTED 3 /* valid in the reference but outside what this build covers */
#define RC_ERR_WORKSPACE 4   /* workspace missing or too small */

#define RC_MAX_LEVELS 8
#define RC_MAX_RADIUS 8

/* arithmetic used for the all-pairs contraction (the `precision` argument of the build) */
#define RC_PREC_FP32 0   /* fp32 FMA on CUDA cores; matches cuBLAS SGEMM to ~1e-6        */
#define RC_PREC_BF16X3 1 /* tcgen05, bf16 hi/lo split operands, 3 MMAs, fp32 accumulate;  */
                         /* normwise error ~1e-5: the fp32-class tensor-core path         */
#define RC_PREC_BF16 2   /* tcgen05, single bf16 pass, fp32 accumulate; ~2e-3 normwise    */
#define RC_PREC_BF16_VOL16 3 /* RC_PREC_BF16 with the pyramid STORED as fp16 (saturating to  */
                             /* +-65504): half the bytes of the build and of every lookup's  */
                             /* reads; tiled layout only (rc_corr_build_pyramid_tiled +       */
                             /* rc_corr_lookup_tiled16_fwd); +2.4e-4 relative storage error   */

#define RC_PREC_FP16_FP8 4 /* tcgen05, fp16 products plus 8-bit compensation: every pixel's D-vector times 2^k = hi + lo */
                          /* with hi fp16 (k from that pixel's largest magnitude, found by the layout prepass itself);   */
                          /* the products hi*hi run in fp16 and the two cross terms hi*lo + lo*hi as ONE e4m3            */
                          /* contraction at twice the rate: 2/3 of BF16X3's tensor work; normwise error ~1e-5            */
                          /* (<= 1e-4 relative whatever the magnitudes); forward only                                    */

int rc_version(void);
const char *rc_last_error(void);

/*
 * Tuning knob, no counterpart in the reference: the L2 fetch granularity of the CURRENT device's
 * context (cudaLimitMaxL2FetchGranularity; 32, 64 or 128 bytes, the driver default is 64).  The
 * lookup reads (2r+2)-float row segments (40 B for r=4) scattered over the volume, so a 64 B
 * granule over-fetches by ~1.5x; 32 B cuts the lookup's DRAM reads.  It is a context-wide hint
 * that also applies to every other kernel of the process, so the library never sets it by
 * itself: the host layer does so only when asked (RAFT_CORR_L2_FETCH=32).
 */
int rc_set_l2_fetch_granularity(int bytes);
int rc_get_l2_fetch_granularity(int *bytes);

/* h_l, w_l of level `level` (h>>level, w>>level); RC_ERR_INVALID if it would be empty. */
int rc_level_dims(int h, int w, int level, int *hl, int *wl);

/* Bytes of scratch rc_corr_build_pyramid needs for this problem (0 for RC_PREC_FP32). */
size_t rc_corr_build_workspace_bytes(int B, int D, int h, int w, int precision);

/*
 * CorrBlock.corr + the pyramid of CorrBlock.__init__ in one pass.
 * Replaces: core/corr.py:52-60 (view, matmul, /sqrt(D)) and core/corr.py:21-27
 * (reshape to (B*h*w,1,h,w), num_levels-1 x F.avg_pool2d).
 *   fmap1, fmap2 : (B, D, h, w) fp32 NCHW
 *   level_ptrs   : HOST array of num_levels DEVICE pointers; level l is written as
 *                  (B*h*w, h>>l, w>>l) fp32, every element exactly once
 *   scale        : multiplies the contraction; pass 0 for the reference's 1/sqrt(D)
 *   workspace    : device scratch of >= rc_corr_build_workspace_bytes(...) (may be NULL when 0)
 */
int rc_corr_build_pyramid(const float *fmap1, const float *fmap2, int B, int D, int h, int w,
                          float *const *level_ptrs, int num_levels, int precision, float scale,
                          void *workspace, size_t workspace_bytes, void *stream);

/*
 * rc_corr_build_pyramid without a caller workspace and without any allocation.
 * No counterpart in the reference (its build is torch.matmul; torch allocates as it likes).
 * It exists for hosts that may not allocate device memory while running: a real-time inference
 * loop (the ROS node's worker thread) that cannot take cudaMalloc's unbounded latency and its
 * device-wide synchronisation, or a process whose whole device memory already belongs to
 * another pool, where cudaMalloc would fail.  The scratch is a fixed array inside the library
 * image: each GPU that loads libraftcorr.so holds RC_STATIC_SCRATCH_BYTES until exit.
 *   - only the tensor-core precisions (RC_PREC_BF16X3, RC_PREC_BF16); RC_PREC_FP32 needs no
 *     scratch and is served by rc_corr_build_pyramid;
 *   - a problem whose rc_corr_build_workspace_bytes(...) exceeds RC_STATIC_SCRATCH_BYTES is
 *     rejected with RC_ERR_WORKSPACE and nothing is launched (32 MiB covers two 440x1024 pairs in
 *     BF16X3, four in BF16);
 *   - warm-up: the first call on a device looks up the array's device address, and -- because
 *     the CUDA runtime loads kernels lazily -- the first launch of each kernel variant (one per
 *     precision; one more for num_levels > 4) may itself make the driver allocate or synchronise
 *     while it loads the code.  So make one call per device AND per precision you are going to use
 *     (any small problem will do) before the phase in which nothing may allocate, and before
 *     beginning any stream capture on that device.  The looked-up address is cached per device
 *     for the life of the process: it does not survive cudaDeviceReset;
 *   - every call on a device uses the same array and finds the previous call's contents in it.
 *     Nothing must be cleared: each call rewrites all it reads.  Calls that overlap on the GPU
 *     race on the array; order them (same stream, or an event between streams).
 * Inputs, outputs, return codes: exactly as rc_corr_build_pyramid.
 */
#define RC_STATIC_SCRATCH_BYTES (32u << 20)
int rc_corr_build_pyramid_noalloc(const float *fmap1, const float *fmap2, int B, int D, int h, int w,
                                  float *const *level_ptrs, int num_levels, int precision, float scale,
                                  void *stream);

/*
 * CorrBlock.__call__: the (2r+1)^2 bilinear window of every level around coords / 2^l.
 * Replaces: core/corr.py:29-50 and, through it, core/utils/utils.py:57-71
 * (bilinear_sampler -> F.grid_sample bilinear, zeros padding, align_corners=True).
 *   level_ptrs : HOST array of num_levels DEVICE pointers (as written by the build)
 *   coords     : (B, 2, h, w) fp32, channel 0 = x, 1 = y (core/utils/utils.py:74-77)
 *   out        : (B, num_levels*(2r+1)^2, h, w) fp32, fully overwritten
 */
int rc_corr_lookup_fwd(const float *const *level_ptrs, const float *coords, float *out, int B,
                       int h, int w, int num_levels, int radius, void *stream);

/*
 * The TILED volume layout (no counterpart in the reference; it is what CorrBlock keeps internally
 * when no gradient is wanted).  Level l of row p = (b, query pixel q) -- h_l x w_l floats in the
 * row layout -- is cut into tiles of 2 rows x 8 floats (64 B = 4 chunks of 16 B, chunk
 * k = 2*(y%2) + (x%8)/4), and the tiles of G = RC_TILE_GROUP = 32 consecutive query pixels are
 * interleaved:
 *     [B][ceil(h*w/G)][hp/2][wp/8][G pixels][4 chunks][4],  hp = h_l rounded up to 2, wp = w_l rounded up to 8,
 * pixel u = q%G of a group keeping chunk k at position k ^ ((u >> 1) & 3).  Element (b, q, y, x):
 *     ((((b*ceil(h*w/G) + q/G) * (hp/2) + y/2) * (wp/8) + x/8) * G + u) * 16 + (k ^ ((u>>1)&3)) * 4 + x%4.
 * rc_corr_tiled_level_elems floats per level, zeros in every padding element a lookup can reach.
 * Why: (1) a 10x10 lookup window covers a few 64 B tiles instead of ten 40 B row segments that
 * straddle DRAM fill granules; (2) neighbouring query pixels look at neighbouring places, so the
 * tiles a group of pixels needs are contiguous and DRAM sees long bursts (lookup 60 -> 47 us);
 * (3) the build's epilogue warp (32 query pixels = one group) holds two whole tiles per pixel = 4 KB
 * of contiguous bytes, which it lays out in shared memory without bank conflicts (hence the chunk
 * swizzle) and hands to the bulk-copy engine -- for any h and w (the row layout degrades to
 * partial-sector writes on ragged widths: 2.3x slower build at 47x156).
 *
 * rc_corr_build_pyramid_tiled : rc_corr_build_pyramid writing that layout.  RC_PREC_BF16X3 /
 *                     RC_PREC_BF16 (/ RC_PREC_BF16_VOL16, below) only, num_levels <= 4, level pointers 64-byte aligned and
 *                     rc_corr_tiled_level_elems floats each.  Same workspace as the row-layout build.
 * rc_corr_lookup_tiled_fwd    : rc_corr_lookup_fwd reading it.  Same coords / out contract;
 *                     radius 3 (raft-small) or 4 (raft-things).
 * rc_corr_relayout_level      : converts one level (hl x wl per row, N = h*w query pixels per batch
 *                     item) between the layouts (to_tiled != 0: rows -> tiled, padding zeroed;
 *                     0: tiled -> rows).
 */
#ifndef RC_TILE_GROUP
#define RC_TILE_GROUP 32 /* fixed: one warp of the build epilogue */
#endif
int rc_corr_tile_group(void); /* the RC_TILE_GROUP the library was built with */
int rc_corr_tiled_level_elems(int B, int h, int w, int level, size_t *elems);
int rc_corr_build_pyramid_tiled(const float *fmap1, const float *fmap2, int B, int D, int h, int w,
                                float *const *level_ptrs, int num_levels, int precision, float scale,
                                void *workspace, size_t workspace_bytes, void *stream);
int rc_corr_lookup_tiled_fwd(const float *const *level_ptrs, const float *coords, float *out, int B,
                             int h, int w, int num_levels, int radius, void *stream);
int rc_corr_relayout_level(const float *src, float *dst, int B, int N, int hl, int wl, int to_tiled,
                           void *stream);

/*
 * Producer-side fusion: the feature encoder's output layer, the 1x1 convolution `conv2` of
 * BasicEncoder / SmallEncoder (core/extractor.py:144,184 -- 128 -> 256; :217,256 -- 96 -> 128), writing
 * its result directly as the K-major bf16 hi / lo operand maps the tensor-core build reads, so that
 * the fp32 (B, 256, h, w) feature maps of core/raft.py:99-103 are never written and the build's own
 * layout prepass never runs:
 *     y[b, p, o] = bias[o] + sum_c weight[o, c] * x[b, c, p],   hi = bf16(y), lo = bf16(y - hi)
 * (error-compensated bf16 products, fp32 accumulation: fp32-class, like RC_PREC_BF16X3).
 *   x      : (Bt, Cin, H, W) fp32 NCHW, the input of conv2 (both images of the pairs: Bt = 2B when the
 *            encoder runs them as one batch, core/raft.py:100); Cin a multiple of 16, <= 128
 *   wprep  : weight (Cout, Cin) fp32 converted once per model by rc_fnet_conv1x1_prep_weight into
 *            rc_fnet_conv1x1_weight_bytes() bytes; Cout a multiple of 64, <= 256
 *   hi, lo : (Bt, H*W, Cout) bf16; lo may be NULL when only RC_PREC_BF16 / _VOL16 builds follow
 * rc_corr_build_pyramid_tiled_prepared is rc_corr_build_pyramid_tiled on such maps (fmap1 = the first
 * image's maps, fmap2 = the second's; D = Cout is also the channel pitch); no workspace is needed.
 */
size_t rc_fnet_conv1x1_weight_bytes(int Cout, int Cin);
int rc_fnet_conv1x1_prep_weight(const float *weight, void *wprep, int Cout, int Cin, void *stream);
int rc_fnet_conv1x1_prep(const float *x, const void *wprep, const float *bias, void *hi, void *lo, int Bt,
                         int Cin, int Cout, int H, int W, void *stream);
int rc_corr_build_pyramid_tiled_prepared(const void *fmap1_hi, const void *fmap1_lo, const void *fmap2_hi,
                                         const void *fmap2_lo, int B, int D, int h, int w,
                                         float *const *level_ptrs, int num_levels, int precision, float scale,
                                         void *stream);

/*
 * The 16-bit flavour of the tiled layout (RC_PREC_BF16_VOL16): the same structure with fp16
 * elements, tiles of 2 rows x 16 halves (still 64 B = 4 chunks of 16 B = 8 halves, chunk
 * k = 2*(y%2) + (x%16)/8), wp = w_l rounded up to 16:
 *     [B][ceil(h*w/32)][hp/2][wp/16][32 pixels][4 chunks][8 halves].
 * rc_corr_tiled16_level_elems : number of fp16 elements of a level.
 * rc_corr_build_pyramid_tiled  with precision = RC_PREC_BF16_VOL16 writes it (level pointers
 *                     then address fp16 data).
 * rc_corr_lookup_tiled16_fwd  : rc_corr_lookup_tiled_fwd reading it (fp32 interpolation and output).
 * rc_corr_relayout16_level    : tiled fp16 -> row-layout fp32 (to_tiled = 0) or back (rounding,
 *                     saturating; padding zeroed).
 */
int rc_corr_tiled16_level_elems(int B, int h, int w, int level, size_t *elems);
int rc_corr_lookup_tiled16_fwd(const void *const *level_ptrs, const float *coords, float *out, int B,
                               int h, int w, int num_levels, int radius, void *stream);
int rc_corr_relayout16_level(const void *src, void *dst, int B, int N, int hl, int wl, int to_tiled,
                             void *stream);

/*
 * The lookup FUSED into its only consumer, the motion encoder's 1x1 convolution convc1 + ReLU
 * (core/update.py:83,90: `cor = F.relu(self.convc1(corr))`; raft-small :66,72), so that the
 * (B, L*(2r+1)^2, h, w) lookup result of core/corr.py:29-50 is never written to or re-read from
 * global memory:
 *     out[b, o, y, x] = relu(bias[o] + sum_c weight[o, c] * CorrBlock.__call__(coords)[b, c, y, x])
 * levels : the TILED pyramid, fp32 (vol_fp16 = 0) or fp16 (vol_fp16 = 1), as for the tiled lookups.
 * wprep  : the convolution's weight (Cout, L*(2r+1)^2) fp32, converted once per model by
 *          rc_corr_convc1_prep_weight into rc_corr_convc1_weight_bytes() bytes of 16-bit operands
 *          (K of every level zero-padded to whole tensor-core K blocks).
 * bf16_ops : 0 = fp16 operands (taps and weights rounded to fp16, saturating -- the arithmetic
 *          the reference's --mixed_precision autocast gives this convolution), 1 = bf16 operands.
 *          One tensor-core pass, fp32 accumulation, fp32 bias / ReLU / output either way.
 * out    : (B, Cout, h, w) fp32.  Cout a multiple of 16, <= 256 (256 raft-things, 96 raft-small);
 *          radius 3 or 4; num_levels <= 4.
 * This is a stated-bounds fast path beside rc_corr_lookup_tiled_fwd (whose fp32 result followed
 * by an fp32 convolution is what the reference computes without autocast), not a replacement.
 */
size_t rc_corr_convc1_weight_bytes(int Cout, int num_levels, int radius);
int rc_corr_convc1_prep_weight(const float *weight, void *wprep, int Cout, int num_levels, int radius,
                               int bf16_ops, void *stream);
int rc_corr_lookup_convc1_fwd(const void *const *level_ptrs, int vol_fp16, const float *coords,
                              const void *wprep, const float *bias, float *out, int B, int h, int w,
                              int num_levels, int radius, int Cout, int bf16_ops, void *stream);

/*
 * Gradient of the lookup w.r.t. the pyramid (coords carry no gradient: they are detached at
 * core/raft.py:123).  Replaces what autograd does through core/corr.py:45 /
 * F.grid_sample's backward, but deterministically: row p of every level is only ever
 * touched by query p, so the scatter needs no atomics.
 *   grad_out        : (B, num_levels*(2r+1)^2, h, w) fp32
 *   grad_level_ptrs : HOST array of DEVICE pointers; gradients are ACCUMULATED (+=) so the
 *                     K iterations of one step add up in stream order; zero them once per step
 */
int rc_corr_lookup_bwd(const float *grad_out, const float *coords, float *const *grad_level_ptrs,
                       int B, int h, int w, int num_levels, int radius, void *stream);

/* Bytes of scratch rc_corr_build_bwd needs for this problem and precision. */
size_t rc_corr_build_bwd_workspace_bytes(int B, int D, int h, int w, int precision);

/*
 * Gradient of the build: folds the grad pyramid back to level 0 (avg_pool2d backward,
 * core/corr.py:26), applies 1/sqrt(D) (:60) and the two contractions of matmul's backward
 * (:58):  dfmap1[d,p] = sum_q dC[p,q] fmap2[d,q],  dfmap2[d,q] = sum_p dC[p,q] fmap1[d,p].
 *   dfmap1, dfmap2 : (B, D, h, w) fp32, fully overwritten
 *   precision      : RC_PREC_FP32 = fp32 FMA on CUDA cores; RC_PREC_BF16X3 / RC_PREC_BF16 = tcgen05
 *                    with bf16 hi/lo split operands (three MMAs, fp32 accumulate: gradients are
 *                    always fp32-class, also when the forward was single-pass bf16)
 *   workspace      : >= rc_corr_build_bwd_workspace_bytes(...), 256 B aligned
 */
int rc_corr_build_bwd(const float *const *grad_level_ptrs, const float *fmap1, const float *fmap2,
                      float *dfmap1, float *dfmap2, int B, int D, int h, int w, int num_levels,
                      int precision, float scale, void *workspace, size_t workspace_bytes, void *stream);

/*
 * alt_cuda_corr.forward: windowed dot products computed on the fly (no volume).
 * Replaces: alt_cuda_corr/correlation.cpp:23-33 -> correlation_kernel.cu:260-286 -> :18-119.
 *   fmap1  : (B, H1, W1, C) fp32 NHWC      fmap2 : (B, H2, W2, C) fp32 NHWC
 *   coords : (B, N, H1, W1, 2) fp32 [x, y]
 *   corr   : (B, N, (2r+1)^2, H1, W1) fp32, fully overwritten (no pre-zeroing needed);
 *            tap index = oy + (2r+1)*ox; NOT scaled by 1/sqrt(C) (the caller does that,
 *            core/corr.py:91)
 */
int rc_alt_corr_fwd(const float *fmap1, const float *fmap2, const float *coords, float *corr,
                    int B, int N, int H1, int W1, int H2, int W2, int C, int radius, void *stream);

/*
 * AlternateCorrBlock on tensor cores: every pyramid level of the on-the-fly correlation in ONE
 * launch, output already in CorrBlock's layout and scaled.
 * Replaces: the per-level loop of AlternateCorrBlock.__call__ (core/corr.py:74-91: 8 NHWC
 * permute-copies, num_levels x alt_cuda_corr.forward, stack, reshape, divide).
 *
 * rc_alt_prep_fmap converts one feature map once (AlternateCorrBlock.__init__ time):
 *   fmap (B, C, H, W) fp32 NCHW  ->  hi, lo : (B, H, W, Cp) bf16, Cp = C rounded up to 64,
 *   x ~= hi + lo (lo may be NULL when only RC_PREC_BF16 will be used).
 * rc_alt_corr_pyramid_fwd then takes fmap1 (hi, lo), HOST arrays of num_levels DEVICE pointers to
 * the prepared fmap2 levels (level l = the (H>>l, W>>l) average-pooled fmap2, core/corr.py:68-72)
 * and
 *   coords : (B, 2, H, W) fp32 [x, y] at level-0 scale (divided by 2^l inside, core/corr.py:85)
 *   out    : (B, num_levels*(2r+1)^2, H, W) fp32, fully overwritten, channel order as
 *            rc_corr_lookup_fwd; multiplied by scale (0 -> 1/sqrt(C), core/corr.py:91)
 *   plan   : device scratch of rc_alt_corr_pyramid_plan_bytes(...) bytes, 16 B aligned
 * radius must be 3 or 4 and num_levels <= 4 (raft-small / raft-things); other shapes use rc_alt_corr_fwd.
 *
 * rc_alt_prep_fmap_level is rc_alt_prep_fmap for pyramid level `level` (0..3) of the map, straight
 * from the level-0 fp32 map: the (H>>level, W>>level) average-pooled map of core/corr.py:68-72 is
 * formed in registers with the same nested 2x2 means (bit-identical to the avg_pool2d chain) and
 * never written as fp32.
 *
 * rc_alt_corr_pyramid_setup / _run split rc_alt_corr_pyramid_fwd into what an AlternateCorrBlock
 * fixes once (core/corr.py:64-72: the feature maps -> tensor maps, geometry) and what changes per call
 * (core/corr.py:74: coords).  `setup` is a HOST buffer of rc_alt_corr_pyramid_setup_bytes() bytes owned
 * by the caller; it holds no pointers to host memory and may be copied.  Here fmap1 may have its own
 * spatial size (H1, W1) -- a band of query rows of a larger frame, coords and out then are (B, 2, H1, W1)
 * and (B, L*(2r+1)^2, H1, W1) -- against the full (H2, W2) fmap2 pyramid: that is how one large pair is
 * split over several GPUs by query rows.
 *
 * RC_PREC_FP16_FP8 on this path (2/3 of BF16X3's tensor work and operand traffic, ~1e-5 normwise): the maps are
 * prepared by rc_alt_prep_fmap_level_f8 -- hi as fp16, and instead of lo the 8-bit pair map (same byte size) -- after
 * rc_alt_absmax has put the bits of max|x| per batch item into amax (B words, ZERO-FILLED by the caller; one call per
 * feature map; the level-0 magnitude serves every pooled level); side = 0 for fmap1, 1 for fmap2.  The setup takes the
 * pair maps in place of the lo maps with precision = RC_PREC_FP16_FP8, then rc_alt_corr_pyramid_set_norm hands it the
 * two amax arrays (device memory, alive as long as the setup is used).
 */
int rc_alt_absmax(const float *fmap, unsigned int *amax, int B, int C, int H, int W, void *stream);
int rc_alt_prep_fmap_level_f8(const float *fmap, void *hi, void *pair, const unsigned int *amax, int side, int B, int C,
                              int H, int W, int level, void *stream);
int rc_alt_corr_pyramid_set_norm(void *setup, const unsigned int *amax1, const unsigned int *amax2);
int rc_alt_prep_fmap(const float *fmap, void *hi, void *lo, int B, int C, int H, int W, void *stream);
int rc_alt_prep_fmap_level(const float *fmap, void *hi, void *lo, int B, int C, int H, int W, int level,
                           void *stream);
size_t rc_alt_corr_pyramid_setup_bytes(void);
int rc_alt_corr_pyramid_setup(void *setup, const void *fmap1_hi, const void *fmap1_lo,
                              const void *const *fmap2_hi, const void *const *fmap2_lo, int B, int H1, int W1,
                              int H2, int W2, int C, int num_levels, int radius, int precision);
int rc_alt_corr_pyramid_run(const void *setup, const float *coords, float *out, float scale, void *plan,
                            size_t plan_bytes, void *stream);
size_t rc_alt_corr_pyramid_plan_bytes(int B, int H, int W, int num_levels);
int rc_alt_corr_pyramid_fwd(const void *fmap1_hi, const void *fmap1_lo, const void *const *fmap2_hi,
                            const void *const *fmap2_lo, const float *coords, float *out, int B, int H, int W,
                            int C, int num_levels, int radius, int precision, float scale, void *plan,
                            size_t plan_bytes, void *stream);

/*
 * alt_cuda_corr.backward.
 * Replaces: alt_cuda_corr/correlation.cpp:36-48 -> correlation_kernel.cu:288-324 -> :122-256.
 *   corr_grad  : (B, N, (2r+1)^2, H1, W1)
 *   fmap1_grad : (B, H1, W1, C)  fully overwritten
 *   fmap2_grad : (B, H2, W2, C)  MUST be zero-filled by the caller; accumulated into
 * coords_grad is identically zero in the reference (:307, never written); the Python shim
 * returns zeros for it and this entry point does not take it.
 */
int rc_alt_corr_bwd(const float *fmap1, const float *fmap2, const float *coords,
                    const float *corr_grad, float *fmap1_grad, float *fmap2_grad, int B, int N,
                    int H1, int W1, int H2, int W2, int C, int radius, void *stream);

/*
 * rc_alt_corr_bwd with a DETERMINISTIC fmap2_grad.  The reference scatters fmap2_grad with float atomicAdd
 * (correlation_kernel.cu:229-238), so its low bits change from run to run; here the contributions are
 * accumulated as 64-bit fixed-point integers (scale chosen on the device from max|corr_grad| * max|fmap1| and
 * the number of contributions so that no sum can overflow; resolution far below fp32), whose sum does not depend
 * on the order of the atomics, and converted to fp32 once.  Same arguments and results (to fp32 rounding) as
 * rc_alt_corr_bwd, except: fmap2_grad need NOT be zeroed by the caller, inputs must be finite, and
 *   workspace : rc_alt_corr_bwd_det_workspace_bytes(B, H2, W2, C) bytes of device scratch, 256 B aligned.
 */
size_t rc_alt_corr_bwd_det_workspace_bytes(int B, int H2, int W2, int C);
int rc_alt_corr_bwd_det(const float *fmap1, const float *fmap2, const float *coords, const float *corr_grad,
                        float *fmap1_grad, float *fmap2_grad, int B, int N, int H1, int W1, int H2, int W2, int C,
                        int radius, void *workspace, size_t workspace_bytes, void *stream);

/*
 * Stride-1 "same" convolution + bias + activation on tensor cores: the convolutions of the per-iteration update
 * (core/update.py: convc2 / convf2 / conv of the motion encoder :84-97, the 1x5 / 5x1 gates of SepConvGRU :33-60, the
 * flow head :6-14 and the mask head :122-125), i.e. what follows rc_corr_lookup_convc1_fwd in an iteration.
 *     out[b, o, y, x] = act(bias[o] + sum_{c, ky, kx} weight[o, c, ky, kx] * in[b, c, y + ky - KH/2, x + kx - KW/2])
 *   x1, C1, x2 : the input is the channel concatenation of x1 (B, C1, H, W) and x2 (B, Cin - C1, H, W), both fp32
 *                NCHW, read in place (the reference's torch.cat copies); x2 = NULL and C1 = Cin for one tensor
 *   wprep      : weight (Cout, Cin, KH, KW) fp32 converted once per model by rc_conv_prep_weight into
 *                rc_conv_weight_bytes() bytes (fp16, [channel block][tap][Cout][64])
 *   bias       : (Cout) fp32 or NULL;   act : RC_ACT_*;   out : (B, Cout, H, W) fp32
 * Cout a multiple of 16, <= 256 (wider layers: call per slice of output channels); KH, KW odd with a halo tile of at
 * most 400 pixels: KH * (128 + KW - 1) for KW > 1, (8 + KH - 1) * 16 for KW = 1 (1x1, 3x3, 1x5, 5x1, 3x5, 1x3, ...).  Operands are rounded to fp16 (saturating), one pass,
 * fp32 accumulation: the class of TF32, which is what cuDNN runs these layers in by torch's default.  Forward only.
 *
 * rc_conv2d_ex_fwd is the same convolution for tensors that are channel slices of wider ones and with the GRU's
 * element-wise steps in the epilogue (core/update.py:43-46,50-53: z, r, r*h, (1-z)*h + z*q), so that a GRU half is two
 * launches and nothing else:
 *   x1_bs, x2_bs, out_bs, aux1_bs, aux2_bs : floats from one batch item to the next (0 = dense, C * H * W)
 *   store_channels : only filters [0, store_channels) are written (0 = all): the zero filters that pad a filter count to
 *                    a multiple of 16 then touch nothing (the motion encoder's 126 filters next to the 2 flow channels)
 *   epilogue = RC_EPI_NONE     : out = act(.)
 *              RC_EPI_MUL_FROM : out[b, o] = act(.) * (o >= split ? aux1[b, o - split] : 1)     z | r*h in one pass
 *              RC_EPI_GRU      : out[b, o] = (1 - aux1[b, o]) * aux2[b, o] + aux1[b, o] * act(.)   aux1 = z, aux2 = h
 *   (out may alias aux2 in RC_EPI_GRU -- every element is read before it is written by the same thread -- but no input)
 */
#define RC_ACT_NONE 0
#define RC_ACT_RELU 1
#define RC_ACT_SIGMOID 2
#define RC_ACT_TANH 3
size_t rc_conv_weight_bytes(int Cout, int Cin, int KH, int KW);
int rc_conv_prep_weight(const float *weight, void *wprep, int Cout, int Cin, int KH, int KW, void *stream);
int rc_conv2d_fwd(const float *x1, int C1, const float *x2, const void *wprep, const float *bias, float *out, int B,
                  int Cin, int Cout, int H, int W, int KH, int KW, int act, void *stream);
#define RC_EPI_NONE 0
#define RC_EPI_MUL_FROM 1
#define RC_EPI_GRU 2
int rc_conv2d_ex_fwd(const float *x1, int C1, long long x1_bs, const float *x2, long long x2_bs, const void *wprep,
                     const float *bias, float *out, long long out_bs, int store_channels, int B, int Cin, int Cout, int H,
                     int W, int KH, int KW, int act, int epilogue, int split, const float *aux1, long long aux1_bs,
                     const float *aux2, long long aux2_bs, void *stream);

/*
 * The two remaining pieces of an iteration (core/update.py:86,92 `flo = F.relu(self.convf1(flow))`; core/raft.py:123,
 * 131-132 `flow = coords1 - coords0 ... coords1 = coords1 + delta_flow`), one launch each:
 * rc_small_conv_relu_fwd : K x K "same" convolution (K odd <= 7) of a map with 1..4 channels (the flow field) + bias
 *                          (+ ReLU if relu != 0) on CUDA cores, fp32: in (B, Cin, H, W), weight (Cout, Cin, K, K),
 *                          out (B, Cout, H, W)
 * rc_flow_update         : coords1 (B, 2, H, W, dense) += delta (B, 2, H, W, batch stride delta_bs floats, 0 = dense);
 *                          flow = coords1 - coords0 is written to flow_out and, if not NULL, flow_out2 (own batch strides)
 */
int rc_small_conv_relu_fwd(const float *in, const float *weight, const float *bias, float *out, int B, int Cin, int Cout,
                           int H, int W, int K, int relu, void *stream);
int rc_flow_update(float *coords1, const float *delta, long long delta_bs, const float *coords0, float *flow_out,
                   long long flow_bs, float *flow_out2, long long flow2_bs, int B, int H, int W, void *stream);

/*
 * RAFT.upsample_flow (core/raft.py:72-83): convex 8x upsampling of the flow field, one launch instead of
 * view / softmax / unfold / mul / sum / permute / reshape with their temporaries:
 *     out[n, c, 8y+i, 8x+j] = sum_k softmax_k(mask[n, k*64 + i*8 + j, y, x]) * 8 * flow[n, c, y + k/3 - 1, x + k%3 - 1]
 *   flow : (N, 2, H, W) fp32      mask : (N, 576, H, W) fp32 (the mask head's output, already scaled by 0.25,
 *   core/update.py:134-135)       out  : (N, 2, 8H, 8W) fp32, 16 B aligned.  Forward only.
 */
int rc_convex_upsample_fwd(const float *flow, const float *mask, float *out, int N, int H, int W, void *stream);

/*
 * bilinear_sampler: pixel-coordinate bilinear sampling with zeros outside.
 * Replaces: core/utils/utils.py:57-71 (normalise to [-1,1], F.grid_sample(..., align_corners=True),
 * optional in-bounds mask at :67-69).
 *   img    : (M, C, H, W)      coords : (M, Ho, Wo, 2) [x, y] in pixels
 *   out    : (M, C, Ho, Wo)    mask   : (M, Ho, Wo, 1) or NULL; 1 where 0<x<W-1 and 0<y<H-1
 */
int rc_bilinear_sampler_fwd(const float *img, const float *coords, float *out, float *mask, int M,
                            int C, int H, int W, int Ho, int Wo, void *stream);

/* Gradient of the sampler w.r.t. img (what autograd produces through F.grid_sample at
 * core/utils/utils.py:65); grad_img (M, C, H, W) MUST be zero-filled by the caller. */
int rc_bilinear_sampler_bwd(const float *grad_out, const float *coords, float *grad_img, int M, int C,
                            int H, int W, int Ho, int Wo, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* RAFTCORR_H_ */
