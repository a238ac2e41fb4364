// tcgen05 / TMA / mbarrier PTX wrappers and the epilogue's row-store helper, shared by the
// tensor-core kernels (build_tc.cu, gemm_tc.cu).  sm_100a only.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace rc {
namespace tc {

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity), "r"(0x989680u) /* suspend-time hint: sleep in hardware instead of spinning */
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// ---- cta_group::2 (CTA pair) variants.  In a cluster launch the 32-bit shared address carries the
// CTA rank in bit 24; clearing it addresses the same offset in the even (leader) CTA of the pair.
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerMask) : "memory");
}
__device__ __forceinline__ void tma2_load_3d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma2_load_4d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t *slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// one thread of the leader CTA issues the pair's 256 x 256 x 16 MMA
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// the same with 8-bit float operands (kind::f8f6f4, K = 32 per instruction)
__device__ __forceinline__ void umma2_f8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all earlier MMAs are done) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma2_commit_both(uint64_t *bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}

__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 operands, fp32 accumulate, issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every previously issued MMA of this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t *u = reinterpret_cast<uint32_t *>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
        "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]),
        "=r"(u[16]), "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]),
        "=r"(u[24]), "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory operand descriptor (rows of 128 B, 8-row atoms 1024 B apart)
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) /*LBO (unused for swizzled K-major)*/ |
         ((uint64_t)(1024 >> 4) << 32) /*SBO*/ | (1ull << 46) /*sm100 descriptor version*/ |
         (2ull << 61) /*SWIZZLE_128B*/;
}

// ---- small shared / global access helpers (explicit state spaces: no generic addressing)
__device__ __forceinline__ void sts128(uint32_t a, float x, float y, float z, float w) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts32(uint32_t a, float x) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(x) : "memory");
}
__device__ __forceinline__ float lds32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void stg128(float *p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void stg64(float *p, float x, float y) {
  asm volatile("st.global.L1::no_allocate.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t *u = reinterpret_cast<uint32_t *>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
        "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
      : "r"(taddr)
      : "memory");
}


// Write this warp's 32 rows x NC columns (thread <-> row, NC values in registers) to global
// memory as contiguous NC-float row segments.  The rows are transposed through a 2 KB
// shared-memory buffer so that a store instruction covers whole segments:
//   VEC path  (16 B aligned rows): 16 B chunks, XOR-swizzled so that both the row-wise writes and
//             the segment-wise reads are bank-conflict free; st.global.v4, 512 B per instruction.
//   scalar path (ragged widths): same idea with 4 B elements.
template <int NC>
__device__ __forceinline__ void store_rows(uint32_t buf, const float *v, int lane, float *__restrict__ g,
                                           size_t pitch, int rows_valid, int cols_valid, bool vec) {
  static_assert(NC == 32 || NC == 16 || NC == 8, "segment width");
  constexpr int CH = NC / 4;          // 16 B chunks per row
  constexpr int RPI = 32 / CH;        // rows per vector store instruction
  constexpr int SH = (NC == 32) ? 0 : (NC == 16) ? 1 : 2;  // log2(rows per 128 B of the buffer)
  if (vec) {
#pragma unroll
    for (int c = 0; c < CH; ++c)
      sts128(buf + lane * (NC * 4) + (((c ^ ((lane >> SH) & (CH - 1)))) << 4), v[4 * c], v[4 * c + 1], v[4 * c + 2],
             v[4 * c + 3]);
    __syncwarp();
    const int rsub = lane / CH, c = lane % CH;
    float *gp = g + (size_t)rsub * pitch + 4 * c;
    const bool cok = 4 * c < cols_valid;
#pragma unroll
    for (int i = 0; i < 32 / RPI; ++i) {
      const int row = i * RPI + rsub;
      const float4 val = lds128(buf + row * (NC * 4) + (((c ^ ((row >> SH) & (CH - 1)))) << 4));
      if (cok && row < rows_valid) stg128(gp, val);
      gp += (size_t)RPI * pitch;
    }
    __syncwarp();
  } else {
#pragma unroll
    for (int i = 0; i < NC; ++i) sts32(buf + ((lane * NC + (i ^ ((lane ^ (lane >> 3)) & (NC - 1)))) << 2), v[i]);
    __syncwarp();
    constexpr int RPS = 32 / NC;  // rows per scalar store instruction
    const int rsub = lane / NC, c = lane % NC;
    float *gp = g + (size_t)rsub * pitch + c;
    const bool cok = c < cols_valid;
#pragma unroll 8
    for (int i = 0; i < 32 / RPS; ++i) {
      const int row = i * RPS + rsub;
      const float val = lds32(buf + ((row * NC + (c ^ ((row ^ (row >> 3)) & (NC - 1)))) << 2));
      if (cok && row < rows_valid) st_stream(gp, val);
      gp += (size_t)RPS * pitch;
    }
    __syncwarp();
  }
}

// ---- stores into the TILED layout (include/raftcorr.h):
//   [group of G = 32 query pixels][tile][pixel u][4 chunks of 16 B], chunk k = 2*(tile row) + (column half),
//   stored at position k ^ ((u >> 1) & 3).
// The warp's 32 rows (query pixels) are one group; `g` points at (group, first tile, pixel 0).
// The chunk swizzle is what lets a warp write its pixels' tiles into shared memory in exactly the
// global order without bank conflicts, so that a plain bulk copy can move them.
__device__ __forceinline__ int tile_sw(int u) { return (u >> 1) & 3; }

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store(void *gdst, uint32_t ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// Fast path: v = two whole tiles per query pixel (32 floats: tile 0 row 0, tile 0 row 1, tile 1
// row 0, tile 1 row 1), all 32 pixels and both tiles valid.  The warp lays the 4 KB out in shared
// memory in global order and lane 0 hands them to the bulk-copy engine: the warp does not wait
// for the store to drain (the caller waits for the buffer to be READ before reusing it).
// (Two 2 KB copies on alternating halves of the buffer, one always outstanding, measured slower.)
__device__ __forceinline__ void bulk_tiles32(uint32_t buf, const float *v, int lane, float *__restrict__ g) {
  const int sw = tile_sw(lane);
#pragma unroll
  for (int t = 0; t < 2; ++t)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      sts128(buf + t * 2048 + lane * 64 + ((k ^ sw) << 4), v[16 * t + 4 * k], v[16 * t + 4 * k + 1], v[16 * t + 4 * k + 2],
             v[16 * t + 4 * k + 3]);
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0) {
    bulk_store(g, buf, 4096);
    bulk_commit();
  }
}

// Edge path (partial group or a single valid tile): same data through per-lane stores.
template <int G>
__device__ __forceinline__ void store_tiles32(uint32_t buf, const float *v, int lane, float *__restrict__ g,
                                              int rows_valid, int tiles_valid) {
  static_assert(G == 32, "one warp = one group");
#pragma unroll
  for (int c = 0; c < 8; ++c)
    sts128(buf + lane * 128 + ((c ^ (lane & 7)) << 4), v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
  __syncwarp();
  const int pix = lane >> 2, ch = lane & 3;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int t = i & 1;
    const int row = (i >> 1) * 8 + pix;
    const float4 val = lds128(buf + row * 128 + ((((t << 2) | ch) ^ pix) << 4));
    if (t < tiles_valid && row < rows_valid) stg128(g + t * (16 * G) + row * 16 + ((ch ^ tile_sw(row)) << 2), val);
  }
  __syncwarp();
}

// v = one 8-float tile row per query pixel (tile row `half`: chunks 2*half, 2*half+1).  One
// instruction stores 16 pixels x 32 B.
template <int G>
__device__ __forceinline__ void store_halftile8(uint32_t buf, const float *v, int lane, float *__restrict__ g, int half,
                                                int rows_valid, bool tile_valid) {
  static_assert(G == 32, "one warp = one group");
  sts128(buf + lane * 32 + ((0 ^ ((lane >> 2) & 1)) << 4), v[0], v[1], v[2], v[3]);
  sts128(buf + lane * 32 + ((1 ^ ((lane >> 2) & 1)) << 4), v[4], v[5], v[6], v[7]);
  __syncwarp();
  const int pixl = lane >> 1, ch = lane & 1;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int row = i * 16 + pixl;
    const float4 val = lds128(buf + row * 32 + ((ch ^ ((row >> 2) & 1)) << 4));
    if (tile_valid && row < rows_valid) stg128(g + row * 16 + (((half * 2 + ch) ^ tile_sw(row)) << 2), val);
  }
  __syncwarp();
}

}  // namespace tc
}  // namespace rc
