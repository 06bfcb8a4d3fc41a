// K9: grouped GEMM on 5th-generation tensor cores (tcgen05.mma), accumulators in TMEM, operands staged by
// TMA with 128-byte swizzle.  See include/fedml_b200.h.
//
//   for every group g:  D_g[M][N] = op(A_g) * op(B_g)     bf16 operands, fp32 accumulate, bf16/fp32 output
//
// Three operand layouts share one kernel template (MODE bit 0: A MN-major, bit 1: B MN-major):
//   NT  (MODE 0): A_g is [M][K], B_g is [N][K], K contiguous                      (forward / Gram / Q*K^T)
//   TN  (MODE 3): A_g is [K][M], B_g is [K][N], M resp. N contiguous              (weight gradients, P^T*dO)
//   NN  (MODE 2): A_g is [M][K], B_g is [K][N]                                    (data gradients, P*V)
// Groups may be two-level (outer, inner) with independent strides (4-D tensor maps), so attention heads are
// addressed in place inside [batch][seq][heads*dh] buffers.
//
// Persistent, warp-specialised CTA of 192 threads (one per SM):
//   warp 0     TMA producer   (one elected lane): global -> smem ring of STAGES k-blocks (BK = 64)
//   warp 1     MMA issuer     (one elected lane): tcgen05.mma 128 x BN x 16 from smem descriptors into one of
//                             two TMEM accumulator buffers; tcgen05.commit releases smem slots / publishes tiles
//   warps 2-5  epilogue       : tcgen05.ld (32 lanes x 16 columns per instruction) -> registers -> bias / ReLU ->
//                             16-byte global stores, overlapped with the main loop of the next tile
// Tiles are (group, m-tile, n-tile) triples walked with a grid-stride; partial tiles rely on TMA zero fill for
// loads and on predication for stores, so any M, N and any K that is a multiple of 8 work.
#include "common.cuh"
#include "../include/fedml_b200.h"
#include <cuda.h>
#include <stdlib.h>

#define GEMM_BM 128
#define GEMM_BK 64
#define GEMM_MAX_STAGES 8
#define GEMM_STAGING_BYTES (4 * 2 * 4096)   /* per epilogue warp: two 4 KiB store buffers */
#define GEMM_THREADS 192
#define GEMM_A_BYTES (GEMM_BM * 128)

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_3d(const CUtensorMap* map, void* smem_dst, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, void* smem_dst, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], kind::f16 (bf16 inputs, fp32 accumulate)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// same with kind::tf32: fp32 operands in shared memory (the tensor core reads sign, exponent and the top 10 mantissa bits)
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// ---- CTA-pair (cta_group::2) forms: two CTAs of a cluster share one 256-row MMA tile
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion is signalled on the LEADER CTA's mbarrier (cluster address), data lands in this CTA
__device__ __forceinline__ void tma_load_4d_pair(const CUtensorMap* map, void* smem_dst, uint32_t leader_bar, int c0, int c1,
                                                 int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// arrive on the barrier at this smem offset in BOTH CTAs of the pair once the MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptor, SWIZZLE_128B, sm_100 version field = 1
//   K-major : rows of 128 B, 8-row groups 1024 B apart (SBO); LBO unused
//   MN-major: 64-element (128 B) MN chunks `lbo_bytes` apart, 8-row K groups 1024 B apart (SBO)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t swizzle_code = 2 /* 2: 128B, 4: 64B, 6: 32B */) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
  d |= (uint64_t)swizzle_code << 61;
  return d;
}
// instruction descriptor: fp32 accumulate, bf16 x bf16, M = 128, N = bn; majors: 0 = K, 1 = MN
__device__ __forceinline__ uint32_t make_idesc(int bn, int a_mn, int b_mn, int m = GEMM_BM, bool tf32 = false) {
  uint32_t d = 0;
  d |= 1u << 4;                       // c_format = F32
  d |= (tf32 ? 2u : 1u) << 7;         // a_format: 1 = BF16, 2 = TF32
  d |= (tf32 ? 2u : 1u) << 10;        // b_format
  d |= (uint32_t)(a_mn & 1) << 15;
  d |= (uint32_t)(b_mn & 1) << 16;
  d |= (uint32_t)(bn >> 3) << 17;     // N / 8
  d |= (uint32_t)(m >> 4) << 24;        // M / 16
  return d;
}

struct GemmArgs {
  void* d;
  const float* bias;
  int64_t d_gs, bias_gs;
  int G, M, N, K;
  int ldd, d_fp32, relu;
  int BN;          // n-tile width (multiple of 16, <= 256)
  int BNpad;       // smem rows/chunks reserved for B (BN, rounded to 64 for MN-major)
  int m_tiles, n_tiles, k_blocks;
  int tmem_cols;
  int stages;       // smem ring depth (2..GEMM_MAX_STAGES), as many as fit
  int tma_store;    // 1: epilogue stages tiles in smem and stores them with TMA (coalesced, clipped at the edges)
  int chunk_cols;   // columns per TMA store box
  int chunk_bytes;  // bytes per staged row (128 => SWIZZLE_128B box, else unswizzled)
  int reduce_add;   // 1: TMA reduce-store (fp32 add) instead of a plain store: split-K slices accumulate in place
  int group_div;    // output group = group / group_div (split-K slices of one problem share an output)
  int gi;           // inner group count of the two-level group index (g = outer * gi + inner); G when one level
  int gi_out;       // same for the output group index
  int64_t d_gs2;    // outer output group stride (pointer-store path)
  // MODE 4 (implicit 3x3 convolution, see fb_conv3x3_bf16): A tiles are gathered by TMA from the NHWC activation
  int cv_B, cv_HoWo, cv_Wo, cv_stride, cv_flip;   // images per group, output plane, output width, conv stride, flipped taps
  int cv_kblk, cv_kcs;                            // channels per k-block (16 / 32 / 64), k-blocks per tap
  int cv_b_inner_off, cv_b_row_off;               // per-tap coordinate offsets into the weight map
  int cv_sbo, cv_swz;                             // 8-row group stride (bytes) and descriptor swizzle code
  int cv_tpt;                                     // MODE 5: taps per n-tile (9, 3 or 1)
  int cv_out_cols;                                // MODE 5: columns of an n-tile that are stored (0 = BN); tf32 with 16 input
                                                  // channels pads every tap to a 32-column tile whose upper half is zero
  // per-column statistics of the output (BatchNorm batch statistics produced by the convolution that feeds it):
  // stats[g][0][n] += sum over rows of D[g][:, n], stats[g][1][n] += sum of squares; group stride stats_gs floats
  float* stats;
  int64_t stats_gs;
};
#define GEMM_STATS_BYTES (4 * 2 * 256 * 4)   /* per epilogue warp: sum and sum of squares for up to 256 tile columns */

// MODE bit 0: A is MN-major ([K][M] in memory), bit 1: B is MN-major ([K][N] in memory); MODE 4: implicit 3x3 conv
// PAIR: the kernel runs as clusters of two CTAs (one TPC) that share 256 x BN tiles through tcgen05 cta_group::2: each CTA
// stages its 128 rows of A and its half of B, the leader CTA issues M = 256 MMAs that read both CTAs' shared memory, each
// CTA drains its own 128 accumulator rows.  Per output tile the pair moves half the B bytes from L2 that two
// independent CTAs would.
// TF32: operands are fp32 in memory and in shared memory, multiplied by tcgen05.mma kind::tf32 (the precision of the
// reference's cuDNN convolutions).  A 128-byte swizzle row then holds 32 elements instead of 64, so a k-block is 32
// elements (K-major) / 32 rows of 32-element chunks (MN-major) and one MMA covers K = 8; byte geometry of the stages,
// descriptors' 32-byte K-major advance and the epilogue are unchanged.
template <int MODE, bool PAIR = false, bool TF32 = false, bool STATS = false>
__global__ void __launch_bounds__(GEMM_THREADS, STATS ? 2 : 1)
grouped_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmD, const GemmArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128B swizzle atoms
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t b_bytes = (uint32_t)(PAIR ? a.BNpad / 2 : a.BNpad) * 128u;   // a pair CTA stages half of B
  const uint32_t stage_bytes = GEMM_A_BYTES + b_bytes;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  constexpr int TILE_M = PAIR ? 2 * GEMM_BM : GEMM_BM;
  constexpr uint32_t ES = TF32 ? 4u : 2u;      // operand element size
  constexpr int BK = 128 / (int)ES;            // elements per k-block (one 128-byte swizzle row); also the MN chunk width
  constexpr int UK = 32 / (int)ES;             // K elements per MMA instruction
  // MN-major operands: 16-bit types use the 128-byte swizzle (pattern period 8 k-rows); 32-bit types only exist in the
  // "128-byte swizzle, 32-byte atom" layout (TMA SWIZZLE_128B_ATOM_32B, descriptor layout type 1), whose pattern repeats
  // every 4 k-rows of 128 bytes
  constexpr uint32_t MN_SBO = TF32 ? 512u : 1024u, MN_SWZ = TF32 ? 1u : 2u;
  const int STAGES = a.stages;
  uint8_t* staging = smem + (size_t)STAGES * stage_bytes;            // 1024-aligned: stage_bytes is a multiple of 2048
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + GEMM_STAGING_BYTES);
  uint64_t* empty_bar = full_bar + GEMM_MAX_STAGES;
  uint64_t* tmem_full = empty_bar + GEMM_MAX_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* stats_acc = reinterpret_cast<float*>(staging + GEMM_STAGING_BYTES + 256);   // only carved when a.stats != nullptr
  // Tile walk.  Default: grid-stride over (group, m-tile, n-tile) with the n-tile fastest.  With output statistics a CTA
  // takes a CONTIGUOUS range of tiles, m-tile fastest, so that consecutive tiles share (group, n-tile) and the column
  // sums stay in shared memory across them: one round of global atomics per CTA and group instead of one per tile.
  const int n_ctas = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x, cta = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tiles_all = a.G * a.m_tiles * a.n_tiles;
  const int per_cta = (tiles_all + n_ctas - 1) / n_ctas;
  const int t_first = STATS ? cta * per_cta : cta;
  const int t_last = STATS ? min(tiles_all, t_first + per_cta) : tiles_all;
  const int t_step = STATS ? 1 : n_ctas;

  constexpr bool CONV = MODE == 4, WGRAD = MODE == 5;   // 5: TN form whose B operand is gathered from the NHWC activation
  constexpr bool A_MN = (MODE & 1) != 0, B_MN = (MODE & 2) != 0 || WGRAD;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = a.G * a.m_tiles * a.n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (a.tma_store) tma_prefetch_desc(&tmD);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tmem_full + i, 1);
      mbar_init(tmem_empty + i, PAIR ? 8 : 4);   // one arrive per epilogue warp (of both CTAs of a pair)
    }
    fence_barrier_init();
  }
  if (PAIR) cluster_sync_all();        // the peer's barriers exist before anything is signalled across CTAs
  if (warp == 1) {
    if (PAIR) tmem_alloc_pair(tmem_slot, (uint32_t)a.tmem_cols);
    else tmem_alloc(tmem_slot, (uint32_t)a.tmem_cols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // programmatic dependent launch: everything above overlapped the tail of the previous kernel; operands (and the
  // output buffer) may only be touched once that kernel has completed
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = t_first; t < t_last; t += t_step) {
        const int nt = STATS ? (t / a.m_tiles) % a.n_tiles : t % a.n_tiles;
      const int mt = STATS ? t % a.m_tiles : (t / a.n_tiles) % a.m_tiles, g = t / (a.n_tiles * a.m_tiles);
        if (CONV) {
          // 128 consecutive output pixels of group g: image n0 (a tile spans 128 / (Ho*Wo) images when the plane is
          // smaller), output rows from ho0.  One k-block per (tap, channel block): the tap's shifted input window
          // is one TMA box (padding = out-of-bounds zero fill, stride = element strides of the map).
          const int p0 = mt * TILE_M + (int)rank * GEMM_BM;       // a pair CTA gathers its own 128 pixels
          const int n0 = g * a.cv_B + p0 / a.cv_HoWo, ho0 = (p0 % a.cv_HoWo) / a.cv_Wo;
          const uint32_t inner_bytes = (uint32_t)a.cv_kblk * ES;
          for (int kb = 0; kb < a.k_blocks; ++kb) {
            const int tap = kb / a.cv_kcs, kc = kb - tap * a.cv_kcs;
            const int kh = tap / 3, kw = tap - kh * 3;
            const int wt = a.cv_flip ? 8 - tap : tap;
            mbar_wait(empty_bar + stage, phase ^ 1);
            uint8_t* sa = smem + (size_t)stage * stage_bytes;
            uint8_t* sb = sa + GEMM_A_BYTES;
            if constexpr (PAIR) {
              const uint32_t lbar = mapa_u32(smem_u32(full_bar + stage), 0);
              const int nh = a.BN / 2;                               // this CTA's half of the filter rows
              if (rank == 0) mbar_expect_tx(full_bar + stage, 2u * (GEMM_BM + (uint32_t)nh) * inner_bytes);
              tma_load_4d_pair(&tmA, sa, lbar, kc * a.cv_kblk, kw - 1, ho0 * a.cv_stride + kh - 1, n0);
              tma_load_4d_pair(&tmB, sb, lbar, kc * a.cv_kblk + wt * a.cv_b_inner_off,
                               nt * a.BN + (int)rank * nh + wt * a.cv_b_row_off, g, 0);
            } else {
              mbar_expect_tx(full_bar + stage, (GEMM_BM + (uint32_t)a.BN) * inner_bytes);
              tma_load_4d(&tmA, sa, full_bar + stage, kc * a.cv_kblk, kw - 1, ho0 * a.cv_stride + kh - 1, n0);
              tma_load_4d(&tmB, sb, full_bar + stage, kc * a.cv_kblk + wt * a.cv_b_inner_off, nt * a.BN + wt * a.cv_b_row_off,
                          g, 0);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          continue;
        }
        for (int kb = 0; kb < a.k_blocks; ++kb) {
          mbar_wait(empty_bar + stage, phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * stage_bytes;
          uint8_t* sb = sa + GEMM_A_BYTES;
          const int g0 = g % a.gi, g1 = g / a.gi;
          if (WGRAD) {
            // weight gradient of a 3x3 convolution: k-block = 64 consecutive output pixels of this (client, K-slice)
            // group; n-tile = (tap, channel chunk); B = the tap's shifted input window, gathered MN-major by TMA
            const uint32_t row_bytes = (uint32_t)a.cv_kblk * ES;
            const int subs = a.BN / a.cv_kblk;                       // BK-row B chunks per k-block (taps x channel blocks)
            const int per_tap = subs / a.cv_tpt;
            const int a_chunks = min(GEMM_BM / BK, (a.M - mt * GEMM_BM + BK - 1) / BK);   // rows of D beyond cout are never stored
            mbar_expect_tx(full_bar + stage, (uint32_t)a_chunks * BK * 128u + (uint32_t)subs * BK * row_bytes);
            for (int i = 0; i < a_chunks; ++i)
              tma_load_4d(&tmA, sa + (size_t)i * BK * 128, full_bar + stage, mt * GEMM_BM + i * BK, kb * BK, g0, g1);
            const int tap0 = a.cv_tpt > 1 ? nt * a.cv_tpt : nt / a.cv_kcs;
            const int cc = a.cv_tpt > 1 ? 0 : nt - tap0 * a.cv_kcs;
            const int p0 = (g % a.group_div) * a.K + kb * BK;
            const int n0 = (g / a.group_div) * a.cv_B + p0 / a.cv_HoWo, ho0 = (p0 % a.cv_HoWo) / a.cv_Wo;
            for (int i = 0; i < subs; ++i) {
              const int tap = tap0 + i / per_tap, sub = i - (i / per_tap) * per_tap;
              const int kh = tap / 3, kw = tap - kh * 3;
              tma_load_4d(&tmB, sb + (size_t)i * BK * row_bytes, full_bar + stage,
                          cc * (a.BN / a.cv_tpt) + sub * a.cv_kblk, kw - 1, ho0 * a.cv_stride + kh - 1, n0);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
            continue;
          }
          if constexpr (PAIR) {
            // this CTA's 128 rows of A and its half of B; every byte of both CTAs is counted on the leader's barrier
            const uint32_t lbar = mapa_u32(smem_u32(full_bar + stage), 0);
            if (rank == 0) mbar_expect_tx(full_bar + stage, 2 * stage_bytes);
            const int m0 = mt * TILE_M + (int)rank * GEMM_BM, nh = a.BN / 2, n0 = nt * a.BN + (int)rank * nh;
            if (!A_MN) {
              tma_load_4d_pair(&tmA, sa, lbar, kb * BK, m0, g0, g1);
            } else {
              for (int i = 0; i < GEMM_BM / BK; ++i)
                tma_load_4d_pair(&tmA, sa + (size_t)i * BK * 128, lbar, m0 + i * BK, kb * BK, g0, g1);
            }
            if (!B_MN) {
              tma_load_4d_pair(&tmB, sb, lbar, kb * BK, n0, g0, g1);
            } else {
              for (int i = 0; i < nh / BK; ++i)
                tma_load_4d_pair(&tmB, sb + (size_t)i * BK * 128, lbar, n0 + i * BK, kb * BK, g0, g1);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
            continue;
          }
          mbar_expect_tx(full_bar + stage, stage_bytes);
          if (!A_MN) {
            tma_load_4d(&tmA, sa, full_bar + stage, kb * BK, mt * GEMM_BM, g0, g1);
          } else {
            for (int i = 0; i < GEMM_BM / BK; ++i)
              tma_load_4d(&tmA, sa + (size_t)i * BK * 128, full_bar + stage, mt * GEMM_BM + i * BK, kb * BK, g0, g1);
          }
          if (!B_MN) {
            tma_load_4d(&tmB, sb, full_bar + stage, kb * BK, nt * a.BN, g0, g1);
          } else {
            for (int i = 0; i < a.BNpad / BK; ++i)
              tma_load_4d(&tmB, sb + (size_t)i * BK * 128, full_bar + stage, nt * a.BN + i * BK, kb * BK, g0, g1);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0 && rank == 0) {
      const uint32_t idesc = make_idesc(a.BN, A_MN, B_MN, TILE_M, TF32);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = t_first; t < t_last; t += t_step) {
        mbar_wait(tmem_empty + acc, acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * a.BN);
        for (int kb = 0; kb < a.k_blocks; ++kb) {
          mbar_wait(full_bar + stage, phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * stage_bytes);
          const uint32_t sb = sa + GEMM_A_BYTES;
          const int k_rem = a.K - kb * BK;
          const int k_steps = CONV ? a.cv_kblk / UK : (k_rem >= BK ? BK / UK : (k_rem + UK - 1) / UK);
          for (int k = 0; k < k_steps; ++k) {
            uint64_t ad, bd;
            if (CONV) {
              ad = make_smem_desc(sa + k * 32, 16, (uint32_t)a.cv_sbo, (uint32_t)a.cv_swz);
              bd = make_smem_desc(sb + k * 32, 16, (uint32_t)a.cv_sbo, (uint32_t)a.cv_swz);
            } else if (WGRAD) {
              const uint32_t row_bytes = (uint32_t)a.cv_kblk * ES;
              ad = make_smem_desc(sa + k * UK * 128, BK * 128, MN_SBO, MN_SWZ);
              bd = make_smem_desc(sb + k * UK * row_bytes, BK * row_bytes, (uint32_t)a.cv_sbo, (uint32_t)a.cv_swz);
            } else {
              // MN-major: UK k-rows of 128 bytes per MMA; the next MN chunk (BK elements) starts BK rows further
              ad = A_MN ? make_smem_desc(sa + k * UK * 128, BK * 128, MN_SBO, MN_SWZ) : make_smem_desc(sa + k * 32, 16, 1024);
              bd = B_MN ? make_smem_desc(sb + k * UK * 128, BK * 128, MN_SBO, MN_SWZ) : make_smem_desc(sb + k * 32, 16, 1024);
            }
            if constexpr (TF32) {
              if constexpr (PAIR) umma_tf32_pair(tmem_d, ad, bd, idesc, (kb | k) ? 1u : 0u);
              else umma_tf32(tmem_d, ad, bd, idesc, (kb | k) ? 1u : 0u);
            } else {
              if constexpr (PAIR) umma_bf16_pair(tmem_d, ad, bd, idesc, (kb | k) ? 1u : 0u);
              else umma_bf16(tmem_d, ad, bd, idesc, (kb | k) ? 1u : 0u);
            }
          }
          // smem slot reusable once these MMAs have read it (in a pair: the slot of both CTAs)
          if constexpr (PAIR) umma_commit_pair(empty_bar + stage);
          else umma_commit(empty_bar + stage);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if constexpr (PAIR) umma_commit_pair(tmem_full + acc);   // accumulator complete -> epilogue (of both CTAs)
        else umma_commit(tmem_full + acc);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================================================== epilogue (4 warps; warp % 4 selects the TMEM lane quarter)
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint8_t* my_stage = staging + (size_t)q * 2 * 4096;
    int sbuf = 0;
    float* my_acc = stats_acc + q * 512;          // [2][256] of this warp
    int acc_key = -1;                             // (group, n-tile) the accumulators belong to
    if (STATS)
      for (int i = lane; i < 512; i += 32) my_acc[i] = 0.f;
    for (int t = t_first; t < t_last; t += t_step) {
      const int nt = STATS ? (t / a.m_tiles) % a.n_tiles : t % a.n_tiles;
      const int mt = STATS ? t % a.m_tiles : (t / a.n_tiles) % a.m_tiles, g = t / (a.n_tiles * a.m_tiles);
      mbar_wait(tmem_full + acc, acc_phase);
      tc_fence_after();
      if (STATS && acc_key != g * a.n_tiles + nt) {
        if (acc_key >= 0) {
          const int pg = acc_key / a.n_tiles, pn = acc_key - pg * a.n_tiles;
          float* dst = a.stats + (size_t)pg * a.stats_gs + pn * a.BN;
          for (int c = lane; c < a.BN; c += 32) {
            if (pn * a.BN + c < a.N) {
              atomicAdd(dst + c, my_acc[c]);
              atomicAdd(dst + a.N + c, my_acc[256 + c]);
            }
            my_acc[c] = 0.f;
            my_acc[256 + c] = 0.f;
          }
          __syncwarp();
        }
        acc_key = g * a.n_tiles + nt;
      }
      const int row0 = mt * TILE_M + (int)rank * GEMM_BM + q * 32;   // first output row of this warp's TMEM lane quarter
      const int row = row0 + lane;
      const bool row_ok = row < a.M;
      const uint32_t taddr = tmem_base + (uint32_t)(acc * a.BN) + ((uint32_t)(q * 32) << 16);
      const float* bias = a.bias ? a.bias + (size_t)g * a.bias_gs : nullptr;
      const int og = g / a.group_div, og0 = og % a.gi_out, og1 = og / a.gi_out;
      if (a.tma_store) {
        // TMEM -> registers -> (bias, ReLU, convert) -> swizzled smem rows -> one TMA store per 32-row x chunk box
        const int CW = a.chunk_cols, RB = a.chunk_bytes;
        const int lds = CW / 16;   // tcgen05.ld.x16 per chunk: 4 (bf16, 64 cols), 2 (fp32, 32 cols) or fewer
        const int out_cols = a.cv_out_cols ? a.cv_out_cols : a.BN;      // columns stored per n-tile = its stride in D
        for (int c0 = 0; c0 < out_cols; c0 += CW) {
          uint32_t r[64];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (i < lds) tmem_ld16(taddr + (uint32_t)(c0 + 16 * i), r + 16 * i);
          tmem_ld_wait();
          const int col0 = nt * out_cols + c0;
          if (col0 >= a.N) continue;
          uint8_t* buf = my_stage + (size_t)sbuf * 4096;
          // the TMA store that last read this buffer (two chunks ago) must have finished reading it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (i < lds) {
              float f[16];
#pragma unroll
              for (int e = 0; e < 16; ++e) f[e] = __uint_as_float(r[16 * i + e]);
              if (bias) {
#pragma unroll
                for (int e = 0; e < 16; ++e)
                  if (col0 + 16 * i + e < a.N) f[e] += __ldg(bias + col0 + 16 * i + e);
              }
              if (a.relu) {
#pragma unroll
                for (int e = 0; e < 16; ++e) f[e] = fmaxf(f[e], 0.f);
              }
              if (a.d_fp32) {
#pragma unroll
                for (int sidx = 0; sidx < 4; ++sidx) {
                  int slot = 4 * i + sidx;
                  if (RB == 128) slot ^= (lane & 7);
                  *reinterpret_cast<float4*>(buf + lane * RB + slot * 16) =
                      make_float4(f[4 * sidx], f[4 * sidx + 1], f[4 * sidx + 2], f[4 * sidx + 3]);
                }
              } else {
#pragma unroll
                for (int sidx = 0; sidx < 2; ++sidx) {
                  uint4 o;
                  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
                  for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(f[8 * sidx + 2 * e], f[8 * sidx + 2 * e + 1]);
                  int slot = 2 * i + sidx;
                  if (RB == 128) slot ^= (lane & 7);
                  *reinterpret_cast<uint4*>(buf + lane * RB + slot * 16) = o;
                }
              }
            }
          }
          fence_async_proxy();   // generic-proxy smem writes -> visible to the TMA (async proxy)
          __syncwarp();
          if (STATS) {
            // column sums of the staged 32 x CW box (the values exactly as stored): lane c walks column c down the rows;
            // at a fixed row the 32 lanes read one (swizzle-permuted) row segment, so the reads are conflict-free
            const int rows_valid = min(32, a.M - row0);
            if (a.d_fp32) {
              if (lane < CW && col0 + lane < a.N) {
                float s1 = 0.f, s2 = 0.f;
                for (int r = 0; r < rows_valid; ++r) {
                  int slot = lane >> 2;
                  if (RB == 128) slot ^= (r & 7);
                  const float v = *reinterpret_cast<const float*>(buf + r * RB + slot * 16 + (lane & 3) * 4);
                  s1 += v;
                  s2 += v * v;
                }
                my_acc[c0 + lane] += s1;
                my_acc[256 + c0 + lane] += s2;
              }
            } else {
              if (2 * lane < CW && col0 + 2 * lane < a.N) {
                float s1a = 0.f, s2a = 0.f, s1b = 0.f, s2b = 0.f;
                for (int r = 0; r < rows_valid; ++r) {
                  int slot = lane >> 2;
                  if (RB == 128) slot ^= (r & 7);
                  const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(buf + r * RB + slot * 16 + (lane & 3) * 4));
                  s1a += v.x; s2a += v.x * v.x;
                  s1b += v.y; s2b += v.y * v.y;
                }
                my_acc[c0 + 2 * lane] += s1a;
                my_acc[256 + c0 + 2 * lane] += s2a;
                my_acc[c0 + 2 * lane + 1] += s1b;
                my_acc[256 + c0 + 2 * lane + 1] += s2b;
              }
            }
          }
          if (lane == 0) {
            if (a.reduce_add) {
              asm volatile("cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                           ::"l"(&tmD), "r"(smem_u32(buf)), "r"(col0), "r"(row0), "r"(og0), "r"(og1)
                           : "memory");
            } else {
              asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(&tmD),
                           "r"(smem_u32(buf)), "r"(col0), "r"(row0), "r"(og0), "r"(og1)
                           : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          sbuf ^= 1;
        }
      } else {
        for (int c0 = 0; c0 < a.BN; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + (uint32_t)c0, r);
          tmem_ld_wait();
          const int col0 = nt * a.BN + c0;
          if (!row_ok || col0 >= a.N) continue;
          float f[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(r[i]);
          if (bias) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (col0 + i < a.N) f[i] += __ldg(bias + col0 + i);
          }
          if (a.relu) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
          }
          if (a.d_fp32) {
            float* dst = reinterpret_cast<float*>(a.d) + (size_t)og1 * a.d_gs2 + (size_t)og0 * a.d_gs + (size_t)row * a.ldd + col0;
            for (int i = 0; i < 16 && col0 + i < a.N; ++i) dst[i] = f[i];
          } else {
            __nv_bfloat16* dst =
                reinterpret_cast<__nv_bfloat16*>(a.d) + (size_t)og1 * a.d_gs2 + (size_t)og0 * a.d_gs + (size_t)row * a.ldd + col0;
            for (int i = 0; i < 16 && col0 + i < a.N; ++i) dst[i] = __float2bfloat16_rn(f[i]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR) mbar_arrive_cluster(mapa_u32(smem_u32(tmem_empty + acc), 0));   // the leader issues the next MMAs
        else mbar_arrive(tmem_empty + acc);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (STATS && acc_key >= 0) {
      const int pg = acc_key / a.n_tiles, pn = acc_key - pg * a.n_tiles;
      float* dst = a.stats + (size_t)pg * a.stats_gs + pn * a.BN;
      for (int c = lane; c < a.BN; c += 32) {
        if (pn * a.BN + c < a.N) {
          atomicAdd(dst + c, my_acc[c]);
          atomicAdd(dst + a.N + c, my_acc[256 + c]);
        }
      }
    }
    if (a.tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // stores complete before exit
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();        // neither CTA frees TMEM / exits while the peer may still signal or read it
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, (uint32_t)a.tmem_cols);
    else tmem_dealloc(tmem_base, (uint32_t)a.tmem_cols);
  }
}

// ---------------------------------------------------------------- host side
// fp32 operands travel as plain FLOAT32 tensor maps; the tensor core reads sign, exponent and the top 10 mantissa bits
// (measured: results match the truncated-operand product to fp32 accumulation order; TFLOAT32 maps gave no accuracy gain)
#define FB_TF32_LOAD_TYPE CU_TENSOR_MAP_DATA_TYPE_FLOAT32
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

// 4-D bf16 tensor map: dims (inner, rows, inner groups, outer groups), box (64, box_rows, 1, 1), 128B swizzle,
// zero OOB fill.  A dimension of extent 1 gets a harmless dense stride.
static int make_map(CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows, uint64_t gi, uint64_t go,
                    uint64_t row_stride_elems, uint64_t gs_elems, uint64_t gs2_elems, uint32_t box_rows, bool tf32 = false,
                    bool mn_major = false) {
  PFN_encodeTiled enc = get_encode();
  FB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  const uint64_t es = tf32 ? 4 : 2;
  cuuint64_t dims[4] = {inner, rows, gi, go};
  const uint64_t dense = row_stride_elems * rows;
  cuuint64_t strides[3] = {row_stride_elems * es, (gi > 1 ? gs_elems : dense) * es, (go > 1 ? gs2_elems : dense * gi) * es};
  cuuint32_t box[4] = {(cuuint32_t)(128 / es), box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  FB_REQUIRE((strides[0] % 16) == 0 && (strides[1] % 16) == 0 && (strides[2] % 16) == 0 && ((uintptr_t)base % 16) == 0,
             "grouped GEMM operands need 16-byte aligned base pointers and strides");
  CUresult r = enc(m, tf32 ? FB_TF32_LOAD_TYPE : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   (tf32 && mn_major) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with code %d", (int)r);
  return FB_OK;
}

// 4-D output map: dims (N, M, inner groups, outer groups), box (chunk_cols, 32, 1, 1); swizzled when a staged row is
// exactly 128 bytes
static int make_store_map(CUtensorMap* m, void* base, int fp32, uint64_t N, uint64_t M, uint64_t gi, uint64_t go, uint64_t ldd,
                          uint64_t gs_elems, uint64_t gs2_elems, uint32_t chunk_cols, bool swizzle128) {
  PFN_encodeTiled enc = get_encode();
  FB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  const uint64_t es = fp32 ? 4 : 2;
  cuuint64_t dims[4] = {N, M, gi, go};
  const uint64_t dense = ldd * M;
  cuuint64_t strides[3] = {ldd * es, (gi > 1 ? gs_elems : dense) * es, (go > 1 ? gs2_elems : dense * gi) * es};
  cuuint32_t box[4] = {chunk_cols, 32, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, strides, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (output) failed with code %d", (int)r);
  return FB_OK;
}

static int launch_gemm(const FbGroupedGemm* p, int mode, fb_stream_t stream, bool tf32 = false) {
  const bool a_mn = (mode & 1) != 0, b_mn = (mode & 2) != 0;
  const int BK = tf32 ? 32 : 64;      // elements per k-block and per MN-major chunk (one 128-byte row)
  FB_REQUIRE(p && p->a && p->b && p->d, "fb_grouped_gemm: null operand");
  FB_REQUIRE(p->G >= 1 && p->M >= 1 && p->N >= 1 && p->K >= 1, "fb_grouped_gemm: empty problem");
  const int group_div = p->out_group_div > 0 ? p->out_group_div : 1;
  FB_REQUIRE(p->accumulate == 0 || p->d_fp32, "fb_grouped_gemm: accumulate needs an fp32 output");
  FB_REQUIRE(p->G % group_div == 0, "fb_grouped_gemm: G must be a multiple of out_group_div");
  FB_REQUIRE(group_div == 1 || p->accumulate, "fb_grouped_gemm: out_group_div > 1 needs accumulate");
  const int gi = p->group_inner > 0 ? p->group_inner : p->G;
  FB_REQUIRE(p->G % gi == 0, "fb_grouped_gemm: G must be a multiple of group_inner");
  FB_REQUIRE(group_div == 1 || gi == p->G, "fb_grouped_gemm: out_group_div and group_inner are mutually exclusive");
  const int go = p->G / gi;
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.d = p->d; a.bias = p->bias; a.d_gs = p->d_gs; a.bias_gs = p->bias_gs;
  a.G = p->G; a.M = p->M; a.N = p->N; a.K = p->K; a.ldd = p->ldd; a.d_fp32 = p->d_fp32; a.relu = p->relu;
  int bn = (p->N + 15) / 16 * 16;
  if (bn > 64) {
    // n-tiles of at most 256 columns, multiples of 64 so that every TMA store box lies inside its own tile.  When the
    // 256-wide tiling leaves the last wave of tiles mostly empty, 128-wide tiles balance the SMs better (cost in
    // column units per SM: waves x (tile width + fixed per-tile overhead)); ties go to the wider tile, which
    // re-reads A less often.
    const int64_t base = (int64_t)p->G * ((p->M + GEMM_BM - 1) / GEMM_BM);
    const int sms = fb_num_sms();
    int64_t cost[2];
    int cand[2];
    for (int i = 0; i < 2; ++i) {
      const int cap = i == 0 ? 256 : 128;
      const int nt = (p->N + cap - 1) / cap;
      cand[i] = ((p->N + nt - 1) / nt + 63) / 64 * 64;
      const int64_t tiles = base * ((p->N + cand[i] - 1) / cand[i]);
      cost[i] = ((tiles + sms - 1) / sms) * (cand[i] + 24);
    }
    bn = (cost[1] * 108 < cost[0] * 100) ? cand[1] : cand[0];
  }
  // CTA pairs (cta_group::2, 256 x bn tiles over two SMs): large problems only; n-tiles are multiples of 128 so that each
  // CTA's half of B is whole 64-column chunks
  static int pair_env = -1;
  if (pair_env < 0) {
    const char* e = getenv("FB_GEMM_PAIR");
    pair_env = e ? atoi(e) : 1;
  }
  const bool pair = pair_env != 0 && p->M >= 512 && p->N >= 128 && p->K >= 128 && group_div == 1 && !p->stats;
  if (pair) {
    const int nt = (p->N + 255) / 256;
    bn = ((p->N + nt - 1) / nt + 127) / 128 * 128;
  }
  a.BN = bn;
  const uint64_t es = p->d_fp32 ? 4 : 2;
  a.gi = gi;
  a.gi_out = gi / group_div;
  a.d_gs2 = p->d_gs2;
  a.tma_store = (((uint64_t)p->ldd * es) % 16 == 0) && (((uint64_t)p->d_gs * es) % 16 == 0 || a.gi_out == 1) &&
                (((uint64_t)p->d_gs2 * es) % 16 == 0 || go == 1) && ((uintptr_t)p->d % 16 == 0);
  a.chunk_cols = p->d_fp32 ? (bn >= 32 ? 32 : 16) : (bn >= 64 ? 64 : bn);
  a.chunk_bytes = a.chunk_cols * (int)es;
  a.reduce_add = p->accumulate ? 1 : 0;
  a.group_div = group_div;
  FB_REQUIRE(!a.reduce_add || a.tma_store, "fb_grouped_gemm: accumulate needs a 16-byte aligned output (TMA reduce-store)");
  a.stats = p->stats;
  a.stats_gs = p->stats_gs;
  FB_REQUIRE(!a.stats || (a.tma_store && group_div == 1 && go == 1),
             "fb_grouped_gemm: output statistics need a 16-byte aligned output and one-level groups without split-K");
  a.BNpad = b_mn ? (bn + BK - 1) / BK * BK : bn;
  a.n_tiles = (p->N + bn - 1) / bn;
  a.m_tiles = pair ? (p->M + 2 * GEMM_BM - 1) / (2 * GEMM_BM) : (p->M + GEMM_BM - 1) / GEMM_BM;
  a.k_blocks = (p->K + BK - 1) / BK;
  int cols = 32;
  while (cols < 2 * bn) cols <<= 1;
  a.tmem_cols = cols;
  CUtensorMap tmA, tmB;
  int rc;
  FB_REQUIRE(p->lda % (tf32 ? 4 : 8) == 0 && p->ldb % (tf32 ? 4 : 8) == 0, "fb_grouped_gemm: lda/ldb must be multiples of 16 bytes");
  if (!a_mn)
    rc = make_map(&tmA, p->a, (uint64_t)p->K, (uint64_t)p->M, (uint64_t)gi, (uint64_t)go, (uint64_t)p->lda, (uint64_t)p->a_gs,
                  (uint64_t)p->a_gs2, GEMM_BM, tf32);
  else
    rc = make_map(&tmA, p->a, (uint64_t)p->M, (uint64_t)p->K, (uint64_t)gi, (uint64_t)go, (uint64_t)p->lda, (uint64_t)p->a_gs,
                  (uint64_t)p->a_gs2, (uint32_t)BK, tf32, true);
  if (rc) return rc;
  if (!b_mn)
    rc = make_map(&tmB, p->b, (uint64_t)p->K, (uint64_t)p->N, (uint64_t)gi, (uint64_t)go, (uint64_t)p->ldb, (uint64_t)p->b_gs,
                  (uint64_t)p->b_gs2, (uint32_t)(pair ? a.BN / 2 : a.BN), tf32);
  else
    rc = make_map(&tmB, p->b, (uint64_t)p->N, (uint64_t)p->K, (uint64_t)gi, (uint64_t)go, (uint64_t)p->ldb, (uint64_t)p->b_gs,
                  (uint64_t)p->b_gs2, (uint32_t)BK, tf32, true);
  if (rc) return rc;
  const size_t stage_bytes = GEMM_A_BYTES + (size_t)(pair ? a.BNpad / 2 : a.BNpad) * 128;
  const size_t fixed = 1024 /*align slack*/ + GEMM_STAGING_BYTES + 256 /*barriers*/ + (a.stats ? GEMM_STATS_BYTES : 0);
  // two co-resident CTAs per SM hide each other's per-tile latencies (TMEM allows it while 2 x tmem_cols <= 512)
  static int force_ctas = -1;
  if (force_ctas < 0) {
    const char* e = getenv("FB_GEMM_CTAS_PER_SM");
    force_ctas = e ? atoi(e) : 0;
  }
  int ctas_per_sm = (a.tmem_cols <= 256) ? 2 : 1;
  if (force_ctas == 1 || force_ctas == 2) ctas_per_sm = (force_ctas == 2 && a.tmem_cols <= 256) ? 2 : 1;
  if (pair) ctas_per_sm = 1;
  int stages = (int)((227 * 1024 / ctas_per_sm - fixed) / stage_bytes);
  if (stages < 3 && ctas_per_sm == 2) {
    ctas_per_sm = 1;
    stages = (int)((227 * 1024 - fixed) / stage_bytes);
  }
  if (stages > GEMM_MAX_STAGES) stages = GEMM_MAX_STAGES;
  FB_REQUIRE(stages >= 2, "fb_grouped_gemm: tile does not fit shared memory");
  a.stages = stages;
  const size_t smem = (size_t)stages * stage_bytes + fixed;
  CUtensorMap tmD;
  memset(&tmD, 0, sizeof(tmD));
  if (a.tma_store) {
    rc = make_store_map(&tmD, p->d, p->d_fp32, (uint64_t)p->N, (uint64_t)p->M, (uint64_t)a.gi_out, (uint64_t)go, (uint64_t)p->ldd,
                        (uint64_t)p->d_gs, (uint64_t)p->d_gs2, (uint32_t)a.chunk_cols, a.chunk_bytes == 128);
    if (rc) return rc;
  }
  const int total = a.G * a.m_tiles * a.n_tiles;
  int grid = fb_num_sms() * ctas_per_sm;
#define FB_LAUNCH_PAIR(MODE_, TF_)                                                                                              \
  do {                                                                                                                            \
    FB_CHECK_CUDA(cudaFuncSetAttribute(grouped_gemm_kernel<MODE_, true, TF_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    FB_CHECK_CUDA(fb_launch_pdl_cluster(grouped_gemm_kernel<MODE_, true, TF_>, dim3(grid), dim3(GEMM_THREADS), (size_t)smem,    \
                                        (cudaStream_t)stream, 2u, tmA, tmB, tmD, a));                                            \
  } while (0)
#define FB_LAUNCH_ONE_ST(MODE_, TF_)                                                                                            \
  do {                                                                                                                            \
    FB_CHECK_CUDA(cudaFuncSetAttribute(grouped_gemm_kernel<MODE_, false, TF_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    FB_CHECK_CUDA(fb_launch_pdl(grouped_gemm_kernel<MODE_, false, TF_, true>, dim3(grid), dim3(GEMM_THREADS), (size_t)smem,     \
                                (cudaStream_t)stream, tmA, tmB, tmD, a));                                                        \
  } while (0)
#define FB_LAUNCH_ONE(MODE_, TF_)                                                                                               \
  do {                                                                                                                            \
    FB_CHECK_CUDA(cudaFuncSetAttribute(grouped_gemm_kernel<MODE_, false, TF_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    FB_CHECK_CUDA(fb_launch_pdl(grouped_gemm_kernel<MODE_, false, TF_>, dim3(grid), dim3(GEMM_THREADS), (size_t)smem,           \
                                (cudaStream_t)stream, tmA, tmB, tmD, a));                                                        \
  } while (0)
  if (pair) {
    grid &= ~1;
    if (grid > 2 * total) grid = 2 * total;
    if (tf32) {
      if (mode == 3) FB_LAUNCH_PAIR(3, true);
      else if (mode == 2) FB_LAUNCH_PAIR(2, true);
      else FB_LAUNCH_PAIR(0, true);
    } else {
      if (mode == 3) FB_LAUNCH_PAIR(3, false);
      else if (mode == 2) FB_LAUNCH_PAIR(2, false);
      else FB_LAUNCH_PAIR(0, false);
    }
    return FB_OK;
  }
  if (grid > total) grid = total;
  if (tf32) {
    if (mode == 3) FB_LAUNCH_ONE(3, true);
    else if (mode == 2) FB_LAUNCH_ONE(2, true);
    else if (a.stats) FB_LAUNCH_ONE_ST(0, true);
    else FB_LAUNCH_ONE(0, true);
  } else {
    if (mode == 3) FB_LAUNCH_ONE(3, false);
    else if (mode == 2) FB_LAUNCH_ONE(2, false);
    else if (a.stats) FB_LAUNCH_ONE_ST(0, false);
    else FB_LAUNCH_ONE(0, false);
  }
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- implicit 3x3 convolution (MODE 4)
static CUtensorMapSwizzle swizzle_for(int inner_bytes) {
  return inner_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : inner_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
}

static int conv3x3_impl(const FbConv3x3* p, fb_stream_t stream, const bool tf32) {
  const int es = tf32 ? 4 : 2, BK = 128 / es;
  FB_REQUIRE(p && p->x && p->w && p->y, "fb_conv3x3: null operand");
  FB_REQUIRE(p->G >= 1 && p->B >= 1 && p->H >= 1 && p->W >= 1, "fb_conv3x3: empty problem");
  FB_REQUIRE(p->stride == 1 || p->stride == 2, "fb_conv3x3: stride must be 1 or 2");
  FB_REQUIRE(!(p->dgrad && p->stride != 1), "fb_conv3x3: the data-gradient form needs stride 1");
  FB_REQUIRE(p->cin * es >= 32 && (p->cin <= BK ? (p->cin & (p->cin - 1)) == 0 : p->cin % BK == 0),
             "fb_conv3x3: input channels must fill 32 or 64 bytes or a multiple of 128 bytes (got %d)", p->cin);
  FB_REQUIRE(p->cout >= 8 && p->cout % 8 == 0, "fb_conv3x3: output channels must be a multiple of 8");
  const int s = p->stride;
  const int Ho = (p->H - 1) / s + 1, Wo = (p->W - 1) / s + 1;
  FB_REQUIRE(Wo <= 128 && (Wo & (Wo - 1)) == 0 && (Ho & (Ho - 1)) == 0 && Wo * s <= 256,
             "fb_conv3x3: output plane must be a power of two per side, at most 128 wide (got %d x %d)", Ho, Wo);
  const int plane = Ho * Wo;
  const int bh = plane >= GEMM_BM ? GEMM_BM / Wo : Ho;          // output rows per tile
  const int bn = plane >= GEMM_BM ? 1 : GEMM_BM / plane;        // images per tile
  FB_REQUIRE(bh >= 1 && bh * s <= 256 && bn <= 256, "fb_conv3x3: unsupported plane shape");
  PFN_encodeTiled enc = get_encode();
  FB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  FB_REQUIRE(((uintptr_t)p->x % 16) == 0 && ((uintptr_t)p->w % 16) == 0 && ((uintptr_t)p->y % 16) == 0 && (p->w_gs * es) % 16 == 0,
             "fb_conv3x3: operands must be 16-byte aligned");

  const int kblk = p->cin < BK ? p->cin : BK;
  const int inner_bytes = kblk * es;
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.d = p->y; a.bias = nullptr;
  a.G = p->G; a.M = p->B * plane; a.N = p->cout; a.K = 9 * p->cin;
  a.ldd = p->cout; a.d_fp32 = tf32 ? 1 : 0; a.relu = 0;
  a.d_gs = (int64_t)a.M * a.N;
  int bn_cols = (p->cout + 15) / 16 * 16;
  if (bn_cols > 64) {
    const int nt = (p->cout + 255) / 256;
    bn_cols = ((p->cout + nt - 1) / nt + 63) / 64 * 64;
  }
  // CTA pairs (cta_group::2): 256 output pixels x bn filters per pair, each CTA gathers its own 128 pixels and stages
  // half of the filter rows.  Wide layers only: the halves must be whole 64-row blocks.  OPT-IN (FB_CONV_PAIR=1): measured
  // on ResNet-18 x 64 clients it is not faster (54.7 vs 54.5 ms/round; these convolutions are pipeline-latency bound, not
  // L2-traffic bound), and only one forward shape (stride 2, 64 -> 128 channels) is checked against F.conv2d; the
  // data-gradient form of the pair path has no oracle test yet.
  static int conv_pair_env = -1;
  if (conv_pair_env < 0) {
    const char* e = getenv("FB_CONV_PAIR");
    conv_pair_env = e ? atoi(e) : 0;
  }
  const bool pair = conv_pair_env != 0 && p->cout >= 128 && (int64_t)p->B * plane >= 512 && inner_bytes == 128 && !p->stats;
  if (pair) {
    const int nt = (p->cout + 255) / 256;
    bn_cols = ((p->cout + nt - 1) / nt + 127) / 128 * 128;
  }
  a.BN = bn_cols; a.BNpad = bn_cols;
  a.gi = p->G; a.gi_out = p->G; a.group_div = 1; a.d_gs2 = 0;
  a.tma_store = ((uint64_t)p->cout * es) % 16 == 0;
  a.chunk_cols = tf32 ? (bn_cols >= 32 ? 32 : 16) : (bn_cols >= 64 ? 64 : bn_cols);
  a.chunk_bytes = a.chunk_cols * es;
  a.n_tiles = (p->cout + bn_cols - 1) / bn_cols;
  a.m_tiles = pair ? (a.M + 2 * GEMM_BM - 1) / (2 * GEMM_BM) : (a.M + GEMM_BM - 1) / GEMM_BM;
  a.cv_B = p->B; a.cv_HoWo = plane; a.cv_Wo = Wo; a.cv_stride = s; a.cv_flip = p->dgrad ? 1 : 0;
  a.cv_kblk = kblk; a.cv_kcs = p->cin / kblk; a.k_blocks = 9 * a.cv_kcs;
  a.cv_sbo = 8 * inner_bytes;
  a.cv_swz = inner_bytes == 32 ? 6 : inner_bytes == 64 ? 4 : 2;
  // weights: forward [G][cout][9*cin] (tap-major columns); data gradient [G][9*cout_fwd][cin_fwd] = per tap a
  // [N = this call's cout][K = this call's cin] matrix, rows tap*cout + n
  a.cv_b_inner_off = p->dgrad ? 0 : p->cin;
  a.cv_b_row_off = p->dgrad ? p->cout : 0;
  int cols = 32;
  while (cols < 2 * bn_cols) cols <<= 1;
  a.tmem_cols = cols;
  a.stats = p->stats;
  a.stats_gs = p->stats_gs;
  FB_REQUIRE(!a.stats || a.tma_store, "fb_conv3x3: output statistics need cout * element size to be a multiple of 16 bytes");

  CUtensorMap tmA, tmB, tmD;
  {
    cuuint64_t dims[4] = {(cuuint64_t)p->cin, (cuuint64_t)p->W, (cuuint64_t)p->H, (cuuint64_t)p->G * p->B};
    cuuint64_t strides[3] = {(cuuint64_t)p->cin * es, (cuuint64_t)p->cin * p->W * es, (cuuint64_t)p->cin * p->W * p->H * es};
    cuuint32_t box[4] = {(cuuint32_t)kblk, (cuuint32_t)(Wo * s), (cuuint32_t)(bh * s), (cuuint32_t)bn};
    cuuint32_t estr[4] = {1, (cuuint32_t)s, (cuuint32_t)s, 1};
    CUresult r = enc(&tmA, tf32 ? FB_TF32_LOAD_TYPE : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p->x), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(inner_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (conv input) failed with code %d", (int)r);
  }
  {
    const uint64_t inner = p->dgrad ? (uint64_t)p->cin : (uint64_t)9 * p->cin;
    const uint64_t rows = p->dgrad ? (uint64_t)9 * p->cout : (uint64_t)p->cout;
    cuuint64_t dims[4] = {inner, rows, (cuuint64_t)p->G, 1};
    cuuint64_t strides[3] = {inner * es, (cuuint64_t)(p->G > 1 ? p->w_gs : inner * rows) * es, inner * rows * p->G * es};
    cuuint32_t box[4] = {(cuuint32_t)kblk, (cuuint32_t)(pair ? bn_cols / 2 : bn_cols), 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tmB, tf32 ? FB_TF32_LOAD_TYPE : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p->w), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(inner_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (conv weights) failed with code %d", (int)r);
  }
  memset(&tmD, 0, sizeof(tmD));
  if (a.tma_store) {
    int rc = make_store_map(&tmD, p->y, tf32 ? 1 : 0, (uint64_t)p->cout, (uint64_t)a.M, (uint64_t)p->G, 1, (uint64_t)p->cout,
                            (uint64_t)a.d_gs, 0, (uint32_t)a.chunk_cols, a.chunk_bytes == 128);
    if (rc) return rc;
  }
  const size_t stage_bytes = GEMM_A_BYTES + (size_t)(pair ? a.BNpad / 2 : a.BNpad) * 128;   // as in the kernel
  const size_t fixed = 1024 + GEMM_STAGING_BYTES + 256 + (a.stats ? GEMM_STATS_BYTES : 0);
  int ctas_per_sm = (a.tmem_cols <= 256) ? 2 : 1;
  if (pair) ctas_per_sm = 1;
  {
    static int conv_ctas = -1;
    if (conv_ctas < 0) {
      const char* e = getenv("FB_CONV_CTAS_PER_SM");
      conv_ctas = e ? atoi(e) : 0;
    }
    if (conv_ctas == 1) ctas_per_sm = 1;
  }
  int stages = (int)((227 * 1024 / ctas_per_sm - fixed) / stage_bytes);
  if (stages < 3 && ctas_per_sm == 2) {
    ctas_per_sm = 1;
    stages = (int)((227 * 1024 - fixed) / stage_bytes);
  }
  if (stages > GEMM_MAX_STAGES) stages = GEMM_MAX_STAGES;
  FB_REQUIRE(stages >= 2, "fb_conv3x3: tile does not fit shared memory");
  a.stages = stages;
  const size_t smem = (size_t)stages * stage_bytes + fixed;
  const int total = a.G * a.m_tiles * a.n_tiles;
  int grid = fb_num_sms() * ctas_per_sm;
  if (grid > total) grid = total;
  if (pair) {
    grid = fb_num_sms() & ~1;
    if (grid > 2 * total) grid = 2 * total;
    if (tf32) FB_LAUNCH_PAIR(4, true);
    else FB_LAUNCH_PAIR(4, false);
    return FB_OK;
  }
  if (a.stats) {
    if (tf32) FB_LAUNCH_ONE_ST(4, true);
    else FB_LAUNCH_ONE_ST(4, false);
  } else if (tf32) FB_LAUNCH_ONE(4, true);
  else FB_LAUNCH_ONE(4, false);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_conv3x3_bf16(const FbConv3x3* p, fb_stream_t stream) { return conv3x3_impl(p, stream, false); }
extern "C" int fb_conv3x3_tf32(const FbConv3x3* p, fb_stream_t stream) { return conv3x3_impl(p, stream, true); }


static int conv3x3_wgrad_impl(const FbConv3x3Wgrad* p, fb_stream_t stream, const bool tf32) {
  const int es = tf32 ? 4 : 2, BK = 128 / es;
  FB_REQUIRE(p && p->x && p->dy && p->dw, "fb_conv3x3_wgrad: null operand");
  FB_REQUIRE(p->G >= 1 && p->B >= 1 && p->H >= 1 && p->W >= 1, "fb_conv3x3_wgrad: empty problem");
  FB_REQUIRE(p->stride == 1 || p->stride == 2, "fb_conv3x3_wgrad: stride must be 1 or 2");
  // fp32 operands fed MN-major exist only in the 128-byte-row layout (see the kernel): whole 32-channel blocks
  // tf32 with 16 input channels: every tap is its own 32-column tile, the TMA box is 32 channels wide on a 16-channel
  // tensor (the upper half is zero-filled) and only 16 columns of the tile are stored.  Correct but slow (nine tiles per
  // client, 320 k-blocks of 32 pixels each: ResNet-56 round 97.8 -> 149.9 ms when the 16-channel layers use it), so the
  // executor keeps im2col + TN GEMM below 64 channels (FB_IMPLICIT_WGRAD_MIN_CIN)
  const bool pad16 = tf32 && p->cin == 16;
  FB_REQUIRE(pad16 || (p->cin % 16 == 0 && (p->cin <= BK ? (!tf32 && (p->cin & (p->cin - 1)) == 0) || p->cin == BK : (p->cin % BK == 0)) &&
                 (p->cin <= 256 || p->cin % 256 == 0)),
             "fb_conv3x3_wgrad: input channels must fill 32 or 64 bytes, a multiple of 128 bytes up to 256 channels, or a "
             "multiple of 256 channels (got %d)", p->cin);
  FB_REQUIRE(p->cout >= 8 && p->cout % 8 == 0, "fb_conv3x3_wgrad: output channels must be a multiple of 8");
  const int s = p->stride;
  const int Ho = (p->H - 1) / s + 1, Wo = (p->W - 1) / s + 1;
  FB_REQUIRE(Wo <= 64 && (Wo & (Wo - 1)) == 0 && (Ho & (Ho - 1)) == 0,
             "fb_conv3x3_wgrad: output plane must be a power of two per side, at most 64 wide (got %d x %d)", Ho, Wo);
  const int plane = Ho * Wo;
  const int S = p->split > 0 ? p->split : 1;
  const int64_t Ktot = (int64_t)p->B * plane;
  FB_REQUIRE(Ktot % ((int64_t)S * BK) == 0, "fb_conv3x3_wgrad: B*Ho*Wo must be a multiple of %d x split", BK);
  const int Ks = (int)(Ktot / S);
  const int bh = plane >= BK ? BK / Wo : Ho;
  const int bn = plane >= BK ? 1 : BK / plane;
  PFN_encodeTiled enc = get_encode();
  FB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  FB_REQUIRE(((uintptr_t)p->x % 16) == 0 && ((uintptr_t)p->dy % 16) == 0 && ((uintptr_t)p->dw % 16) == 0 &&
                 (p->G == 1 || p->dw_gs % 4 == 0),
             "fb_conv3x3_wgrad: operands must be 16-byte aligned");
  const int kblk = pad16 ? BK : (p->cin < BK ? p->cin : BK);
  const int row_bytes = kblk * es;
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.d = p->dw; a.bias = nullptr;
  a.G = p->G * S; a.M = p->cout; a.N = 9 * p->cin; a.K = Ks;
  a.ldd = 9 * p->cin; a.d_fp32 = 1; a.relu = 0;
  a.d_gs = p->dw_gs; a.d_gs2 = 0;
  // taps per n-tile: all nine when they fit one 256-column tile, else three, else one tap (and a channel chunk);
  // dy (the A operand) is then streamed 1 / 3 / 9 times
  const int tpt = pad16 ? 1 : (9 * p->cin <= 256 ? 9 : (3 * p->cin <= 256 ? 3 : 1));
  const int bn_tap = pad16 ? BK : (p->cin < 256 ? p->cin : 256);
  a.cv_tpt = tpt;
  a.cv_out_cols = pad16 ? 16 : 0;
  a.BN = tpt * bn_tap;
  a.BNpad = (a.BN + BK - 1) / BK * BK;
  a.gi = a.G; a.group_div = S; a.gi_out = p->G;
  a.tma_store = 1;
  a.reduce_add = (S > 1 || p->accumulate) ? 1 : 0;
  a.chunk_cols = (a.BN >= 32 && !pad16) ? 32 : 16;
  a.chunk_bytes = a.chunk_cols * 4;
  a.cv_kcs = pad16 ? 1 : p->cin / bn_tap;
  a.n_tiles = (9 / tpt) * a.cv_kcs;
  a.m_tiles = (p->cout + GEMM_BM - 1) / GEMM_BM;
  a.k_blocks = Ks / BK;
  a.cv_B = p->B; a.cv_HoWo = plane; a.cv_Wo = Wo; a.cv_stride = s;
  a.cv_kblk = kblk;
  a.cv_sbo = tf32 ? 512 : 8 * row_bytes;
  a.cv_swz = tf32 ? 1 : (row_bytes == 32 ? 6 : row_bytes == 64 ? 4 : 2);
  int cols = 32;
  while (cols < 2 * a.BN) cols <<= 1;
  a.tmem_cols = cols;
  CUtensorMap tmA, tmB, tmD;
  int rc = make_map(&tmA, p->dy, (uint64_t)p->cout, (uint64_t)Ks, (uint64_t)a.G, 1, (uint64_t)p->cout, (uint64_t)Ks * p->cout, 0,
                    (uint32_t)BK, tf32, true);
  if (rc) return rc;
  {
    cuuint64_t dims[4] = {(cuuint64_t)p->cin, (cuuint64_t)p->W, (cuuint64_t)p->H, (cuuint64_t)p->G * p->B};
    cuuint64_t strides[3] = {(cuuint64_t)p->cin * es, (cuuint64_t)p->cin * p->W * es, (cuuint64_t)p->cin * p->W * p->H * es};
    cuuint32_t box[4] = {(cuuint32_t)kblk, (cuuint32_t)(Wo * s), (cuuint32_t)(bh * s), (cuuint32_t)bn};
    cuuint32_t estr[4] = {1, (cuuint32_t)s, (cuuint32_t)s, 1};
    CUresult r = enc(&tmB, tf32 ? FB_TF32_LOAD_TYPE : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p->x), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, tf32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : swizzle_for(row_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (wgrad input) failed with code %d", (int)r);
  }
  rc = make_store_map(&tmD, p->dw, 1, (uint64_t)a.N, (uint64_t)a.M, (uint64_t)p->G, 1, (uint64_t)a.ldd, (uint64_t)p->dw_gs, 0,
                      (uint32_t)a.chunk_cols, a.chunk_bytes == 128);
  if (rc) return rc;
  const size_t stage_bytes = GEMM_A_BYTES + (size_t)a.BNpad * 128;
  const size_t fixed = 1024 + GEMM_STAGING_BYTES + 256;
  int ctas_per_sm = (a.tmem_cols <= 256) ? 2 : 1;
  {
    static int conv_ctas = -1;
    if (conv_ctas < 0) {
      const char* e = getenv("FB_CONV_CTAS_PER_SM");
      conv_ctas = e ? atoi(e) : 0;
    }
    if (conv_ctas == 1) ctas_per_sm = 1;
  }
  int stages = (int)((227 * 1024 / ctas_per_sm - fixed) / stage_bytes);
  if (stages < 3 && ctas_per_sm == 2) {
    ctas_per_sm = 1;
    stages = (int)((227 * 1024 - fixed) / stage_bytes);
  }
  if (stages > GEMM_MAX_STAGES) stages = GEMM_MAX_STAGES;
  FB_REQUIRE(stages >= 2, "fb_conv3x3_wgrad: tile does not fit shared memory");
  a.stages = stages;
  const size_t smem = (size_t)stages * stage_bytes + fixed;
  const int total = a.G * a.m_tiles * a.n_tiles;
  int grid = fb_num_sms() * ctas_per_sm;
  if (grid > total) grid = total;
  if (tf32) FB_LAUNCH_ONE(5, true);
  else FB_LAUNCH_ONE(5, false);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_conv3x3_wgrad_bf16(const FbConv3x3Wgrad* p, fb_stream_t stream) { return conv3x3_wgrad_impl(p, stream, false); }
extern "C" int fb_conv3x3_wgrad_tf32(const FbConv3x3Wgrad* p, fb_stream_t stream) { return conv3x3_wgrad_impl(p, stream, true); }


extern "C" int fb_grouped_gemm_bf16(const FbGroupedGemm* p, fb_stream_t stream) { return launch_gemm(p, 0, stream); }
extern "C" int fb_grouped_gemm_tn_bf16(const FbGroupedGemm* p, fb_stream_t stream) { return launch_gemm(p, 3, stream); }
extern "C" int fb_grouped_gemm_nn_bf16(const FbGroupedGemm* p, fb_stream_t stream) { return launch_gemm(p, 2, stream); }
/* fp32 operands, tcgen05 kind::tf32 (see the kernel template): same structs, strides in fp32 elements */
extern "C" int fb_grouped_gemm_tf32(const FbGroupedGemm* p, fb_stream_t stream) { return launch_gemm(p, 0, stream, true); }
extern "C" int fb_grouped_gemm_tn_tf32(const FbGroupedGemm* p, fb_stream_t stream) { return launch_gemm(p, 3, stream, true); }
extern "C" int fb_grouped_gemm_nn_tf32(const FbGroupedGemm* p, fb_stream_t stream) { return launch_gemm(p, 2, stream, true); }
