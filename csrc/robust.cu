// K5/K6: coordinate-wise median and trimmed mean; K7 helpers (pack + scores).  See include/fedml_b200.h.
//
// Selection strategy: a tile of TILE coordinates x K clients is staged in shared memory as
// keys[k][t] (thread index t fastest => conflict-free), floats mapped to order-preserving uint32
// keys.  Each coordinate is owned by SPLIT=4 adjacent lanes which cooperatively run a bitwise
// (radix-2, MSB first) selection for the wanted order statistics: 32 passes of K/4 compares per
// lane + 2 shuffles.  No sort, no global scratch, every HBM/NVLink byte is read exactly once.
#include "common.cuh"
#include <stdlib.h>
#include "../include/fedml_b200.h"

#define SEL_THREADS 256
#define SEL_SPLIT 4
#define SEL_TILE (SEL_THREADS / SEL_SPLIT)  // 64 coordinates per block-iteration
#define SEL_LD (SEL_TILE + 8)               // row stride: the 4 lanes of a coordinate (k, k+1, k+2, k+3) hit 4 different banks

__device__ __forceinline__ uint32_t f2key(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

// k-th smallest key (0-based rank) of column `col`; the SPLIT lanes of a coordinate call it together.
__device__ __forceinline__ uint32_t select_rank(const uint32_t* keys, int K, int col, int sub, int rank) {
  uint32_t prefix = 0;  // bits decided so far
  int remaining = rank; // rank inside the candidate set that matches `prefix` on the decided bits
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t mask_hi = bit == 31 ? 0u : (0xffffffffu << (bit + 1));
    int cnt0 = 0;  // candidates whose current bit is 0
    for (int k = sub; k < K; k += SEL_SPLIT) {
      const uint32_t x = keys[(size_t)k * SEL_LD + col];
      const bool cand = ((x ^ prefix) & mask_hi) == 0u;
      cnt0 += (cand && !((x >> bit) & 1u)) ? 1 : 0;
    }
    cnt0 += __shfl_xor_sync(0xffffffffu, cnt0, 1);
    cnt0 += __shfl_xor_sync(0xffffffffu, cnt0, 2);
    if (remaining >= cnt0) {
      remaining -= cnt0;
      prefix |= (1u << bit);
    }
  }
  return prefix;
}

__global__ void __launch_bounds__(SEL_THREADS) robust_select_kernel(const float* const* __restrict__ rows, int K,
                                                                    int64_t j0, int64_t j1, int mode, int trim,
                                                                    float* const* __restrict__ outs, int n_out) {
  extern __shared__ uint32_t keys[];  // [K][SEL_LD]
  const int tid = threadIdx.x;
  const int col = tid / SEL_SPLIT, sub = tid % SEL_SPLIT;
  const int64_t ntiles = (j1 - j0 + SEL_TILE - 1) / SEL_TILE;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = j0 + tile * SEL_TILE;
    const int width = (int)min((int64_t)SEL_TILE, j1 - base);
    __syncthreads();  // previous tile fully consumed
    // stage: consecutive threads read consecutive coordinates of one client row (coalesced 256 B)
    for (int idx = tid; idx < K * SEL_TILE; idx += SEL_THREADS) {
      const int k = idx / SEL_TILE, t = idx % SEL_TILE;
      float x = 0.f;
      if (t < width) {
        const float* src = rows[k] + base + t;
        asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(x) : "l"(src) : "memory");
      }
      keys[(size_t)k * SEL_LD + t] = f2key(x);
    }
    __syncthreads();
    float result;
    if (mode == 0) {
      result = key2f(select_rank(keys, K, col, sub, (K - 1) / 2));
    } else {
      const uint32_t klo = select_rank(keys, K, col, sub, trim);
      const uint32_t khi = select_rank(keys, K, col, sub, K - trim - 1);
      // every lane runs the sums and the shuffles: coordinates whose K values are all equal (klo == khi, e.g. integer
      // buffers that every client holds identically) share warps with coordinates that differ, and a full-mask shuffle
      // inside a branch only some lanes take never completes
      {
        float sum = 0.f;
        int lt_lo = 0, eq_lo = 0, lt_hi = 0;
        for (int k = sub; k < K; k += SEL_SPLIT) {
          const uint32_t x = keys[(size_t)k * SEL_LD + col];
          lt_lo += x < klo;
          eq_lo += x == klo;
          lt_hi += x < khi;
          if (x > klo && x < khi) sum += key2f(x);
        }
#pragma unroll
        for (int o = 1; o < SEL_SPLIT; o <<= 1) {
          sum += __shfl_xor_sync(0xffffffffu, sum, o);
          lt_lo += __shfl_xor_sync(0xffffffffu, lt_lo, o);
          eq_lo += __shfl_xor_sync(0xffffffffu, eq_lo, o);
          lt_hi += __shfl_xor_sync(0xffffffffu, lt_hi, o);
        }
        const int n_lo = lt_lo + eq_lo - trim;       // copies of v_lo with rank >= trim
        const int n_hi = (K - trim) - lt_hi;         // copies of v_hi with rank <  K - trim
        sum += (float)n_lo * key2f(klo) + (float)n_hi * key2f(khi);
        result = klo == khi ? key2f(klo) : sum / (float)(K - 2 * trim);
      }
    }
    if (sub == 0 && col < width) {
      for (int o = 0; o < n_out; ++o) outs[o][base + col] = result;
    }
  }
}

// Register sorting network: for up to 128 clients every thread owns one coordinate, holds the K values in registers
// (padded with +inf to a power of two) and runs a fully unrolled bitonic network of min / max pairs: K log^2 K / 4
// compare-exchanges instead of 32 counting passes over shared memory (7.8 -> ~1 ms for 64 x 11.2 M).
#define SORT_THREADS 128
template <int KP>
__global__ void __launch_bounds__(SORT_THREADS) robust_sort_kernel(const float* const* __restrict__ rows, int K, int64_t j0,
                                                                   int64_t j1, int mode, int trim,
                                                                   float* const* __restrict__ outs, int n_out) {
  for (int64_t j = j0 + (int64_t)blockIdx.x * SORT_THREADS + threadIdx.x; j < j1; j += (int64_t)gridDim.x * SORT_THREADS) {
    float v[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) {
      v[k] = INFINITY;
      if (k < K) {
        const float* src = rows[k] + j;
        asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v[k]) : "l"(src) : "memory");
      }
    }
#pragma unroll
    for (int size = 2; size <= KP; size <<= 1) {
#pragma unroll
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
        for (int i = 0; i < KP; ++i) {
          const int l = i ^ stride;
          if (l > i) {
            const float lo = fminf(v[i], v[l]), hi = fmaxf(v[i], v[l]);
            const bool up = (i & size) == 0;
            v[i] = up ? lo : hi;
            v[l] = up ? hi : lo;
          }
        }
      }
    }
    float result = 0.f;
    if (mode == 0) {
      const int idx = (K - 1) / 2;                       // lower median, as torch.median
#pragma unroll
      for (int k = 0; k < KP; ++k)
        if (k == idx) result = v[k];
    } else {
#pragma unroll
      for (int k = 0; k < KP; ++k)
        if (k >= trim && k < K - trim) result += v[k];
      result /= (float)(K - 2 * trim);
    }
    for (int o = 0; o < n_out; ++o) outs[o][j] = result;
  }
}
template <int KP>
static int launch_robust_sort(const float* const* rows_dev, int K, int64_t j0, int64_t j1, int mode, int trim,
                              float* const* outs_dev, int n_out, cudaStream_t stream) {
  const int64_t nt = (j1 - j0 + SORT_THREADS - 1) / SORT_THREADS;
  int64_t grid = (int64_t)fb_num_sms() * 8;
  if (grid > nt) grid = nt;
  robust_sort_kernel<KP><<<(int)grid, SORT_THREADS, 0, stream>>>(rows_dev, K, j0, j1, mode, trim, outs_dev, n_out);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

extern "C" int fb_robust_select(const float* const* rows_dev, int32_t K, int64_t j0, int64_t j1, int32_t mode,
                                int32_t trim, float* const* outs_dev, int32_t n_out, fb_stream_t stream) {
  FB_REQUIRE(rows_dev && outs_dev && K >= 1 && j1 >= j0 && n_out >= 1, "fb_robust_select: bad arguments");
  FB_REQUIRE(mode == 0 || (trim >= 0 && 2 * trim < K), "fb_robust_select: trim removes every client");
  const size_t smem = (size_t)K * SEL_LD * sizeof(uint32_t);
  if (smem > 220 * 1024) {
    fb_set_error("fb_robust_select: K=%d exceeds the %d clients that fit one shared-memory tile", K,
                 (int)(220 * 1024 / (SEL_LD * sizeof(uint32_t))));
    return FB_ERR_LIMIT;
  }
  if (j1 == j0) return FB_OK;
  static int use_sort = -1;
  if (use_sort < 0) {
    const char* e = getenv("FB_ROBUST_SORT");
    use_sort = (e && atoi(e) == 0) ? 0 : 1;
  }
  if (use_sort && K <= 128) {
    cudaStream_t st = (cudaStream_t)stream;
    if (K <= 8) return launch_robust_sort<8>(rows_dev, K, j0, j1, mode, trim, outs_dev, n_out, st);
    if (K <= 16) return launch_robust_sort<16>(rows_dev, K, j0, j1, mode, trim, outs_dev, n_out, st);
    if (K <= 32) return launch_robust_sort<32>(rows_dev, K, j0, j1, mode, trim, outs_dev, n_out, st);
    if (K <= 64) return launch_robust_sort<64>(rows_dev, K, j0, j1, mode, trim, outs_dev, n_out, st);
    return launch_robust_sort<128>(rows_dev, K, j0, j1, mode, trim, outs_dev, n_out, st);
  }
  FB_CHECK_CUDA(cudaFuncSetAttribute(robust_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, robust_select_kernel, SEL_THREADS, smem);
  if (per_sm < 1) per_sm = 1;
  const int64_t ntiles = (j1 - j0 + SEL_TILE - 1) / SEL_TILE;
  int64_t grid = (int64_t)per_sm * fb_num_sms();
  if (grid > ntiles) grid = ntiles;
  robust_select_kernel<<<(int)grid, SEL_THREADS, smem, (cudaStream_t)stream>>>(rows_dev, K, j0, j1, mode, trim,
                                                                                outs_dev, n_out);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- many clients (K > 128): sorted runs + merge-select
// Step 1 (fb_robust_sort_runs): the rows are cut into runs of `run_len` <= 128 consecutive rows; for every coordinate
// each run is sorted IN PLACE with the register network above (a run's rows live on one GPU, the GPU that trained
// them, so this is local HBM traffic; median / trimmed mean do not care which client a value came from).
// Step 2 (fb_robust_select_runs): the owner of a coordinate shard stages a tile of all K values in shared memory as
// before and finds the wanted order statistic bit by bit, but a bit step now costs one binary search per run
// (sum over runs of lower_bound(run, T) = number of keys < T) instead of a counting pass over all K keys:
// 32 x R x log2(run_len) probes per coordinate instead of 32 x K (K = 512, runs of 64: 1536 instead of 16384).
template <int KP>
__global__ void __launch_bounds__(SORT_THREADS) robust_sort_runs_kernel(float* const* __restrict__ rows, int K, int run_len,
                                                                        int64_t j0, int64_t j1) {
  const int r0 = blockIdx.y * run_len;
  const int len = min(run_len, K - r0);
  for (int64_t j = j0 + (int64_t)blockIdx.x * SORT_THREADS + threadIdx.x; j < j1; j += (int64_t)gridDim.x * SORT_THREADS) {
    float v[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) {
      v[k] = INFINITY;
      if (k < len) v[k] = rows[r0 + k][j];
    }
#pragma unroll
    for (int size = 2; size <= KP; size <<= 1) {
#pragma unroll
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
        for (int i = 0; i < KP; ++i) {
          const int l = i ^ stride;
          if (l > i) {
            const float lo = fminf(v[i], v[l]), hi = fmaxf(v[i], v[l]);
            const bool up = (i & size) == 0;
            v[i] = up ? lo : hi;
            v[l] = up ? hi : lo;
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < KP; ++k)
      if (k < len) rows[r0 + k][j] = v[k];
  }
}
extern "C" int fb_robust_sort_runs(float* const* rows_dev, int32_t K, int32_t run_len, int64_t j0, int64_t j1,
                                   fb_stream_t stream) {
  FB_REQUIRE(rows_dev && K >= 1 && run_len >= 1 && run_len <= 128 && j1 >= j0, "fb_robust_sort_runs: run_len must be in [1, 128]");
  if (j1 == j0) return FB_OK;
  const int runs = (K + run_len - 1) / run_len;
  const int64_t nt = (j1 - j0 + SORT_THREADS - 1) / SORT_THREADS;
  int64_t gx = ((int64_t)fb_num_sms() * 8 + runs - 1) / runs;
  if (gx > nt) gx = nt;
  if (gx < 1) gx = 1;
  const dim3 grid((unsigned)gx, (unsigned)runs);
  cudaStream_t st = (cudaStream_t)stream;
  if (run_len <= 8) robust_sort_runs_kernel<8><<<grid, SORT_THREADS, 0, st>>>(rows_dev, K, run_len, j0, j1);
  else if (run_len <= 16) robust_sort_runs_kernel<16><<<grid, SORT_THREADS, 0, st>>>(rows_dev, K, run_len, j0, j1);
  else if (run_len <= 32) robust_sort_runs_kernel<32><<<grid, SORT_THREADS, 0, st>>>(rows_dev, K, run_len, j0, j1);
  else if (run_len <= 64) robust_sort_runs_kernel<64><<<grid, SORT_THREADS, 0, st>>>(rows_dev, K, run_len, j0, j1);
  else robust_sort_runs_kernel<128><<<grid, SORT_THREADS, 0, st>>>(rows_dev, K, run_len, j0, j1);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// rank-th smallest key of column `col` when rows [r*L, (r+1)*L) of `keys` are sorted ascending for every run r
__device__ __forceinline__ uint32_t select_rank_runs(const uint32_t* keys, int K, int L, int col, int sub, int rank) {
  const int R = (K + L - 1) / L;
  uint32_t prefix = 0;
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t T = prefix | (1u << bit);
    int cnt = 0;                                    // keys < T
    for (int r = sub; r < R; r += SEL_SPLIT) {
      const uint32_t* run = keys + (size_t)r * L * SEL_LD + col;
      int lo = 0, hi = min(L, K - r * L);           // lower_bound over run[0 .. hi)
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (run[(size_t)mid * SEL_LD] < T) lo = mid + 1; else hi = mid;
      }
      cnt += lo;
    }
    cnt += __shfl_xor_sync(0xffffffffu, cnt, 1);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, 2);
    if (rank >= cnt) prefix = T;                    // at most `rank` keys lie below T: the answer is >= T
  }
  return prefix;
}

// K rows are staged; only the K_valid smallest values per coordinate are real (padding rows hold +inf and sort last)
__global__ void __launch_bounds__(SEL_THREADS) robust_select_runs_kernel(const float* const* __restrict__ rows, int K, int K_valid, int L,
                                                                         int64_t j0, int64_t j1, int mode, int trim,
                                                                         float* const* __restrict__ outs, int n_out) {
  extern __shared__ uint32_t keys[];  // [K][SEL_LD]
  const int tid = threadIdx.x;
  const int col = tid / SEL_SPLIT, sub = tid % SEL_SPLIT;
  const int64_t ntiles = (j1 - j0 + SEL_TILE - 1) / SEL_TILE;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = j0 + tile * SEL_TILE;
    const int width = (int)min((int64_t)SEL_TILE, j1 - base);
    __syncthreads();
    if (width == SEL_TILE && ((base & 3) == 0)) {
      // 16-byte loads: a quarter-warp covers the 64 coordinates of one row
      for (int idx = tid; idx < K * (SEL_TILE / 4); idx += SEL_THREADS) {
        const int k = idx / (SEL_TILE / 4), t4 = idx % (SEL_TILE / 4);
        float4 x;
        const float* src = rows[k] + base + t4 * 4;
        asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "l"(src) : "memory");
        uint32_t* dst = keys + (size_t)k * SEL_LD + t4 * 4;
        dst[0] = f2key(x.x); dst[1] = f2key(x.y); dst[2] = f2key(x.z); dst[3] = f2key(x.w);
      }
    } else {
      for (int idx = tid; idx < K * SEL_TILE; idx += SEL_THREADS) {
        const int k = idx / SEL_TILE, t = idx % SEL_TILE;
        float x = 0.f;
        if (t < width) {
          const float* src = rows[k] + base + t;
          asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(x) : "l"(src) : "memory");
        }
        keys[(size_t)k * SEL_LD + t] = f2key(x);
      }
    }
    __syncthreads();
    float result;
    if (mode == 0) {
      result = key2f(select_rank_runs(keys, K, L, col, sub, (K_valid - 1) / 2));
    } else {
      const uint32_t klo = select_rank_runs(keys, K, L, col, sub, trim);
      const uint32_t khi = select_rank_runs(keys, K, L, col, sub, K_valid - trim - 1);
      // every lane runs the sums and the shuffles: coordinates whose K values are all equal (klo == khi, e.g. integer
      // buffers that every client holds identically) share warps with coordinates that differ, and a full-mask shuffle
      // inside a branch only some lanes take never completes
      {
        float sum = 0.f;
        int lt_lo = 0, eq_lo = 0, lt_hi = 0;
        for (int k = sub; k < K; k += SEL_SPLIT) {
          const uint32_t x = keys[(size_t)k * SEL_LD + col];
          lt_lo += x < klo;
          eq_lo += x == klo;
          lt_hi += x < khi;
          if (x > klo && x < khi) sum += key2f(x);
        }
#pragma unroll
        for (int o = 1; o < SEL_SPLIT; o <<= 1) {
          sum += __shfl_xor_sync(0xffffffffu, sum, o);
          lt_lo += __shfl_xor_sync(0xffffffffu, lt_lo, o);
          eq_lo += __shfl_xor_sync(0xffffffffu, eq_lo, o);
          lt_hi += __shfl_xor_sync(0xffffffffu, lt_hi, o);
        }
        const int n_lo = lt_lo + eq_lo - trim;
        const int n_hi = (K_valid - trim) - lt_hi;
        sum += (float)n_lo * key2f(klo) + (float)n_hi * key2f(khi);
        result = klo == khi ? key2f(klo) : sum / (float)(K_valid - 2 * trim);
      }
    }
    if (sub == 0 && col < width) {
      for (int o = 0; o < n_out; ++o) outs[o][base + col] = result;
    }
  }
}
extern "C" int fb_robust_select_runs(const float* const* rows_dev, int32_t K, int32_t K_valid, int32_t run_len, int64_t j0,
                                     int64_t j1, int32_t mode, int32_t trim, float* const* outs_dev, int32_t n_out,
                                     fb_stream_t stream) {
  FB_REQUIRE(rows_dev && outs_dev && K >= 1 && K_valid >= 1 && K_valid <= K && run_len >= 1 && j1 >= j0 && n_out >= 1,
             "fb_robust_select_runs: bad arguments");
  FB_REQUIRE(mode == 0 || (trim >= 0 && 2 * trim < K_valid), "fb_robust_select_runs: trim removes every client");
  const size_t smem = (size_t)K * SEL_LD * sizeof(uint32_t);
  if (smem > 220 * 1024) {
    fb_set_error("fb_robust_select_runs: K=%d exceeds the %d clients that fit one shared-memory tile", K,
                 (int)(220 * 1024 / (SEL_LD * sizeof(uint32_t))));
    return FB_ERR_LIMIT;
  }
  if (j1 == j0) return FB_OK;
  FB_CHECK_CUDA(cudaFuncSetAttribute(robust_select_runs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, robust_select_runs_kernel, SEL_THREADS, smem);
  if (per_sm < 1) per_sm = 1;
  const int64_t ntiles = (j1 - j0 + SEL_TILE - 1) / SEL_TILE;
  int64_t grid = (int64_t)per_sm * fb_num_sms();
  if (grid > ntiles) grid = ntiles;
  robust_select_runs_kernel<<<(int)grid, SEL_THREADS, smem, (cudaStream_t)stream>>>(rows_dev, K, K_valid, run_len, j0, j1, mode,
                                                                                     trim, outs_dev, n_out);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- K7: pack for the Gram GEMM
// dst_hi[k][j-j0] = bf16(x), dst_lo = bf16(x - float(hi)) with x = rows[k][j] - ref[j].
__global__ void __launch_bounds__(256) krum_pack_kernel(const float* const* __restrict__ rows,
                                                        const float* __restrict__ ref, int64_t j0, int64_t n,
                                                        __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo,
                                                        int64_t ld) {
  const int k = blockIdx.y;
  const float* row = rows[k];
  if ((((uintptr_t)(row + j0)) & 15) == 0 && (ref == nullptr || (((uintptr_t)(ref + j0)) & 15) == 0) && (ld & 3) == 0 &&
      (((uintptr_t)hi) & 7) == 0 && (lo == nullptr || (((uintptr_t)lo) & 7) == 0)) {
    // four coordinates per thread: 16-byte peer loads, 8-byte bf16 stores
    const int64_t nv = ld >> 2;
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = v << 2;
      float x[4] = {0.f, 0.f, 0.f, 0.f};
      if (i + 3 < n) {
        asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3])
                     : "l"(row + j0 + i)
                     : "memory");
        if (ref) {
          const float4 r = __ldg(reinterpret_cast<const float4*>(ref + j0 + i));
          x[0] -= r.x; x[1] -= r.y; x[2] -= r.z; x[3] -= r.w;
        }
      } else {
        for (int e = 0; e < 4; ++e)
          if (i + e < n) {
            asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(x[e]) : "l"(row + j0 + i + e) : "memory");
            if (ref) x[e] -= __ldg(ref + j0 + i + e);
          }
      }
      __nv_bfloat16 h[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) h[e] = __float2bfloat16_rn(x[e]);
      uint2 o;
      o.x = (uint32_t)__bfloat16_as_ushort(h[0]) | ((uint32_t)__bfloat16_as_ushort(h[1]) << 16);
      o.y = (uint32_t)__bfloat16_as_ushort(h[2]) | ((uint32_t)__bfloat16_as_ushort(h[3]) << 16);
      *reinterpret_cast<uint2*>(hi + (size_t)k * ld + i) = o;
      if (lo) {
        __nv_bfloat16 l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) l[e] = __float2bfloat16_rn(x[e] - __bfloat162float(h[e]));
        o.x = (uint32_t)__bfloat16_as_ushort(l[0]) | ((uint32_t)__bfloat16_as_ushort(l[1]) << 16);
        o.y = (uint32_t)__bfloat16_as_ushort(l[2]) | ((uint32_t)__bfloat16_as_ushort(l[3]) << 16);
        *reinterpret_cast<uint2*>(lo + (size_t)k * ld + i) = o;
      }
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < ld; i += (int64_t)gridDim.x * blockDim.x) {
    float x = 0.f;
    if (i < n) {
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(x) : "l"(row + j0 + i) : "memory");
      if (ref) x -= __ldg(ref + j0 + i);
    }
    const __nv_bfloat16 h = __float2bfloat16_rn(x);
    hi[(size_t)k * ld + i] = h;
    if (lo) lo[(size_t)k * ld + i] = __float2bfloat16_rn(x - __bfloat162float(h));
  }
}
extern "C" int fb_krum_pack(const float* const* rows_dev, int32_t K, const float* ref, int64_t j0, int64_t j1,
                            void* dst_hi, void* dst_lo, int64_t ld_dst, fb_stream_t stream) {
  FB_REQUIRE(rows_dev && dst_hi && K >= 1 && K <= 65535 && j1 > j0 && ld_dst >= j1 - j0, "fb_krum_pack: bad arguments");
  int gx = (int)((ld_dst + 255) / 256);
  int lim = (fb_num_sms() * 8 + K - 1) / K;
  if (gx > lim) gx = lim < 1 ? 1 : lim;
  krum_pack_kernel<<<dim3(gx, K), 256, 0, (cudaStream_t)stream>>>(rows_dev, ref, j0, j1 - j0, (__nv_bfloat16*)dst_hi,
                                                                  (__nv_bfloat16*)dst_lo, ld_dst);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- K7: scores from the Gram matrix
// one block per client i: d2(i,j) = g_ii + g_jj - 2 g_ij; sum of the nn smallest (j != i).
__global__ void __launch_bounds__(256) krum_scores_kernel(const float* __restrict__ gram, int K, int nn,
                                                          float* __restrict__ scores) {
  extern __shared__ float d2[];  // [K]
  __shared__ float red[33];
  const int i = blockIdx.x, tid = threadIdx.x;
  const float gii = gram[(size_t)i * K + i];
  for (int j = tid; j < K; j += blockDim.x) {
    float v = gii + gram[(size_t)j * K + j] - 2.f * gram[(size_t)i * K + j];
    d2[j] = (j == i) ? INFINITY : fmaxf(v, 0.f);
  }
  __syncthreads();
  // rank-by-counting: element j is among the nn smallest iff fewer than nn elements precede it
  // in (value, index) order — exact with ties, O(K^2 / threads) per block, K <= a few thousand.
  float acc = 0.f;
  for (int j = tid; j < K; j += blockDim.x) {
    const float v = d2[j];
    int before = 0;
    for (int t = 0; t < K; ++t) {
      const float u = d2[t];
      before += (u < v) || (u == v && t < j);
    }
    if (before < nn && j != i) acc += v;
  }
  const float tot = block_sum(acc, red);
  if (tid == 0) scores[i] = tot;
}
extern "C" int fb_krum_scores(const float* gram, int32_t K, int32_t f, float* scores, fb_stream_t stream) {
  FB_REQUIRE(gram && scores && K >= 1 && K <= 8192, "fb_krum_scores: bad arguments");
  int nn = K - f - 2;
  if (nn < 1) nn = 1;
  krum_scores_kernel<<<K, 256, (size_t)K * sizeof(float), (cudaStream_t)stream>>>(gram, K, nn, scores);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
