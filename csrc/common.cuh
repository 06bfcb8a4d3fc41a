// Shared device/host helpers for the fedml_b200 native library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define FB_OK 0
#define FB_ERR_ARG -1
#define FB_ERR_CUDA -2
#define FB_ERR_LIMIT -3
#define FB_ERR_TIMEOUT -4

void fb_set_error(const char* fmt, ...);

#define FB_CHECK_CUDA(expr)                                                         \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      fb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return FB_ERR_CUDA;                                                           \
    }                                                                               \
  } while (0)

#define FB_REQUIRE(cond, ...)                                                       \
  do {                                                                              \
    if (!(cond)) {                                                                  \
      fb_set_error(__VA_ARGS__);                                                    \
      return FB_ERR_ARG;                                                            \
    }                                                                               \
  } while (0)

static inline int fb_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (sms[dev] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    sms[dev] = n > 0 ? n : 148;
  }
  return sms[dev];
}

// ---------------------------------------------------------------- programmatic dependent launch (PDL)
// Kernels of the local training step form one long dependent chain of small launches.  Launched with the
// programmatic-stream-serialization attribute, kernel N+1 is scheduled while kernel N drains: its CTAs run their
// prologue (barrier init, TMEM allocation, tensor-map prefetch, index setup) and then block in pdl_wait() until every
// CTA of kernel N has finished and its memory is visible.  Every kernel launched this way calls pdl_wait() before its
// first global-memory access and pdl_trigger() right after, so the chain never reads stale data.  Stream capture turns
// the attribute into programmatic edges of the CUDA graph.  FB_PDL=0 turns the attribute off (plain launches).
#include <stdlib.h>
#include <utility>
static inline int fb_pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FB_PDL");
    v = (e && atoi(e) == 0) ? 0 : 1;
  }
  return v;
}
template <typename... KArgs, typename... Args>
static inline cudaError_t fb_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                        Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = fb_pdl_enabled();
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
// same, as thread-block clusters of `cluster_x` CTAs (CTA pairs for tcgen05 cta_group::2)
template <typename... KArgs, typename... Args>
static inline cudaError_t fb_launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                                unsigned cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = fb_pdl_enabled();
  at[1].id = cudaLaunchAttributeClusterDimension;
  at[1].val.clusterDim.x = cluster_x;
  at[1].val.clusterDim.y = 1;
  at[1].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum; every thread gets the result. `red` must hold >= 33 floats.
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  if (wid == 0) {
    float t = lane < nw ? red[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

// streaming 128-bit loads/stores that do not pollute L1
__device__ __forceinline__ float4 ld_stream4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
// plain (coherent) 128-bit load: used for memory another GPU may have just written
__device__ __forceinline__ float4 ld_coherent4(const float* p) {
  float4 r;
  asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void st4(float* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// system-scope release/acquire on 32-bit flags living in (possibly peer) global memory
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// ---------------------------------------------------------------- bounded spins on flags written by other ranks / blocks
// Every wait on a flag that another GPU (or another block) writes is bounded by the global nanosecond timer: when the
// writer never shows up (a crashed rank, a launch that was never made) the waiter records {code, who, target, block} in
// a device-side error record, gives up and lets the kernel run to completion on whatever data it has, so the GPU is
// released; every later wait of the same launch bails out at its next timer check.  The host reads (and clears) the
// record after its next synchronisation point (fb_*_spin_status) and turns it into FB_ERR_TIMEOUT.  One record per
// translation unit that spins (round.cu, cheetah.cu); FB_DEFINE_SPIN_API(prefix) emits its two host accessors.
#ifdef __CUDACC__
static __device__ unsigned g_fb_spin_err[4];                                   // code (0 = none), who, target, block
static __device__ unsigned long long g_fb_spin_timeout_ns = 20000000000ull;   // 20 s unless the host sets it
__device__ __forceinline__ unsigned long long fb_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool fb_spin_failed() { return *reinterpret_cast<volatile unsigned*>(&g_fb_spin_err[0]) != 0u; }
// wait until *flag >= target (wrap-around safe); SYS: the flag lives in memory another GPU writes.  false on time-out.
template <bool SYS>
__device__ __forceinline__ bool fb_spin_wait_ge(const uint32_t* flag, uint32_t target, uint32_t code, uint32_t who) {
  uint32_t polls = 0;
  unsigned long long t0 = 0;
  while ((int32_t)((SYS ? ld_acquire_sys(flag) : ld_acquire_gpu(flag)) - target) < 0) {
    if ((++polls & 0xFFu) == 0u) {
      const unsigned long long now = fb_globaltimer();
      if (t0 == 0) t0 = now;
      if (now - t0 > g_fb_spin_timeout_ns || fb_spin_failed()) {
        if (atomicCAS(&g_fb_spin_err[0], 0u, code) == 0u) {
          g_fb_spin_err[1] = who;
          g_fb_spin_err[2] = target;
          g_fb_spin_err[3] = blockIdx.x;
          __threadfence();
        }
        return false;
      }
    }
  }
  return true;
}
#define FB_DEFINE_SPIN_API(prefix)                                                                                          \
  extern "C" int prefix##_set_spin_timeout_ms(int64_t ms) {                                                                 \
    unsigned long long ns = (unsigned long long)(ms < 1 ? 1 : ms) * 1000000ull;                                            \
    FB_CHECK_CUDA(cudaMemcpyToSymbol(g_fb_spin_timeout_ns, &ns, sizeof(ns)));                                              \
    return FB_OK;                                                                                                           \
  }                                                                                                                         \
  /* out4 = {code, who, target, block}; synchronises with the device; clears the record when `clear`; returns             \
     FB_ERR_TIMEOUT when a wait gave up since the last clear */                                                            \
  extern "C" int prefix##_spin_status(uint32_t* out4, int32_t clear) {                                                      \
    unsigned h[4] = {0, 0, 0, 0};                                                                                           \
    FB_CHECK_CUDA(cudaMemcpyFromSymbol(h, g_fb_spin_err, sizeof(h)));                                                      \
    if (out4) for (int i = 0; i < 4; ++i) out4[i] = h[i];                                                                   \
    if (h[0] != 0 && clear) {                                                                                               \
      unsigned z[4] = {0, 0, 0, 0};                                                                                         \
      FB_CHECK_CUDA(cudaMemcpyToSymbol(g_fb_spin_err, z, sizeof(z)));                                                      \
    }                                                                                                                       \
    if (h[0] != 0) {                                                                                                        \
      fb_set_error(#prefix ": a wait on a flag written by %s %u timed out (wait kind %u, expected value %u, block %u): "   \
                   "a peer rank died or never launched its kernel", h[0] == 3 ? "block group" : "rank", h[1], h[0], h[2], h[3]); \
      return FB_ERR_TIMEOUT;                                                                                                \
    }                                                                                                                       \
    return FB_OK;                                                                                                           \
  }
#endif

// ---------------------------------------------------------------- Philox4x32-10 counter RNG
struct Philox4 {
  uint32_t x, y, z, w;
};
__host__ __device__ __forceinline__ uint32_t fb_mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}
__host__ __device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                            uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = fb_mulhi32(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = fb_mulhi32(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  Philox4 o = {c0, c1, c2, c3};
  return o;
}
// four N(0,1) draws from one Philox block (Box-Muller on two uniform pairs)
__device__ __forceinline__ float4 philox_normal4(uint64_t seed, uint32_t stream, uint32_t substream, uint64_t idx4) {
  Philox4 r = philox4x32_10((uint32_t)idx4, (uint32_t)(idx4 >> 32), stream, substream, (uint32_t)seed,
                            (uint32_t)(seed >> 32));
  const float inv = 2.3283064365386963e-10f;  // 2^-32
  float u0 = ((float)r.x + 0.5f) * inv, u1 = ((float)r.y + 0.5f) * inv;
  float u2 = ((float)r.z + 0.5f) * inv, u3 = ((float)r.w + 0.5f) * inv;
  u0 = fminf(fmaxf(u0, 1e-10f), 1.0f);
  u2 = fminf(fmaxf(u2, 1e-10f), 1.0f);
  float r0 = sqrtf(-2.0f * __logf(u0)), r1 = sqrtf(-2.0f * __logf(u2));
  float s0, c0, s1, c1;
  __sincosf(6.283185307179586f * u1, &s0, &c0);
  __sincosf(6.283185307179586f * u3, &s1, &c1);
  return make_float4(r0 * c0, r0 * s0, r1 * c1, r1 * s1);
}
