// K10 helpers: client-stacked BatchNorm (+residual, +ReLU) forward/backward and 3x3 im2col / col2im on
// NHWC bf16 activations.  See include/fedml_b200.h.  All four are bandwidth-bound streaming kernels:
// 16-byte (8 x bf16) vector accesses, fp32 arithmetic, no atomics, deterministic two-stage reductions.
//
// Layout: activations are [C clients][M = B*H*W rows][ch channels], channel fastest.  A CTA of 256 threads
// walks its row range 2048 elements at a time; because 2048 % ch == 0 a thread always sees the same 8
// channels, so per-channel partial sums live in registers for the whole walk.
#include "common.cuh"
#include "../include/fedml_b200.h"
#include <stdlib.h>
#include <mutex>

#define BN_THREADS 256
#define BN_VEC 8
#define BN_MAX_SLICES 128     // slices per client the partial-sum scratch holds (fb_bn_scratch_floats)

// Eight consecutive channels of one pixel: 16 bytes of bf16 or 32 bytes of fp32 (two 16-byte accesses).  Kernels are
// templated on the activation element type T; all arithmetic is fp32 either way.
template <typename T> struct Raw8;
template <> struct Raw8<__nv_bfloat16> { uint4 v; };
template <> struct Raw8<float> { float4 a, b; };
__device__ __forceinline__ Raw8<__nv_bfloat16> ldraw(const __nv_bfloat16* p) {
  Raw8<__nv_bfloat16> r;
  r.v = __ldg(reinterpret_cast<const uint4*>(p));
  return r;
}
__device__ __forceinline__ Raw8<float> ldraw(const float* p) {
  Raw8<float> r;
  r.a = __ldg(reinterpret_cast<const float4*>(p));
  r.b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  return r;
}
template <typename T> __device__ __forceinline__ Raw8<T> zero_raw();
template <> __device__ __forceinline__ Raw8<__nv_bfloat16> zero_raw<__nv_bfloat16>() {
  Raw8<__nv_bfloat16> r;
  r.v = make_uint4(0, 0, 0, 0);
  return r;
}
template <> __device__ __forceinline__ Raw8<float> zero_raw<float>() {
  Raw8<float> r;
  r.a = make_float4(0.f, 0.f, 0.f, 0.f);
  r.b = r.a;
  return r;
}
__device__ __forceinline__ void unpack8(const Raw8<__nv_bfloat16>& r, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r.v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ void unpack8(const Raw8<float>& r, float* f) {
  f[0] = r.a.x; f[1] = r.a.y; f[2] = r.a.z; f[3] = r.a.w;
  f[4] = r.b.x; f[5] = r.b.y; f[6] = r.b.z; f[7] = r.b.w;
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float* f) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = r;
}
__device__ __forceinline__ void store8(float* p, const float* f) {
  reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}
__device__ __forceinline__ void store_raw(__nv_bfloat16* p, const Raw8<__nv_bfloat16>& r) { *reinterpret_cast<uint4*>(p) = r.v; }
__device__ __forceinline__ void store_raw(float* p, const Raw8<float>& r) {
  reinterpret_cast<float4*>(p)[0] = r.a;
  reinterpret_cast<float4*>(p)[1] = r.b;
}

// reduce the 16 per-thread accumulators (8 channels x 2 quantities) over the threads that share a channel
// group and write [2][ch] to `out`.  smem: BN_THREADS*16 floats.
__device__ __forceinline__ void reduce_channel_pairs(const float* a, const float* b, int ch, float* smem, float* out) {
  const int tid = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    smem[tid * 16 + i] = a[i];
    smem[tid * 16 + 8 + i] = b[i];
  }
  __syncthreads();
  const int ng = ch / BN_VEC;  // channel groups; thread t belongs to group t % ng
  for (int t = tid; t < 2 * ch; t += BN_THREADS) {
    const int q = t / ch, c = t % ch;
    const int grp = c / BN_VEC, comp = c % BN_VEC;
    float s = 0.f;
    for (int tt = grp; tt < BN_THREADS; tt += ng) s += smem[tt * 16 + q * 8 + comp];
    out[q * ch + c] = s;
  }
  __syncthreads();
}

// ---------------------------------------------------------------- BN forward
// stage 1: partial[c][s][2][ch] = (sum x, sum x^2) over the rows of slice s
template <typename T>
__device__ __forceinline__ void
bn_stats_body(const T* __restrict__ x, int64_t M, int ch, int S, float* __restrict__ partial, float* smem) {
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const T* base = x + (size_t)c * M * ch;
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += 4 * BN_THREADS) {   // 4 independent 16-byte loads in flight per thread
    Raw8<T> r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      r[u] = (v + u * BN_THREADS < v1) ? ldraw(base + (v + u * BN_THREADS) * BN_VEC) : zero_raw<T>();
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float f[8];
      unpack8(r[u], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a[i] += f[i];
        b[i] += f[i] * f[i];
      }
    }
  }
  reduce_channel_pairs(a, b, ch, smem, partial + ((size_t)c * S + s) * 2 * ch);
}

// BatchNorm: finish the statistics of client blockIdx.y from its SP partial sums (every CTA redundantly, SP is small),
// update the running statistics (slice 0 only) and leave scale / shift per channel in shared memory
__device__ __forceinline__ void
bn_scale_shift(int64_t M, int ch, const float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
               int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r, const float* __restrict__ active,
               float momentum, float eps, int training, float* __restrict__ save_mean, float* __restrict__ save_rstd,
               float* num_batches_tracked, int64_t ld_n, float* scale, float* shift, int SP,
               const float* presum = nullptr /* shared memory [2][ch]: the client's complete sums, already added up */,
               int c = -1, int s = -1 /* client and slice of this CTA's work item (default: blockIdx.y / blockIdx.x) */) {
  const int tid = threadIdx.x;
  if (c < 0) {
    c = blockIdx.y;
    s = blockIdx.x;
  }
  if (training && s == 0 && tid == 0 && num_batches_tracked)       // the integer BN counter lives in the arena as fp32
    num_batches_tracked[(size_t)c * ld_n] += active ? active[c] : 1.f;
  for (int k = tid; k < ch; k += BN_THREADS) {
    float mean, var;
    if (training) {
      float su = 0.f, sq = 0.f;
      if (presum) {
        su = presum[k];
        sq = presum[ch + k];
      } else {
        for (int j = 0; j < SP; ++j) {
          su += __ldcg(partial + ((size_t)c * SP + j) * 2 * ch + k);       // L2: written by sibling CTAs in the fused kernel
          sq += __ldcg(partial + ((size_t)c * SP + j) * 2 * ch + ch + k);
        }
      }
      mean = su / (float)M;
      var = fmaxf(sq / (float)M - mean * mean, 0.f);
      if (s == 0) {
        const float a = active ? active[c] : 1.f;
        if (a != 0.f) {
          const float unbiased = var * ((float)M / (float)max((int64_t)1, M - 1));
          float* rm = running_mean + (size_t)c * ld_r + k;
          float* rv = running_var + (size_t)c * ld_r + k;
          *rm += momentum * a * (mean - *rm);
          *rv += momentum * a * (unbiased - *rv);
        }
      }
    } else {
      mean = running_mean[(size_t)c * ld_r + k];
      var = running_var[(size_t)c * ld_r + k];
    }
    const float rstd = rsqrtf(var + eps);
    if (s == 0 && save_mean) {
      save_mean[(size_t)c * ch + k] = mean;
      save_rstd[(size_t)c * ch + k] = rstd;
    }
    const float g = gamma[(size_t)c * ld_p + k] * rstd;
    scale[k] = g;
    shift[k] = beta[(size_t)c * ld_p + k] - mean * g;
  }
}

// stage 2: finish the statistics (every CTA redundantly, S is small), update running stats, apply
template <typename T>
__device__ __forceinline__ void
bn_apply_body(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                int64_t M, int ch, int S, const float* partial, const float* __restrict__ gamma,
                const float* __restrict__ beta, int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r,
                const float* __restrict__ active, float momentum, float eps, int relu, int training,
                float* __restrict__ save_mean, float* __restrict__ save_rstd, float* num_batches_tracked, int64_t ld_n,
                float* sm /* scale[ch], shift[ch] (+ 2*ch scratch when cpg > 1) */,
                int SP /* slices of `partial` per client (1: complete sums) */,
                int cb = 1 /* rows (blockIdx.y) per parameter set: GroupNorm runs one row per (client, sample) */,
                int cpg = 0 /* > 1: GroupNorm, statistics over the row's pixels x cpg adjacent channels */) {
  float* scale = sm;
  float* shift = sm + ch;
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int cp = c / cb;                 // row of gamma / beta / running statistics
  if (cpg > 1) {
    // GroupNorm (reference model/cv/group_normalization.py:7-93): per-channel sums of this (client, sample) row, then the
    // cpg channels of a group are pooled; no running statistics
    float* t_su = sm + 2 * ch;
    float* t_sq = sm + 3 * ch;
    for (int k = tid; k < ch; k += BN_THREADS) {
      float su = 0.f, sq = 0.f;
      for (int j = 0; j < SP; ++j) {
        su += __ldcg(partial + ((size_t)c * SP + j) * 2 * ch + k);
        sq += __ldcg(partial + ((size_t)c * SP + j) * 2 * ch + ch + k);
      }
      t_su[k] = su;
      t_sq[k] = sq;
    }
    __syncthreads();
    for (int k = tid; k < ch; k += BN_THREADS) {
      const int g0 = k - k % cpg;
      float su = 0.f, sq = 0.f;
      for (int i = 0; i < cpg; ++i) { su += t_su[g0 + i]; sq += t_sq[g0 + i]; }
      const float n = (float)M * (float)cpg;
      const float mean = su / n, var = fmaxf(sq / n - mean * mean, 0.f);
      const float rstd = rsqrtf(var + eps);
      if (s == 0 && save_mean) {
        save_mean[(size_t)c * ch + k] = mean;
        save_rstd[(size_t)c * ch + k] = rstd;
      }
      const float g = gamma[(size_t)cp * ld_p + k] * rstd;
      scale[k] = g;
      shift[k] = beta[(size_t)cp * ld_p + k] - mean * g;
    }
  } else {
    bn_scale_shift(M, ch, partial, gamma, beta, ld_p, running_mean, running_var, ld_r, active, momentum, eps, training,
                   save_mean, save_rstd, num_batches_tracked, ld_n, scale, shift, SP);
  }
  __syncthreads();
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const size_t cbase = (size_t)c * M * ch;
  const int coff = (tid * BN_VEC) % ch;
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = scale[coff + i];
    sh[i] = shift[coff + i];
  }
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += 2 * BN_THREADS) {    // two vectors (x and res each) in flight per thread
    const int64_t vb = v + BN_THREADS;
    const bool two = vb < v1;
    const Raw8<T> xa = ldraw(x + cbase + v * BN_VEC);
    const Raw8<T> xb = two ? ldraw(x + cbase + vb * BN_VEC) : zero_raw<T>();
    Raw8<T> ra = zero_raw<T>(), rb = ra;
    if (res) {
      ra = ldraw(res + cbase + v * BN_VEC);
      if (two) rb = ldraw(res + cbase + vb * BN_VEC);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
      float f[8], r[8];
      unpack8(u ? xb : xa, f);
      unpack8(u ? rb : ra, r);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        f[i] = f[i] * sc[i] + sh[i] + r[i];
        if (relu) f[i] = fmaxf(f[i], 0.f);
      }
      store8(y + cbase + (u ? vb : v) * BN_VEC, f);
    }
  }
}

static int bn_forced_slices(int bwd) {
  static int f[2] = {-1, -1};
  if (f[bwd] < 0) {
    const char* e = getenv(bwd ? "FB_BN_S_BWD" : "FB_BN_S_FWD");
    f[bwd] = e ? atoi(e) : 0;
  }
  return f[bwd];
}
static int bn_slices(int C, int64_t M, int ch) {
  // slices per client; at least four sweeps of 2048 elements per CTA so the per-CTA prologue stays amortised.
  // total CTAs ~ 3.4 per SM: measured on the 100-client bench (ms / round by slices per client, forward x backward):
  // 5x5 69.0, 5x8 69.7, 8x5 69.8, 8x8 70.6, 9x9 73.7, 11x8 72.2, 18x18 77.8 - more slices pay the S x ch partial-sum
  // prologue in every CTA, and 6 or 9 slices x 100 clients leave a nearly empty last wave
  static int milli_per_sm = 0;
  if (!milli_per_sm) {
    const char* e = getenv("FB_BN_CTAS_PER_SM");
    milli_per_sm = e ? (int)(atof(e) * 1000) : 3400;
    if (milli_per_sm < 100) milli_per_sm = 100;
  }
  int64_t want = ((int64_t)milli_per_sm * fb_num_sms() / 1000 + C / 2) / C;
  int64_t max_s = (M * ch + 8191) / 8192;
  if (want > max_s) want = max_s;
  if (want < 1) want = 1;
  if (want > 64) want = 64;
  return (int)want;
}

// ---------------------------------------------------------------- kernels: two-pass and fused single-launch forms
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const T* __restrict__ x, int64_t M, int ch, int S, float* __restrict__ partial) {
  __shared__ float smem[BN_THREADS * 16];
  pdl_wait();
  pdl_trigger();
  bn_stats_body(x, M, ch, S, partial, smem);
}
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                int64_t M, int ch, int S, const float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
                int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r, const float* __restrict__ active,
                float momentum, float eps, int relu, int training, float* __restrict__ save_mean,
                float* __restrict__ save_rstd, float* num_batches_tracked, int64_t ld_n, int SP, int cb, int cpg) {
  extern __shared__ float sm[];
  pdl_wait();
  pdl_trigger();
  bn_apply_body(x, res, y, M, ch, S, partial, gamma, beta, ld_p, running_mean, running_var, ld_r, active, momentum, eps, relu,
                training, save_mean, save_rstd, num_batches_tracked, ld_n, sm, SP, cb, cpg);
}

// Barrier among the S CTAs of one client (blockIdx.y): {arrival count, generation} per client, self-resetting, so
// the state survives from launch to launch (and CUDA-graph replays) without host involvement.  Sound because CTAs
// are dispatched in linear block order (a client's CTAs are adjacent, earlier clients first) and S is far below the
// number of resident CTA slots.  One BN kernel at a time per device (they share this state and the scratch).
static __device__ unsigned g_bn_barrier[2 * 65536];
// BatchNorm calls on different streams (client groups of the simulator) may overlap: each stream gets its own eighth of
// the barrier array (8192 clients).  More than eight streams, or more than 8192 clients in a call, share region 0 - those
// calls must then not overlap.
static int bn_barrier_region(cudaStream_t stream, int C) {
  static std::mutex mu;
  static cudaStream_t seen[8];
  static int n_seen = 0;
  if (C > 8192) return 0;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < n_seen; ++i)
    if (seen[i] == stream) return i * 8192;
  if (n_seen < 8) {
    seen[n_seen] = stream;
    return (n_seen++) * 8192;
  }
  return 0;
}
__device__ __forceinline__ void client_barrier(int S, int bar_base) {
  __syncthreads();
  if (S > 1 && threadIdx.x == 0) {
    unsigned* st = g_bn_barrier + 2 * (bar_base + (int)blockIdx.y);
    __threadfence();                                   // publish this CTA's partial sums
    const unsigned gen = ld_acquire_gpu(st + 1);
    if (atomicAdd(st, 1u) == (unsigned)S - 1) {
      st[0] = 0;
      __threadfence();
      atomicAdd(st + 1, 1u);
    } else {
      while (ld_acquire_gpu(st + 1) == gen) __nanosleep(32);
    }
    __threadfence();
  }
  __syncthreads();
}

// stats -> per-client barrier -> apply in ONE launch: the second read of x comes from L2 (a client's slice was just
// streamed by its sibling CTAs) and a BN layer costs one kernel instead of two
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_fwd_fused_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                    int64_t M, int ch, int S, float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
                    int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r, const float* __restrict__ active,
                    float momentum, float eps, int relu, float* __restrict__ save_mean, float* __restrict__ save_rstd,
                    float* num_batches_tracked, int64_t ld_n, int bar_base) {
  __shared__ float smem[BN_THREADS * 16];
  extern __shared__ float sm[];
  pdl_wait();
  pdl_trigger();
  bn_stats_body(x, M, ch, S, partial, smem);
  client_barrier(S, bar_base);
  bn_apply_body(x, res, y, M, ch, S, partial, gamma, beta, ld_p, running_mean, running_var, ld_r, active, momentum, eps, relu, 1,
                save_mean, save_rstd, num_batches_tracked, ld_n, sm, S);
}

// One launch per BatchNorm (statistics -> per-client barrier -> apply) for every layer size.  FB_BN_FUSED=0 forces the
// two-pass kernels.  Round 1 chose by activation size (>= 48 MB) because the in-kernel barrier lost to a kernel boundary
// on small layers; with the resident kernels and programmatic launch edges the single launch wins everywhere: ResNet-56
// tf32 round at 13 / 25 / 50 / 100 clients 24.6 -> 21.5, 36.1 -> 32.8, 60.1 -> 56.8, 105.7 -> 104.3 ms; bf16 18.5 -> 17.6,
// 25.6 -> 24.6, 39.5 -> 38.5, 66.0 -> 66.0 ms.
static bool bn_fused_enabled(int64_t act_bytes) {
  static int v = -2;
  if (v == -2) {
    const char* e = getenv("FB_BN_FUSED");
    v = e ? (atoi(e) == 0 ? 0 : 1) : -1;
  }
  (void)act_bytes;
  return v >= 0 ? v == 1 : true;
}

__device__ __forceinline__ void
bn_bwd_coeffs(int64_t M, int ch, int S, const float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
              int64_t ld_p, const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dgamma,
              float* __restrict__ dbeta, int64_t ld_g, float* sm, const float* presum = nullptr, int c = -1, int s = -1);
#include "bn_resident.cuh"
#include "bn_persist.cuh"

extern "C" int64_t fb_bn_scratch_floats(int32_t C, int64_t M, int32_t ch) { return (int64_t)C * BN_MAX_SLICES * 2 * ch; }

template <typename T>
static int bn_fwd_impl(const void* x, const void* res, void* y, int32_t C, int64_t M, int32_t ch,
                                 const float* gamma, const float* beta, int64_t ld_p, float* running_mean,
                                 float* running_var, int64_t ld_r, const float* active, float momentum, float eps,
                                 int32_t relu, int32_t training, float* save_mean, float* save_rstd, float* scratch,
                                 float* num_batches_tracked, int64_t ld_n, fb_stream_t stream) {
  FB_REQUIRE(x && y && gamma && beta && C >= 1 && C <= 65535 && M >= 1, "fb_bn_fwd_stacked: bad arguments");
  FB_REQUIRE(ch >= 8 && ch <= 2048 && (2048 % ch) == 0, "fb_bn_fwd_stacked: ch=%d must be a power of two in [8, 2048]", ch);
  FB_REQUIRE(running_mean && running_var, "fb_bn_fwd_stacked: running statistics missing");
  FB_REQUIRE(!training || scratch, "fb_bn_fwd_stacked: training needs the partial-sum scratch");
  int S = bn_slices(C, M, ch);
  if (bn_forced_slices(0) > 0) S = bn_forced_slices(0);
  // training == 2: `scratch` already holds the complete per-client sums [C][2][ch] (sum x, sum x^2), produced by the
  // epilogue of the convolution that wrote x (FbGroupedGemm.stats): one pass over x instead of two
  const bool have_stats = training == 2;
  if (training == 1 && bn_fused_enabled((int64_t)C * M * ch * (int64_t)sizeof(T))) {
    const int bar_base = bn_barrier_region((cudaStream_t)stream, C);
    int Sr = 0;
    uint32_t cap = 0;
    size_t smem = 0;
    {
      static bool chunks_set = false;
      if (!chunks_set) {
        chunks_set = true;
        const char* e = getenv("FB_BN_CHUNKS");
        if (e) {
          unsigned n = (unsigned)atoi(e);
          if (n < 1) n = 1;
          if (n > BNR_MAX_CHUNKS) n = BNR_MAX_CHUNKS;
          FB_CHECK_CUDA(cudaMemcpyToSymbol(g_bn_chunks, &n, sizeof(n)));
        }
      }
    }
    {
      int Sp = 0, grid = 0;
      uint32_t capp = 0;
      size_t smemp = 0;
      if (bn_persist_plan<T>(C, M, ch, res ? 2 : 1, &Sp, &capp, &smemp, &grid)) {
        if (int rc = bn_resident_allow_smem(bn_fwd_persist_kernel<T>, smemp)) return rc;
        FB_CHECK_CUDA(fb_launch_pdl(bn_fwd_persist_kernel<T>, dim3(grid), dim3(BN_THREADS), smemp, (cudaStream_t)stream,
            (const T*)x, (const T*)res, (T*)y, M, ch, Sp, (int)C, capp, scratch, gamma, beta, ld_p, running_mean,
            running_var, ld_r, active, momentum, eps, relu, save_mean, save_rstd, num_batches_tracked, ld_n, bar_base));
        FB_CHECK_CUDA(cudaGetLastError());
        return FB_OK;
      }
    }
    // FB_BN_RES_EXTRAS=1: the residual rides along in shared memory too.  Off by default: twice the slices per client cost more
    // (per-CTA fixed time) than the bulk load saves - fp32 64ch@32x32 forward 197.8 -> 254.5 us, round 105.7 -> 118.3 ms
    int res_smem = (res && bn_resident_extras()) ? 1 : 0;
    bool ok = bn_resident_plan<T>(M, ch, 1 + res_smem, 0, res != nullptr, &Sr, &cap, &smem);
    if (!ok && res_smem) {
      res_smem = 0;
      ok = bn_resident_plan<T>(M, ch, 1, 0, true, &Sr, &cap, &smem);
    }
    if (ok) {       // the slab stays in shared memory: x crosses HBM once
      if (int rc = bn_resident_allow_smem(bn_fwd_resident_kernel<T>, smem)) return rc;
      FB_CHECK_CUDA(fb_launch_pdl(bn_fwd_resident_kernel<T>, dim3(dim3(Sr, C)), dim3(BN_THREADS), smem, (cudaStream_t)stream,
          (const T*)x, (const T*)res, (T*)y, M, ch, Sr, cap, scratch, gamma, beta, ld_p, running_mean,
          running_var, ld_r, active, momentum, eps, relu, save_mean, save_rstd, num_batches_tracked, ld_n, res_smem, bar_base));
      FB_CHECK_CUDA(cudaGetLastError());
      return FB_OK;
    }
    FB_CHECK_CUDA(fb_launch_pdl(bn_fwd_fused_kernel<T>, dim3(dim3(S, C)), dim3(BN_THREADS), (size_t)(2 * ch * sizeof(float)), (cudaStream_t)stream, 
        (const T*)x, (const T*)res, (T*)y, M, ch, S, scratch, gamma, beta, ld_p, running_mean,
        running_var, ld_r, active, momentum, eps, relu, save_mean, save_rstd, num_batches_tracked, ld_n, bar_base));
    FB_CHECK_CUDA(cudaGetLastError());
    return FB_OK;
  }
  if (training == 1) {
    FB_CHECK_CUDA(fb_launch_pdl(bn_stats_kernel<T>, dim3(dim3(S, C)), dim3(BN_THREADS), (size_t)(0), (cudaStream_t)stream, (const T*)x, M, ch, S, scratch));
    FB_CHECK_CUDA(cudaGetLastError());
  }
  FB_CHECK_CUDA(fb_launch_pdl(bn_apply_kernel<T>, dim3(dim3(S, C)), dim3(BN_THREADS), (size_t)(2 * ch * sizeof(float)), (cudaStream_t)stream, 
      (const T*)x, (const T*)res, (T*)y, M, ch, S, scratch, gamma, beta, ld_p,
      running_mean, running_var, ld_r, active, momentum, eps, relu, training ? 1 : 0, save_mean, save_rstd, num_batches_tracked, ld_n,
      have_stats ? 1 : S, 1, 0));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- BN backward
// stage 1: partial[c][s][2][ch] = (sum dy', sum dy' * xhat), dy' = dy masked by the ReLU of the forward output
template <typename T>
__device__ __forceinline__ void
bn_bwd_stats_body(const T* __restrict__ dy, const T* __restrict__ dy2,
                    const T* __restrict__ x,
                    const T* __restrict__ y, int64_t M, int ch, int S, const float* __restrict__ mean,
                    const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                    int64_t ld_p, int relu, float* __restrict__ partial, float* smem, int cb = 1) {
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int cp = c / cb;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const size_t cbase = (size_t)c * M * ch;
  const int coff = (tid * BN_VEC) % ch;
  // mask recompute (no y): relu input = x*g + k with g = gamma*rstd, k = beta - mean*g
  float g[8], k[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float mu = mean[(size_t)c * ch + coff + i], rs = rstd[(size_t)c * ch + coff + i];
    g[i] = gamma[(size_t)cp * ld_p + coff + i] * rs;
    k[i] = (beta ? beta[(size_t)cp * ld_p + coff + i] : 0.f) - mu * g[i];
  }
  // partial sums of dy' and dy'*x (raw x): the apply kernel turns them into dbeta and dgamma = rstd*(S_dx - mean*S_d)
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += 2 * BN_THREADS) {
    const int64_t vb = v + BN_THREADS;
    const bool two = vb < v1;
    const Raw8<T> da = ldraw(dy + cbase + v * BN_VEC), xa = ldraw(x + cbase + v * BN_VEC);
    const Raw8<T> db_ = two ? ldraw(dy + cbase + vb * BN_VEC) : zero_raw<T>();
    const Raw8<T> xb = two ? ldraw(x + cbase + vb * BN_VEC) : zero_raw<T>();
    Raw8<T> ya = zero_raw<T>(), yb = ya;
    if (relu && y) {
      ya = ldraw(y + cbase + v * BN_VEC);
      if (two) yb = ldraw(y + cbase + vb * BN_VEC);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
      float d[8], xv[8];
      unpack8(u ? db_ : da, d);
      unpack8(u ? xb : xa, xv);
      if (dy2) {   // gradient arriving over a second path (residual branch): summed on load instead of by a separate kernel
        float e[8];
        unpack8(ldraw(dy2 + cbase + (u ? vb : v) * BN_VEC), e);
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] += e[i];
      }
      if (relu) {
        if (y) {
          float yv[8];
          unpack8(u ? yb : ya, yv);
#pragma unroll
          for (int i = 0; i < 8; ++i) d[i] = yv[i] > 0.f ? d[i] : 0.f;
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) d[i] = (xv[i] * g[i] + k[i]) > 0.f ? d[i] : 0.f;
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a[i] += d[i];
        b[i] += d[i] * xv[i];
      }
    }
  }
  reduce_channel_pairs(a, b, ch, smem, partial + ((size_t)c * S + s) * 2 * ch);
}

// BatchNorm backward: finish (sum dy', sum dy'*x) of client blockIdx.y from the S partial sums, write dgamma / dbeta (slice 0)
// and leave, per channel, A = g, B = -g*rstd*dgamma/M, Cc = -g*dbeta/M - mean*B, K = beta - mean*g in shared memory
__device__ __forceinline__ void
bn_bwd_coeffs(int64_t M, int ch, int S, const float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
              int64_t ld_p, const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dgamma,
              float* __restrict__ dbeta, int64_t ld_g, float* sm, const float* presum, int c, int s) {
  float* s_A = sm;
  float* s_B = sm + ch;
  float* s_C = sm + 2 * ch;
  float* s_K = sm + 3 * ch;
  const int tid = threadIdx.x;
  if (c < 0) {
    c = blockIdx.y;
    s = blockIdx.x;
  }
  for (int k = tid; k < ch; k += BN_THREADS) {
    float sd = 0.f, sdx = 0.f;
    if (presum) {
      sd = presum[k];
      sdx = presum[ch + k];
    } else {
      for (int j = 0; j < S; ++j) {
        sd += __ldcg(partial + ((size_t)c * S + j) * 2 * ch + k);
        sdx += __ldcg(partial + ((size_t)c * S + j) * 2 * ch + ch + k);
      }
    }
    const float mu = mean[(size_t)c * ch + k], rs = rstd[(size_t)c * ch + k];
    const float dbt = sd, dgm = rs * (sdx - mu * sd);
    if (s == 0) {
      dbeta[(size_t)c * ld_g + k] = dbt;
      dgamma[(size_t)c * ld_g + k] = dgm;
    }
    const float g = gamma[(size_t)c * ld_p + k] * rs;
    const float B = -g * rs * dgm / (float)M;
    s_A[k] = g;
    s_B[k] = B;
    s_C[k] = -g * dbt / (float)M - mu * B;
    s_K[k] = (beta ? beta[(size_t)c * ld_p + k] : 0.f) - mu * g;
  }
}

template <typename T>
__device__ __forceinline__ void
bn_bwd_apply_body(const T* __restrict__ dy, const T* __restrict__ dy2,
                    const T* __restrict__ x,
                    const T* __restrict__ y, T* __restrict__ dx, T* __restrict__ dres,
                    int64_t M, int ch, int S, const float* partial, const float* __restrict__ gamma,
                    const float* __restrict__ beta, int64_t ld_p, const float* __restrict__ mean,
                    const float* __restrict__ rstd, int relu, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    int64_t ld_g, float* sm, int cb = 1, int cpg = 0) {
  // sm, per channel: A = g, B = -g*rstd*dgamma/M, Cc = -g*dbeta/M - mean*B, K = beta - mean*g
  float* s_A = sm;
  float* s_B = sm + ch;
  float* s_C = sm + 2 * ch;
  float* s_K = sm + 3 * ch;
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int cp = c / cb;
  if (cpg > 1) {
    // GroupNorm: dx = rstd * (gamma_k d - mean_g(gamma d) - xhat * mean_g(gamma d xhat)), means over the group's
    // M x cpg elements; per-row dgamma / dbeta go to [rows][ch] and are summed over the samples of a client by the caller
    float* t_a = sm + 4 * ch;      // gamma_k * sum d
    float* t_b = sm + 5 * ch;      // gamma_k * sum d * xhat
    for (int k = tid; k < ch; k += BN_THREADS) {
      float sd = 0.f, sdx = 0.f;
      for (int j = 0; j < S; ++j) {
        sd += __ldcg(partial + ((size_t)c * S + j) * 2 * ch + k);
        sdx += __ldcg(partial + ((size_t)c * S + j) * 2 * ch + ch + k);
      }
      const float mu = mean[(size_t)c * ch + k], rs = rstd[(size_t)c * ch + k];
      const float dxh = rs * (sdx - mu * sd);                       // sum d * xhat
      if (s == 0) {
        dbeta[(size_t)c * ld_g + k] = sd;
        dgamma[(size_t)c * ld_g + k] = dxh;
      }
      const float gm = gamma[(size_t)cp * ld_p + k];
      t_a[k] = gm * sd;
      t_b[k] = gm * dxh;
    }
    __syncthreads();
    for (int k = tid; k < ch; k += BN_THREADS) {
      const int g0 = k - k % cpg;
      float Ag = 0.f, Bg = 0.f;
      for (int i = 0; i < cpg; ++i) { Ag += t_a[g0 + i]; Bg += t_b[g0 + i]; }
      const float n = (float)M * (float)cpg;
      const float mu = mean[(size_t)c * ch + k], rs = rstd[(size_t)c * ch + k];
      const float g = gamma[(size_t)cp * ld_p + k] * rs;
      const float B = -rs * rs * Bg / n;
      s_A[k] = g;
      s_B[k] = B;
      s_C[k] = -rs * Ag / n - mu * B;
      s_K[k] = (beta ? beta[(size_t)cp * ld_p + k] : 0.f) - mu * g;
    }
  } else {
    bn_bwd_coeffs(M, ch, S, partial, gamma, beta, ld_p, mean, rstd, dgamma, dbeta, ld_g, sm);
  }
  __syncthreads();
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const size_t cbase = (size_t)c * M * ch;
  const int coff = (tid * BN_VEC) % ch;
  float A[8], B[8], Cc[8], K[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    A[i] = s_A[coff + i];
    B[i] = s_B[coff + i];
    Cc[i] = s_C[coff + i];
    K[i] = s_K[coff + i];
  }
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += 2 * BN_THREADS) {
    const int64_t vb = v + BN_THREADS;
    const bool two = vb < v1;
    const Raw8<T> da = ldraw(dy + cbase + v * BN_VEC), xa = ldraw(x + cbase + v * BN_VEC);
    const Raw8<T> db_ = two ? ldraw(dy + cbase + vb * BN_VEC) : zero_raw<T>();
    const Raw8<T> xb = two ? ldraw(x + cbase + vb * BN_VEC) : zero_raw<T>();
    Raw8<T> ya = zero_raw<T>(), yb = ya;
    if (relu && y) {
      ya = ldraw(y + cbase + v * BN_VEC);
      if (two) yb = ldraw(y + cbase + vb * BN_VEC);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
      const int64_t vv = u ? vb : v;
      float d[8], xv[8];
      unpack8(u ? db_ : da, d);
      unpack8(u ? xb : xa, xv);
      if (dy2) {   // gradient arriving over a second path (residual branch): summed on load instead of by a separate kernel
        float e[8];
        unpack8(ldraw(dy2 + cbase + (u ? vb : v) * BN_VEC), e);
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] += e[i];
      }
      if (relu) {
        if (y) {
          float yv[8];
          unpack8(u ? yb : ya, yv);
#pragma unroll
          for (int i = 0; i < 8; ++i) d[i] = yv[i] > 0.f ? d[i] : 0.f;
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) d[i] = (xv[i] * A[i] + K[i]) > 0.f ? d[i] : 0.f;
        }
      }
      if (dres) store8(dres + cbase + vv * BN_VEC, d);
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = d[i] * A[i] + xv[i] * B[i] + Cc[i];   // = g*(d - dbeta/M - xhat*dgamma/M)
      store8(dx + cbase + vv * BN_VEC, o);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_stats_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x,
                    const T* __restrict__ y, int64_t M, int ch, int S, const float* __restrict__ mean,
                    const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta, int64_t ld_p,
                    int relu, float* __restrict__ partial, int cb) {
  __shared__ float smem[BN_THREADS * 16];
  pdl_wait();
  pdl_trigger();
  bn_bwd_stats_body(dy, dy2, x, y, M, ch, S, mean, rstd, gamma, beta, ld_p, relu, partial, smem, cb);
}
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x,
                    const T* __restrict__ y, T* __restrict__ dx, T* __restrict__ dres, int64_t M,
                    int ch, int S, const float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
                    int64_t ld_p, const float* __restrict__ mean, const float* __restrict__ rstd, int relu,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, int64_t ld_g, int cb, int cpg) {
  extern __shared__ float sm[];
  pdl_wait();
  pdl_trigger();
  bn_bwd_apply_body(dy, dy2, x, y, dx, dres, M, ch, S, partial, gamma, beta, ld_p, mean, rstd, relu, dgamma, dbeta, ld_g, sm, cb, cpg);
}
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_fused_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x,
                    const T* __restrict__ y, T* __restrict__ dx, T* __restrict__ dres, int64_t M,
                    int ch, int S, float* partial, const float* __restrict__ gamma, const float* __restrict__ beta, int64_t ld_p,
                    const float* __restrict__ mean, const float* __restrict__ rstd, int relu, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, int64_t ld_g, int bar_base) {
  __shared__ float smem[BN_THREADS * 16];
  extern __shared__ float sm[];
  pdl_wait();
  pdl_trigger();
  bn_bwd_stats_body(dy, dy2, x, y, M, ch, S, mean, rstd, gamma, beta, ld_p, relu, partial, smem);
  client_barrier(S, bar_base);
  bn_bwd_apply_body(dy, dy2, x, y, dx, dres, M, ch, S, partial, gamma, beta, ld_p, mean, rstd, relu, dgamma, dbeta, ld_g, sm);
}

template <typename T>
static int bn_bwd_impl(const void* dy, const void* dy2, const void* x, const void* y, void* dx, void* dres, int32_t C,
                                 int64_t M, int32_t ch, const float* gamma, const float* beta, int64_t ld_p,
                                 const float* save_mean,
                                 const float* save_rstd, int32_t relu, float* dgamma, float* dbeta, int64_t ld_g,
                                 float* scratch, fb_stream_t stream) {
  FB_REQUIRE(dy && x && dx && gamma && save_mean && save_rstd && dgamma && dbeta && scratch && C >= 1 && C <= 65535,
             "fb_bn_bwd_stacked: bad arguments");
  FB_REQUIRE(!relu || y || beta, "fb_bn_bwd_stacked: the ReLU mask needs the forward output or beta to recompute it");
  FB_REQUIRE(ch >= 8 && ch <= 2048 && (2048 % ch) == 0, "fb_bn_bwd_stacked: ch=%d must be a power of two in [8, 2048]", ch);
  int S = bn_slices(C, M, ch);
  if (bn_forced_slices(1) > 0) S = bn_forced_slices(1);
  if (bn_fused_enabled((int64_t)C * M * ch * (int64_t)sizeof(T))) {
    const int bar_base = bn_barrier_region((cudaStream_t)stream, C);
    int Sr = 0;
    uint32_t cap = 0;
    size_t smem = 0;
    {
      int Sp = 0, grid = 0;
      uint32_t capp = 0;
      size_t smemp = 0;
      if (bn_persist_plan<T>(C, M, ch, 2 + (dy2 ? 1 : 0) + ((relu && y) ? 1 : 0), &Sp, &capp, &smemp, &grid)) {
        if (int rc = bn_resident_allow_smem(bn_bwd_persist_kernel<T>, smemp)) return rc;
        FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_persist_kernel<T>, dim3(grid), dim3(BN_THREADS), smemp, (cudaStream_t)stream,
            (const T*)dy, (const T*)dy2, (const T*)x, (const T*)y, (T*)dx, (T*)dres, M, ch, Sp, (int)C, capp, scratch, gamma, beta,
            ld_p, save_mean, save_rstd, relu, dgamma, dbeta, ld_g, bar_base));
        FB_CHECK_CUDA(cudaGetLastError());
        return FB_OK;
      }
    }
    int extras = bn_resident_extras() ? ((dy2 ? 1 : 0) | ((relu && y) ? 2 : 0)) : 0;      // see bn_fwd_impl
    bool ok = bn_resident_plan<T>(M, ch, 2 + (extras & 1) + (extras >> 1), 1, false, &Sr, &cap, &smem);
    if (!ok && extras) {
      extras = 0;
      ok = bn_resident_plan<T>(M, ch, 2, 1, false, &Sr, &cap, &smem);
    }
    if (ok) {       // dy and x stay in shared memory between the two passes
      if (int rc = bn_resident_allow_smem(bn_bwd_resident_kernel<T>, smem)) return rc;
      FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_resident_kernel<T>, dim3(dim3(Sr, C)), dim3(BN_THREADS), smem, (cudaStream_t)stream,
          (const T*)dy, (const T*)dy2, (const T*)x, (const T*)y, (T*)dx, (T*)dres, M, ch, Sr, cap, scratch, gamma, beta, ld_p,
          save_mean, save_rstd, relu, dgamma, dbeta, ld_g, extras, bar_base));
      FB_CHECK_CUDA(cudaGetLastError());
      return FB_OK;
    }
    FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_fused_kernel<T>, dim3(dim3(S, C)), dim3(BN_THREADS), (size_t)(4 * ch * sizeof(float)), (cudaStream_t)stream, 
        (const T*)dy, (const T*)dy2, (const T*)x, (const T*)y, (T*)dx,
        (T*)dres, M, ch, S, scratch, gamma, beta, ld_p, save_mean, save_rstd, relu, dgamma, dbeta, ld_g, bar_base));
    FB_CHECK_CUDA(cudaGetLastError());
    return FB_OK;
  }
  FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_stats_kernel<T>, dim3(dim3(S, C)), dim3(BN_THREADS), (size_t)(0), (cudaStream_t)stream, 
      (const T*)dy, (const T*)dy2, (const T*)x, (const T*)y, M, ch, S, save_mean, save_rstd, gamma,
      beta, ld_p, relu, scratch, 1));
  FB_CHECK_CUDA(cudaGetLastError());
  FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_apply_kernel<T>, dim3(dim3(S, C)), dim3(BN_THREADS), (size_t)(4 * ch * sizeof(float)), (cudaStream_t)stream, 
      (const T*)dy, (const T*)dy2, (const T*)x, (const T*)y, (T*)dx,
      (T*)dres, M, ch, S, scratch, gamma, beta, ld_p, save_mean, save_rstd, relu, dgamma, dbeta, ld_g, 1, 0));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- im2col / col2im (3x3, pad 1, NHWC, bf16)
// cols[n][ho][wo][kh][kw][ch]; one thread moves one 16-byte vector (8 channels)
template <typename T>
__global__ void __launch_bounds__(256)
im2col3x3_kernel(const T* __restrict__ x, T* __restrict__ cols, int64_t total_vec, int H, int W,
                 int Ho, int Wo, int ch, int stride) {
  // all index arithmetic in 32 bits (64-bit div/mod per vector made this kernel ALU-bound); the host
  // guarantees total_vec < 2^32
  const uint32_t cv = ch / 8, per_pix = 9u * cv, per_row = per_pix * (uint32_t)Wo, per_img = per_row * (uint32_t)Ho;
  const uint32_t total = (uint32_t)total_vec;
  pdl_wait();
  pdl_trigger();
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t n = i / per_img, r0 = i - n * per_img;
    const uint32_t ho = r0 / per_row, r1 = r0 - ho * per_row;
    const uint32_t wo = r1 / per_pix, r2 = r1 - wo * per_pix;
    const uint32_t tap = r2 / cv, c8 = r2 - tap * cv;
    const int kh = (int)(tap / 3u), kw = (int)(tap - (tap / 3u) * 3u);
    const int h = (int)ho * stride + kh - 1, w = (int)wo * stride + kw - 1;
    Raw8<T> v = zero_raw<T>();
    if (h >= 0 && h < H && w >= 0 && w < W) v = ldraw(x + (((size_t)n * H + h) * W + w) * ch + (size_t)c8 * 8);
    store_raw(cols + (size_t)i * 8, v);
  }
}
template <typename T>
static int im2col_impl(const void* x, void* cols, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                                      int32_t stride, fb_stream_t stream) {
  FB_REQUIRE(x && cols && n_img >= 1 && H >= 1 && W >= 1 && ch >= 8 && ch % 8 == 0 && (stride == 1 || stride == 2),
             "fb_im2col3x3_nhwc: bad arguments (ch must be a multiple of 8)");
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  const int64_t total = (int64_t)n_img * Ho * Wo * 9 * (ch / 8);
  FB_REQUIRE(total < 4294967296LL, "fb_im2col3x3_nhwc: tensor too large for 32-bit indexing");
  int64_t grid = (total + 255) / 256, cap = (int64_t)fb_num_sms() * 16;
  FB_CHECK_CUDA(fb_launch_pdl(im2col3x3_kernel<T>, dim3((int)(grid < cap ? grid : cap)), dim3(256), (size_t)(0), (cudaStream_t)stream, 
      (const T*)x, (T*)cols, total, H, W, Ho, Wo, ch, stride));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// dx[n][h][w][c] = sum over taps of dcols[n][ho][wo][kh][kw][c] with h = ho*stride + kh - 1 (gather: no atomics)
template <typename T>
__global__ void __launch_bounds__(256)
col2im3x3_kernel(const T* __restrict__ cols, T* __restrict__ dx, int64_t total_vec, int H, int W,
                 int Ho, int Wo, int ch, int stride) {
  const uint32_t cv = ch / 8, per_row = cv * (uint32_t)W, per_img = per_row * (uint32_t)H;
  const uint32_t total = (uint32_t)total_vec;
  pdl_wait();
  pdl_trigger();
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t n = i / per_img, r0 = i - n * per_img;
    const int h = (int)(r0 / per_row);
    const uint32_t r1 = r0 - (uint32_t)h * per_row;
    const int w = (int)(r1 / cv);
    const uint32_t c8 = r1 - (uint32_t)w * cv;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int hh = h + 1 - kh;
      if (hh < 0 || hh % stride) continue;
      const int ho = hh / stride;
      if (ho >= Ho) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int ww = w + 1 - kw;
        if (ww < 0 || ww % stride) continue;
        const int wo = ww / stride;
        if (wo >= Wo) continue;
        float f[8];
        unpack8(ldraw(cols + (((((size_t)n * Ho + ho) * Wo + wo) * 3 + kh) * 3 + kw) * ch + (size_t)c8 * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
    }
    store8(dx + (size_t)i * 8, acc);
  }
}
template <typename T>
static int col2im_impl(const void* cols, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                                      int32_t stride, fb_stream_t stream) {
  FB_REQUIRE(cols && dx && n_img >= 1 && H >= 1 && W >= 1 && ch >= 8 && ch % 8 == 0 && (stride == 1 || stride == 2),
             "fb_col2im3x3_nhwc: bad arguments (ch must be a multiple of 8)");
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  const int64_t total = (int64_t)n_img * H * W * (ch / 8);
  FB_REQUIRE(total < 4294967296LL && (int64_t)n_img * Ho * Wo * 9 * (ch / 8) < 4294967296LL,
             "fb_col2im3x3_nhwc: tensor too large for 32-bit indexing");
  int64_t grid = (total + 255) / 256, cap = (int64_t)fb_num_sms() * 16;
  FB_CHECK_CUDA(fb_launch_pdl(col2im3x3_kernel<T>, dim3((int)(grid < cap ? grid : cap)), dim3(256), (size_t)(0), (cudaStream_t)stream, 
      (const T*)cols, (T*)dx, total, H, W, Ho, Wo, ch, stride));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- GroupNorm (+residual, +ReLU) on the same kernels
// A row is one (client, sample): rows = C * B, M = H * W pixels; statistics pool `cpg` adjacent channels of a row
// (reference model/cv/group_normalization.py:7-93, resnet_gn.py:108-196: two channels per group by default); gamma / beta
// are per client (row / rows_per_client).  Two launches per direction (statistics, apply); scratch as for BatchNorm
// (fb_bn_scratch_floats(rows, M, ch)).  save_mean / save_rstd: [rows][ch] (group value repeated per channel).
// Backward writes per-row dgamma / dbeta [rows][ch]; the caller sums them over the samples of a client.
template <typename T>
static int gn_fwd_impl(const void* x, const void* res, void* y, int32_t rows, int64_t M, int32_t ch, const float* gamma,
                       const float* beta, int64_t ld_p, int32_t rows_per_client, int32_t cpg, float eps, int32_t relu,
                       float* save_mean, float* save_rstd, float* scratch, fb_stream_t stream) {
  FB_REQUIRE(x && y && gamma && beta && save_mean && save_rstd && scratch && rows >= 1 && rows <= 65535 && M >= 1,
             "fb_gn_fwd_stacked: bad arguments (at most 65535 rows)");
  FB_REQUIRE(ch >= 8 && ch <= 2048 && (2048 % ch) == 0, "fb_gn_fwd_stacked: ch=%d must be a power of two in [8, 2048]", ch);
  FB_REQUIRE(cpg >= 2 && ch % cpg == 0 && rows_per_client >= 1 && rows % rows_per_client == 0, "fb_gn_fwd_stacked: bad grouping");
  const int S = bn_slices(rows, M, ch);
  FB_CHECK_CUDA(fb_launch_pdl(bn_stats_kernel<T>, dim3(S, rows), dim3(BN_THREADS), (size_t)0, (cudaStream_t)stream, (const T*)x, M, (int)ch, S,
                              scratch));
  FB_CHECK_CUDA(fb_launch_pdl(bn_apply_kernel<T>, dim3(S, rows), dim3(BN_THREADS), (size_t)(4 * ch * sizeof(float)), (cudaStream_t)stream,
                              (const T*)x, (const T*)res, (T*)y, M, (int)ch, S, (const float*)scratch, gamma, beta, ld_p, (float*)nullptr,
                              (float*)nullptr, (int64_t)0, (const float*)nullptr, 0.f, eps, (int)relu, 1, save_mean, save_rstd,
                              (float*)nullptr, (int64_t)0, S, (int)rows_per_client, (int)cpg));
  return FB_OK;
}
template <typename T>
static int gn_bwd_impl(const void* dy, const void* dy2, const void* x, const void* y, void* dx, void* dres, int32_t rows, int64_t M,
                       int32_t ch, const float* gamma, const float* beta, int64_t ld_p, int32_t rows_per_client, int32_t cpg,
                       const float* save_mean, const float* save_rstd, int32_t relu, float* dgamma_rows, float* dbeta_rows,
                       float* scratch, fb_stream_t stream) {
  FB_REQUIRE(dy && x && dx && gamma && save_mean && save_rstd && dgamma_rows && dbeta_rows && scratch && rows >= 1 && rows <= 65535,
             "fb_gn_bwd_stacked: bad arguments (at most 65535 rows)");
  FB_REQUIRE(!relu || y || beta, "fb_gn_bwd_stacked: the ReLU mask needs the forward output or beta to recompute it");
  FB_REQUIRE(ch >= 8 && ch <= 2048 && (2048 % ch) == 0, "fb_gn_bwd_stacked: ch=%d must be a power of two in [8, 2048]", ch);
  FB_REQUIRE(cpg >= 2 && ch % cpg == 0 && rows_per_client >= 1 && rows % rows_per_client == 0, "fb_gn_bwd_stacked: bad grouping");
  const int S = bn_slices(rows, M, ch);
  FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_stats_kernel<T>, dim3(S, rows), dim3(BN_THREADS), (size_t)0, (cudaStream_t)stream, (const T*)dy,
                              (const T*)dy2, (const T*)x, (const T*)y, M, (int)ch, S, save_mean, save_rstd, gamma, beta, ld_p, (int)relu,
                              scratch, (int)rows_per_client));
  FB_CHECK_CUDA(fb_launch_pdl(bn_bwd_apply_kernel<T>, dim3(S, rows), dim3(BN_THREADS), (size_t)(6 * ch * sizeof(float)),
                              (cudaStream_t)stream, (const T*)dy, (const T*)dy2, (const T*)x, (const T*)y, (T*)dx, (T*)dres, M, (int)ch, S,
                              (const float*)scratch, gamma, beta, ld_p, save_mean, save_rstd, (int)relu, dgamma_rows, dbeta_rows,
                              (int64_t)ch, (int)rows_per_client, (int)cpg));
  return FB_OK;
}
extern "C" int fb_gn_fwd_stacked(const void* x, const void* res, void* y, int32_t act_fp32, int32_t rows, int64_t M, int32_t ch,
                                 const float* gamma, const float* beta, int64_t ld_p, int32_t rows_per_client, int32_t cpg, float eps,
                                 int32_t relu, float* save_mean, float* save_rstd, float* scratch, fb_stream_t stream) {
  if (act_fp32)
    return gn_fwd_impl<float>(x, res, y, rows, M, ch, gamma, beta, ld_p, rows_per_client, cpg, eps, relu, save_mean, save_rstd, scratch, stream);
  return gn_fwd_impl<__nv_bfloat16>(x, res, y, rows, M, ch, gamma, beta, ld_p, rows_per_client, cpg, eps, relu, save_mean, save_rstd, scratch,
                                    stream);
}
extern "C" int fb_gn_bwd_stacked(const void* dy, const void* dy2, const void* x, const void* y, void* dx, void* dres, int32_t act_fp32,
                                 int32_t rows, int64_t M, int32_t ch, const float* gamma, const float* beta, int64_t ld_p,
                                 int32_t rows_per_client, int32_t cpg, const float* save_mean, const float* save_rstd, int32_t relu,
                                 float* dgamma_rows, float* dbeta_rows, float* scratch, fb_stream_t stream) {
  if (act_fp32)
    return gn_bwd_impl<float>(dy, dy2, x, y, dx, dres, rows, M, ch, gamma, beta, ld_p, rows_per_client, cpg, save_mean, save_rstd, relu,
                              dgamma_rows, dbeta_rows, scratch, stream);
  return gn_bwd_impl<__nv_bfloat16>(dy, dy2, x, y, dx, dres, rows, M, ch, gamma, beta, ld_p, rows_per_client, cpg, save_mean, save_rstd,
                                    relu, dgamma_rows, dbeta_rows, scratch, stream);
}

// ---------------------------------------------------------------- C ABI: bf16 and fp32 activation variants
#define FB_BN_FWD_ARGS                                                                                                      \
  const void *x, const void *res, void *y, int32_t C, int64_t M, int32_t ch, const float *gamma, const float *beta,        \
      int64_t ld_p, float *running_mean, float *running_var, int64_t ld_r, const float *active, float momentum, float eps, \
      int32_t relu, int32_t training, float *save_mean, float *save_rstd, float *scratch, float *num_batches_tracked,      \
      int64_t ld_n, fb_stream_t stream
#define FB_BN_FWD_PASS                                                                                                      \
  x, res, y, C, M, ch, gamma, beta, ld_p, running_mean, running_var, ld_r, active, momentum, eps, relu, training,          \
      save_mean, save_rstd, scratch, num_batches_tracked, ld_n, stream
extern "C" int fb_bn_fwd_stacked(FB_BN_FWD_ARGS) { return bn_fwd_impl<__nv_bfloat16>(FB_BN_FWD_PASS); }
extern "C" int fb_bn_fwd_stacked_f32(FB_BN_FWD_ARGS) { return bn_fwd_impl<float>(FB_BN_FWD_PASS); }
#define FB_BN_BWD_ARGS                                                                                                      \
  const void *dy, const void *dy2, const void *x, const void *y, void *dx, void *dres, int32_t C, int64_t M, int32_t ch,   \
      const float *gamma, const float *beta, int64_t ld_p, const float *save_mean, const float *save_rstd, int32_t relu,   \
      float *dgamma, float *dbeta, int64_t ld_g, float *scratch, fb_stream_t stream
#define FB_BN_BWD_PASS dy, dy2, x, y, dx, dres, C, M, ch, gamma, beta, ld_p, save_mean, save_rstd, relu, dgamma, dbeta, ld_g, scratch, stream
extern "C" int fb_bn_bwd_stacked(FB_BN_BWD_ARGS) { return bn_bwd_impl<__nv_bfloat16>(FB_BN_BWD_PASS); }
extern "C" int fb_bn_bwd_stacked_f32(FB_BN_BWD_ARGS) { return bn_bwd_impl<float>(FB_BN_BWD_PASS); }
extern "C" int fb_im2col3x3_nhwc_bf16(const void* x, void* cols, int32_t n_img, int32_t H, int32_t W, int32_t ch, int32_t stride,
                                      fb_stream_t stream) {
  return im2col_impl<__nv_bfloat16>(x, cols, n_img, H, W, ch, stride, stream);
}
extern "C" int fb_im2col3x3_nhwc_f32(const void* x, void* cols, int32_t n_img, int32_t H, int32_t W, int32_t ch, int32_t stride,
                                     fb_stream_t stream) {
  return im2col_impl<float>(x, cols, n_img, H, W, ch, stride, stream);
}
extern "C" int fb_col2im3x3_nhwc_bf16(const void* cols, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch, int32_t stride,
                                      fb_stream_t stream) {
  return col2im_impl<__nv_bfloat16>(cols, dx, n_img, H, W, ch, stride, stream);
}
extern "C" int fb_col2im3x3_nhwc_f32(const void* cols, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch, int32_t stride,
                                     fb_stream_t stream) {
  return col2im_impl<float>(cols, dx, n_img, H, W, ch, stride, stream);
}

// ---------------------------------------------------------------- 2x2 max pooling (stride 2) on NHWC, optional window offset
// y[n][i][j][:] = max over x[n][off_h + 2i + {0,1}][off_w + 2j + {0,1}][:]: the offset lets the pool read the interior of a
// "same"-padded convolution output (a valid convolution is its centre crop) without a cropping copy.  Replaces
// nn.MaxPool2d(2, stride=2) of the reference CNNs (model/cv/cnn.py:74-142).
template <typename T>
__global__ void __launch_bounds__(256)
maxpool2x2_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint32_t total_vec, int H, int W, int ch, int off_h, int off_w,
                      int Ho, int Wo) {
  const uint32_t cv = ch / 8, per_row = cv * (uint32_t)Wo, per_img = per_row * (uint32_t)Ho;
  pdl_wait();
  pdl_trigger();
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += gridDim.x * blockDim.x) {
    const uint32_t n = i / per_img, r0 = i - n * per_img;
    const uint32_t ho = r0 / per_row, r1 = r0 - ho * per_row;
    const uint32_t wo = r1 / cv, c8 = r1 - wo * cv;
    const size_t base = (((size_t)n * H + off_h + 2 * ho) * W + off_w + 2 * wo) * ch + (size_t)c8 * 8;
    float m[8], f[8];
    unpack8(ldraw(x + base), m);
    const size_t o[3] = {(size_t)ch, (size_t)W * ch, (size_t)W * ch + ch};
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      unpack8(ldraw(x + base + o[t]), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) m[e] = fmaxf(m[e], f[e]);
    }
    store8(y + (size_t)i * 8, m);
  }
}
// dx (full [H][W] map, zero outside the pooled windows): dy goes to the first maximal element of its window in scan order
// (the argmax rule of the reference's pooling backward); relu_mask also zeroes it where x <= 0 (pool input = ReLU output)
template <typename T>
__global__ void __launch_bounds__(256)
maxpool2x2_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, uint32_t total_vec, int H, int W,
                      int ch, int off_h, int off_w, int Ho, int Wo, int relu_mask) {
  const uint32_t cv = ch / 8, per_row = cv * (uint32_t)W, per_img = per_row * (uint32_t)H;
  pdl_wait();
  pdl_trigger();
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += gridDim.x * blockDim.x) {
    const uint32_t n = i / per_img, r0 = i - n * per_img;
    const int h = (int)(r0 / per_row);
    const uint32_t r1 = r0 - (uint32_t)h * per_row;
    const int w = (int)(r1 / cv);
    const uint32_t c8 = r1 - (uint32_t)w * cv;
    float out[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int hh = h - off_h, ww = w - off_w;
    if (hh >= 0 && ww >= 0 && hh < 2 * Ho && ww < 2 * Wo) {
      const int ho = hh >> 1, wo = ww >> 1, me = ((hh & 1) << 1) | (ww & 1);      // my position in the window, scan order
      const size_t base = (((size_t)n * H + off_h + 2 * ho) * W + off_w + 2 * wo) * ch + (size_t)c8 * 8;
      const size_t o[4] = {0, (size_t)ch, (size_t)W * ch, (size_t)W * ch + ch};
      float v[4][8], d[8];
#pragma unroll
      for (int t = 0; t < 4; ++t) unpack8(ldraw(x + base + o[t]), v[t]);
      unpack8(ldraw(dy + ((((size_t)n * Ho + ho) * Wo + wo) * ch + (size_t)c8 * 8)), d);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        int arg = 0;
        float best = v[0][e];
#pragma unroll
        for (int t = 1; t < 4; ++t)
          if (v[t][e] > best) {
            best = v[t][e];
            arg = t;
          }
        out[e] = (arg == me && (!relu_mask || best > 0.f)) ? d[e] : 0.f;
      }
    }
    store8(dx + (size_t)i * 8, out);
  }
}
template <typename T>
static int maxpool_impl(const void* x, void* y, const void* dy, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                        int32_t off_h, int32_t off_w, int32_t Ho, int32_t Wo, int32_t relu_mask, fb_stream_t stream) {
  FB_REQUIRE(x && n_img >= 1 && ch >= 8 && ch % 8 == 0 && Ho >= 1 && Wo >= 1 && off_h >= 0 && off_w >= 0 && off_h + 2 * Ho <= H &&
                 off_w + 2 * Wo <= W,
             "fb_maxpool2x2: bad arguments (ch a multiple of 8, windows inside the map)");
  const int64_t total = dx ? (int64_t)n_img * H * W * (ch / 8) : (int64_t)n_img * Ho * Wo * (ch / 8);
  FB_REQUIRE(total < 4294967296LL, "fb_maxpool2x2: tensor too large for 32-bit indexing");
  int64_t grid = (total + 255) / 256, cap = (int64_t)fb_num_sms() * 16;
  if (dx) {
    FB_REQUIRE(dy, "fb_maxpool2x2_bwd: dy missing");
    FB_CHECK_CUDA(fb_launch_pdl(maxpool2x2_bwd_kernel<T>, dim3((int)(grid < cap ? grid : cap)), dim3(256), (size_t)0, (cudaStream_t)stream,
        (const T*)dy, (const T*)x, (T*)dx, (uint32_t)total, H, W, ch, off_h, off_w, Ho, Wo, relu_mask));
  } else {
    FB_REQUIRE(y, "fb_maxpool2x2_fwd: y missing");
    FB_CHECK_CUDA(fb_launch_pdl(maxpool2x2_fwd_kernel<T>, dim3((int)(grid < cap ? grid : cap)), dim3(256), (size_t)0, (cudaStream_t)stream,
        (const T*)x, (T*)y, (uint32_t)total, H, W, ch, off_h, off_w, Ho, Wo));
  }
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_maxpool2x2_fwd(const void* x, void* y, int32_t act_fp32, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                                 int32_t off_h, int32_t off_w, int32_t Ho, int32_t Wo, fb_stream_t stream) {
  return act_fp32 ? maxpool_impl<float>(x, y, nullptr, nullptr, n_img, H, W, ch, off_h, off_w, Ho, Wo, 0, stream)
                  : maxpool_impl<__nv_bfloat16>(x, y, nullptr, nullptr, n_img, H, W, ch, off_h, off_w, Ho, Wo, 0, stream);
}
extern "C" int fb_maxpool2x2_bwd(const void* dy, const void* x, void* dx, int32_t act_fp32, int32_t n_img, int32_t H, int32_t W,
                                 int32_t ch, int32_t off_h, int32_t off_w, int32_t Ho, int32_t Wo, int32_t relu_mask,
                                 fb_stream_t stream) {
  return act_fp32 ? maxpool_impl<float>(x, nullptr, dy, dx, n_img, H, W, ch, off_h, off_w, Ho, Wo, relu_mask, stream)
                  : maxpool_impl<__nv_bfloat16>(x, nullptr, dy, dx, n_img, H, W, ch, off_h, off_w, Ho, Wo, relu_mask, stream);
}

// debugging aid: per-CTA timeline of the resident BatchNorm forward kernel (NULL turns it off); see tools/trace_bn.py
extern "C" int fb_bn_set_trace(void* buf_u64) {
  unsigned long long* p = (unsigned long long*)buf_u64;
  FB_CHECK_CUDA(cudaMemcpyToSymbol(g_bn_trace, &p, sizeof(p)));
  return FB_OK;
}
