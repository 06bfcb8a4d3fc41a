// K12 helpers: client-stacked transformer-encoder kernels (DistilBERT family) around the tcgen05 grouped GEMM.
// See include/fedml_b200.h.  The GEMM-shaped work (QKV / output / FFN projections, Q*K^T, P*V and every
// gradient product) runs on fb_grouped_gemm_{bf16,tn_bf16,nn_bf16} with two-level (batch, head) groups; the
// kernels here are the bandwidth-bound glue, each a single pass over bf16 activations with fp32 arithmetic:
//
//   embed_gather / embed_bwd     word + position embedding rows of the fp32 arena -> bf16, and its scatter-add
//   add_ln_fwd / ln_bwd          (residual add +) LayerNorm, one warp per token row, statistics in registers
//   softmax_fwd / softmax_bwd    scaled, key-masked softmax over fp32 score rows -> bf16 probabilities
//   act_fwd / act_bwd            exact (erf) GELU and ReLU
//   dropout                      Philox-keyed inverted dropout (the backward pass re-derives the same mask)
//   colsum                       bias gradients (column sums of bf16 gradient matrices), deterministic
//
// Layout: activations are [C clients][R = B*S token rows][D], D fastest; per-client parameters are rows of the
// stacked fp32 arena (row stride ld).  Everything is deterministic except embed_bwd's word-row scatter, which
// uses fp32 atomics (repeated tokens collide).
#include "common.cuh"
#include "../include/fedml_b200.h"
#include <stdlib.h>

#define TF_THREADS 256
#define TF_WARPS 8
#define LN_MAXV 4      /* 8-element vectors per lane: D <= 32 * 8 * 4 = 1024 */

namespace {

__device__ __forceinline__ void unpack8(const uint4& r, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// eight consecutive elements of a bf16 or fp32 tensor <-> fp32 registers (kernels templated on the activation type)
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float* f) { unpack8(ldg16(p), f); }
__device__ __forceinline__ void ld8(const float* p, float* f) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float* f) { *reinterpret_cast<uint4*>(p) = pack8(f); }
__device__ __forceinline__ void st8(float* p, const float* f) {
  reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}
__device__ __forceinline__ void ld8f(const float* p, float* f) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------- embeddings
// out[c][r][:] = word[c][ids[c][r]][:] + pos[c][r % S][:]
__global__ void __launch_bounds__(TF_THREADS)
embed_gather_kernel(const int64_t* __restrict__ ids, const float* __restrict__ word, const float* __restrict__ pos,
                    int64_t ld, __nv_bfloat16* __restrict__ out, int64_t R, int S, int D, int vocab) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, c = blockIdx.y;
  const int64_t r = (int64_t)blockIdx.x * TF_WARPS + warp;
  if (r >= R) return;
  int64_t id = ids[(size_t)c * R + r];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const float* w = word + (size_t)c * ld + (size_t)id * D;
  const float* p = pos + (size_t)c * ld + (size_t)(r % S) * D;
  __nv_bfloat16* o = out + ((size_t)c * R + r) * D;
  for (int k = lane * 8; k < D; k += 256) {
    float a[8], b[8];
    ld8f(w + k, a);
    ld8f(p + k, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += b[e];
    *reinterpret_cast<uint4*>(o + k) = pack8(a);
  }
}

// gword[c][ids[c][r]][:] += dz[c][r][:]   (fp32 atomics: tokens repeat)
__global__ void __launch_bounds__(TF_THREADS)
embed_word_bwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dz, float* __restrict__ gword,
                      int64_t ld, int64_t R, int D, int vocab, const float* __restrict__ client_scale) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, c = blockIdx.y;
  const int64_t r = (int64_t)blockIdx.x * TF_WARPS + warp;
  if (r >= R) return;
  int64_t id = ids[(size_t)c * R + r];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  float* g = gword + (size_t)c * ld + (size_t)id * D;
  const __nv_bfloat16* d = dz + ((size_t)c * R + r) * D;
  const float sc = client_scale ? client_scale[c] : 1.f;
  if (sc == 0.f) return;
  for (int k = lane * 8; k < D; k += 256) {
    float f[8];
    unpack8(ldg16(d + k), f);
#pragma unroll
    for (int e = 0; e < 8; ++e) atomicAdd(g + k + e, f[e] * sc);
  }
}

// gpos[c][s][:] = sum_b dz[c][b*S + s][:]   (deterministic: one thread owns 8 columns of one position)
__global__ void __launch_bounds__(TF_THREADS)
embed_pos_bwd_kernel(const __nv_bfloat16* __restrict__ dz, float* __restrict__ gpos, int64_t ld, int B, int S, int D) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int c = blockIdx.y;
  const int64_t v = (int64_t)blockIdx.x * TF_THREADS + threadIdx.x;     // vector index inside [S][D/8]
  const int nv = D / 8;
  if (v >= (int64_t)S * nv) return;
  const int s = (int)(v / nv), k = (int)(v % nv) * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int b = 0; b < B; ++b) {
    float f[8];
    unpack8(ldg16(dz + (((size_t)c * B + b) * S + s) * D + k), f);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] += f[e];
  }
  float* g = gpos + (size_t)c * ld + (size_t)s * D + k;
  *reinterpret_cast<float4*>(g) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4*>(g + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

// ---------------------------------------------------------------- (add +) LayerNorm forward
// z = bf16(a + b)  (b optional);  y = (z - mean) * rstd * gamma + beta.  One warp per token row; the row lives in
// registers (LN_MAXV vectors of 8 per lane), statistics are exact two-pass over the bf16-rounded z so that the
// backward pass, which re-reads z, sees the same x-hat.
__global__ void __launch_bounds__(TF_THREADS)
add_ln_fwd_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b, __nv_bfloat16* __restrict__ z,
                  __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out,
                  const float* __restrict__ gamma, const float* __restrict__ beta, int64_t ld_p, int64_t R, int D, float eps) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, c = blockIdx.y;
  const int64_t r = (int64_t)blockIdx.x * TF_WARPS + warp;
  if (r >= R) return;
  const int nv = D >> 3;
  const size_t base = ((size_t)c * R + r) * D;
  float v[LN_MAXV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int idx = i * 32 + lane;
    if (idx < nv) {
      unpack8(ldg16(a + base + idx * 8), v[i]);
      if (b) {
        float t[8];
        unpack8(ldg16(b + base + idx * 8), t);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[i][e] = bf16_round(v[i][e] + t[e]);
        if (z) *reinterpret_cast<uint4*>(z + base + idx * 8) = pack8(v[i]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) sum += v[i][e];
    }
  }
  const float mean = warp_sum(sum) / (float)D;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (i * 32 + lane < nv) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = v[i][e] - mean;
        sq += d * d;
      }
    }
  const float rstd = rsqrtf(warp_sum(sq) / (float)D + eps);
  if (lane == 0) {
    mean_out[(size_t)c * R + r] = mean;
    rstd_out[(size_t)c * R + r] = rstd;
  }
  const float* g = gamma + (size_t)c * ld_p;
  const float* bt = beta + (size_t)c * ld_p;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int idx = i * 32 + lane;
    if (idx < nv) {
      float gg[8], bb[8], o[8];
      ld8f(g + idx * 8, gg);
      ld8f(bt + idx * 8, bb);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (v[i][e] - mean) * rstd * gg[e] + bb[e];
      *reinterpret_cast<uint4*>(y + base + idx * 8) = pack8(o);
    }
  }
}

// ---------------------------------------------------------------- LayerNorm backward
// g = (dy + dy2) * gamma;  dz = rstd * (g - mean_D(g) - xhat * mean_D(g * xhat));  per-CTA partial sums of
// dgamma = sum_rows (dy + dy2) * xhat and dbeta = sum_rows (dy + dy2) go to partial[c][slice][2][D].
// The running dgamma / dbeta sums of a warp live in its private slice of shared memory (lane-contiguous float4
// slots, conflict-free), not in registers: the row itself (NV vectors of 8 per lane, twice) is all a thread holds,
// which keeps three CTAs resident per SM.
template <int NV>
__global__ void __launch_bounds__(TF_THREADS, 2)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ dy2, const __nv_bfloat16* __restrict__ z,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ gamma,
              int64_t ld_p, __nv_bfloat16* __restrict__ dz, float* __restrict__ partial, int64_t R, int D, int slices) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  extern __shared__ float4 acc4[];                 // [warp][quantity q][vector i][half][lane] float4
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, c = blockIdx.y, s = blockIdx.x;
  const int nv = D >> 3;
  const int64_t rows_per = (R + slices - 1) / slices;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(R, r0 + rows_per);
  float4* mine = acc4 + (size_t)warp * (2 * NV * 2 * 32);
#pragma unroll
  for (int j = 0; j < 2 * NV * 2; ++j) mine[j * 32 + lane] = make_float4(0.f, 0.f, 0.f, 0.f);
  const float* gam = gamma + (size_t)c * ld_p;
  const float invD = 1.f / (float)D;
  for (int64_t r = r0 + warp; r < r1; r += TF_WARPS) {
    const size_t base = ((size_t)c * R + r) * D;
    const float mean = mean_in[(size_t)c * R + r], rstd = rstd_in[(size_t)c * R + r];
    float g[NV][8], xh[NV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = i * 32 + lane;
      if (idx < nv) {
        float gm[8];
        unpack8(ldg16(dy + base + idx * 8), g[i]);
        if (dy2) {
          float t[8];
          unpack8(ldg16(dy2 + base + idx * 8), t);
#pragma unroll
          for (int e = 0; e < 8; ++e) g[i][e] += t[e];
        }
        unpack8(ldg16(z + base + idx * 8), xh[i]);
        ld8f(gam + idx * 8, gm);
        float4 a0 = mine[((0 * NV + i) * 2 + 0) * 32 + lane], a1 = mine[((0 * NV + i) * 2 + 1) * 32 + lane];
        float4 b0 = mine[((1 * NV + i) * 2 + 0) * 32 + lane], b1 = mine[((1 * NV + i) * 2 + 1) * 32 + lane];
        float* pa0 = &a0.x; float* pa1 = &a1.x; float* pb0 = &b0.x; float* pb1 = &b1.x;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          xh[i][e] = (xh[i][e] - mean) * rstd;
          const float d = g[i][e];
          if (e < 4) { pa0[e] += d * xh[i][e]; pb0[e] += d; }
          else { pa1[e - 4] += d * xh[i][e]; pb1[e - 4] += d; }
          g[i][e] = d * gm[e];
          s1 += g[i][e];
          s2 += g[i][e] * xh[i][e];
        }
        mine[((0 * NV + i) * 2 + 0) * 32 + lane] = a0; mine[((0 * NV + i) * 2 + 1) * 32 + lane] = a1;
        mine[((1 * NV + i) * 2 + 0) * 32 + lane] = b0; mine[((1 * NV + i) * 2 + 1) * 32 + lane] = b1;
      }
    }
    s1 = warp_sum(s1) * invD;
    s2 = warp_sum(s2) * invD;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = i * 32 + lane;
      if (idx < nv) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = rstd * (g[i][e] - s1 - xh[i][e] * s2);
        *reinterpret_cast<uint4*>(dz + base + idx * 8) = pack8(o);
      }
    }
  }
  __syncthreads();
  // fixed-order sum over the 8 warps; slot j of a warp's slice holds column ((i*32 + l)*8 + half*4 + e)
  const float* accf = reinterpret_cast<const float*>(acc4);
  constexpr int PER_WARP = 2 * NV * 2 * 32 * 4;
  float* out = partial + ((size_t)c * slices + s) * 2 * D;
  for (int j = threadIdx.x; j < PER_WARP; j += TF_THREADS) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < TF_WARPS; ++w) t += accf[w * PER_WARP + j];
    const int e = j & 3, l = (j >> 2) & 31, half = (j >> 7) & 1, qi = j >> 8;   // qi = q * NV + i
    const int q = qi / NV, i = qi - q * NV;
    const int col = (i * 32 + l) * 8 + half * 4 + e;
    if (col < D) out[q * D + col] = t;
  }
}

__global__ void __launch_bounds__(TF_THREADS)
ln_bwd_finish_kernel(const float* __restrict__ partial, int slices, int D, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     int64_t ld_g) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int c = blockIdx.y, k = blockIdx.x * TF_THREADS + threadIdx.x;
  if (k >= 2 * D) return;
  float acc = 0.f;
  for (int s = 0; s < slices; ++s) acc += partial[((size_t)c * slices + s) * 2 * D + k];
  if (k < D) dgamma[(size_t)c * ld_g + k] = acc;
  else dbeta[(size_t)c * ld_g + (k - D)] = acc;
}

// ---------------------------------------------------------------- dropout mask
// Eight keep/drop decisions for the 8-element vector `v` of a tensor: one Philox block keyed (seed, step, site),
// one 16-bit uniform per element.  Forward and backward passes regenerate identical masks from the same key.
struct DropKey {
  uint64_t seed;
  uint32_t site, step, thresh;
  float keep_scale;
};
__device__ __forceinline__ DropKey make_drop_key(float p, uint64_t seed, uint32_t site, const uint32_t* step_dev) {
  DropKey k;
  k.seed = seed;
  k.site = site;
  k.step = step_dev ? *step_dev : 0u;
  k.thresh = (uint32_t)(p * 65536.f);
  k.keep_scale = 1.f / (1.f - p);
  return k;
}
__device__ __forceinline__ void drop_apply8(const DropKey& k, int64_t v, float* f) {
  const Philox4 r = philox4x32_10((uint32_t)v, (uint32_t)((uint64_t)v >> 32), k.site, k.step, (uint32_t)k.seed,
                                  (uint32_t)(k.seed >> 32));
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const uint32_t u = (w[e >> 1] >> ((e & 1) * 16)) & 0xFFFFu;
    f[e] = u >= k.thresh ? f[e] * k.keep_scale : 0.f;
  }
}

// ---------------------------------------------------------------- attention softmax
// P[row][j] = softmax_j(scale * scores[row][j] + (key_mask[j] ? 0 : -inf));  Pd = dropout(P) when p > 0.
// rows = groups * S with group = (client*batch) * H + head; key_mask is [client*batch][S] (1 = attend) or null.
// LPR lanes own one row (32 / LPR rows per warp, so short sequences keep every lane busy); a lane holds VPL
// vectors of 8 consecutive keys.  S % 8 == 0, S <= 8 * LPR * VPL.
template <int LPR>
__device__ __forceinline__ float sub_sum(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <int LPR>
__device__ __forceinline__ float sub_max(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <int LPR, int VPL>
__global__ void __launch_bounds__(TF_THREADS)
softmax_fwd_kernel(const float* __restrict__ scores, __nv_bfloat16* __restrict__ P, __nv_bfloat16* __restrict__ Pd,
                   const float* __restrict__ key_mask, int64_t rows, int S, int H, float scale, float p, uint64_t seed,
                   uint32_t site, const uint32_t* __restrict__ step_dev) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  constexpr int RPW = 32 / LPR;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, l = lane % LPR;
  const int64_t row = ((int64_t)blockIdx.x * TF_WARPS + warp) * RPW + lane / LPR;
  const bool valid = row < rows;
  const int nv = S >> 3;
  const float* src = scores + (size_t)row * S;
  const float* mk = key_mask ? key_mask + (size_t)(row / S / H) * S : nullptr;
  float v[VPL][8];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int idx = i * LPR + l;
#pragma unroll
    for (int e = 0; e < 8; ++e) v[i][e] = -INFINITY;
    if (valid && idx < nv) {
      float m[8];
      ld8f(src + idx * 8, v[i]);
      if (mk) ld8f(mk + idx * 8, m);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        v[i][e] = (mk && m[e] == 0.f) ? -INFINITY : v[i][e] * scale;
        mx = fmaxf(mx, v[i][e]);
      }
    }
  }
  mx = sub_max<LPR>(mx);
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      v[i][e] = (mx == -INFINITY) ? 0.f : __expf(v[i][e] - mx);
      sum += v[i][e];
    }
  sum = sub_sum<LPR>(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  const DropKey key = make_drop_key(p, seed, site, step_dev);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int idx = i * LPR + l;
    if (valid && idx < nv) {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[i][e] = bf16_round(v[i][e] * inv);
      *reinterpret_cast<uint4*>(P + (size_t)row * S + idx * 8) = pack8(v[i]);
      if (p > 0.f) {
        drop_apply8(key, (int64_t)row * nv + idx, v[i]);
        *reinterpret_cast<uint4*>(Pd + (size_t)row * S + idx * 8) = pack8(v[i]);
      }
    }
  }
}

// dS = scale * P * (dP' - sum_j dP'_j P_j)   with dP' = dropout-backward(dP) when p > 0
template <int LPR, int VPL>
__global__ void __launch_bounds__(TF_THREADS)
softmax_bwd_kernel(const float* __restrict__ dP, const __nv_bfloat16* __restrict__ P, __nv_bfloat16* __restrict__ dS,
                   int64_t rows, int S, float scale, float p, uint64_t seed, uint32_t site,
                   const uint32_t* __restrict__ step_dev) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  constexpr int RPW = 32 / LPR;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, l = lane % LPR;
  const int64_t row = ((int64_t)blockIdx.x * TF_WARPS + warp) * RPW + lane / LPR;
  const bool valid = row < rows;
  const int nv = S >> 3;
  const DropKey key = make_drop_key(p, seed, site, step_dev);
  float pv[VPL][8], gv[VPL][8];
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int idx = i * LPR + l;
#pragma unroll
    for (int e = 0; e < 8; ++e) pv[i][e] = gv[i][e] = 0.f;
    if (valid && idx < nv) {
      unpack8(ldg16(P + (size_t)row * S + idx * 8), pv[i]);
      ld8f(dP + (size_t)row * S + idx * 8, gv[i]);
      if (p > 0.f) drop_apply8(key, (int64_t)row * nv + idx, gv[i]);
#pragma unroll
      for (int e = 0; e < 8; ++e) dot += pv[i][e] * gv[i][e];
    }
  }
  dot = sub_sum<LPR>(dot);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int idx = i * LPR + l;
    if (valid && idx < nv) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = scale * pv[i][e] * (gv[i][e] - dot);
      *reinterpret_cast<uint4*>(dS + (size_t)row * S + idx * 8) = pack8(o);
    }
  }
}

// pick (lanes per row, vectors per lane) for a sequence length; invokes LAUNCH(LPR, VPL)
#define SM_DISPATCH(S, LAUNCH)          \
  do {                                  \
    const int nv_ = (S) >> 3;           \
    if (nv_ > 32) { LAUNCH(32, 2); }    \
    else if (nv_ > 16) { LAUNCH(32, 1); } \
    else if (nv_ > 8) { LAUNCH(16, 1); }  \
    else if (nv_ > 4) { LAUNCH(8, 1); }   \
    else { LAUNCH(4, 1); }              \
  } while (0)
#define SM_MAX_S 512

// ---------------------------------------------------------------- activations / dropout / bias gradients
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad(float x) {
  return 0.5f * (1.f + erff(x * 0.70710678118654752f)) + x * 0.3989422804014327f * __expf(-0.5f * x * x);
}

// mode 0: ReLU, 1: GELU (erf)
template <typename T>
__global__ void __launch_bounds__(TF_THREADS)
act_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t nvec, int mode) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  for (int64_t v = (int64_t)blockIdx.x * TF_THREADS + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * TF_THREADS) {
    float f[8];
    ld8(x + v * 8, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = mode ? gelu_f(f[e]) : fmaxf(f[e], 0.f);
    st8(y + v * 8, f);
  }
}
// dx = dy * act'(x)   (x = pre-activation)
template <typename T>
__global__ void __launch_bounds__(TF_THREADS)
act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t nvec, int mode) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  for (int64_t v = (int64_t)blockIdx.x * TF_THREADS + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * TF_THREADS) {
    float f[8], g[8];
    ld8(x + v * 8, f);
    ld8(dy + v * 8, g);
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] *= mode ? gelu_grad(f[e]) : (f[e] > 0.f ? 1.f : 0.f);
    st8(dx + v * 8, g);
  }
}

// y = dropout(x (+ x2)): inverted dropout, mask from drop_apply8 over the flattened tensor
template <typename T>
__global__ void __launch_bounds__(TF_THREADS)
dropout_kernel(const T* __restrict__ x, const T* __restrict__ x2, T* __restrict__ y,
               int64_t nvec, float p, uint64_t seed, uint32_t site, const uint32_t* __restrict__ step_dev) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const DropKey key = make_drop_key(p, seed, site, step_dev);
  for (int64_t v = (int64_t)blockIdx.x * TF_THREADS + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * TF_THREADS) {
    float f[8];
    ld8(x + v * 8, f);
    if (x2) {
      float t[8];
      ld8(x2 + v * 8, t);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] += t[e];
    }
    drop_apply8(key, v, f);
    st8(y + v * 8, f);
  }
}

// out[c][n] = sum_r x[c][r][n]: 32 row lanes x 8 column vectors per CTA, fixed summation order
template <typename T>
__global__ void __launch_bounds__(TF_THREADS)
colsum_kernel(const T* __restrict__ x, int64_t R, int N, int64_t ldx, int64_t cs, float* __restrict__ out,
              int64_t ld_o) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  __shared__ float red[32][65];
  const int c = blockIdx.y, tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
  const int col = blockIdx.x * 64 + tx * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (col < N) {
    const T* base = x + (size_t)c * cs + col;
    for (int64_t r = ty; r < R; r += 32) {
      float f[8];
      ld8(base + (size_t)r * ldx, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += f[e];
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[ty][tx * 8 + e] = acc[e];
  __syncthreads();
  if (threadIdx.x < 64) {
    float s = 0.f;
    for (int r = 0; r < 32; ++r) s += red[r][threadIdx.x];
    const int n = blockIdx.x * 64 + threadIdx.x;
    if (n < N) out[(size_t)c * ld_o + n] = s;
  }
}

static int ew_grid(int64_t nvec) {
  int64_t g = (nvec + TF_THREADS - 1) / TF_THREADS;
  const int64_t cap = (int64_t)fb_num_sms() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

// ================================================================ C ABI
#define TF_ALIGNED16(p) (((uintptr_t)(p) & 15) == 0)

extern "C" int fb_embed_gather(const int64_t* ids, const float* word, const float* pos, int64_t ld, void* out_bf16,
                               int32_t C, int64_t R, int32_t S, int32_t D, int32_t vocab, fb_stream_t stream) {
  FB_REQUIRE(ids && word && pos && out_bf16, "fb_embed_gather: null pointer");
  FB_REQUIRE(C >= 1 && C <= 65535 && R >= 1 && S >= 1 && R % S == 0 && D >= 8 && D % 8 == 0 && vocab >= 1,
             "fb_embed_gather: bad shape");
  FB_REQUIRE(ld % 4 == 0 && TF_ALIGNED16(word) && TF_ALIGNED16(pos) && TF_ALIGNED16(out_bf16),
             "fb_embed_gather: tables and output must be 16-byte aligned");
  FB_CHECK_CUDA(fb_launch_pdl(embed_gather_kernel, dim3(dim3((unsigned)((R + TF_WARPS - 1) / TF_WARPS), C)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, 
      ids, word, pos, ld, (__nv_bfloat16*)out_bf16, R, S, D, vocab));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

extern "C" int fb_embed_bwd(const int64_t* ids, const void* dz_bf16, float* gword, float* gpos, int64_t ld, int32_t C,
                            int64_t R, int32_t S, int32_t D, int32_t vocab, const float* word_scale, fb_stream_t stream) {
  FB_REQUIRE(ids && dz_bf16 && gword && gpos, "fb_embed_bwd: null pointer");
  FB_REQUIRE(C >= 1 && C <= 65535 && R >= 1 && S >= 1 && R % S == 0 && D >= 8 && D % 8 == 0 && vocab >= 1,
             "fb_embed_bwd: bad shape");
  FB_REQUIRE(ld % 4 == 0 && TF_ALIGNED16(gpos) && TF_ALIGNED16(dz_bf16), "fb_embed_bwd: buffers must be 16-byte aligned");
  FB_CHECK_CUDA(fb_launch_pdl(embed_word_bwd_kernel, dim3(dim3((unsigned)((R + TF_WARPS - 1) / TF_WARPS), C)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, 
      ids, (const __nv_bfloat16*)dz_bf16, gword, ld, R, D, vocab, word_scale));
  const int64_t nv = (int64_t)S * (D / 8);
  FB_CHECK_CUDA(fb_launch_pdl(embed_pos_bwd_kernel, dim3(dim3((unsigned)((nv + TF_THREADS - 1) / TF_THREADS), C)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, 
      (const __nv_bfloat16*)dz_bf16, gpos, ld, (int)(R / S), S, D));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

extern "C" int fb_add_ln_fwd(const void* a, const void* b, void* z, void* y, float* mean, float* rstd, const float* gamma,
                             const float* beta, int64_t ld_p, int32_t C, int64_t R, int32_t D, float eps,
                             fb_stream_t stream) {
  FB_REQUIRE(a && y && mean && rstd && gamma && beta, "fb_add_ln_fwd: null pointer");
  FB_REQUIRE(C >= 1 && C <= 65535 && R >= 1 && D >= 8 && D % 8 == 0 && D <= 256 * LN_MAXV,
             "fb_add_ln_fwd: D must be a multiple of 8 and at most %d", 256 * LN_MAXV);
  FB_REQUIRE(ld_p % 4 == 0 && TF_ALIGNED16(gamma) && TF_ALIGNED16(beta) && TF_ALIGNED16(a) && TF_ALIGNED16(y),
             "fb_add_ln_fwd: buffers must be 16-byte aligned");
  FB_CHECK_CUDA(fb_launch_pdl(add_ln_fwd_kernel, dim3(dim3((unsigned)((R + TF_WARPS - 1) / TF_WARPS), C)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, 
      (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)z, (__nv_bfloat16*)y, mean, rstd, gamma, beta, ld_p, R,
      D, eps));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

static int ln_slices(int C, int64_t R) {
  static int rows_per_warp = 0, ctas_per_sm = 0;
  if (!rows_per_warp) {
    const char* e = getenv("FB_LN_ROWS_PER_WARP");
    rows_per_warp = e ? atoi(e) : 4;
    if (rows_per_warp < 1) rows_per_warp = 1;
    e = getenv("FB_LN_CTAS_PER_SM");
    ctas_per_sm = e ? atoi(e) : 4;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
  }
  int64_t want = ((int64_t)ctas_per_sm * fb_num_sms() + C - 1) / C;
  const int64_t max_s = (R + TF_WARPS * rows_per_warp - 1) / (TF_WARPS * rows_per_warp);
  if (want > max_s) want = max_s;
  if (want < 1) want = 1;
  if (want > 256) want = 256;
  return (int)want;
}
extern "C" int64_t fb_ln_bwd_scratch_floats(int32_t C, int64_t R, int32_t D) { return (int64_t)C * ln_slices(C, R) * 2 * D; }

extern "C" int fb_ln_bwd(const void* dy, const void* dy2, const void* z, const float* mean, const float* rstd,
                         const float* gamma, int64_t ld_p, void* dz, float* dgamma, float* dbeta, int64_t ld_g,
                         float* scratch, int32_t C, int64_t R, int32_t D, fb_stream_t stream) {
  FB_REQUIRE(dy && z && mean && rstd && gamma && dz && dgamma && dbeta && scratch, "fb_ln_bwd: null pointer");
  FB_REQUIRE(C >= 1 && C <= 65535 && R >= 1 && D >= 8 && D % 8 == 0 && D <= 256 * LN_MAXV,
             "fb_ln_bwd: D must be a multiple of 8 and at most %d", 256 * LN_MAXV);
  FB_REQUIRE(ld_p % 4 == 0 && TF_ALIGNED16(gamma) && TF_ALIGNED16(dy) && TF_ALIGNED16(z) && TF_ALIGNED16(dz),
             "fb_ln_bwd: buffers must be 16-byte aligned");
  const int S = ln_slices(C, R);
  const int NV = (D + 255) / 256;
#define FB_LN_BWD(NVV)                                                                                                      \
  do {                                                                                                                      \
    const int smem = TF_WARPS * 2 * NVV * 2 * 32 * 16;                                                                      \
    FB_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<NVV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));              \
    ln_bwd_kernel<NVV><<<dim3(S, C), TF_THREADS, smem, (cudaStream_t)stream>>>(                                              \
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)dy2, (const __nv_bfloat16*)z, mean, rstd, gamma, ld_p,              \
        (__nv_bfloat16*)dz, scratch, R, D, S);                                                                              \
  } while (0)
  if (NV == 1) FB_LN_BWD(1);
  else if (NV == 2) FB_LN_BWD(2);
  else if (NV == 3) FB_LN_BWD(3);
  else FB_LN_BWD(4);
#undef FB_LN_BWD
  FB_CHECK_CUDA(fb_launch_pdl(ln_bwd_finish_kernel, dim3(dim3((2 * D + TF_THREADS - 1) / TF_THREADS, C)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, scratch, S, D, dgamma,
                                                                                                            dbeta, ld_g));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

extern "C" int fb_attn_softmax_fwd(const float* scores, void* P_bf16, void* Pdrop_bf16, const float* key_mask, int64_t groups,
                                   int32_t S, int32_t H, float scale, float p, uint64_t seed, uint32_t site,
                                   const uint32_t* step_dev, fb_stream_t stream) {
  FB_REQUIRE(scores && P_bf16 && groups >= 1 && H >= 1, "fb_attn_softmax_fwd: bad arguments");
  FB_REQUIRE(S >= 8 && S % 8 == 0 && S <= SM_MAX_S, "fb_attn_softmax_fwd: S must be a multiple of 8 and at most %d",
             SM_MAX_S);
  FB_REQUIRE(p >= 0.f && p < 1.f && (p == 0.f || Pdrop_bf16), "fb_attn_softmax_fwd: dropout needs 0 <= p < 1 and a Pdrop buffer");
  FB_REQUIRE(TF_ALIGNED16(scores) && TF_ALIGNED16(P_bf16) && TF_ALIGNED16(Pdrop_bf16) && TF_ALIGNED16(key_mask),
             "fb_attn_softmax_fwd: buffers must be 16-byte aligned");
  const int64_t rows = groups * S;
#define FB_SM_FWD(LPR, VPL)                                                                                               \
  softmax_fwd_kernel<LPR, VPL><<<(unsigned)((rows + TF_WARPS * (32 / LPR) - 1) / (TF_WARPS * (32 / LPR))), TF_THREADS, 0,    \
                                 (cudaStream_t)stream>>>(scores, (__nv_bfloat16*)P_bf16, (__nv_bfloat16*)Pdrop_bf16, key_mask, \
                                                         rows, S, H, scale, p, seed, site, step_dev)
  SM_DISPATCH(S, FB_SM_FWD);
#undef FB_SM_FWD
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

extern "C" int fb_attn_softmax_bwd(const float* dP, const void* P_bf16, void* dS_bf16, int64_t groups, int32_t S, float scale,
                                   float p, uint64_t seed, uint32_t site, const uint32_t* step_dev, fb_stream_t stream) {
  FB_REQUIRE(dP && P_bf16 && dS_bf16 && groups >= 1, "fb_attn_softmax_bwd: bad arguments");
  FB_REQUIRE(S >= 8 && S % 8 == 0 && S <= SM_MAX_S, "fb_attn_softmax_bwd: S must be a multiple of 8 and at most %d",
             SM_MAX_S);
  FB_REQUIRE(p >= 0.f && p < 1.f, "fb_attn_softmax_bwd: 0 <= p < 1");
  FB_REQUIRE(TF_ALIGNED16(dP) && TF_ALIGNED16(P_bf16) && TF_ALIGNED16(dS_bf16), "fb_attn_softmax_bwd: buffers must be 16-byte aligned");
  const int64_t rows = groups * S;
#define FB_SM_BWD(LPR, VPL)                                                                                            \
  softmax_bwd_kernel<LPR, VPL><<<(unsigned)((rows + TF_WARPS * (32 / LPR) - 1) / (TF_WARPS * (32 / LPR))), TF_THREADS, 0, \
                                 (cudaStream_t)stream>>>(dP, (const __nv_bfloat16*)P_bf16, (__nv_bfloat16*)dS_bf16, rows, S, \
                                                         scale, p, seed, site, step_dev)
  SM_DISPATCH(S, FB_SM_BWD);
#undef FB_SM_BWD
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

template <typename T>
static int act_fwd_impl(const void* x, void* y, int64_t n, int32_t mode, fb_stream_t stream) {
  FB_REQUIRE(x && y && n >= 8 && n % 8 == 0 && (mode == 0 || mode == 1) && TF_ALIGNED16(x) && TF_ALIGNED16(y),
             "fb_act_fwd: n must be a multiple of 8, buffers 16-byte aligned, mode 0 (ReLU) or 1 (GELU)");
  FB_CHECK_CUDA(fb_launch_pdl(act_fwd_kernel<T>, dim3(ew_grid(n / 8)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, (const T*)x, (T*)y, n / 8, mode));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_act_fwd(const void* x, void* y, int64_t n, int32_t mode, fb_stream_t stream) {
  return act_fwd_impl<__nv_bfloat16>(x, y, n, mode, stream);
}
extern "C" int fb_act_fwd_f32(const void* x, void* y, int64_t n, int32_t mode, fb_stream_t stream) {
  return act_fwd_impl<float>(x, y, n, mode, stream);
}

template <typename T>
static int act_bwd_impl(const void* dy, const void* x, void* dx, int64_t n, int32_t mode, fb_stream_t stream) {
  FB_REQUIRE(dy && x && dx && n >= 8 && n % 8 == 0 && (mode == 0 || mode == 1) && TF_ALIGNED16(x) && TF_ALIGNED16(dy) &&
                 TF_ALIGNED16(dx),
             "fb_act_bwd: n must be a multiple of 8, buffers 16-byte aligned, mode 0 (ReLU) or 1 (GELU)");
  FB_CHECK_CUDA(fb_launch_pdl(act_bwd_kernel<T>, dim3(ew_grid(n / 8)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, (const T*)dy, (const T*)x,
                              (T*)dx, n / 8, mode));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_act_bwd(const void* dy, const void* x, void* dx, int64_t n, int32_t mode, fb_stream_t stream) {
  return act_bwd_impl<__nv_bfloat16>(dy, x, dx, n, mode, stream);
}
extern "C" int fb_act_bwd_f32(const void* dy, const void* x, void* dx, int64_t n, int32_t mode, fb_stream_t stream) {
  return act_bwd_impl<float>(dy, x, dx, n, mode, stream);
}

template <typename T>
static int dropout_impl(const void* x, const void* x2, void* y, int64_t n, float p, uint64_t seed, uint32_t site,
                        const uint32_t* step_dev, fb_stream_t stream) {
  FB_REQUIRE(x && y && n >= 8 && n % 8 == 0 && p >= 0.f && p < 1.f && TF_ALIGNED16(x) && TF_ALIGNED16(x2) && TF_ALIGNED16(y),
             "fb_dropout: n must be a multiple of 8, buffers 16-byte aligned, 0 <= p < 1");
  FB_CHECK_CUDA(fb_launch_pdl(dropout_kernel<T>, dim3(ew_grid(n / 8)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, (const T*)x, (const T*)x2,
                              (T*)y, n / 8, p, seed, site, step_dev));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_dropout_bf16(const void* x, const void* x2, void* y, int64_t n, float p, uint64_t seed, uint32_t site,
                               const uint32_t* step_dev, fb_stream_t stream) {
  return dropout_impl<__nv_bfloat16>(x, x2, y, n, p, seed, site, step_dev, stream);
}
extern "C" int fb_dropout_f32(const void* x, const void* x2, void* y, int64_t n, float p, uint64_t seed, uint32_t site,
                              const uint32_t* step_dev, fb_stream_t stream) {
  return dropout_impl<float>(x, x2, y, n, p, seed, site, step_dev, stream);
}

template <typename T>
static int colsum_impl(const void* x, int32_t C, int64_t R, int32_t N, int64_t ldx, int64_t client_stride, float* out,
                       int64_t ld_o, fb_stream_t stream) {
  constexpr int A = 16 / (int)sizeof(T);      // elements per 16 bytes
  FB_REQUIRE(x && out && C >= 1 && C <= 65535 && R >= 1 && N >= 8 && N % 8 == 0 && ldx % A == 0 && client_stride % A == 0 &&
                 TF_ALIGNED16(x),
             "fb_colsum: N a multiple of 8, row and client strides multiples of 16 bytes, x 16-byte aligned");
  FB_CHECK_CUDA(fb_launch_pdl(colsum_kernel<T>, dim3(dim3((N + 63) / 64, C)), dim3(TF_THREADS), (size_t)(0), (cudaStream_t)stream, (const T*)x, R, N, ldx, client_stride,
                              out, ld_o));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_colsum_bf16(const void* x, int32_t C, int64_t R, int32_t N, int64_t ldx, int64_t client_stride, float* out,
                              int64_t ld_o, fb_stream_t stream) {
  return colsum_impl<__nv_bfloat16>(x, C, R, N, ldx, client_stride, out, ld_o, stream);
}
extern "C" int fb_colsum_f32(const void* x, int32_t C, int64_t R, int32_t N, int64_t ldx, int64_t client_stride, float* out,
                             int64_t ld_o, fb_stream_t stream) {
  return colsum_impl<float>(x, C, R, N, ldx, client_stride, out, ld_o, stream);
}
