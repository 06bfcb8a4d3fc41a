// K10 helpers: client-stacked BatchNorm (+residual, +ReLU) forward/backward and 3x3 im2col / col2im on
// NHWC bf16 activations.  See include/fedml_b200.h.  All four are bandwidth-bound streaming kernels:
// 16-byte (8 x bf16) vector accesses, fp32 arithmetic, no atomics, deterministic two-stage reductions.
//
// Layout: activations are [C clients][M = B*H*W rows][ch channels], channel fastest.  A CTA of 256 threads
// walks its row range 2048 elements at a time; because 2048 % ch == 0 a thread always sees the same 8
// channels, so per-channel partial sums live in registers for the whole walk.
#include "common.cuh"
#include "../include/fedml_b200.h"

#define BN_THREADS 256
#define BN_VEC 8

struct bf16x8 {
  uint4 raw;
};
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
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, int64_t M, int ch, int S, float* __restrict__ partial) {
  __shared__ float smem[BN_THREADS * 16];
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const __nv_bfloat16* base = x + (size_t)c * M * ch;
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += BN_THREADS) {
    float f[8];
    unpack8(ldg16(base + v * BN_VEC), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] += f[i];
      b[i] += f[i] * f[i];
    }
  }
  reduce_channel_pairs(a, b, ch, smem, partial + ((size_t)c * S + s) * 2 * ch);
}

// stage 2: finish the statistics (every CTA redundantly, S is small), update running stats, apply
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ y,
                int64_t M, int ch, int S, const float* __restrict__ partial, const float* __restrict__ gamma,
                const float* __restrict__ beta, int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r,
                const float* __restrict__ active, float momentum, float eps, int relu, int training,
                float* __restrict__ save_mean, float* __restrict__ save_rstd) {
  extern __shared__ float sm[];  // scale[ch], shift[ch]
  float* scale = sm;
  float* shift = sm + ch;
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  for (int k = tid; k < ch; k += BN_THREADS) {
    float mean, var;
    if (training) {
      float su = 0.f, sq = 0.f;
      for (int j = 0; j < S; ++j) {
        su += partial[((size_t)c * S + j) * 2 * ch + k];
        sq += partial[((size_t)c * S + j) * 2 * ch + ch + k];
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
  for (int64_t v = v0 + tid; v < v1; v += BN_THREADS) {
    float f[8];
    unpack8(ldg16(x + cbase + v * BN_VEC), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = f[i] * sc[i] + sh[i];
    if (res) {
      float r[8];
      unpack8(ldg16(res + cbase + v * BN_VEC), r);
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] += r[i];
    }
    if (relu) {
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = fmaxf(f[i], 0.f);
    }
    *reinterpret_cast<uint4*>(y + cbase + v * BN_VEC) = pack8(f);
  }
}

static int bn_slices(int C, int64_t M, int ch) {
  // enough CTAs to fill the machine twice, but at least 2048 elements (one sweep) per CTA iteration
  int64_t want = (2 * (int64_t)fb_num_sms() + C - 1) / C;
  int64_t max_s = (M * ch + 8191) / 8192;
  if (want > max_s) want = max_s;
  if (want < 1) want = 1;
  if (want > 64) want = 64;
  return (int)want;
}

extern "C" int64_t fb_bn_scratch_floats(int32_t C, int64_t M, int32_t ch) { return (int64_t)C * 64 * 2 * ch; }

extern "C" int fb_bn_fwd_stacked(const void* x, const void* res, void* y, int32_t C, int64_t M, int32_t ch,
                                 const float* gamma, const float* beta, int64_t ld_p, float* running_mean,
                                 float* running_var, int64_t ld_r, const float* active, float momentum, float eps,
                                 int32_t relu, int32_t training, float* save_mean, float* save_rstd, float* scratch,
                                 fb_stream_t stream) {
  FB_REQUIRE(x && y && gamma && beta && C >= 1 && C <= 65535 && M >= 1, "fb_bn_fwd_stacked: bad arguments");
  FB_REQUIRE(ch >= 8 && ch <= 2048 && (2048 % ch) == 0, "fb_bn_fwd_stacked: ch=%d must be a power of two in [8, 2048]", ch);
  FB_REQUIRE(running_mean && running_var, "fb_bn_fwd_stacked: running statistics missing");
  FB_REQUIRE(!training || scratch, "fb_bn_fwd_stacked: training needs the partial-sum scratch");
  const int S = bn_slices(C, M, ch);
  if (training) {
    bn_stats_kernel<<<dim3(S, C), BN_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, M, ch, S, scratch);
    FB_CHECK_CUDA(cudaGetLastError());
  }
  bn_apply_kernel<<<dim3(S, C), BN_THREADS, 2 * ch * sizeof(float), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)res, (__nv_bfloat16*)y, M, ch, S, scratch, gamma, beta, ld_p,
      running_mean, running_var, ld_r, active, momentum, eps, relu, training, save_mean, save_rstd);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- BN backward
// stage 1: partial[c][s][2][ch] = (sum dy', sum dy' * xhat), dy' = dy masked by the ReLU of the forward output
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_stats_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                    const __nv_bfloat16* __restrict__ y, int64_t M, int ch, int S, const float* __restrict__ mean,
                    const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                    int64_t ld_p, int relu, float* __restrict__ partial) {
  __shared__ float smem[BN_THREADS * 16];
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const size_t cbase = (size_t)c * M * ch;
  const int coff = (tid * BN_VEC) % ch;
  float mu[8], rs[8], ga[8], be[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    mu[i] = mean[(size_t)c * ch + coff + i];
    rs[i] = rstd[(size_t)c * ch + coff + i];
    ga[i] = gamma[(size_t)c * ld_p + coff + i];
    be[i] = beta ? beta[(size_t)c * ld_p + coff + i] : 0.f;
  }
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += BN_THREADS) {
    float d[8], xv[8];
    unpack8(ldg16(dy + cbase + v * BN_VEC), d);
    unpack8(ldg16(x + cbase + v * BN_VEC), xv);
    if (relu) {
      if (y) {   // residual block output: the mask needs the stored forward output
        float yv[8];
        unpack8(ldg16(y + cbase + v * BN_VEC), yv);
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] = yv[i] > 0.f ? d[i] : 0.f;
      } else {   // plain BN+ReLU: recompute the sign from x instead of streaming y again
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] = ((xv[i] - mu[i]) * rs[i] * ga[i] + be[i]) > 0.f ? d[i] : 0.f;
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] += d[i];
      b[i] += d[i] * (xv[i] - mu[i]) * rs[i];
    }
  }
  reduce_channel_pairs(a, b, ch, smem, partial + ((size_t)c * S + s) * 2 * ch);
}

__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                    const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres,
                    int64_t M, int ch, int S, const float* __restrict__ partial, const float* __restrict__ gamma,
                    const float* __restrict__ beta, int64_t ld_p, const float* __restrict__ mean,
                    const float* __restrict__ rstd, int relu, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    int64_t ld_g) {
  extern __shared__ float sm[];  // k1[ch] = gamma*rstd, k2[ch] = dbeta/M, k3[ch] = dgamma/M * rstd ... see below
  float* s_g = sm;               // gamma * rstd
  float* s_db = sm + ch;         // dbeta / M
  float* s_dg = sm + 2 * ch;     // dgamma / M
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  for (int k = tid; k < ch; k += BN_THREADS) {
    float db = 0.f, dg = 0.f;
    for (int j = 0; j < S; ++j) {
      db += partial[((size_t)c * S + j) * 2 * ch + k];
      dg += partial[((size_t)c * S + j) * 2 * ch + ch + k];
    }
    if (s == 0) {
      dbeta[(size_t)c * ld_g + k] = db;
      dgamma[(size_t)c * ld_g + k] = dg;
    }
    s_g[k] = gamma[(size_t)c * ld_p + k] * rstd[(size_t)c * ch + k];
    s_db[k] = db / (float)M;
    s_dg[k] = dg / (float)M;
  }
  __syncthreads();
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = (int64_t)s * rows_per, r1 = min(M, r0 + rows_per);
  const size_t cbase = (size_t)c * M * ch;
  const int coff = (tid * BN_VEC) % ch;
  float g[8], db[8], dg[8], mu[8], rs[8], be[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    g[i] = s_g[coff + i];
    db[i] = s_db[coff + i];
    dg[i] = s_dg[coff + i];
    mu[i] = mean[(size_t)c * ch + coff + i];
    rs[i] = rstd[(size_t)c * ch + coff + i];
    be[i] = beta ? beta[(size_t)c * ld_p + coff + i] : 0.f;
  }
  const int64_t v0 = r0 * ch / BN_VEC, v1 = r1 * ch / BN_VEC;
  for (int64_t v = v0 + tid; v < v1; v += BN_THREADS) {
    float d[8], xv[8];
    unpack8(ldg16(dy + cbase + v * BN_VEC), d);
    unpack8(ldg16(x + cbase + v * BN_VEC), xv);
    if (relu) {
      if (y) {
        float yv[8];
        unpack8(ldg16(y + cbase + v * BN_VEC), yv);
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] = yv[i] > 0.f ? d[i] : 0.f;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] = ((xv[i] - mu[i]) * g[i] + be[i]) > 0.f ? d[i] : 0.f;
      }
    }
    if (dres) *reinterpret_cast<uint4*>(dres + cbase + v * BN_VEC) = pack8(d);
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = g[i] * (d[i] - db[i] - (xv[i] - mu[i]) * rs[i] * dg[i]);
    *reinterpret_cast<uint4*>(dx + cbase + v * BN_VEC) = pack8(o);
  }
}

extern "C" int fb_bn_bwd_stacked(const void* dy, const void* x, const void* y, void* dx, void* dres, int32_t C,
                                 int64_t M, int32_t ch, const float* gamma, const float* beta, int64_t ld_p,
                                 const float* save_mean,
                                 const float* save_rstd, int32_t relu, float* dgamma, float* dbeta, int64_t ld_g,
                                 float* scratch, fb_stream_t stream) {
  FB_REQUIRE(dy && x && dx && gamma && save_mean && save_rstd && dgamma && dbeta && scratch && C >= 1 && C <= 65535,
             "fb_bn_bwd_stacked: bad arguments");
  FB_REQUIRE(!relu || y || beta, "fb_bn_bwd_stacked: the ReLU mask needs the forward output or beta to recompute it");
  FB_REQUIRE(ch >= 8 && ch <= 2048 && (2048 % ch) == 0, "fb_bn_bwd_stacked: ch=%d must be a power of two in [8, 2048]", ch);
  const int S = bn_slices(C, M, ch);
  bn_bwd_stats_kernel<<<dim3(S, C), BN_THREADS, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, M, ch, S, save_mean, save_rstd, gamma,
      beta, ld_p, relu, scratch);
  FB_CHECK_CUDA(cudaGetLastError());
  bn_bwd_apply_kernel<<<dim3(S, C), BN_THREADS, 3 * ch * sizeof(float), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (__nv_bfloat16*)dx,
      (__nv_bfloat16*)dres, M, ch, S, scratch, gamma, beta, ld_p, save_mean, save_rstd, relu, dgamma, dbeta, ld_g);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- im2col / col2im (3x3, pad 1, NHWC, bf16)
// cols[n][ho][wo][kh][kw][ch]; one thread moves one 16-byte vector (8 channels)
__global__ void __launch_bounds__(256)
im2col3x3_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols, int64_t total_vec, int H, int W,
                 int Ho, int Wo, int ch, int stride) {
  const int cv = ch / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t t = i;
    const int c8 = (int)(t % cv); t /= cv;
    const int kw = (int)(t % 3); t /= 3;
    const int kh = (int)(t % 3); t /= 3;
    const int wo = (int)(t % Wo); t /= Wo;
    const int ho = (int)(t % Ho); t /= Ho;
    const int64_t n = t;
    const int h = ho * stride + kh - 1, w = wo * stride + kw - 1;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (h >= 0 && h < H && w >= 0 && w < W) v = ldg16(x + (((size_t)n * H + h) * W + w) * ch + (size_t)c8 * 8);
    *reinterpret_cast<uint4*>(cols + (size_t)i * 8) = v;
  }
}
extern "C" int fb_im2col3x3_nhwc_bf16(const void* x, void* cols, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                                      int32_t stride, fb_stream_t stream) {
  FB_REQUIRE(x && cols && n_img >= 1 && H >= 1 && W >= 1 && ch >= 8 && ch % 8 == 0 && (stride == 1 || stride == 2),
             "fb_im2col3x3_nhwc_bf16: bad arguments (ch must be a multiple of 8)");
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  const int64_t total = (int64_t)n_img * Ho * Wo * 9 * (ch / 8);
  int64_t grid = (total + 255) / 256, cap = (int64_t)fb_num_sms() * 16;
  im2col3x3_kernel<<<(int)(grid < cap ? grid : cap), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (__nv_bfloat16*)cols, total, H, W, Ho, Wo, ch, stride);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// dx[n][h][w][c] = sum over taps of dcols[n][ho][wo][kh][kw][c] with h = ho*stride + kh - 1 (gather: no atomics)
__global__ void __launch_bounds__(256)
col2im3x3_kernel(const __nv_bfloat16* __restrict__ cols, __nv_bfloat16* __restrict__ dx, int64_t total_vec, int H, int W,
                 int Ho, int Wo, int ch, int stride) {
  const int cv = ch / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t t = i;
    const int c8 = (int)(t % cv); t /= cv;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); t /= H;
    const int64_t n = t;
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
        unpack8(ldg16(cols + (((((size_t)n * Ho + ho) * Wo + wo) * 3 + kh) * 3 + kw) * ch + (size_t)c8 * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
    }
    *reinterpret_cast<uint4*>(dx + (size_t)i * 8) = pack8(acc);
  }
}
extern "C" int fb_col2im3x3_nhwc_bf16(const void* cols, void* dx, int32_t n_img, int32_t H, int32_t W, int32_t ch,
                                      int32_t stride, fb_stream_t stream) {
  FB_REQUIRE(cols && dx && n_img >= 1 && H >= 1 && W >= 1 && ch >= 8 && ch % 8 == 0 && (stride == 1 || stride == 2),
             "fb_col2im3x3_nhwc_bf16: bad arguments (ch must be a multiple of 8)");
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  const int64_t total = (int64_t)n_img * H * W * (ch / 8);
  int64_t grid = (total + 255) / 256, cap = (int64_t)fb_num_sms() * 16;
  col2im3x3_kernel<<<(int)(grid < cap ? grid : cap), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)cols, (__nv_bfloat16*)dx, total, H, W, Ho, Wo, ch, stride);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
