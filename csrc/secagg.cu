// K8: secure-aggregation arithmetic in GF(p), p < 2^31.  See include/fedml_b200.h.
// Reference maths: turboaggregate/mpc_function.py:4-281 (NumPy int64 `% p` loops on the CPU).
#include "common.cuh"
#include "../include/fedml_b200.h"

static inline int ff_grid(int64_t n) {
  int64_t want = (n + 255) / 256, cap = (int64_t)fb_num_sms() * 8;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

__global__ void __launch_bounds__(256) ff_quantize_kernel(const float* __restrict__ x, int64_t n, float scale,
                                                          uint32_t p, uint32_t* __restrict__ q) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const long long r = __float2ll_rn(x[i] * scale);
    long long m = r % (long long)p;
    if (m < 0) m += p;
    q[i] = (uint32_t)m;
  }
}
__global__ void __launch_bounds__(256) ff_dequantize_kernel(const uint32_t* __restrict__ q, int64_t n, float inv_scale,
                                                            uint32_t p, float* __restrict__ x) {
  const uint32_t half = p >> 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t v = q[i];
    const float s = v > half ? -(float)(p - v) : (float)v;  // values above p/2 are negatives
    x[i] = s * inv_scale;
  }
}
extern "C" int fb_ff_quantize(const float* x, int64_t n, float scale, uint32_t p, uint32_t* q, fb_stream_t stream) {
  FB_REQUIRE(x && q && n > 0 && p > 1 && p < 0x80000000u, "fb_ff_quantize: bad arguments");
  ff_quantize_kernel<<<ff_grid(n), 256, 0, (cudaStream_t)stream>>>(x, n, scale, p, q);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_ff_dequantize(const uint32_t* q, int64_t n, float inv_scale, uint32_t p, float* x,
                                fb_stream_t stream) {
  FB_REQUIRE(x && q && n > 0 && p > 1 && p < 0x80000000u, "fb_ff_dequantize: bad arguments");
  ff_dequantize_kernel<<<ff_grid(n), 256, 0, (cudaStream_t)stream>>>(q, n, inv_scale, p, x);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// out[i][d] = sum_j U[i][j] * X[j][d] mod p.  One thread per (d); U row staged in smem; the running sum
// is kept below 2^63 by reducing every 2 products (each product < 2^62).
#define FF_MAX_IN 1024
__global__ void __launch_bounds__(256) ff_matmul_kernel(const uint32_t* __restrict__ U, int n_in,
                                                        const uint32_t* __restrict__ X, int64_t d, int64_t ldx,
                                                        uint32_t p, uint32_t* __restrict__ out, int64_t ldo) {
  __shared__ uint32_t u_row[FF_MAX_IN];
  const int i = blockIdx.y;
  for (int j = threadIdx.x; j < n_in; j += blockDim.x) u_row[j] = U[(size_t)i * n_in + j] % p;
  __syncthreads();
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < d; c += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long acc = 0ull;
    int j = 0;
    for (; j + 2 <= n_in; j += 2) {
      acc += (unsigned long long)u_row[j] * X[(size_t)j * ldx + c];
      acc += (unsigned long long)u_row[j + 1] * X[(size_t)(j + 1) * ldx + c];
      acc %= p;
    }
    if (j < n_in) acc = (acc + (unsigned long long)u_row[j] * X[(size_t)j * ldx + c]) % p;
    out[(size_t)i * ldo + c] = (uint32_t)acc;
  }
}
extern "C" int fb_ff_matmul(const uint32_t* U, int32_t n_out, int32_t n_in, const uint32_t* X, int64_t d, int64_t ldx,
                            uint32_t p, uint32_t* out, int64_t ldo, fb_stream_t stream) {
  FB_REQUIRE(U && X && out && n_out >= 1 && n_out <= 65535 && n_in >= 1 && d > 0 && ldx >= d && ldo >= d,
             "fb_ff_matmul: bad arguments");
  FB_REQUIRE(p > 1 && p < 0x80000000u, "fb_ff_matmul: p must be below 2^31");
  if (n_in > FF_MAX_IN) {
    fb_set_error("fb_ff_matmul: n_in=%d exceeds the supported %d", n_in, FF_MAX_IN);
    return FB_ERR_LIMIT;
  }
  int gx = ff_grid(d);
  int lim = (fb_num_sms() * 8 + n_out - 1) / n_out;
  if (gx > lim) gx = lim < 1 ? 1 : lim;
  ff_matmul_kernel<<<dim3(gx, n_out), 256, 0, (cudaStream_t)stream>>>(U, n_in, X, d, ldx, p, out, ldo);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// y = x + sum_s sign_s * PRG(seed_s, nonce) mod p.  PRG word i of seed s = Philox(counter=(i/4, stream tag, nonce),
// key=seed)[i%4] mod p.  The nonce (the FL round) makes every round's masks different: with a fixed mask the difference
// of two uploads of one client would reveal the difference of its two quantised updates.
__global__ void __launch_bounds__(256) ff_mask_kernel(const uint32_t* __restrict__ x, int64_t n,
                                                      const uint64_t* __restrict__ seeds,
                                                      const int32_t* __restrict__ signs, int n_seeds, uint32_t p,
                                                      uint32_t nonce, uint32_t* __restrict__ y) {
  const int64_t nvec = (n + 3) / 4;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    long long acc[4] = {0, 0, 0, 0};
    for (int s = 0; s < n_seeds; ++s) {
      const uint64_t seed = seeds[s];
      const Philox4 r = philox4x32_10((uint32_t)v, (uint32_t)(v >> 32), 0x5ec0a66u, nonce, (uint32_t)seed,
                                      (uint32_t)(seed >> 32));
      const long long sg = signs[s];
      acc[0] += sg * (long long)(r.x % p);
      acc[1] += sg * (long long)(r.y % p);
      acc[2] += sg * (long long)(r.z % p);
      acc[3] += sg * (long long)(r.w % p);
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int64_t i = v * 4 + e;
      if (i < n) {
        long long t = ((long long)x[i] + acc[e]) % (long long)p;
        if (t < 0) t += p;
        y[i] = (uint32_t)t;
      }
    }
  }
}
extern "C" int fb_ff_mask(const uint32_t* x, int64_t n, const uint64_t* seeds_dev, const int32_t* signs_dev,
                          int32_t n_seeds, uint32_t p, uint32_t nonce, uint32_t* y, fb_stream_t stream) {
  FB_REQUIRE(x && y && n > 0 && n_seeds >= 0 && (n_seeds == 0 || (seeds_dev && signs_dev)), "fb_ff_mask: bad arguments");
  FB_REQUIRE(p > 1 && p < 0x80000000u, "fb_ff_mask: p must be below 2^31");
  ff_mask_kernel<<<ff_grid((n + 3) / 4), 256, 0, (cudaStream_t)stream>>>(x, n, seeds_dev, signs_dev, n_seeds, p, nonce, y);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

__global__ void __launch_bounds__(256) ff_add_kernel(uint32_t* __restrict__ acc, const uint32_t* __restrict__ x,
                                                     int64_t n, uint32_t p) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t s = acc[i] + x[i];  // both < 2^31: no overflow
    acc[i] = s >= p ? s - p : s;
  }
}
extern "C" int fb_ff_add(uint32_t* acc, const uint32_t* x, int64_t n, uint32_t p, fb_stream_t stream) {
  FB_REQUIRE(acc && x && n > 0 && p > 1 && p < 0x80000000u, "fb_ff_add: bad arguments");
  ff_add_kernel<<<ff_grid(n), 256, 0, (cudaStream_t)stream>>>(acc, x, n, p);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
