// Persistent, software-pipelined BatchNorm (included by conv.cu after bn_resident.cuh).
//
// The resident kernels move the minimum bytes but spend ~6.5 us per CTA outside of data movement (launch, TMA latency,
// per-client barrier, coefficient prologue, store completion; fitted from tools/bench_bn.py), so HBM idles: 3.4-3.9 TB/s
// of min-bytes throughput.  Here one CTA per SM walks work items (client c, slice s) = blockIdx.x + i * gridDim.x with a
// ring of NB = 3 shared-memory buffers filled by bulk TMA:
//   A(i): wait for the slice, per-channel partial sums from shared memory, publish, ARRIVE at the client's barrier
//   B(i): WAIT for the client's barrier, add up the partials, coefficients, apply from shared memory, 16-byte stores to HBM,
//         then refill the buffer with item i + NB
// issued in the order A(i+1), B(i): a CTA arrives for its next item before it waits for the current one, so the barrier
// and the TMA latency of later items hide behind the passes over earlier ones.  A client's S slices sit on S consecutive
// CTAs (S <= 128 < number of SMs), all CTAs are resident (grid <= SMs, one CTA per SM), and every CTA arrives at item
// w's barrier before it can block on any item >= w, so the lowest unfinished client always completes: no deadlock.
#pragma once

#define BNP_NB 3

__device__ __forceinline__ void bnp_arrive(int c, int S, unsigned* saved_gen) {
  // thread 0 only, after the partial sums of this CTA have been written by all threads (__syncthreads before the call)
  unsigned* st = g_bn_barrier + 2 * c;
  __threadfence();
  *saved_gen = ld_acquire_gpu(st + 1);          // read BEFORE arriving: the bump we wait for is the one after this value
  if (S > 1) {
    if (atomicAdd(st, 1u) == (unsigned)S - 1) {
      st[0] = 0;
      __threadfence();
      atomicAdd(st + 1, 1u);
    }
  }
}
__device__ __forceinline__ void bnp_wait(int c, int S, unsigned saved_gen) {
  if (S > 1) {
    const unsigned* st = g_bn_barrier + 2 * c;
    while (ld_acquire_gpu(st + 1) == saved_gen) __nanosleep(20);
    __threadfence();
  }
}

struct BnpItem {
  int c, s;
  uint32_t bytes;     // bytes of one array's slice
  size_t goff;        // element offset of the slice in the [C][M][ch] tensors
};
template <typename T>
__device__ __forceinline__ BnpItem bnp_item(int i, int S, int64_t M, int ch) {
  const int w = blockIdx.x + i * gridDim.x;
  BnpItem it;
  it.c = w / S;
  it.s = w - it.c * S;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = min(M, (int64_t)it.s * rows_per), r1 = min(M, r0 + rows_per);
  it.bytes = (uint32_t)((r1 - r0) * ch * (int64_t)sizeof(T));
  it.goff = ((size_t)it.c * M + r0) * ch;
  return it;
}

// dynamic shared memory: NB buffers of `arrays` slices (slice_cap bytes each), then coefficients [4*ch], complete sums
// [2*ch], reduction scratch [BN_THREADS*2*N]
template <typename T>
__global__ void __launch_bounds__(BN_THREADS, 1)
bn_fwd_persist_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, int64_t M, int ch, int S, int C,
                      uint32_t slice_cap, float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
                      int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r, const float* __restrict__ active,
                      float momentum, float eps, int relu, float* __restrict__ save_mean, float* __restrict__ save_rstd,
                      float* num_batches_tracked, int64_t ld_n, int bar_base) {
  constexpr int N = Vec16<T>::N;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t full[BNP_NB];
  __shared__ unsigned s_gen[BNP_NB];
  const int arrays = res ? 2 : 1;
  const size_t buf_bytes = (size_t)arrays * slice_cap;
  float* coef = reinterpret_cast<float*>(dsm + BNP_NB * buf_bytes);
  float* sums = coef + 4 * ch;
  float* red = coef + 6 * ch;
  const int tid = threadIdx.x;
  const int n_items = C * S;
  const int my_n = ((int)blockIdx.x < n_items) ? (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (tid == 0)
    for (int k = 0; k < BNP_NB; ++k) bnr_mbar_init(full + k, 1);
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  const int coff = (tid * N) % ch;

  auto issue = [&](int i) {            // warp 0: bulk loads of item i into buffer i % NB
    const BnpItem it = bnp_item<T>(i, S, M, ch);
    unsigned char* b = dsm + (size_t)(i % BNP_NB) * buf_bytes;
    if (tid == 0) bnr_mbar_expect_tx(full + i % BNP_NB, it.bytes * (uint32_t)arrays);
    __syncwarp();
    bnr_load_slice(b, x + it.goff, it.bytes, full + i % BNP_NB);
    if (res) bnr_load_slice(b + slice_cap, res + it.goff, it.bytes, full + i % BNP_NB);
  };
  auto phase_a = [&](int i) {          // statistics of item i, publish, arrive
    const BnpItem it = bnp_item<T>(i, S, M, ch);
    const uint4* xs = reinterpret_cast<const uint4*>(dsm + (size_t)(i % BNP_NB) * buf_bytes);
    bnr_mbar_wait(full + i % BNP_NB, (uint32_t)(i / BNP_NB) & 1u);
    const int nvec = (int)(it.bytes / 16u);
    float a[N], b[N];
#pragma unroll
    for (int k = 0; k < N; ++k) a[k] = b[k] = 0.f;
    for (int v = tid; v < nvec; v += BN_THREADS) {
      float f[N];
      unpack16(xs[v], f, (const T*)nullptr);
#pragma unroll
      for (int k = 0; k < N; ++k) {
        a[k] += f[k];
        b[k] += f[k] * f[k];
      }
    }
    bnr_reduce<N>(a, b, ch, red, partial + ((size_t)it.c * S + it.s) * 2 * ch);      // ends with __syncthreads
    if (tid == 0) {
      unsigned g;
      bnp_arrive(bar_base + it.c, S, &g);
      s_gen[i % BNP_NB] = g;
    }
  };

  if (tid < 32)
    for (int k = 0; k < BNP_NB && k < my_n; ++k) issue(k);
  if (my_n > 0) phase_a(0);
  __syncthreads();                                     // s_gen[0] visible to the waiting thread
  for (int i = 0; i < my_n; ++i) {
    if (i + 1 < my_n) phase_a(i + 1);
    // thread 0 is still arriving for item i + 1 while thread 32 polls the barrier of item i (s_gen[i % NB] was written one
    // iteration ago, several __syncthreads back): the two latencies overlap
    const BnpItem it = bnp_item<T>(i, S, M, ch);
    if (tid == 32) bnp_wait(bar_base + it.c, S, s_gen[i % BNP_NB]);
    __syncthreads();
    bnr_sum_partials(partial + (size_t)it.c * S * 2 * ch, S, ch, red, sums);
    bn_scale_shift(M, ch, partial, gamma, beta, ld_p, running_mean, running_var, ld_r, active, momentum, eps, 1, save_mean,
                   save_rstd, num_batches_tracked, ld_n, coef, coef + ch, S, sums, it.c, it.s);
    __syncthreads();
    float sc[N], sh[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      sc[k] = coef[coff + k];
      sh[k] = coef[ch + coff + k];
    }
    const uint4* xs = reinterpret_cast<const uint4*>(dsm + (size_t)(i % BNP_NB) * buf_bytes);
    const uint4* rs = reinterpret_cast<const uint4*>(dsm + (size_t)(i % BNP_NB) * buf_bytes + slice_cap);
    uint4* yg = reinterpret_cast<uint4*>(y + it.goff);
    const int nvec = (int)(it.bytes / 16u);
    for (int v = tid; v < nvec; v += BN_THREADS) {
      float f[N];
      unpack16(xs[v], f, (const T*)nullptr);
      if (res) {
        float q[N];
        unpack16(rs[v], q, (const T*)nullptr);
#pragma unroll
        for (int k = 0; k < N; ++k) f[k] = f[k] * sc[k] + sh[k] + q[k];
      } else {
#pragma unroll
        for (int k = 0; k < N; ++k) f[k] = f[k] * sc[k] + sh[k];
      }
      if (relu) {
#pragma unroll
        for (int k = 0; k < N; ++k) f[k] = fmaxf(f[k], 0.f);
      }
      yg[v] = pack16(f, (const T*)nullptr);
    }
    __syncthreads();                                   // every read of buffer i % NB is done: refill it
    if (tid < 32 && i + BNP_NB < my_n) issue(i + BNP_NB);
  }
}

// backward: buffers hold dy (rewritten in place as d' = mask(dy + dy2)), x, and - when present - dy2 and the forward output y
template <typename T>
__global__ void __launch_bounds__(BN_THREADS, 1)
bn_bwd_persist_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x, const T* __restrict__ y,
                      T* __restrict__ dx, T* __restrict__ dres, int64_t M, int ch, int S, int C, uint32_t slice_cap,
                      float* partial, const float* __restrict__ gamma, const float* __restrict__ beta, int64_t ld_p,
                      const float* __restrict__ mean, const float* __restrict__ rstd, int relu, float* __restrict__ dgamma,
                      float* __restrict__ dbeta, int64_t ld_g, int bar_base) {
  constexpr int N = Vec16<T>::N;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t full[BNP_NB];
  __shared__ unsigned s_gen[BNP_NB];
  const bool has_d2 = dy2 != nullptr, has_y = relu && y != nullptr;
  const int arrays = 2 + (has_d2 ? 1 : 0) + (has_y ? 1 : 0);
  const size_t buf_bytes = (size_t)arrays * slice_cap;
  const size_t off_d2 = 2 * (size_t)slice_cap, off_y = (size_t)(2 + (has_d2 ? 1 : 0)) * slice_cap;
  float* coef = reinterpret_cast<float*>(dsm + BNP_NB * buf_bytes);
  float* sums = coef + 4 * ch;
  float* red = coef + 6 * ch;
  const int tid = threadIdx.x;
  const int n_items = C * S;
  const int my_n = ((int)blockIdx.x < n_items) ? (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (tid == 0)
    for (int k = 0; k < BNP_NB; ++k) bnr_mbar_init(full + k, 1);
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  const int coff = (tid * N) % ch;

  auto issue = [&](int i) {
    const BnpItem it = bnp_item<T>(i, S, M, ch);
    unsigned char* b = dsm + (size_t)(i % BNP_NB) * buf_bytes;
    uint64_t* bar = full + i % BNP_NB;
    if (tid == 0) bnr_mbar_expect_tx(bar, it.bytes * (uint32_t)arrays);
    __syncwarp();
    bnr_load_slice(b, dy + it.goff, it.bytes, bar);
    bnr_load_slice(b + slice_cap, x + it.goff, it.bytes, bar);
    if (has_d2) bnr_load_slice(b + off_d2, dy2 + it.goff, it.bytes, bar);
    if (has_y) bnr_load_slice(b + off_y, y + it.goff, it.bytes, bar);
  };
  auto phase_a = [&](int i) {
    const BnpItem it = bnp_item<T>(i, S, M, ch);
    unsigned char* b = dsm + (size_t)(i % BNP_NB) * buf_bytes;
    uint4* ds = reinterpret_cast<uint4*>(b);
    const uint4* xs = reinterpret_cast<const uint4*>(b + slice_cap);
    const uint4* d2s = reinterpret_cast<const uint4*>(b + off_d2);
    const uint4* ys = reinterpret_cast<const uint4*>(b + off_y);
    // ReLU mask recomputed from x when the forward output was not kept: relu input = x*g + k
    float g[N], kk[N];
    if (relu && !has_y) {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        const float mu = mean[(size_t)it.c * ch + coff + k], rs = rstd[(size_t)it.c * ch + coff + k];
        g[k] = gamma[(size_t)it.c * ld_p + coff + k] * rs;
        kk[k] = (beta ? beta[(size_t)it.c * ld_p + coff + k] : 0.f) - mu * g[k];
      }
    }
    bnr_mbar_wait(full + i % BNP_NB, (uint32_t)(i / BNP_NB) & 1u);
    const int nvec = (int)(it.bytes / 16u);
    float a[N], bb[N];
#pragma unroll
    for (int k = 0; k < N; ++k) a[k] = bb[k] = 0.f;
    for (int v = tid; v < nvec; v += BN_THREADS) {
      float d[N], xv[N];
      unpack16(ds[v], d, (const T*)nullptr);
      unpack16(xs[v], xv, (const T*)nullptr);
      if (has_d2) {
        float q[N];
        unpack16(d2s[v], q, (const T*)nullptr);
#pragma unroll
        for (int k = 0; k < N; ++k) d[k] += q[k];
      }
      if (relu) {
        if (has_y) {
          float q[N];
          unpack16(ys[v], q, (const T*)nullptr);
#pragma unroll
          for (int k = 0; k < N; ++k) d[k] = q[k] > 0.f ? d[k] : 0.f;
        } else {
#pragma unroll
          for (int k = 0; k < N; ++k) d[k] = (xv[k] * g[k] + kk[k]) > 0.f ? d[k] : 0.f;
        }
      }
#pragma unroll
      for (int k = 0; k < N; ++k) {
        a[k] += d[k];
        bb[k] += d[k] * xv[k];
      }
      if (has_d2 || relu) ds[v] = pack16(d, (const T*)nullptr);
    }
    bnr_reduce<N>(a, bb, ch, red, partial + ((size_t)it.c * S + it.s) * 2 * ch);
    if (tid == 0) {
      unsigned gen;
      bnp_arrive(bar_base + it.c, S, &gen);
      s_gen[i % BNP_NB] = gen;
    }
  };

  if (tid < 32)
    for (int k = 0; k < BNP_NB && k < my_n; ++k) issue(k);
  if (my_n > 0) phase_a(0);
  __syncthreads();
  for (int i = 0; i < my_n; ++i) {
    if (i + 1 < my_n) phase_a(i + 1);
    const BnpItem it = bnp_item<T>(i, S, M, ch);
    if (tid == 32) bnp_wait(bar_base + it.c, S, s_gen[i % BNP_NB]);
    __syncthreads();
    bnr_sum_partials(partial + (size_t)it.c * S * 2 * ch, S, ch, red, sums);
    bn_bwd_coeffs(M, ch, S, partial, gamma, beta, ld_p, mean, rstd, dgamma, dbeta, ld_g, coef, sums, it.c, it.s);
    __syncthreads();
    float A[N], B[N], Cc[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      A[k] = coef[coff + k];
      B[k] = coef[ch + coff + k];
      Cc[k] = coef[2 * ch + coff + k];
    }
    const unsigned char* b = dsm + (size_t)(i % BNP_NB) * buf_bytes;
    const uint4* ds = reinterpret_cast<const uint4*>(b);
    const uint4* xs = reinterpret_cast<const uint4*>(b + slice_cap);
    uint4* dxg = reinterpret_cast<uint4*>(dx + it.goff);
    uint4* drg = dres ? reinterpret_cast<uint4*>(dres + it.goff) : nullptr;
    const int nvec = (int)(it.bytes / 16u);
    for (int v = tid; v < nvec; v += BN_THREADS) {
      const uint4 dr = ds[v];
      float d[N], xv[N], o[N];
      unpack16(dr, d, (const T*)nullptr);
      unpack16(xs[v], xv, (const T*)nullptr);
#pragma unroll
      for (int k = 0; k < N; ++k) o[k] = d[k] * A[k] + xv[k] * B[k] + Cc[k];     // = g*(d - dbeta/M - xhat*dgamma/M)
      dxg[v] = pack16(o, (const T*)nullptr);
      if (drg) drg[v] = dr;
    }
    __syncthreads();
    if (tid < 32 && i + BNP_NB < my_n) issue(i + BNP_NB);
  }
}

// ---- host side
// FB_BN_PERSIST=1 selects these kernels.  Off by default: with NB x arrays slices sharing one SM's shared memory the slices
// are small (S = 53-105 per client on the fp32 stage-1 layers) and every item pays the S x 2*ch partial-sum read after its
// barrier; measured 394 vs 200 us (fp32 64ch@32x32 forward with residual), round 129.2 vs 105.7 ms
// (profiles/r2/bn_resident.md).  Kept as the tested starting point for a version with a hierarchical partial reduction.
static int bn_persist_mode() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FB_BN_PERSIST");
    v = e ? (atoi(e) != 0) : 0;
  }
  return v;
}
// S (slices per client, <= BN_MAX_SLICES), slice capacity and dynamic shared memory for NB buffers of `arrays` slices
template <typename T>
static bool bn_persist_plan(int32_t C, int64_t M, int ch, int arrays, int* S_out, uint32_t* cap_out, size_t* smem_out, int* grid_out) {
  constexpr int N = Vec16<T>::N;
  if (!bn_persist_mode() || (BN_THREADS * N) % ch != 0) return false;
  const int64_t row_bytes = (int64_t)ch * (int64_t)sizeof(T);
  const int64_t extra = (int64_t)(6 * ch + BN_THREADS * 2 * N) * 4;
  static int budget_kb = 0;
  if (!budget_kb) {
    const char* e = getenv("FB_BN_PERSIST_KB");
    budget_kb = e ? atoi(e) : 216;
    if (budget_kb < 32) budget_kb = 32;
    if (budget_kb > 222) budget_kb = 222;
  }
  const int64_t budget = (int64_t)budget_kb * 1024 - extra;
  const int64_t rows_fit = budget / ((int64_t)BNP_NB * arrays * row_bytes);
  if (rows_fit < 1) return false;
  int64_t S = (M + rows_fit - 1) / rows_fit;
  if (S > BN_MAX_SLICES || S > fb_num_sms()) return false;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t cap = (rows_per * row_bytes + 127) / 128 * 128;
  *S_out = (int)S;
  *cap_out = (uint32_t)cap;
  *smem_out = (size_t)(BNP_NB * arrays * cap + extra);
  const int64_t items = (int64_t)C * S;
  *grid_out = (int)(items < fb_num_sms() ? items : fb_num_sms());
  return *smem_out <= (size_t)223 * 1024 && (int64_t)C * S < (int64_t)1 << 30;
}
