// Slab-resident BatchNorm (included by conv.cu after the fused kernels).
//
// The fused kernels read every activation twice (statistics, then apply) and count on L2 for the second read; with 100
// clients in flight the fp32 stage-1 tensors of the bottleneck ResNet (262 MB each) do not fit and the second read comes
// from HBM (ncu: 524 MB read for a 262 MB input).  Here a CTA keeps its slice of the client's slab in shared memory:
//   forward : x -> smem (1-D bulk TMA), statistics from smem, per-client barrier, y computed in place, bulk store
//   backward: dy, x -> smem, d' = mask(dy + dy2) written back in place while the sums are taken, barrier,
//             dx computed in place over x, dx (and d' as the residual gradient) bulk-stored
// so every tensor crosses HBM once.  Slices are cut so that two CTAs fit on an SM (one loads while the other computes).
// Elements move as 16-byte vectors (8 bf16 / 4 fp32): conflict-free in shared memory.
#pragma once

#define BNR_CHUNK 16384u       // bytes per bulk copy
#define BNR_MAX_CHUNKS 8        // chunks (and mbarriers) a slice is cut into by the chunk-pipelined forward kernel

__device__ __forceinline__ uint32_t bnr_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bnr_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bnr_smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bnr_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bnr_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bnr_mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = bnr_smem_u32(bar);
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
// global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bnr_bulk_load(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   bnr_smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(bnr_smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bnr_bulk_store(void* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(bnr_smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
// the first warp moves `bytes` (a multiple of 16) in BNR_CHUNK pieces, one piece per lane and round
__device__ __forceinline__ void bnr_load_slice(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  const uint32_t lane = threadIdx.x;
  for (uint32_t off = lane * BNR_CHUNK; off < bytes; off += 32u * BNR_CHUNK)
    bnr_bulk_load((char*)dst_smem + off, (const char*)src + off, min(BNR_CHUNK, bytes - off), bar);
}
__device__ __forceinline__ void bnr_store_slice(void* dst, const void* src_smem, uint32_t bytes) {
  const uint32_t lane = threadIdx.x;
  for (uint32_t off = lane * BNR_CHUNK; off < bytes; off += 32u * BNR_CHUNK)
    bnr_bulk_store((char*)dst + off, (const char*)src_smem + off, min(BNR_CHUNK, bytes - off));
}

// Optional per-CTA timeline (tools/trace_bn.py): when a trace buffer is registered, thread 0 of every forward CTA stores
// globaltimer stamps [start, loaded, stats, barrier, coeffs, applied, stored, smid] at row blockIdx.y * gridDim.x + blockIdx.x.
static __device__ unsigned long long* g_bn_trace = nullptr;
// FB_BN_CHUNKS: chunks per slice.  ResNet-56 tf32 round with 1 / 2 / 4 / 8 chunks: 99.6 / 99.1 / 99.8 / 101.0 ms
static __device__ unsigned g_bn_chunks = 2;
__device__ __forceinline__ void bnr_stamp(int slot) {
  if (g_bn_trace && threadIdx.x == 0) {
    unsigned long long* row = g_bn_trace + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 8;
    if (slot == 7) {
      unsigned sm;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
      row[7] = sm;
    } else {
      row[slot] = fb_globaltimer();
    }
  }
}

template <typename T> struct Vec16 { static constexpr int N = 16 / (int)sizeof(T); };
__device__ __forceinline__ void unpack16(const uint4& r, float* f, const __nv_bfloat16*) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ void unpack16(const uint4& r, float* f, const float*) {
  f[0] = __uint_as_float(r.x); f[1] = __uint_as_float(r.y); f[2] = __uint_as_float(r.z); f[3] = __uint_as_float(r.w);
}
__device__ __forceinline__ uint4 pack16(const float* f, const __nv_bfloat16*) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
__device__ __forceinline__ uint4 pack16(const float* f, const float*) {
  return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
}

// per-channel sums over the threads that share a channel group: a[N], b[N] per thread -> out[2][ch].  red: BN_THREADS*2*N floats
template <int N>
__device__ __forceinline__ void bnr_reduce(const float* a, const float* b, int ch, float* red, float* out) {
  const int tid = threadIdx.x;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    red[tid * 2 * N + i] = a[i];
    red[tid * 2 * N + N + i] = b[i];
  }
  __syncthreads();
  const int ng = ch / N;
  for (int t = tid; t < 2 * ch; t += BN_THREADS) {
    const int q = t / ch, c = t % ch;
    const int grp = c / N, comp = c % N;
    float s = 0.f;
    for (int tt = grp; tt < BN_THREADS; tt += ng) s += red[tt * 2 * N + q * N + comp];
    out[q * ch + c] = s;
  }
  __syncthreads();
}

// Barrier among the S CTAs of a client, split into arrive / wait, one global round trip each on the critical path
// (the generation barrier of the fused kernels costs ~4.9 us per CTA in tools/trace_bn.py: fence, generation load, atomic
// with return, polling with sleep, fence).  {count, seen} per client: a CTA arrives with a release reduction (no return
// value), waits until count >= S, and when it is done with the barrier adds itself to `seen`; the CTA that makes
// seen == S resets both words, so the state is zero again when the kernel ends (the next launch is stream-ordered).
static __device__ unsigned g_bn_barrier2[2 * 65536];
__device__ __forceinline__ void client_arrive(int S, int idx) {          // one thread, after the CTA's partials are written
  if (S > 1) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(g_bn_barrier2 + 2 * idx) : "memory");
}
__device__ __forceinline__ void client_wait(int S, int idx) {            // one thread
  if (S > 1) {
    const unsigned* st = g_bn_barrier2 + 2 * idx;
    while (ld_acquire_gpu(st) < (unsigned)S) {
    }
  }
}
__device__ __forceinline__ void client_leave(int S, int idx) {           // one thread, any time after its client_wait
  if (S > 1) {
    unsigned* st = g_bn_barrier2 + 2 * idx;
    if (atomicAdd(st + 1, 1u) == (unsigned)S - 1) {
      st[0] = 0;
      st[1] = 0;
    }
  }
}

// The client's S partial rows [S][2*ch] (L2, written by the sibling CTAs) -> complete sums [2*ch] in shared memory, with
// every thread loading: a column is shared by BN_THREADS / (2*ch) threads that take every other row, eight loads in flight
// each (the per-channel loop of the fused kernels walks S rows one after the other: 55 dependent L2 round trips at S = 55).
// red: BN_THREADS floats.  Deterministic: fixed row order per thread, fixed order across threads.
__device__ __forceinline__ void bnr_sum_partials(const float* part, int S, int ch, float* red, float* sums) {
  // rows are taken sixteen at a time with the tail masked, so S <= 16 * groups rows cost ONE round trip to L2 (the tail
  // loops of the first version made it 3-4 rounds at S = 27 and 6 at S = 54); the summation order per thread is fixed
  const int tid = threadIdx.x, cols = 2 * ch;
  if (cols >= BN_THREADS) {
    for (int col = tid; col < cols; col += BN_THREADS) {
      float acc = 0.f;
      for (int j = 0; j < S; j += 16) {
        float v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) v[u] = (j + u < S) ? __ldcg(part + (size_t)(j + u) * cols + col) : 0.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) acc += v[u];
      }
      sums[col] = acc;
    }
  } else {
    const int groups = BN_THREADS / cols, col = tid % cols, grp = tid / cols;
    float acc = 0.f;
    for (int j = grp; j < S; j += 16 * groups) {
      float v[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) v[u] = (j + u * groups < S) ? __ldcg(part + (size_t)(j + u * groups) * cols + col) : 0.f;
#pragma unroll
      for (int u = 0; u < 16; ++u) acc += v[u];
    }
    red[tid] = acc;
    __syncthreads();
    if (tid < cols) {
      float t = 0.f;
      for (int g2 = 0; g2 < groups; ++g2) t += red[g2 * cols + tid];
      sums[tid] = t;
    }
  }
  __syncthreads();
}

// dynamic shared memory: [slices ...][coefficients 4*ch floats][complete sums 2*ch floats][reduction BN_THREADS*2*N floats]
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_fwd_resident_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, int64_t M, int ch, int S,
                       uint32_t slice_cap, float* partial, const float* __restrict__ gamma, const float* __restrict__ beta,
                       int64_t ld_p, float* running_mean, float* running_var, int64_t ld_r, const float* __restrict__ active,
                       float momentum, float eps, int relu, float* __restrict__ save_mean, float* __restrict__ save_rstd,
                       float* num_batches_tracked, int64_t ld_n, int res_smem /* the residual slice is resident too */, int bar_base) {
  constexpr int N = Vec16<T>::N;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t bar[BNR_MAX_CHUNKS];
  uint4* xs = reinterpret_cast<uint4*>(dsm);
  uint4* rs = reinterpret_cast<uint4*>(dsm + slice_cap);
  float* coef = reinterpret_cast<float*>(dsm + (size_t)(1 + res_smem) * slice_cap);
  float* red = coef + 6 * ch;
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = min(M, (int64_t)s * rows_per), r1 = min(M, r0 + rows_per);
  const uint32_t bytes = (uint32_t)((r1 - r0) * ch * (int64_t)sizeof(T));
  const size_t goff = ((size_t)c * M + r0) * ch;
  // the slice moves in <= BNR_MAX_CHUNKS chunks of a multiple of 4 KiB (256 threads x 16 bytes, so a thread keeps its
  // channels from chunk to chunk): statistics start on chunk 0 while the rest is still in flight, and chunk i of the
  // output is stored while chunk i + 1 is being computed
  const uint32_t want_chunks = g_bn_chunks;
  const uint32_t chunk = ((bytes + want_chunks - 1) / want_chunks + 4095u) / 4096u * 4096u;
  const int nchunks = bytes ? (int)((bytes + chunk - 1) / chunk) : 0;
  if (tid < BNR_MAX_CHUNKS) bnr_mbar_init(bar + tid, 1);
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  bnr_stamp(0);
  bnr_stamp(7);
  if (tid < nchunks) {
    const uint32_t off = (uint32_t)tid * chunk, nb = min(chunk, bytes - off);
    bnr_mbar_expect_tx(bar + tid, nb * (uint32_t)(1 + res_smem));
    bnr_bulk_load((char*)xs + off, (const char*)(x + goff) + off, nb, bar + tid);
    if (res_smem) bnr_bulk_load((char*)rs + off, (const char*)(res + goff) + off, nb, bar + tid);
  }
  const int nvec = (int)(bytes / 16u), cvec = (int)(chunk / 16u);
  const int coff = (tid * N) % ch;
  float a[N], b[N];
#pragma unroll
  for (int i = 0; i < N; ++i) a[i] = b[i] = 0.f;
  for (int k = 0; k < nchunks; ++k) {
    bnr_mbar_wait(bar + k, 0);
    const int v1 = min(nvec, (k + 1) * cvec);
    for (int v = k * cvec + tid; v < v1; v += BN_THREADS) {
      float f[N];
      unpack16(xs[v], f, (const T*)nullptr);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        a[i] += f[i];
        b[i] += f[i] * f[i];
      }
    }
  }
  bnr_stamp(1);
  bnr_reduce<N>(a, b, ch, red, partial + ((size_t)c * S + s) * 2 * ch);
  bnr_stamp(2);
  if (tid == 0) client_arrive(S, bar_base + c);
  // nothing below depends on the barrier until the partial sums are read: fetch the per-channel inputs of the
  // coefficients and the first residual vectors while the siblings arrive
  const bool one_k = ch <= BN_THREADS, mine = one_k && tid < ch;
  float pg = 0.f, pb = 0.f, prm = 0.f, prv = 0.f, pact = 1.f;
  if (mine) {
    pg = gamma[(size_t)c * ld_p + tid];
    pb = beta[(size_t)c * ld_p + tid];
    pact = active ? active[c] : 1.f;
    if (s == 0) {
      prm = running_mean[(size_t)c * ld_r + tid];
      prv = running_var[(size_t)c * ld_r + tid];
    }
  }
  const bool res_g = res && !res_smem;
  const uint4* rg = reinterpret_cast<const uint4*>(res + goff);
  constexpr int RU = 8;                       // residual vectors per thread and round: two rounds' worth (2 x 8 x 16 B per thread) in flight
  uint4 rq[RU];
#pragma unroll
  for (int u = 0; u < RU; ++u) rq[u] = make_uint4(0, 0, 0, 0);
  if (res_g) {
#pragma unroll
    for (int u = 0; u < RU; ++u) {
      const int v = tid + u * BN_THREADS;
      if (v < min(nvec, cvec) && u * BN_THREADS < cvec) rq[u] = __ldg(rg + v);
    }
  }
  if (tid == 0) client_wait(S, bar_base + c);
  __syncthreads();
  bnr_stamp(3);
  bnr_sum_partials(partial + (size_t)c * S * 2 * ch, S, ch, red, coef + 2 * ch);
  if (one_k) {
    if (mine) {
      const float su = coef[2 * ch + tid], sq = coef[3 * ch + tid];
      const float mean = su / (float)M, var = fmaxf(sq / (float)M - mean * mean, 0.f);
      const float rstd = rsqrtf(var + eps);
      if (s == 0) {
        if (pact != 0.f) {
          const float unbiased = var * ((float)M / (float)max((int64_t)1, M - 1));
          running_mean[(size_t)c * ld_r + tid] = prm + momentum * pact * (mean - prm);
          running_var[(size_t)c * ld_r + tid] = prv + momentum * pact * (unbiased - prv);
        }
        if (save_mean) {
          save_mean[(size_t)c * ch + tid] = mean;
          save_rstd[(size_t)c * ch + tid] = rstd;
        }
        if (tid == 0 && num_batches_tracked) num_batches_tracked[(size_t)c * ld_n] += pact;
      }
      const float g = pg * rstd;
      coef[tid] = g;
      coef[ch + tid] = pb - mean * g;
    }
  } else {
    bn_scale_shift(M, ch, partial, gamma, beta, ld_p, running_mean, running_var, ld_r, active, momentum, eps, 1, save_mean,
                   save_rstd, num_batches_tracked, ld_n, coef, coef + ch, S, coef + 2 * ch);
  }
  __syncthreads();
  bnr_stamp(4);
  float sc[N], sh[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    sc[i] = coef[coff + i];
    sh[i] = coef[ch + coff + i];
  }
  for (int k = 0; k < nchunks; ++k) {
    const int v0 = k * cvec, v1 = min(nvec, v0 + cvec);
    for (int vb = v0; vb < v1; vb += RU * BN_THREADS) {        // one round = RU vectors per thread
      uint4 rn[RU];
#pragma unroll
      for (int u = 0; u < RU; ++u) rn[u] = make_uint4(0, 0, 0, 0);
      if (res_g) {                                               // residual of the next round: in flight during this one
        int nb = vb + RU * BN_THREADS, ne = v1;
        if (nb >= v1) {
          nb = v1;                                               // first round of the next chunk
          ne = min(nvec, v1 + cvec);
        }
#pragma unroll
        for (int u = 0; u < RU; ++u) {
          const int v = nb + tid + u * BN_THREADS;
          if (v < ne) rn[u] = __ldg(rg + v);
        }
      }
#pragma unroll
      for (int u = 0; u < RU; ++u) {
        const int v = vb + tid + u * BN_THREADS;
        if (v < v1) {
          float f[N];
          unpack16(xs[v], f, (const T*)nullptr);
          if (res) {
            float q[N];
            unpack16(res_smem ? rs[v] : rq[u], q, (const T*)nullptr);
#pragma unroll
            for (int i = 0; i < N; ++i) f[i] = f[i] * sc[i] + sh[i] + q[i];
          } else {
#pragma unroll
            for (int i = 0; i < N; ++i) f[i] = f[i] * sc[i] + sh[i];
          }
          if (relu) {
#pragma unroll
            for (int i = 0; i < N; ++i) f[i] = fmaxf(f[i], 0.f);
          }
          xs[v] = pack16(f, (const T*)nullptr);
        }
      }
#pragma unroll
      for (int u = 0; u < RU; ++u) rq[u] = rn[u];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes visible to the bulk store
    __syncthreads();
    if (tid == 0) {
      const uint32_t off = (uint32_t)k * chunk;
      bnr_bulk_store((char*)(y + goff) + off, (const char*)xs + off, min(chunk, bytes - off));
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  bnr_stamp(5);
  if (tid == 0) {
    client_leave(S, bar_base + c);
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");    // shared memory must outlive the reads of the bulk stores
    bnr_stamp(6);
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_resident_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x, const T* __restrict__ y,
                       T* __restrict__ dx, T* __restrict__ dres, int64_t M, int ch, int S, uint32_t slice_cap, float* partial,
                       const float* __restrict__ gamma, const float* __restrict__ beta, int64_t ld_p,
                       const float* __restrict__ mean, const float* __restrict__ rstd, int relu, float* __restrict__ dgamma,
                       float* __restrict__ dbeta, int64_t ld_g, int extras /* bit 0: dy2 resident, bit 1: y resident */, int bar_base) {
  constexpr int N = Vec16<T>::N;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t bar;
  const int e_d2 = extras & 1, e_y = (extras >> 1) & 1;
  uint4* ds = reinterpret_cast<uint4*>(dsm);                  // dy, then d' = mask(dy + dy2)
  uint4* xs = reinterpret_cast<uint4*>(dsm + slice_cap);      // x, then dx
  uint4* d2s = reinterpret_cast<uint4*>(dsm + 2 * (size_t)slice_cap);
  uint4* ys = reinterpret_cast<uint4*>(dsm + (size_t)(2 + e_d2) * slice_cap);
  float* coef = reinterpret_cast<float*>(dsm + (size_t)(2 + e_d2 + e_y) * slice_cap);
  float* red = coef + 6 * ch;
  const int c = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const int64_t rows_per = (M + S - 1) / S;
  const int64_t r0 = min(M, (int64_t)s * rows_per), r1 = min(M, r0 + rows_per);
  const uint32_t bytes = (uint32_t)((r1 - r0) * ch * (int64_t)sizeof(T));
  const size_t goff = ((size_t)c * M + r0) * ch;
  if (tid == 0) bnr_mbar_init(&bar, 1);
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  if (tid < 32) {
    if (tid == 0) bnr_mbar_expect_tx(&bar, (uint32_t)(2 + e_d2 + e_y) * bytes);
    __syncwarp();
    bnr_load_slice(ds, dy + goff, bytes, &bar);
    bnr_load_slice(xs, x + goff, bytes, &bar);
    if (e_d2) bnr_load_slice(d2s, dy2 + goff, bytes, &bar);
    if (e_y) bnr_load_slice(ys, y + goff, bytes, &bar);
  }
  const int nvec = (int)(bytes / 16u);
  const int coff = (tid * N) % ch;
  // ReLU mask recomputed from x when the forward output was not kept: relu input = x*g + k
  float g[N], k[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float mu = mean[(size_t)c * ch + coff + i], rs = rstd[(size_t)c * ch + coff + i];
    g[i] = gamma[(size_t)c * ld_p + coff + i] * rs;
    k[i] = (beta ? beta[(size_t)c * ld_p + coff + i] : 0.f) - mu * g[i];
  }
  const uint4* d2g = dy2 ? reinterpret_cast<const uint4*>(dy2 + goff) : nullptr;
  const uint4* yg = (relu && y) ? reinterpret_cast<const uint4*>(y + goff) : nullptr;
  float a[N], b[N];
#pragma unroll
  for (int i = 0; i < N; ++i) a[i] = b[i] = 0.f;
  bnr_mbar_wait(&bar, 0);
  for (int v = tid; v < nvec; v += 4 * BN_THREADS) {
    uint4 e[4], m[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const bool in = v + u * BN_THREADS < nvec;
      e[u] = (d2g && in) ? (e_d2 ? d2s[v + u * BN_THREADS] : __ldg(d2g + v + u * BN_THREADS)) : make_uint4(0, 0, 0, 0);
      m[u] = (yg && in) ? (e_y ? ys[v + u * BN_THREADS] : __ldg(yg + v + u * BN_THREADS)) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int vv = v + u * BN_THREADS;
      if (vv >= nvec) break;
      float d[N], xv[N];
      unpack16(ds[vv], d, (const T*)nullptr);
      unpack16(xs[vv], xv, (const T*)nullptr);
      if (d2g) {
        float q[N];
        unpack16(e[u], q, (const T*)nullptr);
#pragma unroll
        for (int i = 0; i < N; ++i) d[i] += q[i];
      }
      if (relu) {
        if (yg) {
          float q[N];
          unpack16(m[u], q, (const T*)nullptr);
#pragma unroll
          for (int i = 0; i < N; ++i) d[i] = q[i] > 0.f ? d[i] : 0.f;
        } else {
#pragma unroll
          for (int i = 0; i < N; ++i) d[i] = (xv[i] * g[i] + k[i]) > 0.f ? d[i] : 0.f;
        }
      }
#pragma unroll
      for (int i = 0; i < N; ++i) {
        a[i] += d[i];
        b[i] += d[i] * xv[i];
      }
      if (d2g || relu) ds[vv] = pack16(d, (const T*)nullptr);
    }
  }
  bnr_reduce<N>(a, b, ch, red, partial + ((size_t)c * S + s) * 2 * ch);
  if (tid == 0) client_arrive(S, bar_base + c);
  // per-channel inputs of the coefficients: fetched while the siblings arrive
  const bool one_k = ch <= BN_THREADS, mine = one_k && tid < ch;
  float pg = 0.f, pb = 0.f, pmu = 0.f, prs = 0.f;
  if (mine) {
    pg = gamma[(size_t)c * ld_p + tid];
    pb = beta ? beta[(size_t)c * ld_p + tid] : 0.f;
    pmu = mean[(size_t)c * ch + tid];
    prs = rstd[(size_t)c * ch + tid];
  }
  if (tid == 0) client_wait(S, bar_base + c);
  __syncthreads();
  bnr_sum_partials(partial + (size_t)c * S * 2 * ch, S, ch, red, coef + 4 * ch);
  if (one_k) {
    if (mine) {
      // A = g, B = -g*rstd*dgamma/M, Cc = -g*dbeta/M - mean*B, K = beta - mean*g (bn_bwd_coeffs)
      const float sd = coef[4 * ch + tid], sdx = coef[5 * ch + tid];
      const float dbt = sd, dgm = prs * (sdx - pmu * sd);
      if (s == 0) {
        dbeta[(size_t)c * ld_g + tid] = dbt;
        dgamma[(size_t)c * ld_g + tid] = dgm;
      }
      const float gg = pg * prs;
      const float Bk = -gg * prs * dgm / (float)M;
      coef[tid] = gg;
      coef[ch + tid] = Bk;
      coef[2 * ch + tid] = -gg * dbt / (float)M - pmu * Bk;
      coef[3 * ch + tid] = pb - pmu * gg;
    }
  } else {
    bn_bwd_coeffs(M, ch, S, partial, gamma, beta, ld_p, mean, rstd, dgamma, dbeta, ld_g, coef, coef + 4 * ch);
  }
  __syncthreads();
  float A[N], B[N], Cc[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    A[i] = coef[coff + i];
    B[i] = coef[ch + coff + i];
    Cc[i] = coef[2 * ch + coff + i];
  }
  for (int v = tid; v < nvec; v += BN_THREADS) {
    float d[N], xv[N], o[N];
    unpack16(ds[v], d, (const T*)nullptr);
    unpack16(xs[v], xv, (const T*)nullptr);
#pragma unroll
    for (int i = 0; i < N; ++i) o[i] = d[i] * A[i] + xv[i] * B[i] + Cc[i];     // = g*(d - dbeta/M - xhat*dgamma/M)
    xs[v] = pack16(o, (const T*)nullptr);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (tid < 32) {
    bnr_store_slice(dx + goff, xs, bytes);
    if (dres) bnr_store_slice(dres + goff, ds, bytes);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (tid == 0) client_leave(S, bar_base + c);
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");    // shared memory must outlive the reads of the bulk stores
  }
}

// ---- host side: slice count and shared-memory budget
// FB_BN_RESIDENT=0 turns the resident kernels off (fused kernels instead); default: whenever the slab fits (first version:
// tf32 round 114.0 -> 106.9 ms, bf16 backward only; with the arrive/wait barrier and two bf16 CTAs per SM the bf16 forward
// pass wins as well: 64.28 -> 63.16 ms; profiles/r2/bn_resident.md)
static int bn_resident_mode() {
  static int v = -2;
  if (v == -2) {
    const char* e = getenv("FB_BN_RESIDENT");
    v = e ? (atoi(e) != 0) : -1;
  }
  return v;
}
static int bn_resident_extras() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FB_BN_RES_EXTRAS");
    v = e ? (atoi(e) != 0) : 0;
  }
  return v;
}
static int bn_resident_budget_kb(int bwd) {     // bytes of slices per CTA; the default leaves room for two CTAs per SM
  static int kb[2] = {0, 0};
  if (!kb[bwd]) {
    const char* e = getenv(bwd ? "FB_BN_RES_KB_BWD" : "FB_BN_RES_KB_FWD");
    kb[bwd] = e ? atoi(e) : 96;
    if (kb[bwd] < 8) kb[bwd] = 8;
    if (kb[bwd] > 200) kb[bwd] = 200;
  }
  return kb[bwd];
}
// Picks S (slices per client, <= 64: the partial-sum scratch holds 64) and the per-array slice capacity for `arrays`
// resident arrays; returns false when a client's slab does not fit (the fused / two-pass kernels take over).
template <typename T>
static bool bn_resident_plan(int64_t M, int ch, int arrays, int bwd, bool has_res, int* S_out, uint32_t* cap_out,
                             size_t* smem_out) {
  constexpr int N = Vec16<T>::N;
  const int mode = bn_resident_mode();
  if (!mode || (BN_THREADS * N) % ch != 0) return false;
  (void)has_res;
  (void)bwd;
  const int64_t row_bytes = (int64_t)ch * (int64_t)sizeof(T);
  const int64_t extra = (int64_t)(6 * ch + BN_THREADS * 2 * N) * 4;
  // FB_BN_RES_TOTAL_KB (default 110): slices + scratch together stay below this many KB per CTA, so that two CTAs share an SM
  // whatever the scratch size (bf16: 17.9 KB); 0 = the slice budget alone.  bf16 round 64.28 -> 64.03 ms, with the bf16
  // forward pass resident as well 63.16 ms
  static int total_kb = -1;
  if (total_kb < 0) {
    const char* e = getenv("FB_BN_RES_TOTAL_KB");
    total_kb = e ? atoi(e) : 110;
  }
  for (int pass = 0; pass < 2; ++pass) {
    int64_t budget = (pass == 0 ? (int64_t)bn_resident_budget_kb(bwd) : 200) * 1024 - (pass == 0 ? 0 : extra);
    // only when two CTAs would otherwise not fit on an SM (fp32 keeps its 96 KB slices: 98.24 vs 98.91 ms per round)
    if (pass == 0 && total_kb > 0 && budget + extra > (int64_t)113 * 1024) budget = (int64_t)total_kb * 1024 - extra;
    const int64_t rows_fit = budget / (arrays * row_bytes);
    if (rows_fit < 1) continue;
    int64_t S = (M + rows_fit - 1) / rows_fit;
    if (S > 64) continue;
    const int64_t rows_per = (M + S - 1) / S;
    const int64_t cap = (rows_per * row_bytes + 127) / 128 * 128;
    *S_out = (int)S;
    *cap_out = (uint32_t)cap;
    *smem_out = (size_t)(arrays * cap + extra);
    return *smem_out <= (size_t)220 * 1024;
  }
  return false;
}
template <typename K>
static int bn_resident_allow_smem(K kernel, size_t bytes) {
  // once per kernel instantiation and device: raise the dynamic shared-memory limit to the hardware maximum
  static bool done[64] = {};
  int dev = 0;
  FB_CHECK_CUDA(cudaGetDevice(&dev));
  (void)bytes;
  if (dev < 64 && !done[dev]) {
    FB_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    done[dev] = true;
  }
  return FB_OK;
}
