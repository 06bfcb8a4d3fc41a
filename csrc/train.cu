// K11 / K15: fused multi-client optimizer step, softmax cross-entropy fwd+bwd, fused evaluation counters.
// See include/fedml_b200.h.
#include "common.cuh"
#include "../include/fedml_b200.h"

static inline int ew_grid(int64_t nvec) {
  int64_t want = (nvec + 255) / 256, cap = (int64_t)fb_num_sms() * 8;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

// ---------------------------------------------------------------- stacked SGD (+wd, momentum, FedProx, FedNova)
__global__ void __launch_bounds__(256)
sgd_step_stacked_kernel(float* __restrict__ w, const float* __restrict__ g, int64_t ld, int C, int64_t nvec_row,
                        const float* __restrict__ lr_dev, float wd, float momentum, float dampening, int nesterov,
                        float* __restrict__ mom, int first_step, float mu, const float* __restrict__ anchor,
                        float* __restrict__ cum) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int64_t total = nvec_row * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / nvec_row);
    const int64_t j = (i - (int64_t)c * nvec_row) * 4;
    const float lr = __ldg(lr_dev + c);
    if (lr == 0.f) continue;  // frozen client (finished its local steps)
    const size_t off = (size_t)c * ld + j;
    float4 ww = *reinterpret_cast<const float4*>(w + off);
    float4 d = ld_stream4(g + off);
    if (wd != 0.f) { d.x += wd * ww.x; d.y += wd * ww.y; d.z += wd * ww.z; d.w += wd * ww.w; }
    if (momentum != 0.f) {
      float4 b;
      if (first_step) {
        b = d;
      } else {
        b = *reinterpret_cast<const float4*>(mom + off);
        const float k = 1.f - dampening;
        b.x = momentum * b.x + k * d.x; b.y = momentum * b.y + k * d.y;
        b.z = momentum * b.z + k * d.z; b.w = momentum * b.w + k * d.w;
      }
      *reinterpret_cast<float4*>(mom + off) = b;
      if (nesterov) { d.x += momentum * b.x; d.y += momentum * b.y; d.z += momentum * b.z; d.w += momentum * b.w; }
      else d = b;
    }
    if (mu != 0.f) {
      const float4 a = *reinterpret_cast<const float4*>(anchor + j);
      d.x += mu * (ww.x - a.x); d.y += mu * (ww.y - a.y); d.z += mu * (ww.z - a.z); d.w += mu * (ww.w - a.w);
    }
    if (cum) {
      float4 cg = *reinterpret_cast<const float4*>(cum + off);
      cg.x += lr * d.x; cg.y += lr * d.y; cg.z += lr * d.z; cg.w += lr * d.w;
      *reinterpret_cast<float4*>(cum + off) = cg;
    }
    ww.x -= lr * d.x; ww.y -= lr * d.y; ww.z -= lr * d.z; ww.w -= lr * d.w;
    *reinterpret_cast<float4*>(w + off) = ww;
  }
}

extern "C" int fb_sgd_step_stacked(float* w, const float* g, int64_t ld, int32_t C, int64_t n, const float* lr_dev,
                                   float weight_decay, float momentum, float dampening, int32_t nesterov,
                                   float* mom_buf, int32_t first_step, float mu, const float* anchor, float* cum_grad,
                                   fb_stream_t stream) {
  FB_REQUIRE(w && g && lr_dev && C >= 1 && n > 0 && n % 4 == 0 && ld % 4 == 0 && ld >= n, "fb_sgd_step_stacked: bad shape");
  FB_REQUIRE(momentum == 0.f || mom_buf, "fb_sgd_step_stacked: momentum needs mom_buf");
  FB_REQUIRE(mu == 0.f || anchor, "fb_sgd_step_stacked: mu needs the anchor (global) model");
  FB_CHECK_CUDA(fb_launch_pdl(sgd_step_stacked_kernel, dim3(ew_grid(n / 4 * C)), dim3(256), (size_t)(0), (cudaStream_t)stream, 
      w, g, ld, C, n / 4, lr_dev, weight_decay, momentum, dampening, nesterov, mom_buf, first_step, mu, anchor, cum_grad));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

__global__ void __launch_bounds__(256)
sgd_step_stacked_bf16g_kernel(float* __restrict__ w, const __nv_bfloat16* __restrict__ g, int64_t ld, int64_t ld_g,
                              int C, int64_t nvec_row, const float* __restrict__ lr_dev, float wd, float mu,
                              const float* __restrict__ anchor) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int64_t total = nvec_row * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / nvec_row);
    const int64_t j = (i - (int64_t)c * nvec_row) * 4;
    const float lr = __ldg(lr_dev + c);
    if (lr == 0.f) continue;
    const size_t off = (size_t)c * ld + j;
    float4 ww = *reinterpret_cast<const float4*>(w + off);
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(g + (size_t)c * ld_g + j));
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
    float4 d = make_float4(a.x, a.y, b.x, b.y);
    if (wd != 0.f) { d.x += wd * ww.x; d.y += wd * ww.y; d.z += wd * ww.z; d.w += wd * ww.w; }
    if (mu != 0.f) {
      const float4 an = *reinterpret_cast<const float4*>(anchor + j);
      d.x += mu * (ww.x - an.x); d.y += mu * (ww.y - an.y); d.z += mu * (ww.z - an.z); d.w += mu * (ww.w - an.w);
    }
    ww.x -= lr * d.x; ww.y -= lr * d.y; ww.z -= lr * d.z; ww.w -= lr * d.w;
    *reinterpret_cast<float4*>(w + off) = ww;
  }
}
extern "C" int fb_sgd_step_stacked_bf16g(float* w, const void* g_bf16, int64_t ld, int64_t ld_g, int32_t C, int64_t n,
                                         const float* lr_dev, float weight_decay, float mu, const float* anchor,
                                         fb_stream_t stream) {
  FB_REQUIRE(w && g_bf16 && lr_dev && C >= 1 && n > 0 && n % 4 == 0 && ld % 4 == 0 && ld_g % 4 == 0,
             "fb_sgd_step_stacked_bf16g: bad shape");
  FB_REQUIRE(mu == 0.f || anchor, "fb_sgd_step_stacked_bf16g: mu needs the anchor (global) model");
  FB_CHECK_CUDA(fb_launch_pdl(sgd_step_stacked_bf16g_kernel, dim3(ew_grid(n / 4 * C)), dim3(256), (size_t)(0), (cudaStream_t)stream, 
      w, (const __nv_bfloat16*)g_bf16, ld, ld_g, C, n / 4, lr_dev, weight_decay, mu, anchor));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- softmax cross-entropy (one warp per row)
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, int64_t R, int ncls,
                    int64_t ignore_index, int pre_sigmoid, const float* __restrict__ row_scale, int64_t rows_per_group,
                    float* __restrict__ loss_rows, float* __restrict__ dlogits) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < R; r += nwarps) {
    const float* row = logits + (size_t)r * ncls;
    const int64_t y = labels[r];
    const bool ignored = (y == ignore_index) || y < 0 || y >= ncls;
    float mx = -INFINITY;
    for (int c = lane; c < ncls; c += 32) {
      float z = row[c];
      if (pre_sigmoid) z = sigmoidf_(z);
      mx = fmaxf(mx, z);
    }
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < ncls; c += 32) {
      float z = row[c];
      if (pre_sigmoid) z = sigmoidf_(z);
      se += __expf(z - mx);
    }
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    if (loss_rows && lane == 0) {
      float zy = 0.f;
      if (!ignored) {
        zy = row[y];
        if (pre_sigmoid) zy = sigmoidf_(zy);
      }
      loss_rows[r] = ignored ? 0.f : (lse - zy);
    }
    if (dlogits) {
      const float sc = ignored ? 0.f : (row_scale ? __ldg(row_scale + r / rows_per_group) : 1.f / (float)R);
      for (int c = lane; c < ncls; c += 32) {
        float z = row[c], s = 0.f;
        if (pre_sigmoid) { s = sigmoidf_(z); z = s; }
        float gsm = (__expf(z - lse) - (c == y ? 1.f : 0.f)) * sc;
        if (pre_sigmoid) gsm *= s * (1.f - s);
        dlogits[(size_t)r * ncls + c] = gsm;
      }
    }
  }
}
extern "C" int fb_softmax_xent(const float* logits, const int64_t* labels, int64_t R, int32_t ncls,
                               int64_t ignore_index, int32_t pre_sigmoid, const float* row_scale,
                               int64_t rows_per_group, float* loss_rows, float* dlogits, fb_stream_t stream) {
  FB_REQUIRE(logits && labels && R > 0 && ncls > 0, "fb_softmax_xent: bad shape");
  if (rows_per_group < 1) rows_per_group = 1;
  int64_t want = (R * 32 + 255) / 256, cap = (int64_t)fb_num_sms() * 8;
  int grid = (int)(want < cap ? want : cap);
  FB_CHECK_CUDA(fb_launch_pdl(softmax_xent_kernel, dim3(grid < 1 ? 1 : grid), dim3(256), (size_t)(0), (cudaStream_t)stream, 
      logits, labels, R, ncls, ignore_index, pre_sigmoid, row_scale, rows_per_group, loss_rows, dlogits));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- fused evaluation counters
// counts = [correct, total, tp[ncls], fp[ncls], fn[ncls]]
__global__ void __launch_bounds__(256)
eval_counts_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, int64_t R, int ncls,
                   int pre_sigmoid, uint32_t* __restrict__ counts, float* __restrict__ loss_sum) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  __shared__ float red[33];
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  float loss_acc = 0.f;
  uint32_t correct = 0, total = 0;
  for (int64_t r = warp0; r < R; r += nwarps) {
    const float* row = logits + (size_t)r * ncls;
    const int64_t y = labels[r];
    float mx = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = lane; c < ncls; c += 32) {
      float z = row[c];
      if (pre_sigmoid) z = sigmoidf_(z);
      if (z > mx) { mx = z; arg = c; }
    }
    // warp arg-max, lowest index wins ties (torch.max / argmax behaviour on the first maximal element)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    float se = 0.f;
    for (int c = lane; c < ncls; c += 32) {
      float z = row[c];
      if (pre_sigmoid) z = sigmoidf_(z);
      se += __expf(z - mx);
    }
    se = warp_sum(se);
    if (lane == 0 && y >= 0 && y < ncls) {
      float zy = row[y];
      if (pre_sigmoid) zy = sigmoidf_(zy);
      loss_acc += mx + __logf(se) - zy;
      total += 1;
      if (arg == (int)y) {
        correct += 1;
        atomicAdd(counts + 2 + y, 1u);
      } else {
        atomicAdd(counts + 2 + ncls + arg, 1u);       // false positive for the predicted class
        atomicAdd(counts + 2 + 2 * ncls + y, 1u);     // false negative for the true class
      }
    }
  }
  const float bl = block_sum(loss_acc, red);
  const float bc = block_sum((float)correct, red);
  const float bt = block_sum((float)total, red);
  if (threadIdx.x == 0) {
    atomicAdd(loss_sum, bl);
    atomicAdd(counts + 0, (uint32_t)(bc + 0.5f));
    atomicAdd(counts + 1, (uint32_t)(bt + 0.5f));
  }
}
extern "C" int fb_eval_counts(const float* logits, const int64_t* labels, int64_t R, int32_t ncls, int32_t pre_sigmoid,
                              uint32_t* counts, float* loss_sum, fb_stream_t stream) {
  FB_REQUIRE(logits && labels && counts && loss_sum && R > 0 && ncls > 0, "fb_eval_counts: bad shape");
  int64_t want = (R * 32 + 255) / 256, cap = (int64_t)fb_num_sms() * 4;
  int grid = (int)(want < cap ? want : cap);
  FB_CHECK_CUDA(fb_launch_pdl(eval_counts_kernel, dim3(grid < 1 ? 1 : grid), dim3(256), (size_t)(0), (cudaStream_t)stream, logits, labels, R, ncls, pre_sigmoid, counts,
                                                                            loss_sum));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- conv / linear weight staging (one launch per step)
// For every listed layer: fp32 master weights [cout][cin][kh][kw] of every client (rows of the stacked arena) ->
//   Wm  bf16 [cout][k*k*cin_pad]   ((kh,kw,cin) order = the im2col column order, zero-padded input channels)
//   WmT bf16 [k*k*cin_pad][cout]   (K-major operand for the dgrad GEMM), optional
// Segment table (int64 x 8 per layer): src_off, cout, cin, cin_pad, k, dst_off, dstT_off (-1: none), n = cout*k*k*cin_pad
// staged GEMM operands are bf16 (kind::f16 path) or fp32 (kind::tf32 path): ST is the staged element type
__device__ __forceinline__ void stage_st1(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
__device__ __forceinline__ void stage_st1(float* p, float v) { *p = v; }
__device__ __forceinline__ void stage_st4(__nv_bfloat16* p, float a, float b, float c, float d) {
  uint2 o;
  *reinterpret_cast<__nv_bfloat162*>(&o.x) = __floats2bfloat162_rn(a, b);
  *reinterpret_cast<__nv_bfloat162*>(&o.y) = __floats2bfloat162_rn(c, d);
  *reinterpret_cast<uint2*>(p) = o;
}
__device__ __forceinline__ void stage_st4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void stage_st8(__nv_bfloat16* p, const float4& a, const float4& b) {
  uint4 o;
  __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&o);
  h2[0] = __floats2bfloat162_rn(a.x, a.y);
  h2[1] = __floats2bfloat162_rn(a.z, a.w);
  h2[2] = __floats2bfloat162_rn(b.x, b.y);
  h2[3] = __floats2bfloat162_rn(b.z, b.w);
  *reinterpret_cast<uint4*>(p) = o;
}
__device__ __forceinline__ void stage_st8(float* p, const float4& a, const float4& b) {
  reinterpret_cast<float4*>(p)[0] = a;
  reinterpret_cast<float4*>(p)[1] = b;
}

template <typename ST>
__global__ void __launch_bounds__(256)
stage_weights_kernel(const float* __restrict__ arena, int64_t ld, ST* __restrict__ stage, int64_t stage_ld,
                     const int64_t* __restrict__ table, int n_layers) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  const int c = blockIdx.y;
  const float* row = arena + (size_t)c * ld;
  ST* out = stage + (size_t)c * stage_ld;
  for (int l = blockIdx.z; l < n_layers; l += gridDim.z) {
    const int64_t* t = table + (size_t)l * 8;
    const int64_t src = t[0], dst = t[5], dstT = t[6];
    const int cout = (int)t[1], cin = (int)t[2], cinp = (int)t[3], k = (int)t[4], n = (int)t[7];
    const int kk = k * k, kdim = kk * cinp;
    if (k == 1 && cin == cinp && dstT < 0 && (n & 7) == 0 && (src & 3) == 0 && (dst & 7) == 0 && (ld & 3) == 0 &&
        (stage_ld & 7) == 0) {
      // plain fp32 -> bf16 cast of a linear layer / 1x1 conv: 32-byte loads, 16-byte stores
      const float4* s4 = reinterpret_cast<const float4*>(row + src);
      for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < (n >> 3); v += gridDim.x * blockDim.x) {
        const float4 a = __ldg(s4 + 2 * v), b = __ldg(s4 + 2 * v + 1);
        stage_st8(out + dst + (size_t)v * 8, a, b);
      }
      continue;
    }
    if (k == 3) {
      // 3 x 3 weights: (cout 32) x (cin 32) x 9 tiles through shared memory, so the fp32 rows are read in contiguous
      // runs and both bf16 layouts are written in contiguous runs (the element-wise form below strides by 9 floats on
      // the read side and by cout on the transposed write side: 8x / 16x sector amplification on big layers).  All
      // index arithmetic divides by compile-time constants.
      __shared__ float tile[32 * 289];   // [cout 32][cin 32][tap 9] + 1 pad per row: every access pattern below is conflict-free
      constexpr int KK = 9, RUN = 32 * KK;
      const int n_cb = (cout + 31) / 32, n_ib = (cinp + 31) / 32;
      const bool full = cin == cinp && (cin & 31) == 0 && (cout & 31) == 0 && (src & 3) == 0 && (dst & 3) == 0 &&
                        (dstT < 0 || (dstT & 3) == 0) && (ld & 3) == 0 && (stage_ld & 3) == 0;
      for (int t = blockIdx.x; t < n_cb * n_ib; t += gridDim.x) {
        const int co0 = (t / n_ib) * 32, ci0 = (t % n_ib) * 32;
        __syncthreads();
        if (full) {                                                    // vectorised whole tile, as in sgd_restage_kernel
          const float* wbase = row + src + ((size_t)co0 * cin + ci0) * KK;
#pragma unroll 3
          for (int i = threadIdx.x; i < 32 * (RUN / 4); i += 256) {
            const int co = i / (RUN / 4), v4 = i - co * (RUN / 4);
            const float4 v = __ldg(reinterpret_cast<const float4*>(wbase + (size_t)co * cin * KK + v4 * 4));
            float* tt = tile + co * 289 + v4 * 4;
            tt[0] = v.x; tt[1] = v.y; tt[2] = v.z; tt[3] = v.w;
          }
          __syncthreads();
#pragma unroll 3
          for (int i = threadIdx.x; i < 32 * KK * 8; i += 256) {
            const int ci4 = i & 7, q = i >> 3;
            const int co = q / KK, tap = q - co * KK;
            const float* tt = tile + co * 289 + ci4 * 36 + tap;
            stage_st4(out + dst + ((size_t)(co0 + co) * KK + tap) * cinp + ci0 + ci4 * 4, tt[0], tt[9], tt[18], tt[27]);
          }
          if (dstT >= 0) {
#pragma unroll 3
            for (int i = threadIdx.x; i < KK * 32 * 8; i += 256) {
              const int co4 = i & 7, q = i >> 3;
              const int ci = q & 31, tap = q >> 5;
              const float* tt = tile + co4 * 4 * 289 + ci * 9 + tap;
              stage_st4(out + dstT + ((size_t)tap * cinp + ci0 + ci) * cout + co0 + co4 * 4, tt[0], tt[289], tt[578], tt[867]);
            }
          }
          continue;
        }
#pragma unroll 12
        for (int i = threadIdx.x; i < 32 * RUN; i += 256) {            // 36 independent loads per thread: keep them in flight
          const int co = i / RUN, r = i - co * RUN;
          const int ci = r / KK, tap = r - ci * KK;
          float v = 0.f;
          if (co0 + co < cout && ci0 + ci < cin) v = __ldg(row + src + ((size_t)(co0 + co) * cin + ci0 + ci) * KK + tap);
          tile[co * 289 + r] = v;
        }
        __syncthreads();
        const int lane = threadIdx.x & 31;
#pragma unroll 12
        for (int q = threadIdx.x >> 5; q < 32 * KK; q += 8) {                 // Wm: a warp writes 32 input channels
          const int co = q / KK, tap = q - co * KK;
          if (co0 + co < cout && ci0 + lane < cinp)
            stage_st1(out + dst + ((size_t)(co0 + co) * KK + tap) * cinp + ci0 + lane, tile[co * 289 + lane * 9 + tap]);
        }
        if (dstT >= 0) {
#pragma unroll 12
          for (int q = threadIdx.x >> 5; q < 32 * KK; q += 8) {               // WmT: a warp writes 32 output channels
            const int tap = q >> 5, ci = q & 31;
            if (co0 + lane < cout && ci0 + ci < cinp)
              stage_st1(out + dstT + ((size_t)tap * cinp + ci0 + ci) * cout + co0 + lane, tile[lane * 289 + ci * 9 + tap]);
          }
        }
      }
      continue;
    }
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
      const int co = e / kdim, r = e - co * kdim;
      const int tap = r / cinp, ci = r - tap * cinp;
      const float v = ci < cin ? row[src + ((size_t)co * cin + ci) * kk + tap] : 0.f;
      stage_st1(out + dst + e, v);
      if (dstT >= 0) stage_st1(out + dstT + (size_t)r * cout + co, v);
    }
  }
}
template <typename ST>
static int stage_conv_weights_impl(const float* arena, int64_t ld, int32_t C, void* stage, int64_t stage_ld,
                                   const int64_t* table_dev, int32_t n_layers, int32_t max_n, fb_stream_t stream) {
  FB_REQUIRE(arena && stage && table_dev && C >= 1 && C <= 65535 && n_layers >= 1, "fb_stage_conv_weights: bad arguments");
  int gx = (max_n + 255) / 256;
  const int gx_cap = max_n >= (1 << 18) ? 64 : 16;   // big linear layers: more CTAs per (client, layer) slice
  if (gx > gx_cap) gx = gx_cap;
  if (gx < 1) gx = 1;
  int gz = n_layers < 64 ? n_layers : 64;
  FB_CHECK_CUDA(fb_launch_pdl(stage_weights_kernel<ST>, dim3(dim3(gx, C, gz)), dim3(256), (size_t)(0), (cudaStream_t)stream, arena, ld, (ST*)stage, stage_ld,
                                                                         table_dev, n_layers));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_stage_conv_weights(const float* arena, int64_t ld, int32_t C, void* stage, int64_t stage_ld,
                                     const int64_t* table_dev, int32_t n_layers, int32_t max_n, fb_stream_t stream) {
  return stage_conv_weights_impl<__nv_bfloat16>(arena, ld, C, stage, stage_ld, table_dev, n_layers, max_n, stream);
}
extern "C" int fb_stage_conv_weights_f32(const float* arena, int64_t ld, int32_t C, void* stage, int64_t stage_ld,
                                         const int64_t* table_dev, int32_t n_layers, int32_t max_n, fb_stream_t stream) {
  return stage_conv_weights_impl<float>(arena, ld, C, stage, stage_ld, table_dev, n_layers, max_n, stream);
}

// ---------------------------------------------------------------- fused optimizer step + re-staging (plain SGD)
// One launch per local step does, for every trainable segment of every client,  w -= lr[c] * g  AND refreshes the bf16
// GEMM operands of the next step, so the separate staging pass (read w again, write bf16) and the gradient un-staging
// pass (GEMM layout -> parameter layout) disappear: k x k weight gradients are read straight from the GEMM-layout
// scratch the weight-gradient kernels wrote.  Table (int64 x 10 per segment):
//   w_off, cout, cin, cin_pad, k, dst (stage offset or -1: no bf16 copy), dstT (or -1), n = cout*k*k*cin_pad,
//   g_off, g_kind (0: gradient in the arena at w_off, parameter layout; 1: in the scratch stack at g_off, GEMM layout)
// Plain segments (BN affine, biases, fc) are listed as cout = numel, cin = k = 1, dst = -1.
#define RS_MAX_SEGS 1024
#define RS_VEC_TILE 4096      /* elements per tile of the vector path */
#define RS_GEN_TILE 2048      /* elements per tile of the element-wise path */
struct RsSeg {
  int64_t src, dst, dstT, goff;
  int cout, cin, cinp, k, n, gkind, path;   // path 0: vector, 1: 3x3 tiles, 2: element-wise
};
__device__ __forceinline__ RsSeg rs_load(const int64_t* t, int64_t ld, int64_t stage_ld) {
  RsSeg s;
  s.src = t[0]; s.cout = (int)t[1]; s.cin = (int)t[2]; s.cinp = (int)t[3]; s.k = (int)t[4];
  s.dst = t[5]; s.dstT = t[6]; s.n = (int)t[7]; s.goff = t[8]; s.gkind = (int)t[9];
  if (s.k == 1 && s.cin == s.cinp && s.dstT < 0 && s.gkind == 0 && (s.n & 7) == 0 && (s.src & 3) == 0 &&
      (s.dst < 0 || (s.dst & 7) == 0) && (ld & 3) == 0 && (stage_ld & 7) == 0)
    s.path = 0;
  else if (s.k == 3 && s.gkind == 1 && s.dst >= 0)
    s.path = 1;
  else
    s.path = 2;
  return s;
}
__device__ __forceinline__ int rs_tiles(const RsSeg& s) {
  if (s.path == 0) return (s.n + RS_VEC_TILE - 1) / RS_VEC_TILE;
  if (s.path == 1) return ((s.cout + 31) / 32) * ((s.cinp + 31) / 32);
  return (s.n + RS_GEN_TILE - 1) / RS_GEN_TILE;
}

// grid (X, C): the CTAs of a client walk ONE flat list of tiles over all segments (prefix of per-segment tile counts in
// shared memory), so small segments cost a tile, not a grid slice of mostly empty CTAs
template <typename ST>
__global__ void __launch_bounds__(256)
sgd_restage_kernel(float* __restrict__ arena, const float* __restrict__ grads, int64_t ld, const float* __restrict__ scratch,
                   int64_t scratch_ld, ST* __restrict__ stage, int64_t stage_ld, const int64_t* __restrict__ table,
                   int n_layers, const float* __restrict__ lr_dev) {
  __shared__ float tile[32 * 289];
  __shared__ int tstart[RS_MAX_SEGS + 1];
  for (int l = threadIdx.x; l < n_layers; l += blockDim.x) {          // the table is host-written: readable before the wait
    const RsSeg sg = rs_load(table + (size_t)l * 10, ld, stage_ld);
    tstart[l + 1] = rs_tiles(sg);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    tstart[0] = 0;
    for (int l = 0; l < n_layers; ++l) tstart[l + 1] += tstart[l];
  }
  __syncthreads();
  const int total = tstart[n_layers];
  pdl_wait();
  pdl_trigger();
  const int c = blockIdx.y;
  const float lr = lr_dev[c];
  float* row = arena + (size_t)c * ld;
  const float* grow = grads + (size_t)c * ld;
  const float* srow = scratch ? scratch + (size_t)c * scratch_ld : nullptr;
  ST* out = stage ? stage + (size_t)c * stage_ld : nullptr;
  for (int gt = blockIdx.x; gt < total; gt += gridDim.x) {
    int lo = 0, hi = n_layers - 1;                                      // last segment with tstart <= gt
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tstart[mid] <= gt) lo = mid; else hi = mid - 1;
    }
    const RsSeg sg = rs_load(table + (size_t)lo * 10, ld, stage_ld);
    const int tl = gt - tstart[lo];
    const int64_t src = sg.src, dst = sg.dst, dstT = sg.dstT, goff = sg.goff;
    const int cout = sg.cout, cin = sg.cin, cinp = sg.cinp, n = sg.n;
    if (sg.path == 0) {
      float4* w4 = reinterpret_cast<float4*>(row + src);
      const float4* g4 = reinterpret_cast<const float4*>(grow + src);
      ST* d8 = dst >= 0 ? out + dst : nullptr;
      const int v0 = tl * (RS_VEC_TILE / 8), v1 = min(n >> 3, v0 + RS_VEC_TILE / 8);
      for (int v = v0 + threadIdx.x; v < v1; v += 256) {
        float4 a = w4[2 * v], b = w4[2 * v + 1];
        const float4 ga = __ldg(g4 + 2 * v), gb = __ldg(g4 + 2 * v + 1);
        a.x -= lr * ga.x; a.y -= lr * ga.y; a.z -= lr * ga.z; a.w -= lr * ga.w;
        b.x -= lr * gb.x; b.y -= lr * gb.y; b.z -= lr * gb.z; b.w -= lr * gb.w;
        if (lr != 0.f) { w4[2 * v] = a; w4[2 * v + 1] = b; }
        if (d8) stage_st8(d8 + (size_t)v * 8, a, b);
      }
    } else if (sg.path == 1) {
      constexpr int KK = 9, RUN = 32 * KK;
      const int n_ib = (cinp + 31) / 32;
      const int co0 = (tl / n_ib) * 32, ci0 = (tl % n_ib) * 32;
      const int lane = threadIdx.x & 31;
      __syncthreads();
      const bool full = cin == cinp && (cin & 31) == 0 && (cout & 31) == 0 && (src & 3) == 0 && (goff & 3) == 0 && (dst & 3) == 0 &&
                        (dstT < 0 || (dstT & 3) == 0) && (ld & 3) == 0 && (scratch_ld & 3) == 0 && (stage_ld & 3) == 0;
      if (full) {
        // whole 32 x 32 x 9 tile, everything 16-byte aligned: four elements per thread and access (float4 / 4 x bf16), a
        // quarter of the index arithmetic of the element-wise passes below (which issue-bound this kernel: ncu IPC 2.4,
        // DRAM 28 %)
        float* wbase = row + src + ((size_t)co0 * cin + ci0) * KK;
#pragma unroll 3
        for (int i = threadIdx.x; i < 32 * (RUN / 4); i += 256) {            // w tile: 32 rows x 72 float4
          const int co = i / (RUN / 4), v4 = i - co * (RUN / 4);
          const float4 v = *reinterpret_cast<const float4*>(wbase + (size_t)co * cin * KK + v4 * 4);
          float* t = tile + co * 289 + v4 * 4;
          t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
        }
        __syncthreads();
#pragma unroll 3
        for (int i = threadIdx.x; i < 32 * KK * 8; i += 256) {                // GEMM order: (cout, tap, 4 input channels)
          const int ci4 = i & 7, q = i >> 3;
          const int co = q / KK, tap = q - co * KK;
          const size_t gi = ((size_t)(co0 + co) * KK + tap) * cinp + ci0 + ci4 * 4;
          const float4 g = __ldg(reinterpret_cast<const float4*>(srow + goff + gi));
          float* t = tile + co * 289 + ci4 * 36 + tap;
          const float w0 = t[0] - lr * g.x, w1 = t[9] - lr * g.y, w2 = t[18] - lr * g.z, w3 = t[27] - lr * g.w;
          t[0] = w0; t[9] = w1; t[18] = w2; t[27] = w3;
          stage_st4(out + dst + gi, w0, w1, w2, w3);
        }
        __syncthreads();
        if (dstT >= 0) {
#pragma unroll 3
          for (int i = threadIdx.x; i < KK * 32 * 8; i += 256) {              // transposed copy: (tap, cin, 4 output channels)
            const int co4 = i & 7, q = i >> 3;
            const int ci = q & 31, tap = q >> 5;
            const float* t = tile + co4 * 4 * 289 + ci * 9 + tap;
            stage_st4(out + dstT + ((size_t)tap * cinp + ci0 + ci) * cout + co0 + co4 * 4, t[0], t[289], t[578], t[867]);
          }
        }
        if (lr != 0.f) {
#pragma unroll 3
          for (int i = threadIdx.x; i < 32 * (RUN / 4); i += 256) {          // updated weights back
            const int co = i / (RUN / 4), v4 = i - co * (RUN / 4);
            const float* t = tile + co * 289 + v4 * 4;
            *reinterpret_cast<float4*>(wbase + (size_t)co * cin * KK + v4 * 4) = make_float4(t[0], t[1], t[2], t[3]);
          }
        }
        continue;
      }
#pragma unroll 12
      for (int i = threadIdx.x; i < 32 * RUN; i += 256) {                   // w tile, parameter layout runs
        const int co = i / RUN, r = i - co * RUN;
        const int ci = r / KK;
        float v = 0.f;
        if (co0 + co < cout && ci0 + ci < cin) v = row[src + ((size_t)(co0 + co) * cin + ci0) * KK + r];
        tile[co * 289 + r] = v;
      }
      __syncthreads();
#pragma unroll 12
      for (int q = threadIdx.x >> 5; q < 32 * KK; q += 8) {                 // GEMM-layout pass: gradient, update, Wm
        const int co = q / KK, tap = q - co * KK;
        if (co0 + co < cout && ci0 + lane < cinp) {
          const size_t gi = ((size_t)(co0 + co) * KK + tap) * cinp + ci0 + lane;
          float w = tile[co * 289 + lane * 9 + tap];
          if (ci0 + lane < cin) {
            w -= lr * __ldg(srow + goff + gi);
            tile[co * 289 + lane * 9 + tap] = w;
          }
          stage_st1(out + dst + gi, w);
        }
      }
      __syncthreads();
      if (dstT >= 0) {
#pragma unroll 12
        for (int q = threadIdx.x >> 5; q < 32 * KK; q += 8) {               // WmT: a warp writes 32 output channels
          const int tap = q >> 5, ci = q & 31;
          if (co0 + lane < cout && ci0 + ci < cinp)
            stage_st1(out + dstT + ((size_t)tap * cinp + ci0 + ci) * cout + co0 + lane, tile[lane * 289 + ci * 9 + tap]);
        }
      }
      if (lr != 0.f) {
#pragma unroll 12
        for (int i = threadIdx.x; i < 32 * RUN; i += 256) {                 // updated weights back, parameter layout runs
          const int co = i / RUN, r = i - co * RUN;
          const int ci = r / KK;
          if (co0 + co < cout && ci0 + ci < cin) row[src + ((size_t)(co0 + co) * cin + ci0) * KK + r] = tile[co * 289 + r];
        }
      }
    } else {
      // element-wise form (tiny or odd-sized segments, channel-padded 1x1, k x k without scratch gradients)
      const int kk = sg.k * sg.k, kdim = kk * cinp;
      const int e0 = tl * RS_GEN_TILE, e1 = min(n, e0 + RS_GEN_TILE);
      for (int e = e0 + threadIdx.x; e < e1; e += 256) {
        const int co = e / kdim, r = e - co * kdim;
        const int tap = r / cinp, ci = r - tap * cinp;
        float w = 0.f;
        if (ci < cin) {
          const size_t wi = src + ((size_t)co * cin + ci) * kk + tap;
          w = row[wi] - lr * (sg.gkind ? srow[goff + e] : grow[wi]);
          if (lr != 0.f) row[wi] = w;
        }
        if (dst >= 0) {
          stage_st1(out + dst + e, w);
          if (dstT >= 0) stage_st1(out + dstT + (size_t)r * cout + co, w);
        }
      }
    }
  }
}
template <typename ST>
static int sgd_restage_impl(float* arena, const float* grads, int64_t ld, int32_t C, const float* scratch, int64_t scratch_ld,
                            void* stage, int64_t stage_ld, const int64_t* table_dev, int32_t n_layers, int32_t total_elems,
                            const float* lr_dev, fb_stream_t stream) {
  FB_REQUIRE(arena && grads && table_dev && lr_dev && C >= 1 && C <= 65535 && n_layers >= 1, "fb_sgd_restage: bad arguments");
  FB_REQUIRE(n_layers <= RS_MAX_SEGS, "fb_sgd_restage: at most %d segments", RS_MAX_SEGS);
  // CTAs per client: enough to put ~16 CTAs on every SM overall, at least one per ~8K elements, at most 256
  int gx = (16 * fb_num_sms() + C - 1) / C;
  const int by_work = total_elems / 8192 + 1;
  if (gx > by_work) gx = by_work;
  if (gx > 256) gx = 256;
  if (gx < 1) gx = 1;
  FB_CHECK_CUDA(fb_launch_pdl(sgd_restage_kernel<ST>, dim3(gx, C), dim3(256), (size_t)0, (cudaStream_t)stream, arena, grads, ld, scratch,
                              scratch_ld, (ST*)stage, stage_ld, table_dev, n_layers, lr_dev));
  return FB_OK;
}
extern "C" int fb_sgd_restage(float* arena, const float* grads, int64_t ld, int32_t C, const float* scratch, int64_t scratch_ld,
                              void* stage, int64_t stage_ld, const int64_t* table_dev, int32_t n_layers, int32_t total_elems,
                              const float* lr_dev, fb_stream_t stream) {
  return sgd_restage_impl<__nv_bfloat16>(arena, grads, ld, C, scratch, scratch_ld, stage, stage_ld, table_dev, n_layers, total_elems,
                                         lr_dev, stream);
}
extern "C" int fb_sgd_restage_f32(float* arena, const float* grads, int64_t ld, int32_t C, const float* scratch, int64_t scratch_ld,
                                  void* stage, int64_t stage_ld, const int64_t* table_dev, int32_t n_layers, int32_t total_elems,
                                  const float* lr_dev, fb_stream_t stream) {
  return sgd_restage_impl<float>(arena, grads, ld, C, scratch, scratch_ld, stage, stage_ld, table_dev, n_layers, total_elems, lr_dev,
                                 stream);
}

// ---------------------------------------------------------------- weight-gradient un-staging (one launch per step)
// The k x k weight-gradient GEMMs produce dW in the GEMM layout fp32 [cout][(kh,kw,cin_pad)] (scratch rows); the
// gradient arena wants the parameter layout [cout][cin][kh][kw].  Table (int64 x 8 per layer): src_off (scratch),
// cout, cin, cin_pad, k, dst_off (arena), unused, unused.  Same shared-memory tiling as the staging kernel, the other
// way round.  Replaces one strided permute-copy launch per convolution per step.
__global__ void __launch_bounds__(256)
unstage_grads_kernel(const float* __restrict__ scratch, int64_t scratch_ld, float* __restrict__ grads, int64_t ld,
                     const int64_t* __restrict__ table, int n_layers) {
  pdl_wait();      // programmatic dependent launch: see common.cuh
  pdl_trigger();
  __shared__ float tile[32 * 289];   // [cout 32][cin 32][tap 9] + 1 pad per row: every access pattern below is conflict-free
  const int c = blockIdx.y;
  const float* in = scratch + (size_t)c * scratch_ld;
  float* out = grads + (size_t)c * ld;
  for (int l = blockIdx.z; l < n_layers; l += gridDim.z) {
    const int64_t* t = table + (size_t)l * 8;
    const int64_t src = t[0], dst = t[5];
    const int cout = (int)t[1], cin = (int)t[2], cinp = (int)t[3], k = (int)t[4];
    const int kk = k * k;
    const int n_cb = (cout + 31) / 32, n_ib = (cin + 31) / 32;
    if (k != 3) {                                                     // generic (rare): element-wise
      const int kdim = kk * cinp;
      for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < cout * kk * cin; e += gridDim.x * blockDim.x) {
        const int co = e / (kk * cin), r = e - co * (kk * cin);
        const int ci = r / kk, tap = r - ci * kk;
        out[dst + e] = in[src + (size_t)co * kdim + (size_t)tap * cinp + ci];
      }
      continue;
    }
    constexpr int KK = 9, RUN = 32 * KK;
    for (int tl = blockIdx.x; tl < n_cb * n_ib; tl += gridDim.x) {
      const int co0 = (tl / n_ib) * 32, ci0 = (tl % n_ib) * 32;
      const int lane = threadIdx.x & 31;
      __syncthreads();
#pragma unroll 12
      for (int q = threadIdx.x >> 5; q < 32 * KK; q += 8) {                   // a warp reads 32 input channels (36 loads in flight)
        const int co = q / KK, tap = q - co * KK;
        float v = 0.f;
        if (co0 + co < cout && ci0 + lane < cin) v = __ldg(in + src + ((size_t)(co0 + co) * KK + tap) * cinp + ci0 + lane);
        tile[co * 289 + lane * 9 + tap] = v;
      }
      __syncthreads();
#pragma unroll 12
      for (int i = threadIdx.x; i < 32 * RUN; i += 256) {                     // contiguous (ci, tap) runs per cout
        const int co = i / RUN, r = i - co * RUN;
        const int ci = r / KK, tap = r - ci * KK;
        if (co0 + co < cout && ci0 + ci < cin) out[dst + ((size_t)(co0 + co) * cin + ci0 + ci) * KK + tap] = tile[co * 289 + r];
      }
    }
  }
}
extern "C" int fb_unstage_conv_grads(const float* scratch, int64_t scratch_ld, float* grads, int64_t ld, int32_t C,
                                     const int64_t* table_dev, int32_t n_layers, int32_t max_tiles, fb_stream_t stream) {
  FB_REQUIRE(scratch && grads && table_dev && C >= 1 && C <= 65535 && n_layers >= 1, "fb_unstage_conv_grads: bad arguments");
  int gx = max_tiles < 1 ? 1 : (max_tiles > 64 ? 64 : max_tiles);
  int gz = (8 * fb_num_sms()) / (gx * C);
  if (gz > n_layers) gz = n_layers;
  if (gz > 64) gz = 64;
  if (gz < 1) gz = 1;
  FB_CHECK_CUDA(fb_launch_pdl(unstage_grads_kernel, dim3(dim3(gx, C, gz)), dim3(256), (size_t)(0), (cudaStream_t)stream, scratch, scratch_ld, grads, ld, table_dev, n_layers));
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- fused classifier head: pool -> fc -> loss -> backward
// One launch replaces the tail of the forward pass and the head of the backward pass of a client-stacked ResNet:
// global average pool, fc (+bias), softmax cross-entropy, d(logits), fc weight / bias gradients, d(features) and its
// broadcast back over the H x W positions.  One CTA per client; everything between the two big tensors (h in, dh out)
// lives in shared memory, the fc weights are read in place from the parameter arena (fp32) and all arithmetic is fp32.
// This also takes the 10-column fc products off the library GEMM path (their K / N are not multiples of 16 bytes).
//   h  [C][B][HW][ch] (T)      W [C][ncls][ch] fp32, client stride ld_w      bias [C][ncls], stride ld_b
//   labels int64 [C][B] (negative or >= ncls: padding row)       row_scale [C] (NULL: 1 / B)
//   loss [C] = row_scale * sum_b CE_b      dh [C][B][HW][ch] (T) = d loss / d h
//   gW [C][ncls][ch] (stride ld_gw), gb [C][ncls] (stride ld_gb): gradients, overwritten
#define HEAD_THREADS 512
template <typename T> __device__ __forceinline__ float head_ld(const T* p);
template <> __device__ __forceinline__ float head_ld<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float head_ld<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void head_st4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void head_st4(__nv_bfloat16* p, float a, float b, float c, float d) { stage_st4(p, a, b, c, d); }

template <typename T>
__global__ void __launch_bounds__(HEAD_THREADS)
head_fwd_bwd_kernel(const T* __restrict__ h, int B, int HW, int ch, const float* __restrict__ W, int64_t ld_w,
                    const float* __restrict__ bias, int64_t ld_b, int ncls, const int64_t* __restrict__ labels,
                    const float* __restrict__ row_scale, float* __restrict__ loss, T* __restrict__ dh, float* __restrict__ gW,
                    int64_t ld_gw, float* __restrict__ gb, int64_t ld_gb) {
  extern __shared__ float sm[];
  float* feat = sm;                        // [B][ch]  pooled features, later d(features) / HW
  float* dlog = sm + (size_t)B * ch;       // [B][ncls] logits, then d(logits)
  float* lrow = dlog + (size_t)B * ncls;   // [B] per-row loss
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = HEAD_THREADS / 32;
  pdl_wait();
  pdl_trigger();
  const T* hc = h + (size_t)c * B * HW * ch;
  const float* Wc = W + (size_t)c * ld_w;
  const float inv_hw = 1.f / (float)HW;
  // ---- average pool: thread <-> (image, channel); consecutive threads read consecutive channels
  for (int idx = tid; idx < B * ch; idx += HEAD_THREADS) {
    const int b = idx / ch, k = idx - b * ch;
    const T* src = hc + (size_t)b * HW * ch + k;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int p = 0;
    for (; p + 4 <= HW; p += 4) {
      s0 += head_ld<T>(src + (size_t)p * ch);
      s1 += head_ld<T>(src + (size_t)(p + 1) * ch);
      s2 += head_ld<T>(src + (size_t)(p + 2) * ch);
      s3 += head_ld<T>(src + (size_t)(p + 3) * ch);
    }
    for (; p < HW; ++p) s0 += head_ld<T>(src + (size_t)p * ch);
    feat[idx] = (s0 + s1 + s2 + s3) * inv_hw;
  }
  __syncthreads();
  // ---- logits: one warp per (image, class)
  for (int pr = warp; pr < B * ncls; pr += nwarps) {
    const int b = pr / ncls, k = pr - b * ncls;
    const float* w = Wc + (size_t)k * ch;
    const float* f = feat + (size_t)b * ch;
    float s = 0.f;
    for (int j = lane; j < ch; j += 32) s += f[j] * __ldg(w + j);
    s = warp_sum(s);
    if (lane == 0) dlog[pr] = s + __ldg(bias + (size_t)c * ld_b + k);
  }
  __syncthreads();
  // ---- softmax cross-entropy and d(logits): one thread per image
  const float scale = row_scale ? __ldg(row_scale + c) : 1.f / (float)B;
  if (tid < B) {
    float* z = dlog + (size_t)tid * ncls;
    const int64_t y = labels[(size_t)c * B + tid];
    const bool ignored = y < 0 || y >= ncls;
    float mx = -INFINITY;
    for (int k = 0; k < ncls; ++k) mx = fmaxf(mx, z[k]);
    float se = 0.f;
    for (int k = 0; k < ncls; ++k) se += __expf(z[k] - mx);
    const float lse = mx + __logf(se);
    lrow[tid] = ignored ? 0.f : lse - z[y];
    const float sc = ignored ? 0.f : scale;
    for (int k = 0; k < ncls; ++k) z[k] = (__expf(z[k] - lse) - (k == (int)y ? 1.f : 0.f)) * sc;
  }
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int b = 0; b < B; ++b) t += lrow[b];
    loss[c] = t * scale;
  }
  // ---- fc gradients: gW[k][j] = sum_b dlog[b][k] * feat[b][j], gb[k] = sum_b dlog[b][k]
  for (int idx = tid; idx < ncls * ch; idx += HEAD_THREADS) {
    const int k = idx / ch, j = idx - k * ch;
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dlog[b * ncls + k] * feat[(size_t)b * ch + j];
    gW[(size_t)c * ld_gw + idx] = s;
  }
  if (tid < ncls) {
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dlog[b * ncls + tid];
    gb[(size_t)c * ld_gb + tid] = s;
  }
  __syncthreads();
  // ---- d(features) / HW, in place over feat
  for (int idx = tid; idx < B * ch; idx += HEAD_THREADS) {
    const int b = idx / ch, j = idx - b * ch;
    float s = 0.f;
    for (int k = 0; k < ncls; ++k) s += dlog[b * ncls + k] * __ldg(Wc + (size_t)k * ch + j);
    feat[idx] = s * inv_hw;
  }
  __syncthreads();
  // ---- broadcast over the H x W positions: dh[b][p][:] = dfeat[b][:]   (ch % 4 == 0: four channels per store)
  T* dc = dh + (size_t)c * B * HW * ch;
  const int ch4 = ch >> 2;
  const int64_t total4 = (int64_t)B * HW * ch4;
  for (int64_t i = tid; i < total4; i += HEAD_THREADS) {
    const int j4 = (int)(i % ch4);
    const int b = (int)(i / ((int64_t)HW * ch4));
    const float* f = feat + (size_t)b * ch + j4 * 4;
    head_st4(dc + i * 4, f[0], f[1], f[2], f[3]);
  }
}

template <typename T>
static int head_impl(const void* h, int32_t C, int32_t B, int32_t HW, int32_t ch, const float* W, int64_t ld_w, const float* bias,
                     int64_t ld_b, int32_t ncls, const int64_t* labels, const float* row_scale, float* loss, void* dh, float* gW,
                     int64_t ld_gw, float* gb, int64_t ld_gb, fb_stream_t stream) {
  FB_REQUIRE(h && W && bias && labels && loss && dh && gW && gb, "fb_head_fwd_bwd: null argument");
  FB_REQUIRE(C >= 1 && B >= 1 && B <= HEAD_THREADS && HW >= 1 && ch >= 4 && ch % 4 == 0 && ncls >= 1 && ncls <= HEAD_THREADS,
             "fb_head_fwd_bwd: bad shape (B, ncls <= %d, ch a multiple of 4)", HEAD_THREADS);
  const size_t smem = ((size_t)B * ch + (size_t)B * ncls + B) * sizeof(float);
  FB_REQUIRE(smem <= 200 * 1024, "fb_head_fwd_bwd: B x (ch + ncls) = %d x (%d + %d) floats exceed shared memory", B, ch, ncls);
  FB_CHECK_CUDA(cudaFuncSetAttribute(head_fwd_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  FB_CHECK_CUDA(fb_launch_pdl(head_fwd_bwd_kernel<T>, dim3(C), dim3(HEAD_THREADS), smem, (cudaStream_t)stream, (const T*)h, B, HW, ch, W,
                              ld_w, bias, ld_b, ncls, labels, row_scale, loss, (T*)dh, gW, ld_gw, gb, ld_gb));
  return FB_OK;
}
extern "C" int fb_head_fwd_bwd(const void* h, int32_t act_fp32, int32_t C, int32_t B, int32_t HW, int32_t ch, const float* W,
                               int64_t ld_w, const float* bias, int64_t ld_b, int32_t ncls, const int64_t* labels,
                               const float* row_scale, float* loss, void* dh, float* gW, int64_t ld_gw, float* gb, int64_t ld_gb,
                               fb_stream_t stream) {
  if (act_fp32)
    return head_impl<float>(h, C, B, HW, ch, W, ld_w, bias, ld_b, ncls, labels, row_scale, loss, dh, gW, ld_gw, gb, ld_gb, stream);
  return head_impl<__nv_bfloat16>(h, C, B, HW, ch, W, ld_w, bias, ld_b, ncls, labels, row_scale, loss, dh, gW, ld_gw, gb, ld_gb, stream);
}
