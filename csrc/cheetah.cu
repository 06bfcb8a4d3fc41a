// K13: Cheetah data-parallel step — gradient all-reduce over peer memory fused with the SGD(+momentum) update
// and the redistribution of the new parameters.  See include/fedml_b200.h.
//
// Rank r owns shard r of the flat bucket.  One launch per rank:
//   phase 0  per-block arrival handshake: block b of every rank has reached the kernel, hence (stream order)
//            every rank's backward has finished writing its gradient bucket;
//   phase 1  block b pulls its slice of shard r from all G gradient buckets (fixed rank order), averages,
//            applies weight decay / momentum / lr to the fp32 master slice and pushes the new parameters
//            into every rank's parameter bucket;
//   phase 2  completion handshake: all pushes into my bucket have landed before the kernel retires.
// Replaces DDP's bucketed NCCL all-reduce + one torch.optim kernel per tensor.
#include "common.cuh"
#include "../include/fedml_b200.h"

#define DP_THREADS 256

__device__ __forceinline__ void dp_signal_wait(uint32_t* const* peer_flags, int phase, int rank, int G, int block,
                                               uint32_t epoch) {
  __syncthreads();
  if ((int)threadIdx.x < G) {
    const int q = threadIdx.x;
    __threadfence_system();
    st_release_sys(peer_flags[q] + ((size_t)phase * G + rank) * FB_ROUND_MAX_BLOCKS + block, epoch);
    const uint32_t* mine = peer_flags[rank] + ((size_t)phase * G + q) * FB_ROUND_MAX_BLOCKS + block;
    fb_spin_wait_ge<true>(mine, epoch, 1u + (uint32_t)phase, (uint32_t)q);     // bounded: see common.cuh
  }
  __syncthreads();
}
FB_DEFINE_SPIN_API(fb_dp)

__device__ void dp_body(const FbDpStepParams& p, int block, int nblocks) {
  const int G = p.G, rank = p.rank;
  const int64_t off = p.offset;          // this launch covers elements [offset, offset + n) of the buckets (one gradient bucket)
  const int64_t nvec = p.n / 4;
  const int64_t S4 = (nvec + G - 1) / G;
  const int64_t lo = (int64_t)rank * S4, hi = min(lo + S4, nvec);
  if (G > 1) dp_signal_wait(p.peer_flags, 0, rank, G, block, p.epoch);
  for (int64_t v = lo + (int64_t)block * blockDim.x + threadIdx.x; v < hi; v += (int64_t)nblocks * blockDim.x) {
    const int64_t j = off + v * 4;
    float4 g = make_float4(0, 0, 0, 0);
    if (G <= 8) {
      // all peer loads are issued before the first add: G NVLink round trips overlap instead of queueing
      float4 xs[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (q < G) {
          const float* src = p.peer_grad[q] + j;
          asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(xs[q].x), "=f"(xs[q].y), "=f"(xs[q].z), "=f"(xs[q].w) : "l"(src) : "memory");
        }
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (q < G) { g.x += xs[q].x; g.y += xs[q].y; g.z += xs[q].z; g.w += xs[q].w; }
      }
    } else {
      for (int q = 0; q < G; ++q) {
        float4 x;
        const float* src = p.peer_grad[q] + j;
        asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "l"(src) : "memory");
        g.x += x.x; g.y += x.y; g.z += x.z; g.w += x.w;
      }
    }
    g.x *= p.inv_world; g.y *= p.inv_world; g.z *= p.inv_world; g.w *= p.inv_world;
    float4 w = *reinterpret_cast<const float4*>(p.peer_param[rank] + j);
    if (p.weight_decay != 0.f) {
      g.x += p.weight_decay * w.x; g.y += p.weight_decay * w.y; g.z += p.weight_decay * w.z; g.w += p.weight_decay * w.w;
    }
    if (p.momentum != 0.f && p.mom) {
      float4 b = *reinterpret_cast<const float4*>(p.mom + j);
      b.x = p.momentum * b.x + g.x; b.y = p.momentum * b.y + g.y; b.z = p.momentum * b.z + g.z; b.w = p.momentum * b.w + g.w;
      *reinterpret_cast<float4*>(p.mom + j) = b;
      if (p.nesterov) { g.x += p.momentum * b.x; g.y += p.momentum * b.y; g.z += p.momentum * b.z; g.w += p.momentum * b.w; }
      else g = b;
    }
    w.x -= p.lr * g.x; w.y -= p.lr * g.y; w.z -= p.lr * g.z; w.w -= p.lr * g.w;
    for (int q = 0; q < G; ++q) st4(p.peer_param[q] + j, w);
  }
  if (G > 1) dp_signal_wait(p.peer_flags, 1, rank, G, block, p.epoch);
}

__global__ void __launch_bounds__(DP_THREADS) dp_kernel(const FbDpStepParams p) { dp_body(p, blockIdx.x, gridDim.x); }
__global__ void __launch_bounds__(DP_THREADS) dp_kernel_emulated(const FbDpStepParams* __restrict__ ps) {
  __shared__ FbDpStepParams sp;
  if (threadIdx.x == 0) sp = ps[blockIdx.y];
  __syncthreads();
  dp_body(sp, blockIdx.x, gridDim.x);
}

static int dp_grid(int64_t n, int G, int emulated, int max_blocks = 0) {
  const int64_t S4 = (n / 4 + G - 1) / G;
  int64_t want = (S4 + DP_THREADS - 1) / DP_THREADS;
  int cap = 2 * fb_num_sms();
  if (max_blocks > 0 && cap > max_blocks) cap = max_blocks;     // overlapped buckets leave most SMs to the backward pass
  if (emulated > 1) cap /= emulated;
  if (cap > FB_ROUND_MAX_BLOCKS) cap = FB_ROUND_MAX_BLOCKS;
  if (cap < 1) cap = 1;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

static int dp_check(const FbDpStepParams* p) {
  FB_REQUIRE(p && p->peer_grad && p->peer_param && p->n > 0 && p->n % 4 == 0 && p->offset >= 0 && p->offset % 4 == 0,
             "fb_dp_allreduce_step: bad bucket");
  FB_REQUIRE(p->G >= 1 && p->G <= 64 && p->rank >= 0 && p->rank < p->G, "fb_dp_allreduce_step: bad rank/G");
  FB_REQUIRE(p->G == 1 || p->peer_flags, "fb_dp_allreduce_step: flags missing");
  return FB_OK;
}

extern "C" int fb_dp_allreduce_step(const FbDpStepParams* p, fb_stream_t stream) {
  int rc = dp_check(p);
  if (rc) return rc;
  FbDpStepParams local = *p;
  void* args[] = {&local};
  FB_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)dp_kernel, dim3(dp_grid(p->n, p->G, 1, p->max_blocks)), dim3(DP_THREADS), args, 0,
                                            (cudaStream_t)stream));
  return FB_OK;
}
extern "C" int fb_dp_allreduce_step_emulated(const FbDpStepParams* params_dev, int32_t G, int64_t n, fb_stream_t stream) {
  FB_REQUIRE(params_dev && G >= 1 && G <= 16 && n > 0 && n % 4 == 0, "fb_dp_allreduce_step_emulated: bad arguments");
  void* args[] = {&params_dev};
  FB_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)dp_kernel_emulated, dim3(dp_grid(n, G, G), G), dim3(DP_THREADS),
                                            args, 0, (cudaStream_t)stream));
  return FB_OK;
}
