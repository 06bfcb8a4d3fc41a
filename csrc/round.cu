// K1+K2+K3+K4: the fused FL round kernel and its stand-alone pieces.  See include/fedml_b200.h.
//
// Work decomposition (identical on every rank, which is what makes per-block cross-rank flags
// sufficient — no grid-wide or device-wide barrier is needed on the FedAvg/FedOpt path):
//   the model is cut into G shards of S floats; block b owns, inside EVERY shard, the float4
//   lanes v with (v / blockDim) % nblocks == b.  Phase 1 (local weighted accumulate into
//   `partial`) walks all shards; phase 3 (cross-rank reduce + epilogue) walks all shards in
//   one-shot mode and only shard `rank` in two-shot mode.  Either way the data block b pulls
//   from rank q in phase 3 was produced by block b of rank q in phase 1.
#include "common.cuh"
#include "../include/fedml_b200.h"
#include <cooperative_groups.h>
#include <stdarg.h>

namespace cg = cooperative_groups;

static thread_local char g_err[512] = "";
void fb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* fb_last_error(void) { return g_err; }
extern "C" int fb_version(void) { return 10001; }
extern "C" int fb_device_sm_count(void) { return fb_num_sms(); }

#define ROUND_THREADS 256

__device__ __forceinline__ float4 ld_relaxed_sys4(const float* p) {
  float4 r;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// ---------------------------------------------------------------- server-optimizer epilogue
struct OptCfg {
  int opt, nesterov;
  float lr, momentum, beta1, beta2, eps, bc1, bc2, grad_scale;
};

__device__ __forceinline__ float opt_apply(const OptCfg& o, float w, float avg, float* m, float* v) {
  // m, v point at this element's moment slots (may be nullptr when unused)
  float g = (w - avg) * o.grad_scale;
  switch (o.opt) {
    case FB_OPT_SGD: {
      if (o.momentum != 0.f) {
        float buf = o.momentum * (*m) + g;
        *m = buf;
        g = o.nesterov ? g + o.momentum * buf : buf;
      }
      return w - o.lr * g;
    }
    case FB_OPT_ADAGRAD: {
      float vv = *v + g * g;
      *v = vv;
      return w - o.lr * g / (sqrtf(vv) + o.eps);
    }
    case FB_OPT_ADAM:
    case FB_OPT_YOGI: {
      float mm = o.beta1 * (*m) + (1.f - o.beta1) * g;
      float g2 = g * g, vo = *v, vv;
      if (o.opt == FB_OPT_ADAM) {
        vv = o.beta2 * vo + (1.f - o.beta2) * g2;
      } else {
        float d = vo - g2;
        float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        vv = vo - (1.f - o.beta2) * g2 * sgn;
      }
      *m = mm;
      *v = vv;
      float denom = sqrtf(vv) * rsqrtf(o.bc2) + o.eps;
      return w - (o.lr / o.bc1) * mm / denom;
    }
    default:
      return avg;
  }
}

__device__ __forceinline__ float4 opt_apply4(const OptCfg& o, float4 w, float4 avg, float* m, float* v, int64_t j) {
  if (o.opt == FB_OPT_NONE) return avg;
  float4 mm = make_float4(0, 0, 0, 0), vv = make_float4(0, 0, 0, 0);
  const bool use_m = m != nullptr, use_v = v != nullptr;
  if (use_m) mm = *reinterpret_cast<const float4*>(m + j);
  if (use_v) vv = *reinterpret_cast<const float4*>(v + j);
  float4 r;
  r.x = opt_apply(o, w.x, avg.x, &mm.x, &vv.x);
  r.y = opt_apply(o, w.y, avg.y, &mm.y, &vv.y);
  r.z = opt_apply(o, w.z, avg.z, &mm.z, &vv.z);
  r.w = opt_apply(o, w.w, avg.w, &mm.w, &vv.w);
  if (use_m) *reinterpret_cast<float4*>(m + j) = mm;
  if (use_v) *reinterpret_cast<float4*>(v + j) = vv;
  return r;
}

// ---------------------------------------------------------------- cross-rank per-block flags
// flags layout on every rank: uint32 [2 phases][G senders][FB_ROUND_MAX_BLOCKS]
// Race-detection aid (tests/test_gpu_round_hardening.py): with fb_round_set_skew(seed, max_ns) every (rank, block, phase)
// stalls a pseudo-random time below max_ns before it signals, so ranks and blocks reach the handshakes in shuffled order;
// the flag protocol must give the same bits whatever the interleaving.  max_ns = 0 (default): one load per handshake.
static __device__ unsigned g_fb_round_skew[2] = {0u, 0u};     // seed, max_ns
__device__ __forceinline__ void skew_delay(int phase, int rank, int block) {
  const unsigned max_ns = *reinterpret_cast<volatile unsigned*>(&g_fb_round_skew[1]);
  if (max_ns == 0u) return;
  if (threadIdx.x == 0) {
    const Philox4 r = philox4x32_10((uint32_t)block, (uint32_t)rank, (uint32_t)phase, 0x5EEDu, g_fb_round_skew[0], 0x9E3779B9u);
    const unsigned long long wait = r.x % max_ns, t0 = fb_globaltimer();
    while (fb_globaltimer() - t0 < wait) {
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void block_signal_and_wait(uint32_t* const* peer_flags, int phase, int rank, int G,
                                                      int block, uint32_t epoch) {
  skew_delay(phase, rank, block);
  __syncthreads();  // all of this block's data stores are issued
  if ((int)threadIdx.x < G) {
    const int q = threadIdx.x;
    __threadfence_system();
    uint32_t* remote = peer_flags[q] + ((size_t)phase * G + rank) * FB_ROUND_MAX_BLOCKS + block;
    st_release_sys(remote, epoch);
    const uint32_t* mine = peer_flags[rank] + ((size_t)phase * G + q) * FB_ROUND_MAX_BLOCKS + block;
    fb_spin_wait_ge<true>(mine, epoch, 1u + (uint32_t)phase, (uint32_t)q);     // bounded: see common.cuh
  }
  __syncthreads();
}

// intra-rank barrier over the blocks of ONE rank (monotonic counter, no reset needed)
__device__ __forceinline__ void rank_grid_barrier(uint32_t* counter, uint32_t target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    fb_spin_wait_ge<false>(counter, target, 3u, 0u);
  }
  __syncthreads();
}

// sum_c wt_c * contribution_c[j..j+3] for this rank's clients (phase 1 of the round)
__device__ __forceinline__ float4 local_accumulate(const FbRoundParams& p, const int64_t j, const int64_t v,
                                                   const bool dp, const bool clip, const bool noise,
                                                   const float* wts, const float* wts_b, const float* s_scale) {
  const int C = p.C;
  const int64_t Pt = p.P_train;
  const bool train = j < Pt;
  const float* wt = train ? wts : wts_b;
  float4 acc = make_float4(0, 0, 0, 0);
  if (!dp || !train) {
    int c = 0;
    for (; c + 4 <= C; c += 4) {  // 4 independent 128-bit loads in flight
      float4 x0 = ld_stream4(p.clients + (size_t)(c + 0) * p.ld + j);
      float4 x1 = ld_stream4(p.clients + (size_t)(c + 1) * p.ld + j);
      float4 x2 = ld_stream4(p.clients + (size_t)(c + 2) * p.ld + j);
      float4 x3 = ld_stream4(p.clients + (size_t)(c + 3) * p.ld + j);
      float w0 = __ldg(wt + c), w1 = __ldg(wt + c + 1), w2 = __ldg(wt + c + 2), w3 = __ldg(wt + c + 3);
      acc.x += w0 * x0.x; acc.y += w0 * x0.y; acc.z += w0 * x0.z; acc.w += w0 * x0.w;
      acc.x += w1 * x1.x; acc.y += w1 * x1.y; acc.z += w1 * x1.z; acc.w += w1 * x1.w;
      acc.x += w2 * x2.x; acc.y += w2 * x2.y; acc.z += w2 * x2.z; acc.w += w2 * x2.w;
      acc.x += w3 * x3.x; acc.y += w3 * x3.y; acc.z += w3 * x3.z; acc.w += w3 * x3.w;
    }
    for (; c < C; ++c) {
      float4 x = ld_stream4(p.clients + (size_t)c * p.ld + j);
      float w = __ldg(wt + c);
      acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
    }
  } else {
    const float4 g = *reinterpret_cast<const float4*>(p.global_model + j);
    for (int c0 = 0; c0 < C; c0 += 4) {
      const int nc = min(4, C - c0);
      float4 x[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (u < nc) x[u] = ld_stream4(p.clients + (size_t)(c0 + u) * p.ld + j);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u < nc) {
          const int c = c0 + u;
          const float sc = clip ? s_scale[c] : 1.f;
          float4 y = make_float4(g.x + sc * (x[u].x - g.x), g.y + sc * (x[u].y - g.y), g.z + sc * (x[u].z - g.z),
                                 g.w + sc * (x[u].w - g.w));
          if (noise) {
            float4 z = philox_normal4(p.seed, p.round, (uint32_t)(p.client_base + c), (uint64_t)v);
            y.x += p.noise_std * z.x; y.y += p.noise_std * z.y; y.z += p.noise_std * z.z; y.w += p.noise_std * z.w;
          }
          const float w = __ldg(wt + c);
          acc.x += w * y.x; acc.y += w * y.y; acc.z += w * y.z; acc.w += w * y.w;
        }
      }
    }
  }
  return acc;
}

// ---------------------------------------------------------------- the round body
__device__ void round_body(const FbRoundParams& p, const int block, const int nblocks, float* red /*smem 33+*/,
                           float* s_scale /*smem [C]*/) {
  const int G = p.G, rank = p.rank, C = p.C;
  const int64_t P = p.P, Pt = p.P_train;
  const int64_t nvec = P / 4;
  const int64_t S4 = (nvec + G - 1) / G;  // float4 lanes per shard
  const int tid = threadIdx.x, T = blockDim.x;
  const int64_t stride = (int64_t)nblocks * T;
  const bool clip = p.clip_bound > 0.f, noise = p.noise_std > 0.f;
  const bool dp = clip || noise;
  const float* wts = p.weights;
  const float* wts_b = p.weights_buf ? p.weights_buf : p.weights;

  // ---- phase 0: per-client clip factors s_c = 1 / max(1, ||x_c - w||_2 / bound) over [0, Pt)
  if (clip && C > 0) {
    for (int c0 = 0; c0 < C; c0 += 4) {     // four client rows per sweep: four independent 128-bit loads in flight
      const int nc = min(4, C - c0);
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int s = 0; s < G; ++s) {
        const int64_t lo = (int64_t)s * S4, hi = min(lo + S4, nvec);
        for (int64_t v = lo + (int64_t)block * T + tid; v < hi; v += stride) {
          const int64_t j = v * 4;
          if (j < Pt) {
            const float4 w = *reinterpret_cast<const float4*>(p.global_model + j);
            float4 x[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (u < nc) x[u] = ld_stream4(p.clients + (size_t)(c0 + u) * p.ld + j);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              if (u < nc) {
                const float a = x[u].x - w.x, b = x[u].y - w.y, cc = x[u].z - w.z, d = x[u].w - w.w;
                acc[u] += a * a + b * b + cc * cc + d * d;
              }
            }
          }
        }
      }
      for (int u = 0; u < nc; ++u) {
        const float tot = block_sum(acc[u], red);
        if (tid == 0) p.sqnorm_part[(size_t)block * C + c0 + u] = tot;
      }
    }
    rank_grid_barrier(p.grid_bar, p.grid_bar_target);
    for (int c = tid; c < C; c += T) {
      float tot = 0.f;
      for (int b = 0; b < nblocks; ++b) tot += __ldcg(p.sqnorm_part + (size_t)b * C + c);  // fixed order
      float sc = 1.f / fmaxf(1.f, sqrtf(tot) / p.clip_bound);
      s_scale[c] = sc;
      if (block == 0 && p.clip_scale_out) p.clip_scale_out[c] = sc;
    }
    __syncthreads();
  }

  // ---- phase 1: partial[j] = sum_c wt_c * contribution_c[j]   (skipped when G == 1: fused into phase 3)
  if (G > 1) {
    for (int s = 0; s < G; ++s) {
      const int64_t lo = (int64_t)s * S4, hi = min(lo + S4, nvec);
      for (int64_t v = lo + (int64_t)block * T + tid; v < hi; v += stride) {
        const int64_t j = v * 4;
        st4(p.partial + j, local_accumulate(p, j, v, dp, clip, noise, wts, wts_b, s_scale));
      }
    }
  }

  // ---- phase 2: block-level cross-rank handshake
  if (G > 1) block_signal_and_wait(p.peer_flags, 0, rank, G, block, p.epoch);

  // ---- phase 3: reduce over ranks (fixed order => bit-identical on every rank), epilogue, store
  OptCfg o;
  o.opt = p.opt; o.nesterov = p.nesterov; o.lr = p.lr; o.momentum = p.momentum; o.beta1 = p.beta1; o.beta2 = p.beta2;
  o.eps = p.eps; o.bc1 = p.bias_corr1; o.bc2 = p.bias_corr2; o.grad_scale = p.grad_scale;
  float* mom = (p.opt == FB_OPT_SGD && p.momentum == 0.f) ? nullptr : p.m;
  float* sec = (p.opt == FB_OPT_ADAM || p.opt == FB_OPT_YOGI || p.opt == FB_OPT_ADAGRAD) ? p.v : nullptr;
  if (p.opt == FB_OPT_ADAGRAD) mom = nullptr;
  const int s_begin = p.two_shot ? rank : 0, s_end = p.two_shot ? rank + 1 : G;
  const bool use_mc = G > 1 && p.two_shot && p.mc_partial != nullptr && p.mc_global != nullptr;
  const float* pp[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) pp[q] = (G > 1 && q < G) ? p.peer_partial[q] : nullptr;
  // averaged sum -> (aggregate noise) -> server optimizer -> value to store, for the float4 lane v
  auto finish = [&](const int64_t v, const float4 tot) -> float4 {
    const int64_t j = v * 4;
    const bool train = j < Pt;
    const float inv = train ? p.inv_total : p.inv_total_buf;
    float4 avg = make_float4(tot.x * inv, tot.y * inv, tot.z * inv, tot.w * inv);
    if (train && p.agg_noise_std > 0.f) {
      // one Gaussian draw on the aggregate: identical in distribution to summing the per-client draws
      const float4 z = philox_normal4(p.seed, p.round, 0x00A66E6u, (uint64_t)v);
      avg.x += p.agg_noise_std * z.x; avg.y += p.agg_noise_std * z.y; avg.z += p.agg_noise_std * z.z; avg.w += p.agg_noise_std * z.w;
    }
    float4 out = avg;
    if (train && (p.opt != FB_OPT_NONE || p.self_coef != 0.f)) {
      const float4 w = *reinterpret_cast<const float4*>(p.global_model + j);
      if (p.self_coef != 0.f) {
        avg.x += p.self_coef * w.x; avg.y += p.self_coef * w.y; avg.z += p.self_coef * w.z; avg.w += p.self_coef * w.w;
      }
      out = opt_apply4(o, w, avg, mom, sec, j);
    }
    return out;
  };
  for (int s = s_begin; s < s_end; ++s) {
    const int64_t lo = (int64_t)s * S4, hi = min(lo + S4, nvec);
    if (use_mc) {
      // NVLS: the switch reduces the G replicas of a 16-byte word and returns the sum (one load instead of G) and
      // multimem.st writes the result into all G replicas.  Four independent reductions are in flight per thread before
      // the first epilogue: a reduction is a full NVLink round trip (~ 3-4 us), so bytes in flight, not issue rate, set
      // the bandwidth (2 CTAs x 256 threads x 4 x 16 B = 32 KiB per SM).
      constexpr int U = 4;
      for (int64_t v0 = lo + (int64_t)block * T + tid; v0 < hi; v0 += (int64_t)U * stride) {
        float4 t[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t v = v0 + (int64_t)u * stride;
          if (v < hi)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(t[u].x), "=f"(t[u].y), "=f"(t[u].z), "=f"(t[u].w)
                         : "l"(p.mc_partial + v * 4)
                         : "memory");
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t v = v0 + (int64_t)u * stride;
          if (v < hi) {
            const float4 out = finish(v, t[u]);
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p.mc_global + v * 4), "f"(out.x),
                         "f"(out.y), "f"(out.z), "f"(out.w)
                         : "memory");
          }
        }
      }
      continue;
    }
    for (int64_t v = lo + (int64_t)block * T + tid; v < hi; v += stride) {
      const int64_t j = v * 4;
      float4 tot = make_float4(0, 0, 0, 0);
      if (G == 1) {
        tot = local_accumulate(p, j, v, dp, clip, noise, wts, wts_b, s_scale);
      } else if (G <= 8) {
        // all peer loads are issued before the first add so that G NVLink round trips overlap
        float4 xs[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < G) xs[q] = (q == rank) ? *reinterpret_cast<const float4*>(p.partial + j) : ld_relaxed_sys4(pp[q] + j);
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < G) { tot.x += xs[q].x; tot.y += xs[q].y; tot.z += xs[q].z; tot.w += xs[q].w; }
        }
      } else {
        for (int q = 0; q < G; ++q) {
          float4 x = (q == rank) ? *reinterpret_cast<const float4*>(p.partial + j)
                                 : ld_relaxed_sys4(p.peer_partial[q] + j);
          tot.x += x.x; tot.y += x.y; tot.z += x.z; tot.w += x.w;
        }
      }
      const float4 out = finish(v, tot);
      if (p.two_shot && G > 1) {
        for (int q = 0; q < G; ++q) st4((q == rank ? p.global_model : p.peer_global[q]) + j, out);
      } else {
        st4(p.global_model + j, out);
      }
    }
  }

  // ---- phase 4: completion handshake.  Two-shot: peers' pushes into my global model must have
  // landed before this kernel retires.  One-shot: peers must be done *reading* my partial before a
  // later launch overwrites it.
  if (G > 1) block_signal_and_wait(p.peer_flags, 1, rank, G, block, p.epoch);
}

__global__ void __launch_bounds__(ROUND_THREADS, 2) round_kernel(const FbRoundParams p) {
  extern __shared__ float smem[];
  round_body(p, blockIdx.x, gridDim.x, smem, smem + 64);
}

__global__ void __launch_bounds__(ROUND_THREADS, 2) round_kernel_emulated(const FbRoundParams* __restrict__ ps) {
  extern __shared__ float smem[];
  __shared__ FbRoundParams sp;
  if (threadIdx.x == 0) sp = ps[blockIdx.y];
  __syncthreads();
  round_body(sp, blockIdx.x, gridDim.x, smem, smem + 64);
}

static int round_max_blocks(const void* fn, size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, ROUND_THREADS, smem) != cudaSuccess || per_sm < 1)
    per_sm = 1;
  return per_sm * fb_num_sms();
}

extern "C" int fb_round_grid(int64_t P, int32_t G, int32_t emulated_ranks) {
  if (G < 1) G = 1;
  const int64_t nvec = P / 4;
  const int64_t S4 = (nvec + G - 1) / G;
  int64_t want = (S4 + ROUND_THREADS - 1) / ROUND_THREADS;  // one float4 per thread per shard
  int cap = 2 * fb_num_sms();                                // 2 CTAs of 256 threads per SM is always resident
  if (emulated_ranks > 1) cap = cap / emulated_ranks;
  if (cap > FB_ROUND_MAX_BLOCKS) cap = FB_ROUND_MAX_BLOCKS;
  if (cap < 1) cap = 1;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

extern "C" int64_t fb_round_flag_bytes(int32_t G) {
  return (int64_t)2 * (G < 1 ? 1 : G) * FB_ROUND_MAX_BLOCKS * sizeof(uint32_t);
}

static int round_check(const FbRoundParams* p) {
  FB_REQUIRE(p != nullptr, "fb_round: null params");
  FB_REQUIRE(p->P > 0 && p->P % 4 == 0, "fb_round: P=%lld must be a positive multiple of 4", (long long)p->P);
  FB_REQUIRE(p->P_train >= 0 && p->P_train <= p->P && p->P_train % 4 == 0, "fb_round: bad P_train");
  FB_REQUIRE(p->C >= 0 && (p->C == 0 || (p->clients && p->weights)), "fb_round: clients/weights missing");
  FB_REQUIRE(p->C == 0 || (p->ld >= p->P && p->ld % 4 == 0), "fb_round: ld must be >= P and a multiple of 4");
  FB_REQUIRE(p->C <= 4096, "fb_round: at most 4096 clients per rank");
  FB_REQUIRE(p->global_model && p->partial, "fb_round: global/partial missing");
  FB_REQUIRE(p->G >= 1 && p->G <= 64 && p->rank >= 0 && p->rank < p->G, "fb_round: bad rank/G");
  FB_REQUIRE(p->G == 1 || (p->peer_partial && p->peer_global && p->peer_flags), "fb_round: peer tables missing");
  FB_REQUIRE(p->clip_bound <= 0.f || p->C == 0 || (p->sqnorm_part && p->grid_bar),
             "fb_round: clipping needs sqnorm_part and grid_bar scratch");
  if (p->opt == FB_OPT_ADAM || p->opt == FB_OPT_YOGI) FB_REQUIRE(p->m && p->v, "fb_round: adam/yogi need m and v");
  if (p->opt == FB_OPT_ADAGRAD) FB_REQUIRE(p->v, "fb_round: adagrad needs v");
  if (p->opt == FB_OPT_SGD && p->momentum != 0.f) FB_REQUIRE(p->m, "fb_round: sgd momentum needs m");
  return FB_OK;
}

extern "C" int fb_round_launch(const FbRoundParams* p, fb_stream_t stream) {
  int rc = round_check(p);
  if (rc) return rc;
  const int grid = fb_round_grid(p->P, p->G, 1);
  const size_t smem = (64 + (size_t)p->C) * sizeof(float);
  FbRoundParams local = *p;
  void* args[] = {&local};
  // cooperative launch: guarantees all blocks are co-resident, which the spin-waits require
  FB_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)round_kernel, dim3(grid), dim3(ROUND_THREADS), args, smem,
                                            (cudaStream_t)stream));
  return FB_OK;
}

extern "C" int fb_round_launch_emulated(const FbRoundParams* params_dev, int32_t G, int64_t P, fb_stream_t stream) {
  FB_REQUIRE(params_dev && G >= 1 && G <= 16, "fb_round_launch_emulated: bad arguments");
  const int grid = fb_round_grid(P, G, G);
  const size_t smem = (64 + 4096) * sizeof(float);
  FB_REQUIRE(grid * G <= round_max_blocks((const void*)round_kernel_emulated, smem),
             "fb_round_launch_emulated: grid does not fit co-resident");
  void* args[] = {&params_dev};
  FB_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)round_kernel_emulated, dim3(grid, G), dim3(ROUND_THREADS),
                                            args, smem, (cudaStream_t)stream));
  return FB_OK;
}

// ---------------------------------------------------------------- K4 stand-alone
template <bool BF16>
__global__ void __launch_bounds__(256) accumulate_kernel(const void* __restrict__ clients, int64_t ld, int C,
                                                         const float* __restrict__ weights, float scale, float beta,
                                                         float* __restrict__ out, int64_t nvec) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = v * 4;
    float4 acc = make_float4(0, 0, 0, 0);
    if (!BF16) {
      const float* base = reinterpret_cast<const float*>(clients);
      int c = 0;
      for (; c + 4 <= C; c += 4) {
        float4 x0 = ld_stream4(base + (size_t)(c + 0) * ld + j), x1 = ld_stream4(base + (size_t)(c + 1) * ld + j);
        float4 x2 = ld_stream4(base + (size_t)(c + 2) * ld + j), x3 = ld_stream4(base + (size_t)(c + 3) * ld + j);
        float w0 = __ldg(weights + c), w1 = __ldg(weights + c + 1), w2 = __ldg(weights + c + 2), w3 = __ldg(weights + c + 3);
        acc.x += w0 * x0.x; acc.y += w0 * x0.y; acc.z += w0 * x0.z; acc.w += w0 * x0.w;
        acc.x += w1 * x1.x; acc.y += w1 * x1.y; acc.z += w1 * x1.z; acc.w += w1 * x1.w;
        acc.x += w2 * x2.x; acc.y += w2 * x2.y; acc.z += w2 * x2.z; acc.w += w2 * x2.w;
        acc.x += w3 * x3.x; acc.y += w3 * x3.y; acc.z += w3 * x3.z; acc.w += w3 * x3.w;
      }
      for (; c < C; ++c) {
        float4 x = ld_stream4(base + (size_t)c * ld + j);
        float w = __ldg(weights + c);
        acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
      }
    } else {
      const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(clients);
      for (int c = 0; c < C; ++c) {
        const uint2 raw = __ldg(reinterpret_cast<const uint2*>(base + (size_t)c * ld + j));
        const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
        const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
        const float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
        float w = __ldg(weights + c);
        acc.x += w * a.x; acc.y += w * a.y; acc.z += w * b.x; acc.w += w * b.y;
      }
    }
    float4 r = make_float4(acc.x * scale, acc.y * scale, acc.z * scale, acc.w * scale);
    if (beta != 0.f) {
      float4 o = *reinterpret_cast<const float4*>(out + j);
      r.x += beta * o.x; r.y += beta * o.y; r.z += beta * o.z; r.w += beta * o.w;
    }
    st4(out + j, r);
  }
}

static inline int stream_grid(int64_t nvec, int threads) {
  int64_t want = (nvec + threads - 1) / threads;
  int64_t cap = (int64_t)fb_num_sms() * 8;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

extern "C" int fb_weighted_accumulate(const float* clients, int64_t ld, int32_t C, const float* weights, float scale,
                                      float beta, float* out, int64_t P, fb_stream_t stream) {
  FB_REQUIRE(P > 0 && P % 4 == 0 && ld % 4 == 0 && C >= 0 && out, "fb_weighted_accumulate: bad shape");
  accumulate_kernel<false><<<stream_grid(P / 4, 256), 256, 0, (cudaStream_t)stream>>>(clients, ld, C, weights, scale,
                                                                                      beta, out, P / 4);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
extern "C" int fb_weighted_accumulate_bf16(const void* clients, int64_t ld, int32_t C, const float* weights,
                                           float scale, float beta, float* out, int64_t P, fb_stream_t stream) {
  FB_REQUIRE(P > 0 && P % 4 == 0 && ld % 4 == 0 && C >= 0 && out, "fb_weighted_accumulate_bf16: bad shape");
  accumulate_kernel<true><<<stream_grid(P / 4, 256), 256, 0, (cudaStream_t)stream>>>(clients, ld, C, weights, scale,
                                                                                     beta, out, P / 4);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

__global__ void __launch_bounds__(256) broadcast_rows_kernel(const float* __restrict__ src, float* __restrict__ rows,
                                                             int64_t ld, int C, int64_t nvec) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const float4 x = *reinterpret_cast<const float4*>(src + v * 4);
    for (int c = 0; c < C; ++c) st4(rows + (size_t)c * ld + v * 4, x);
  }
}
extern "C" int fb_broadcast_rows(const float* src, float* rows, int64_t ld, int32_t C, int64_t P, fb_stream_t stream) {
  FB_REQUIRE(P > 0 && P % 4 == 0 && ld % 4 == 0 && ld >= P && C >= 0, "fb_broadcast_rows: bad shape");
  if (C == 0) return FB_OK;
  broadcast_rows_kernel<<<stream_grid(P / 4, 256), 256, 0, (cudaStream_t)stream>>>(src, rows, ld, C, P / 4);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- K3 stand-alone
__global__ void __launch_bounds__(256) server_opt_kernel(float* __restrict__ w, const float* __restrict__ avg,
                                                         float* m, float* v, int64_t nvec, OptCfg o) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = i * 4;
    float4 ww = *reinterpret_cast<const float4*>(w + j);
    float4 aa = *reinterpret_cast<const float4*>(avg + j);
    st4(w + j, opt_apply4(o, ww, aa, m, v, j));
  }
}
extern "C" int fb_server_opt(float* w, const float* avg, float* m, float* v, int64_t n, int32_t opt, int32_t nesterov,
                             float lr, float momentum, float beta1, float beta2, float eps, float bias_corr1,
                             float bias_corr2, float grad_scale, fb_stream_t stream) {
  FB_REQUIRE(w && avg && n > 0 && n % 4 == 0, "fb_server_opt: bad shape");
  OptCfg o;
  o.opt = opt; o.nesterov = nesterov; o.lr = lr; o.momentum = momentum; o.beta1 = beta1; o.beta2 = beta2; o.eps = eps;
  o.bc1 = bias_corr1; o.bc2 = bias_corr2; o.grad_scale = grad_scale;
  float* mm = (opt == FB_OPT_SGD && momentum == 0.f) || opt == FB_OPT_ADAGRAD || opt == FB_OPT_NONE ? nullptr : m;
  float* vv = (opt == FB_OPT_ADAM || opt == FB_OPT_YOGI || opt == FB_OPT_ADAGRAD) ? v : nullptr;
  if (opt == FB_OPT_ADAM || opt == FB_OPT_YOGI) FB_REQUIRE(mm && vv, "fb_server_opt: adam/yogi need m and v");
  server_opt_kernel<<<stream_grid(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(w, avg, mm, vv, n / 4, o);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- K2 stand-alone pieces
__global__ void __launch_bounds__(256) sqnorm_diff_kernel(const float* __restrict__ clients, int64_t ld,
                                                          const float* __restrict__ ref, int64_t nvec,
                                                          float* __restrict__ out) {
  __shared__ float red[33];
  const int c = blockIdx.y;
  const float* row = clients + (size_t)c * ld;
  float acc = 0.f;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    float4 x = ld_stream4(row + v * 4);
    float4 w = *reinterpret_cast<const float4*>(ref + v * 4);
    float a = x.x - w.x, b = x.y - w.y, cc = x.z - w.z, d = x.w - w.w;
    acc += a * a + b * b + cc * cc + d * d;
  }
  float tot = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(out + c, tot);
}
extern "C" int fb_sqnorm_diff(const float* clients, int64_t ld, int32_t C, const float* ref, int64_t n, float* out_sq,
                              fb_stream_t stream) {
  FB_REQUIRE(clients && ref && out_sq && n > 0 && n % 4 == 0 && C > 0 && C <= 65535, "fb_sqnorm_diff: bad shape");
  FB_CHECK_CUDA(cudaMemsetAsync(out_sq, 0, sizeof(float) * C, (cudaStream_t)stream));
  int gx = stream_grid(n / 4, 256);
  int lim = (fb_num_sms() * 8 + C - 1) / C;
  if (gx > lim) gx = lim < 1 ? 1 : lim;
  sqnorm_diff_kernel<<<dim3(gx, C), 256, 0, (cudaStream_t)stream>>>(clients, ld, ref, n / 4, out_sq);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

__global__ void __launch_bounds__(256) philox_normal_kernel(float* out, int64_t nvec, uint64_t seed, uint32_t s0,
                                                            uint32_t s1) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x)
    st4(out + v * 4, philox_normal4(seed, s0, s1, (uint64_t)v));
}
extern "C" int fb_philox_normal(float* out, int64_t n, uint64_t seed, uint32_t stream_id, uint32_t substream,
                                fb_stream_t stream) {
  FB_REQUIRE(out && n > 0 && n % 4 == 0, "fb_philox_normal: n must be a positive multiple of 4");
  philox_normal_kernel<<<stream_grid(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(out, n / 4, seed, stream_id,
                                                                                  substream);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- stand-alone cross-rank barrier
// flags layout: uint32 [8 slots][G]
__global__ void peer_barrier_kernel(uint32_t* const* peer_flags, int rank, int G, uint32_t epoch, int slot) {
  if ((int)threadIdx.x < G) {
    const int q = threadIdx.x;
    __threadfence_system();
    st_release_sys(peer_flags[q] + (size_t)slot * G + rank, epoch);
    const uint32_t* mine = peer_flags[rank] + (size_t)slot * G + q;
    fb_spin_wait_ge<true>(mine, epoch, 4u, (uint32_t)q);
  }
}
FB_DEFINE_SPIN_API(fb_round)

extern "C" int fb_round_set_skew(uint32_t seed, uint32_t max_ns) {
  const unsigned v[2] = {seed, max_ns};
  FB_CHECK_CUDA(cudaMemcpyToSymbol(g_fb_round_skew, v, sizeof(v)));
  return FB_OK;
}
extern "C" int64_t fb_peer_barrier_flag_bytes(int32_t G) { return (int64_t)8 * (G < 1 ? 1 : G) * sizeof(uint32_t); }
extern "C" int fb_peer_barrier(uint32_t* const* peer_flags, int32_t rank, int32_t G, uint32_t epoch, int32_t slot,
                               fb_stream_t stream) {
  FB_REQUIRE(peer_flags && G >= 1 && G <= 64 && rank >= 0 && rank < G && slot >= 0 && slot < 8,
             "fb_peer_barrier: bad arguments");
  peer_barrier_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(peer_flags, rank, G, epoch, slot);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}

// ---------------------------------------------------------------- allocation-free single-GPU round
// Same kernel as fb_round_launch on one rank, but the scratch it needs for clipping (per-block
// squared-norm partials and the intra-grid barrier counter) lives in a fixed device array that is
// part of this library's image instead of being supplied by the caller.  See the declaration of
// fb_round_single_noalloc in include/fedml_b200.h for who this is for and what the limits are.
#define FB_NOALLOC_MAX_CLIENTS 64
#define FB_NOALLOC_COUNTER_FLOATS 64  /* the barrier counter gets its own 256-byte line at the front */
#define FB_NOALLOC_SCRATCH_FLOATS (FB_NOALLOC_COUNTER_FLOATS + FB_ROUND_MAX_BLOCKS * FB_NOALLOC_MAX_CLIENTS)

static __device__ __align__(256) float g_noalloc_scratch[FB_NOALLOC_SCRATCH_FLOATS];  // 256.25 KiB on every GPU that loads the library

// Device address of g_noalloc_scratch per device.  In host code the symbol name is only a handle for
// the runtime; the address a kernel can dereference has to be asked for, with that device current.
static float* g_noalloc_ptr[64] = {nullptr};

static int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  FB_CHECK_CUDA(cudaGetDevice(&dev));
  FB_REQUIRE(dev >= 0 && dev < 64, "fb_round_single_noalloc: device ordinal %d not supported", dev);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    FB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    FB_REQUIRE(p != nullptr, "fb_round_single_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return FB_OK;
}

extern "C" int fb_round_single_noalloc_max_clients(void) { return FB_NOALLOC_MAX_CLIENTS; }

extern "C" int fb_round_single_noalloc(const float* clients, int64_t ld, int32_t C, const float* weights,
                                       float inv_total, float clip_bound, float* global_model, int64_t P,
                                       int64_t P_train, float* clip_scale_out, fb_stream_t stream) {
  if (C > FB_NOALLOC_MAX_CLIENTS) {
    fb_set_error("fb_round_single_noalloc: %d clients exceed the built-in scratch limit of %d; use fb_round_launch "
                 "with caller-provided scratch for larger rounds", C, FB_NOALLOC_MAX_CLIENTS);
    return FB_ERR_LIMIT;
  }
  FB_REQUIRE(C >= 1, "fb_round_single_noalloc: need at least one client");
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  FbRoundParams p;
  memset(&p, 0, sizeof(p));
  p.clients = clients; p.ld = ld; p.C = C; p.weights = weights;
  p.P = P; p.P_train = P_train;
  p.inv_total = inv_total; p.inv_total_buf = inv_total;
  p.global_model = global_model;
  p.partial = global_model;   // never touched when G == 1; only has to be non-null for the argument check
  p.rank = 0; p.G = 1; p.epoch = 1;
  p.clip_bound = clip_bound;
  p.opt = FB_OPT_NONE; p.grad_scale = 1.f; p.bias_corr1 = 1.f; p.bias_corr2 = 1.f;
  p.grid_bar = reinterpret_cast<uint32_t*>(scratch);
  p.sqnorm_part = scratch + FB_NOALLOC_COUNTER_FLOATS;
  p.clip_scale_out = clip_scale_out;
  rc = round_check(&p);
  if (rc) return rc;
  const int grid = fb_round_grid(P, 1, 1);
  FB_REQUIRE(grid <= FB_ROUND_MAX_BLOCKS, "fb_round_single_noalloc: grid exceeds the scratch layout");
  p.grid_bar_target = (uint32_t)grid;
  // The array remembers the previous call: the barrier counter has to start from zero every time.  The
  // per-block partials are fully rewritten by the kernel before it reads them, so they need no clearing.
  FB_CHECK_CUDA(cudaMemsetAsync(scratch, 0, FB_NOALLOC_COUNTER_FLOATS * sizeof(float), (cudaStream_t)stream));
  const size_t smem = (64 + (size_t)C) * sizeof(float);
  void* args[] = {&p};
  FB_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)round_kernel, dim3(grid), dim3(ROUND_THREADS), args, smem,
                                            (cudaStream_t)stream));
  return FB_OK;
}
