// K14: HS-FedAvg amplitude normalisation (reference single_process/hs_fedavg/hs_fft.py:8-84, NumPy complex128 on
// the CPU): per image channel a 2-D DFT, the amplitude spectrum is replaced by the (running) batch-mean amplitude,
// the phase is kept, inverse DFT.  See include/fedml_b200.h.
//
// Hand-written radix-2 FFTs in shared memory (no cuFFT): a 2-D transform is a row pass and a column pass.
//   fft_rows_kernel   one CTA transforms 512 / W rows at a time; real input is widened to complex on load, the
//                     inverse pass writes only the real part, scaled by 1 / (H*W)
//   fft_cols_kernel   one CTA transforms a tile of 8 adjacent columns (64-byte row segments keep the strided
//                     column traffic sector-aligned), in place in the complex workspace
//   amp_kernel        batch-mean amplitude per (channel, y, x), running-amplitude update, spectrum rebuild
// H and W are powers of two in [8, 512]; twiddles come from one sincospif table per CTA.
#include "common.cuh"
#include "../include/fedml_b200.h"

#define FFT_THREADS 256
#define FFT_COL_TILE 8

namespace {

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// tw[k] = exp(sign * 2*pi*i * k / N), k < N/2
__device__ __forceinline__ void make_twiddles(float2* tw, int N, float sign) {
  for (int k = threadIdx.x; k < N / 2; k += blockDim.x) {
    float s, c;
    sincospif(2.f * (float)k / (float)N, &s, &c);
    tw[k] = make_float2(c, sign * s);
  }
}

// nb independent length-N transforms stored bit-reversed in a[b * stride + i]; iterative radix-2 butterflies
__device__ __forceinline__ void fft_smem(float2* a, int nb, int stride, int N, int logN, const float2* tw) {
  const int half = N >> 1, total = nb * half;
  for (int s = 1; s <= logN; ++s) {
    const int m2 = 1 << (s - 1);          // butterfly span
    const int tstep = N >> s;             // twiddle stride
    __syncthreads();
    for (int t = threadIdx.x; t < total; t += blockDim.x) {
      const int b = t / half, q = t - b * half;
      const int grp = q >> (s - 1), k = q & (m2 - 1);
      const int i = b * stride + (grp << s) + k, j = i + m2;
      const float2 u = a[i], v = cmul(a[j], tw[k * tstep]);
      a[i] = make_float2(u.x + v.x, u.y + v.y);
      a[j] = make_float2(u.x - v.x, u.y - v.y);
    }
  }
  __syncthreads();
}

__device__ __forceinline__ int bitrev(int v, int bits) { return (int)(__brev((unsigned)v) >> (32 - bits)); }

// rows: forward reads real x [n][H][W] and writes complex F; inverse reads F and writes real y * scale
__global__ void __launch_bounds__(FFT_THREADS)
fft_rows_kernel(const float* __restrict__ x, float2* __restrict__ F, float* __restrict__ y, int64_t total_rows, int W, int logW,
                int rows_per_cta, int inverse, float scale) {
  extern __shared__ float2 sm[];
  float2* tw = sm;                       // [W/2]
  float2* a = sm + W / 2;                // [rows_per_cta][W + 1]
  const int stride = W + 1;
  make_twiddles(tw, W, inverse ? 1.f : -1.f);
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int nb = (int)min((int64_t)rows_per_cta, total_rows - row0);
  for (int t = threadIdx.x; t < nb * W; t += blockDim.x) {
    const int b = t / W, i = t - b * W;
    const int64_t g = (row0 + b) * W + i;
    a[b * stride + bitrev(i, logW)] = inverse ? F[g] : make_float2(x[g], 0.f);
  }
  fft_smem(a, nb, stride, W, logW, tw);
  for (int t = threadIdx.x; t < nb * W; t += blockDim.x) {
    const int b = t / W, i = t - b * W;
    const int64_t g = (row0 + b) * W + i;
    if (inverse) y[g] = a[b * stride + i].x * scale;
    else F[g] = a[b * stride + i];
  }
}

// columns, in place: CTA (tile, image) transforms FFT_COL_TILE adjacent columns of one H x W plane
__global__ void __launch_bounds__(FFT_THREADS)
fft_cols_kernel(float2* __restrict__ F, int H, int W, int logH, int inverse) {
  extern __shared__ float2 sm[];
  float2* tw = sm;                       // [H/2]
  float2* a = sm + H / 2;                // [FFT_COL_TILE][H + 1]
  const int stride = H + 1;
  make_twiddles(tw, H, inverse ? 1.f : -1.f);
  float2* plane = F + (size_t)blockIdx.y * H * W + (size_t)blockIdx.x * FFT_COL_TILE;
  for (int t = threadIdx.x; t < H * FFT_COL_TILE; t += blockDim.x) {
    const int r = t / FFT_COL_TILE, c = t - r * FFT_COL_TILE;
    a[c * stride + bitrev(r, logH)] = plane[(size_t)r * W + c];
  }
  fft_smem(a, FFT_COL_TILE, stride, H, logH, tw);
  for (int t = threadIdx.x; t < H * FFT_COL_TILE; t += blockDim.x) {
    const int r = t / FFT_COL_TILE, c = t - r * FFT_COL_TILE;
    plane[(size_t)r * W + c] = a[c * stride + r];
  }
}

// per (channel, y, x): batch-mean amplitude, running update, F <- amp_new * F / |F|
__global__ void __launch_bounds__(FFT_THREADS)
amp_kernel(float2* __restrict__ F, int B, int ch, int64_t plane, float* __restrict__ amp_run, float* __restrict__ amp_batch,
           float momentum, int update_running, int have_running) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;     // over [ch][H*W]
  if (idx >= (int64_t)ch * plane) return;
  float acc = 0.f;
  for (int b = 0; b < B; ++b) {
    const float2 f = F[(size_t)b * ch * plane + idx];
    acc += sqrtf(f.x * f.x + f.y * f.y);
  }
  const float mean = acc / (float)B;
  if (amp_batch) amp_batch[idx] = mean;
  const float amp_new = have_running ? (1.f - momentum) * amp_run[idx] + momentum * mean : mean;
  if (update_running) amp_run[idx] = amp_new;
  for (int b = 0; b < B; ++b) {
    float2 f = F[(size_t)b * ch * plane + idx];
    const float mag = sqrtf(f.x * f.x + f.y * f.y);
    f = mag > 0.f ? make_float2(f.x / mag * amp_new, f.y / mag * amp_new) : make_float2(amp_new, 0.f);   // angle(0) = 0
    F[(size_t)b * ch * plane + idx] = f;
  }
}

static int ilog2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}
static bool pow2_in_range(int v) { return v >= 8 && v <= 512 && (v & (v - 1)) == 0; }

}  // namespace

extern "C" int64_t fb_amp_normalize_workspace_bytes(int32_t n_img, int32_t ch, int32_t H, int32_t W) {
  return (int64_t)n_img * ch * H * W * (int64_t)sizeof(float2);
}

extern "C" int fb_amp_normalize(const float* x, float* y, int32_t n_img, int32_t ch, int32_t H, int32_t W, float* amp_run,
                                float* amp_batch_mean, float momentum, int32_t update_running, int32_t have_running,
                                void* workspace, fb_stream_t stream) {
  FB_REQUIRE(x && y && workspace && n_img >= 1 && ch >= 1, "fb_amp_normalize: bad arguments");
  FB_REQUIRE(pow2_in_range(H) && pow2_in_range(W), "fb_amp_normalize: H and W must be powers of two in [8, 512] (got %d x %d)", H, W);
  FB_REQUIRE(amp_run || !(have_running || update_running), "fb_amp_normalize: running amplitude buffer required");
  FB_REQUIRE((int64_t)n_img * ch <= 65535, "fb_amp_normalize: at most 65535 image planes per call");
  cudaStream_t st = (cudaStream_t)stream;
  float2* F = reinterpret_cast<float2*>(workspace);
  const int logW = ilog2(W), logH = ilog2(H);
  const int64_t rows = (int64_t)n_img * ch * H;
  const int rpc = 512 / W > 0 ? 512 / W : 1;
  const size_t smem_r = ((size_t)W / 2 + (size_t)rpc * (W + 1)) * sizeof(float2);
  const size_t smem_c = ((size_t)H / 2 + (size_t)FFT_COL_TILE * (H + 1)) * sizeof(float2);
  const unsigned grid_r = (unsigned)((rows + rpc - 1) / rpc);
  const dim3 grid_c(W / FFT_COL_TILE, n_img * ch);
  const int64_t plane = (int64_t)H * W;
  fft_rows_kernel<<<grid_r, FFT_THREADS, smem_r, st>>>(x, F, nullptr, rows, W, logW, rpc, 0, 1.f);
  fft_cols_kernel<<<grid_c, FFT_THREADS, smem_c, st>>>(F, H, W, logH, 0);
  amp_kernel<<<(unsigned)((ch * plane + FFT_THREADS - 1) / FFT_THREADS), FFT_THREADS, 0, st>>>(
      F, n_img, ch, plane, amp_run, amp_batch_mean, momentum, update_running, have_running);
  fft_cols_kernel<<<grid_c, FFT_THREADS, smem_c, st>>>(F, H, W, logH, 1);
  fft_rows_kernel<<<grid_r, FFT_THREADS, smem_r, st>>>(nullptr, F, y, rows, W, logW, rpc, 1, 1.f / (float)plane);
  FB_CHECK_CUDA(cudaGetLastError());
  return FB_OK;
}
