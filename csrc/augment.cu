// Train-time image augmentation for the client-batched input path: RandomCrop(H, padding) + horizontal flip + Cutout on
// uint8 NHWC images, all random choices from one Philox4x32-10 block per sample (see fedml_b200/data/augment.py, which
// holds the CPU oracle with the same arithmetic).  Reference: the torchvision transform chain of the CIFAR loaders
// (data/cifar10/data_loader.py:58-103), run there per sample on the host inside DataLoader workers.
#include "common.cuh"
#include "../include/fedml_b200.h"

#define AUG_STREAM_TAG 0xA06A06u

// one thread per output pixel (C <= 4 channels as one 32-bit word when C == 4, else byte stores)
__global__ void __launch_bounds__(256)
augment_cifar_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, int64_t n_img, int H, int W, int C, uint64_t seed,
                     uint32_t nonce, const int64_t* __restrict__ nonce_dev, int pad, int cutout, uint32_t fill_rgba,
                     int64_t first_index) {
  pdl_wait();
  pdl_trigger();
  const uint32_t nn = nonce_dev ? (uint32_t)(*nonce_dev) : nonce;
  const int64_t total = n_img * H * W;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
    const int64_t img = p / (H * W);
    const int rem = (int)(p - img * H * W);
    const int oy = rem / W, ox = rem - oy * W;
    const uint64_t idx = (uint64_t)(first_index + img);
    const Philox4 r = philox4x32_10((uint32_t)idx, (uint32_t)(idx >> 32), AUG_STREAM_TAG, nn, (uint32_t)seed, (uint32_t)(seed >> 32));
    const uint32_t span = 2u * (uint32_t)pad + 1u;
    const int dy = (int)(r.x % span), dx = (int)((r.x >> 16) % span);
    const bool flip = (r.y & 1u) != 0u;
    const int cy = (int)(r.z % (uint32_t)H), cx = (int)(r.w % (uint32_t)W);
    uint8_t* dst = y + (size_t)p * C;
    const int h2 = cutout / 2;
    const bool cut = cutout > 0 && oy >= max(0, cy - h2) && oy < min(H, cy + h2) && ox >= max(0, cx - h2) && ox < min(W, cx + h2);
    if (cut) {
      for (int c = 0; c < C; ++c) dst[c] = (uint8_t)(fill_rgba >> (8 * c));
      continue;
    }
    const int sx_out = flip ? (W - 1 - ox) : ox;          // flip acts on the cropped image
    const int sy = oy + dy - pad, sx = sx_out + dx - pad;
    if (sy < 0 || sy >= H || sx < 0 || sx >= W) {
      for (int c = 0; c < C; ++c) dst[c] = 0;              // crop padding is black
    } else {
      const uint8_t* src = x + (((size_t)img * H + sy) * W + sx) * C;
      for (int c = 0; c < C; ++c) dst[c] = src[c];
    }
  }
}

extern "C" int fb_augment_cifar(const uint8_t* x, uint8_t* y, int64_t n_img, int32_t H, int32_t W, int32_t C, uint64_t seed,
                                uint32_t nonce, const int64_t* nonce_dev, int32_t pad, int32_t cutout, uint32_t fill_rgba,
                                int64_t first_index, fb_stream_t stream) {
  FB_REQUIRE(x && y && x != y && n_img >= 1 && H >= 1 && W >= 1 && C >= 1 && C <= 4 && pad >= 0 && pad < 32768 && cutout >= 0,
             "fb_augment_cifar: bad arguments (out of place, 1..4 channels)");
  const int64_t total = n_img * H * W;
  int64_t grid = (total + 255) / 256, cap = (int64_t)fb_num_sms() * 16;
  if (grid > cap) grid = cap;
  FB_CHECK_CUDA(fb_launch_pdl(augment_cifar_kernel, dim3((unsigned)grid), dim3(256), (size_t)0, (cudaStream_t)stream, x, y, n_img,
                              (int)H, (int)W, (int)C, seed, nonce, nonce_dev, (int)pad, (int)cutout, fill_rgba, first_index));
  return FB_OK;
}
