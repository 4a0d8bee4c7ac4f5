// Image augmentation on the device: uint8 in, normalised compute-dtype NCHW out (SURVEY §8 K16 "cast / normalise on device").
//
// The reference trains CIFAR-10 through RandomCrop(32, padding=4) + RandomHorizontalFlip + ToTensor + Normalize on PIL
// images, one sample at a time on the host (src/utils/functions.py:5-12, main.py:12-28), then uploads fp32.  Here the dataset
// stays uint8 (as CIFAR ships it, [N, H, W, C], or [N, C, H, W]); a batch is uploaded as uint8 (4x fewer bytes over PCIe) and
// ONE kernel does, per sample: zero-pad by `pad`, crop an out_h x out_w window at a random offset, flip with probability 1/2,
// scale by 1/255, subtract the channel mean and divide by the channel std - the same arithmetic in the same order as the
// torchvision pipeline (padding pixels are 0 BEFORE normalisation, so they come out as -mean/std).
//
// Randomness is counter-based (Philox4x32-10 keyed by the seed, counter = position of the sample in the training stream), so a
// run is reproducible and the host can state the parameters of any sample (mlt_augment_params) - that is the test oracle.
// Evaluation (train = 0): centre window, no flip.
#include "common.cuh"
#include "../include/mlt_b200.h"

namespace mlt {

__host__ __device__ inline uint32_t mulhi32(uint32_t a, uint32_t b) { return (uint32_t)(((unsigned long long)a * b) >> 32); }

__host__ __device__ inline void philox4x32_10_hd(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = mulhi32(M0, c[0]), lo0 = M0 * c[0];
    const uint32_t hi1 = mulhi32(M1, c[2]), lo1 = M1 * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
    k0 += W0; k1 += W1;
  }
}

// (top, left, flip) of the sample at stream position `pos`
__host__ __device__ inline void augment_params(unsigned long long seed, unsigned long long pos, int pad, int do_flip, int& top, int& left,
                                               int& flip) {
  uint32_t c[4] = {(uint32_t)pos, (uint32_t)(pos >> 32), 0x61756730u /* "aug0" */, 0u};
  philox4x32_10_hd(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  const uint32_t span = (uint32_t)(2 * pad + 1);
  top = (int)(mulhi32(c[0], span));            // uniform in [0, 2 pad] (multiply-shift, bias < 2^-28)
  left = (int)(mulhi32(c[1], span));
  flip = do_flip ? (int)(c[2] >> 31) : 0;
}

struct AugArgs {
  int C, H, W, out_h, out_w, pad, flip, hwc;
  float mean[4], inv_std[4];
  unsigned long long seed, pos0;
  int train;
};

// one CTA per sample; threads walk the OUTPUT in NCHW order (coalesced stores); the 3 KB source sample is L1/L2 resident
template <typename T>
__global__ void __launch_bounds__(256)
augment_u8_kernel(const unsigned char* __restrict__ src, T* __restrict__ dst, const AugArgs a) {
  const int b = blockIdx.x;
  int top, left, flip;
  if (a.train) {
    augment_params(a.seed, a.pos0 + (unsigned long long)b, a.pad, a.flip, top, left, flip);
  } else {
    top = (a.H + 2 * a.pad - a.out_h) / 2; left = (a.W + 2 * a.pad - a.out_w) / 2; flip = 0;
  }
  const unsigned char* s = src + (size_t)b * a.C * a.H * a.W;
  T* d = dst + (size_t)b * a.C * a.out_h * a.out_w;
  const int n = a.C * a.out_h * a.out_w;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int c = i / (a.out_h * a.out_w);
    const int r = i - c * a.out_h * a.out_w;
    const int oy = r / a.out_w, ox = r - oy * a.out_w;
    const int sx = (flip ? a.out_w - 1 - ox : ox) + left - a.pad;      // position in the unpadded source image
    const int sy = oy + top - a.pad;
    float v = 0.f;
    if (sx >= 0 && sx < a.W && sy >= 0 && sy < a.H)
      v = (float)(a.hwc ? s[(sy * a.W + sx) * a.C + c] : s[(c * a.H + sy) * a.W + sx]) * (1.0f / 255.0f);
    d[i] = from_float<T>((v - a.mean[c]) * a.inv_std[c]);
  }
}

}  // namespace mlt

using namespace mlt;

static int fill_aug(AugArgs& a, const mlt_augment_cfg* cfg, unsigned long long pos0, int train) {
  MLT_REQUIRE(cfg != nullptr, "augment: configuration is required");
  MLT_REQUIRE(cfg->C >= 1 && cfg->C <= 4 && cfg->H > 0 && cfg->W > 0, "augment: bad image shape (C=%d, H=%d, W=%d)", cfg->C, cfg->H, cfg->W);
  MLT_REQUIRE(cfg->pad >= 0 && cfg->out_h > 0 && cfg->out_w > 0 && cfg->out_h <= cfg->H + 2 * cfg->pad && cfg->out_w <= cfg->W + 2 * cfg->pad,
              "augment: the %dx%d window does not fit the padded %dx%d image", cfg->out_h, cfg->out_w, cfg->H + 2 * cfg->pad, cfg->W + 2 * cfg->pad);
  a.C = cfg->C; a.H = cfg->H; a.W = cfg->W; a.out_h = cfg->out_h; a.out_w = cfg->out_w; a.pad = cfg->pad; a.flip = cfg->flip; a.hwc = cfg->hwc;
  for (int c = 0; c < 4; ++c) { a.mean[c] = cfg->mean[c]; a.inv_std[c] = cfg->std[c] != 0.f ? 1.0f / cfg->std[c] : 1.0f; }
  a.seed = cfg->seed; a.pos0 = pos0; a.train = train;
  return 0;
}

// src: [B, H, W, C] (cfg->hwc) or [B, C, H, W] uint8 on the device; dst: [B, C, out_h, out_w] fp32 or bf16.
// pos0: position of sample 0 of this batch in the training stream (sample b uses pos0 + b).
extern "C" int mlt_augment_u8(const void* src_u8, void* dst, int dst_bf16, int B, const mlt_augment_cfg* cfg, unsigned long long pos0,
                              int train, void* stream) {
  AugArgs a;
  if (fill_aug(a, cfg, pos0, train)) return 1;
  MLT_REQUIRE(B > 0, "augment: empty batch");
  // the random offset range [0, 2 pad] assumes the torchvision case out = in (crop back to the original size after padding)
  MLT_REQUIRE(train == 0 || (a.out_h == a.H && a.out_w == a.W), "augment: random crops keep the image size (out %dx%d, in %dx%d)", a.out_h,
              a.out_w, a.H, a.W);
  cudaStream_t st = (cudaStream_t)stream;
  if (dst_bf16) augment_u8_kernel<__nv_bfloat16><<<B, 256, 0, st>>>((const unsigned char*)src_u8, (__nv_bfloat16*)dst, a);
  else augment_u8_kernel<float><<<B, 256, 0, st>>>((const unsigned char*)src_u8, (float*)dst, a);
  MLT_LAUNCH_CHECK();
  return 0;
}

// Host-side statement of the random draw of stream position `pos`: out3 = {top, left, flip} (the oracle of the tests, and what
// a CPU data path uses to produce the same images).
extern "C" int mlt_augment_params(const mlt_augment_cfg* cfg, unsigned long long pos, int* out3) {
  MLT_REQUIRE(cfg != nullptr && out3 != nullptr, "mlt_augment_params: null argument");
  augment_params(cfg->seed, pos, cfg->pad, cfg->flip, out3[0], out3[1], out3[2]);
  return 0;
}
