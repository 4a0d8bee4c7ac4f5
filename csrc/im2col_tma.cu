// TMA im2col tensor maps (cuTensorMapEncodeIm2col) for NHWC bf16 activations, shared by the implicit-GEMM convolution
// path of gemm_tcgen05.cu, plus a probe kernel that loads one [pixels x 64 channels] tile exactly the way the GEMM producer
// does and writes it back un-swizzled - the test for the bounding-box / offset semantics (tests/test_gpu_im2col_tma.py).
#include "common.cuh"
#include "ptx.cuh"
#include "im2col_tma.h"
#include "../include/mlt_b200.h"
#include <mutex>

namespace mlt {

using namespace ptx;

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_im2col_tmap(CUtensorMap* out, const void* x, int B, int H, int W, int C, int KH, int KW, int stride, int pad,
                     int channels_per_pixel, int pixels_per_column) {
  static EncodeIm2colFn enc = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fp, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      enc = reinterpret_cast<EncodeIm2colFn>(fp);
  });
  MLT_REQUIRE(enc != nullptr, "cuTensorMapEncodeIm2col is not available from the driver");
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { MLT_CHECK_CUDA(cudaFree(nullptr)); ctx_bound = true; }
  MLT_REQUIRE(((uintptr_t)x % 16) == 0 && (C * 2) % 16 == 0, "im2col map: base and channel pitch must be 16-byte multiples");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  // base pixels range over [-pad, dim + upper - 1]; a filter tap adds its offset on top.  upper = pad - (K - 1).
  int lower[2] = {-pad, -pad};
  int upper[2] = {pad - (KW - 1), pad - (KH - 1)};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, lower, upper,
                   (cuuint32_t)channels_per_pixel, (cuuint32_t)pixels_per_column, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MLT_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeIm2col failed with CUresult %d (B=%d H=%d W=%d C=%d K=%dx%d stride=%d pad=%d)", (int)r,
              B, H, W, C, KH, KW, stride, pad);
  return 0;
}

__global__ void __launch_bounds__(128)
im2col_probe_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* out, int pixels, int c0, int w0, int h0, int n0, int off_w,
                    int off_h) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, pixels * 128);
    tma_load_im2col_4d(smem, &tm, &bar, c0, w0, h0, n0, off_w, off_h);
  }
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < pixels * 8; i += blockDim.x) {
    const int row = i >> 3, chunk = i & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(smem + row * 128 + ((chunk ^ (row & 7)) << 4));
    *reinterpret_cast<uint4*>(out + (size_t)row * 64 + chunk * 8) = v;
  }
}

}  // namespace mlt

using namespace mlt;

// Loads the [pixels x 64] patch tile for output pixels m0 .. m0+pixels-1, filter tap (r, s), channels c0 .. c0+63.
extern "C" int mlt_im2col_probe(const void* x, int B, int H, int W, int C, int KH, int KW, int stride, int pad, long long m0, int r, int s,
                                int c0, int pixels, void* out, void* stream) {
  MLT_REQUIRE(pixels > 0 && pixels <= 256 && C % 64 == 0, "mlt_im2col_probe: pixels in 1..256, C a multiple of 64");
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  CUtensorMap tm;
  if (make_im2col_tmap(&tm, x, B, H, W, C, KH, KW, stride, pad, 64, pixels)) return 1;
  const int q0 = (int)(m0 % OW), p0 = (int)((m0 / OW) % OH), n0 = (int)(m0 / ((long long)OW * OH));
  MLT_CHECK_CUDA(cudaFuncSetAttribute(im2col_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 128 + 1024));
  im2col_probe_kernel<<<1, 128, 256 * 128 + 1024, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, pixels, c0, q0 * stride - pad,
                                                                          p0 * stride - pad, n0, s, r);
  MLT_LAUNCH_CHECK();
  return 0;
}
