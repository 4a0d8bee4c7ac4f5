// TMA im2col tensor map for an NHWC bf16 activation tensor (see im2col_tma.cu).
#pragma once
#include <cuda.h>

namespace mlt {
int make_im2col_tmap(CUtensorMap* out, const void* x, int B, int H, int W, int C, int KH, int KW, int stride, int pad,
                     int channels_per_pixel, int pixels_per_column);
}
