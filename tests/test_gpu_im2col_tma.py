"""TMA im2col tensor maps: the tile the implicit-GEMM convolution producer loads == the explicit patch matrix."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _patches(x, KH, KW, stride, pad):
    """[B*OH*OW, KH, KW, C] fp32 patch tensor of an NHWC input."""
    B, H, W, C = x.shape
    OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    xp = torch.zeros(B, H + 2 * pad, W + 2 * pad, C, dtype=x.dtype, device=x.device)
    xp[:, pad:pad + H, pad:pad + W] = x
    out = torch.empty(B, OH, OW, KH, KW, C, dtype=x.dtype, device=x.device)
    for r in range(KH):
        for s in range(KW):
            out[:, :, :, r, s] = xp[:, r:r + stride * OH:stride, s:s + stride * OW:stride]
    return out.view(B * OH * OW, KH, KW, C)


@pytest.mark.parametrize("B,H,W,C,K,stride,pad", [(2, 8, 8, 64, 3, 1, 1), (3, 14, 14, 128, 3, 2, 1), (2, 10, 12, 64, 1, 2, 0),
                                                   (2, 9, 7, 64, 3, 1, 1), (1, 16, 16, 192, 3, 1, 1)])
def test_im2col_tma_tiles(B, H, W, C, K, stride, pad):
    from ml_trainer_b200.ops import _lib
    from ml_trainer_b200.ops._lib import check, ptr

    torch.manual_seed(0)
    x = torch.randn(B, H, W, C, device=DEV).to(torch.bfloat16)
    ref = _patches(x, K, K, stride, pad)
    M = ref.shape[0]
    L = _lib.lib()
    for pixels in (128, 64):
        for m0 in range(0, M, pixels):
            for (r, s, c0) in {(0, 0, 0), (K - 1, K - 1, C - 64), (K // 2, 0, 0), (0, K - 1, C - 64)}:
                out = torch.full((pixels, 64), 7.0, dtype=torch.bfloat16, device=DEV)
                check(L.mlt_im2col_probe(ptr(x), B, H, W, C, K, K, stride, pad, ctypes.c_longlong(m0), r, s, c0, pixels, ptr(out), None),
                      "mlt_im2col_probe")
                torch.cuda.synchronize()
                n = min(pixels, M - m0)
                want = ref[m0:m0 + n, r, s, c0:c0 + 64]
                assert torch.equal(out[:n], want), (pixels, m0, r, s, c0, (out[:n].float() - want.float()).abs().max().item())
                if n < pixels:      # rows past the last image: zero-filled
                    assert torch.count_nonzero(out[n:]).item() == 0
