"""The allocation-free conv backward entry point, driven through the C ABI.

Some users of the library may not allocate device memory where they call it: a plug-in inside an
application that has already claimed the whole GPU for its own pool, or a loop that cannot afford the
device-wide synchronisation a ``cudaMalloc`` brings.  ``mlt_conv2d_direct_bwd_noalloc`` serves them by
keeping the kernel's reduction scratch in a fixed 256 KiB array inside the library instead of taking it
from the caller.  These tests call it with ctypes exactly as such a C caller would (raw pointers, a
stream handle, an int status) and hold it to the ordinary entry point's results.

The file name sorts last on purpose: if this test ever fails, ``pytest -x`` has already run everything else.
It must pass both on its own in a fresh process and after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

C_INT, C_LL, C_VP = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p


def _lib():
    from ml_trainer_b200.ops import _lib as L

    lib = L.lib()
    lib.mlt_conv2d_direct_bwd_scratch_bytes.restype = C_LL
    lib.mlt_last_error.restype = ctypes.c_char_p
    return lib


def _p(t):
    return C_VP(0 if t is None else t.data_ptr())


def _problem(seed, B, Cin, H, Cout, K, dev):
    """Forward through the library (pooled output + arg-max indices), then an upstream gradient."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, Cin, H, H, generator=g).to(dev)
    w = (torch.randn(Cout, Cin, K, K, generator=g) * 0.2).to(dev)
    b = torch.randn(Cout, generator=g).to(dev)
    P = (H - K + 1) // 2
    y = torch.empty(B, Cout, P, P, device=dev)
    idx = torch.empty(B, Cout, P, P, dtype=torch.uint8, device=dev)
    st = C_VP(torch.cuda.current_stream().cuda_stream)
    assert _lib().mlt_conv2d_direct_fwd(_p(x), _p(w), _p(b), _p(y), _p(idx), C_INT(B), C_INT(Cin), C_INT(H), C_INT(H),
                                        C_INT(Cout), C_INT(K), C_INT(K), C_INT(1), C_INT(1), st) == 0
    dy = torch.randn(B, Cout, P, P, generator=g).to(dev)
    return x, w, y, idx, dy


def _bwd(entry, x, w, y, idx, dy, scratch=None):
    lib = _lib()
    B, Cin, H, _ = x.shape
    Cout, _, K, _ = w.shape
    dx, dw, db = torch.empty_like(x), torch.empty_like(w), torch.empty(Cout, device=x.device)
    st = C_VP(torch.cuda.current_stream().cuda_stream)
    dims = (C_INT(B), C_INT(Cin), C_INT(H), C_INT(H), C_INT(Cout), C_INT(K), C_INT(K), C_INT(1), C_INT(1), C_INT(0), st)
    if entry == "noalloc":
        rc = lib.mlt_conv2d_direct_bwd_noalloc(_p(x), _p(w), _p(dy), _p(y), _p(idx), _p(dx), _p(dw), _p(db), *dims)
    else:
        rc = lib.mlt_conv2d_direct_bwd(_p(x), _p(w), _p(dy), _p(y), _p(idx), _p(dx), _p(dw), _p(db), _p(scratch),
                                       C_LL(scratch.numel()), *dims)
    assert rc == 0, lib.mlt_last_error()
    torch.cuda.synchronize()
    return dx, dw, db


@pytest.mark.parametrize("B,Cin,H,Cout,K", [(32, 3, 32, 6, 5), (32, 6, 14, 16, 5)])
def test_noalloc_matches_regular_entry_point_twice_in_a_row(B, Cin, H, Cout, K):
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = _lib()
    nbytes = lib.mlt_conv2d_direct_bwd_scratch_bytes(C_INT(B), C_INT(Cin), C_INT(Cout), C_INT(K), C_INT(K))
    assert 0 < nbytes <= 256 * 1024
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device=dev)          # the caller-provided scratch of the regular path
    for call, seed in enumerate((11, 12)):                                # two calls in a row, different data
        x, w, y, idx, dy = _problem(seed, B, Cin, H, Cout, K, dev)
        ref = _bwd("regular", x, w, y, idx, dy, scratch)
        got = _bwd("noalloc", x, w, y, idx, dy)
        for name, a, e in zip(("dx", "dw", "db"), got, ref):
            assert torch.allclose(a, e, rtol=1e-3, atol=1e-3), (call, name, (a - e).abs().max().item())


def test_noalloc_rejects_problems_beyond_its_buffer():
    dev = torch.device("cuda", 0)
    lib = _lib()
    B, Cin, H, Cout, K = 64, 16, 20, 64, 7
    assert lib.mlt_conv2d_direct_bwd_scratch_bytes(C_INT(B), C_INT(Cin), C_INT(Cout), C_INT(K), C_INT(K)) > 256 * 1024
    t = torch.zeros(16, device=dev)
    st = C_VP(torch.cuda.current_stream().cuda_stream)
    rc = lib.mlt_conv2d_direct_bwd_noalloc(_p(t), _p(t), _p(t), _p(t), _p(t), _p(None), _p(t), _p(t), C_INT(B), C_INT(Cin),
                                           C_INT(H), C_INT(H), C_INT(Cout), C_INT(K), C_INT(K), C_INT(1), C_INT(1),
                                           C_INT(0), st)
    assert rc != 0
    msg = lib.mlt_last_error().decode()
    assert "built-in buffer" in msg and "caller-provided scratch" in msg
