"""tcgen05 GEMM front end (csrc/gemm_tcgen05.cu) and the tensor-core Linear autograd function.

``gemm(a, b, a_mn, b_mn)`` computes ``D[M,N] = A @ B^T`` where each operand is given in its *storage* form:
``a_mn=False`` -> ``a`` is ``[M,K]``; ``a_mn=True`` -> ``a`` is ``[K,M]`` (the reduction dim is the slow one).  Likewise
``b`` is ``[N,K]`` or ``[K,N]``.  All three GEMMs of a linear layer map onto it without any transpose:

    forward   Y  = X  W^T        gemm(X,  W,  False, False)
    dgrad     dX = dY W          gemm(dY, W,  False, True)
    wgrad     dW = dY^T X        gemm(dY, X,  True,  True)
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn.functional as TF

from . import _lib
from ._lib import check, launches, ptr, stream_handle

_c_int, _c_ll = ctypes.c_int, ctypes.c_longlong
ACT_IDS = {None: 0, "none": 0, "relu": 1, "gelu": 2, "tanh": 3}
SPLITK_WS_BYTES = 256 << 20          # fp32 partials of up to 16 K-slices of the largest weight gradient
SPLITK_MAX_OUT = 8 << 20             # outputs beyond 8M elements have plenty of tiles already


def gemm_supported(a: torch.Tensor, b: torch.Tensor) -> bool:
    return (a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
            and a.stride(1) == 1 and b.stride(1) == 1 and a.stride(0) % 8 == 0 and b.stride(0) % 8 == 0
            and a.data_ptr() % 16 == 0 and b.data_ptr() % 16 == 0)


def gemm(a: torch.Tensor, b: torch.Tensor, a_mn: bool = False, b_mn: bool = False, out: Optional[torch.Tensor] = None,
         out_dtype: torch.dtype = torch.bfloat16, bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
         aux_mode: int = 0, aux: Optional[torch.Tensor] = None, block_n: int = 0, split_k: bool = True) -> torch.Tensor:
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    if K != Kb:
        raise ValueError(f"gemm: reduction dims differ ({K} vs {Kb})")
    if out is None:
        n_pad = (N + 7) // 8 * 8             # keep the row pitch a multiple of 16 bytes
        full = torch.empty((M, n_pad), dtype=out_dtype, device=a.device)
        out = full[:, :N] if n_pad != N else full
    ws = None
    if split_k and bias is None and act in (None, "none") and aux_mode == 0 and M * N <= SPLITK_MAX_OUT and K >= 512:
        from .functional import scratch

        ws = scratch(a.device, "gemm_splitk", SPLITK_WS_BYTES)
    check(_lib.lib().mlt_gemm_bf16(ptr(a), _c_ll(a.stride(0)), _c_int(int(a_mn)), ptr(b), _c_ll(b.stride(0)),
                                   _c_int(int(b_mn)), ptr(out), _c_ll(out.stride(0)), _c_int(int(out.dtype == torch.float32)),
                                   _c_int(M), _c_int(N), _c_int(K), ptr(bias),
                                   _c_int(int(bias is not None and bias.dtype == torch.float32)), _c_int(ACT_IDS[act]),
                                   _c_int(aux_mode), ptr(aux), _c_ll(aux.stride(0) if aux is not None else 0),
                                   _c_int(block_n), ptr(ws), _c_ll(ws.numel() if ws is not None else 0), stream_handle()),
          "mlt_gemm_bf16")
    launches.bump("gemm_bf16_tcgen05")
    return out


def gemm_batched(a, lda, a_s1, a_s2, a_mn, b, ldb, b_s1, b_s2, b_mn, out, ldo, o_s1, o_s2, M, N, K, nb1, nb2, block_n: int = 0):
    """``nb1*nb2`` independent ``D = A B^T`` problems addressed by two batch strides each (elements); see csrc/gemm_tcgen05.cu."""
    check(_lib.lib().mlt_gemm_bf16_batched(ptr(a), _c_ll(lda), _c_ll(a_s1), _c_ll(a_s2), _c_int(int(a_mn)), ptr(b), _c_ll(ldb),
                                           _c_ll(b_s1), _c_ll(b_s2), _c_int(int(b_mn)), ptr(out), _c_ll(ldo), _c_ll(o_s1),
                                           _c_ll(o_s2), _c_int(int(out.dtype == torch.float32)), _c_int(M), _c_int(N), _c_int(K),
                                           _c_int(nb1), _c_int(nb2), _c_int(block_n), stream_handle()), "mlt_gemm_bf16_batched")
    launches.bump("gemm_bf16_tcgen05_batched")
    return out


def gemm_reference(a, b, a_mn=False, b_mn=False, bias=None, act=None):
    """fp32 PyTorch statement of :func:`gemm` (oracle)."""
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    d = A @ B.t()
    if bias is not None:
        d = d + bias.float()
    if act == "relu":
        d = TF.relu(d)
    elif act == "gelu":
        d = TF.gelu(d)
    elif act == "tanh":
        d = torch.tanh(d)
    return d


class _LinearTCFn(torch.autograd.Function):
    """``act(x W^T + b)`` on tcgen05; backward = two more tcgen05 GEMMs + a column-sum for the bias."""

    @staticmethod
    def forward(ctx, x2, weight, bias, act, wparam, bparam):
        need_pre = act in ("gelu", "tanh")
        aux = torch.empty((x2.shape[0], weight.shape[0]), dtype=torch.bfloat16, device=x2.device) if need_pre else None
        y = gemm(x2, weight, False, False, bias=bias, act=act, aux_mode=1 if need_pre else 0, aux=aux)
        ctx.save_for_backward(x2, weight, aux if need_pre else y)
        ctx.cfg = (act, bias is not None, need_pre)
        ctx.params = (wparam, bparam)
        return y

    @staticmethod
    def backward(ctx, dy):
        from .functional import _deliver, _sink, act_backward, colsum

        x2, weight, saved = ctx.saved_tensors
        act, has_bias, need_pre = ctx.cfg
        wparam, bparam = ctx.params
        dy = dy.contiguous()
        dz = act_backward(dy, saved, act, from_pre=need_pre) if act not in (None, "none") else dy
        dx = gemm(dz, weight, False, True) if ctx.needs_input_grad[0] else None
        wsink = _sink(wparam)
        dw = gemm(dz, x2, True, True, out=wsink if (wsink is not None and wsink.dtype in (torch.bfloat16, torch.float32)) else None,
                  out_dtype=weight.dtype if weight.dtype == torch.bfloat16 else torch.float32)
        gw = _deliver(wparam, dw, wsink)
        gb = None
        if has_bias:
            bsink = _sink(bparam)
            db = colsum(dz, out=bsink)
            gb = _deliver(bparam, db, bsink)
        return dx, gw, gb, None, None, None


def _wgrad_into(dy, x2, wparam, weight):
    """dW = dy^T x2, written into the parameter's arena sink when there is one."""
    from .functional import _deliver, _sink

    wsink = _sink(wparam)
    dw = gemm(dy, x2, True, True, out=wsink if (wsink is not None and wsink.dtype in (torch.bfloat16, torch.float32)) else None,
              out_dtype=weight.dtype if weight.dtype == torch.bfloat16 else torch.float32)
    return _deliver(wparam, dw, wsink)


def _bgrad_into(dy, bparam):
    from .functional import _deliver, _sink, colsum

    bsink = _sink(bparam)
    return _deliver(bparam, colsum(dy, out=bsink), bsink)


class _FFNFn(torch.autograd.Function):
    """``act(x W1^T + b1) W2^T + b2`` as one autograd node (the transformer MLP block).

    The first GEMM's epilogue applies bias + activation and also stores the activation's derivative act'(z1); the
    dgrad GEMM of the second layer multiplies by it in ITS epilogue, so the backward of the activation costs no kernel
    and dz1 is produced directly: 6 tcgen05 GEMMs + 2 column sums per block, no elementwise passes."""

    AUX_FWD, AUX_BWD = 3, 4          # store act'(z1) / multiply by it;  (1, 2) = store z1 / multiply by act'(z1)

    @staticmethod
    def forward(ctx, x2, w1, b1, w2, b2, act, params, skip=None):
        ctx.skip = skip
        dact = torch.empty((x2.shape[0], w1.shape[0]), dtype=torch.bfloat16, device=x2.device)
        f = gemm(x2, w1, False, False, bias=b1, act=act, aux_mode=_FFNFn.AUX_FWD, aux=dact)
        y = gemm(f, w2, False, False, bias=b2)
        ctx.save_for_backward(x2, w1, w2, f, dact)
        ctx.act = act
        ctx.params = params
        ctx.has_bias = (b1 is not None, b2 is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, w1, w2, f, dact = ctx.saved_tensors
        w1p, b1p, w2p, b2p = ctx.params
        dy = dy.contiguous()
        dz = gemm(dy, w2, False, True, act=ctx.act, aux_mode=_FFNFn.AUX_BWD, aux=dact)   # (dy W2) * act'(z1)
        gw2 = _wgrad_into(dy, f, w2p, w2)
        gb2 = _bgrad_into(dy, b2p) if ctx.has_bias[1] else None
        dx = _dgrad_plus_skip(dz, w1, ctx.skip) if ctx.needs_input_grad[0] else None
        if ctx.skip is not None:
            ctx.skip.dres = None
        gw1 = _wgrad_into(dz, x2, w1p, w1)
        gb1 = _bgrad_into(dz, b1p) if ctx.has_bias[0] else None
        return dx, gw1, gb1, gw2, gb2, None, None, None


def _dgrad_plus_skip(dy, w, skip):
    """``dy @ w`` (the data gradient of a linear layer), plus the gradient an identity skip connection parked in ``skip``
    (ops/functional.py::SkipGrad) - added in the GEMM's epilogue instead of by a separate elementwise kernel."""
    parked = skip.dres if skip is not None else None
    if parked is None:
        return gemm(dy, w, False, True)
    skip.dres = None
    if parked.dtype == torch.bfloat16 and parked.dim() == 2 and parked.stride(1) == 1 and parked.data_ptr() % 16 == 0 \
            and parked.stride(0) % 8 == 0 and parked.shape == (dy.shape[0], w.shape[1]):
        return gemm(dy, w, False, True, aux_mode=5, aux=parked)
    return gemm(dy, w, False, True) + parked.reshape(dy.shape[0], w.shape[1])


def ffn_fast_path_ok(x, w1, b1, w2, b2, act: str = "gelu") -> bool:
    x2 = x.reshape(-1, x.shape[-1])
    return bool(gemm_supported(x2, w1) and w2.is_cuda and w2.dtype == torch.bfloat16 and w2.stride(1) == 1 and w2.stride(0) % 8 == 0
                and w1.shape[0] % 8 == 0 and w2.shape[0] % 8 == 0 and act in ("gelu", "relu", "tanh")
                and all(b is None or b.dtype == torch.bfloat16 for b in (b1, b2)))


def ffn_tc(x, w1, b1, w2, b2, act: str = "gelu", skip=None):
    """Transformer MLP block on the tensor cores; falls back to two :func:`linear_act_tc` calls off the fast path.
    ``skip``: a SkipGrad shared with the LayerNorm that adds this block's input back (only on the fast path, with grad)."""
    from .functional import _has_sink

    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    ok = ffn_fast_path_ok(x, w1, b1, w2, b2, act)
    assert skip is None or (ok and torch.is_grad_enabled()), "ffn_tc: a SkipGrad needs the fast path"
    if not ok:
        return linear_act_tc(linear_act_tc(x, w1, b1, act), w2, b2)
    if not torch.is_grad_enabled():
        return gemm(gemm(x2, w1, False, False, bias=b1, act=act), w2, False, False, bias=b2).reshape(*lead, w2.shape[0])
    det = lambda t: None if t is None else t.detach()
    if _has_sink(w1, b1) and _has_sink(w2, b2):
        y = _FFNFn.apply(x2, det(w1), det(b1), det(w2), det(b2), act, (w1, b1, w2, b2), skip)
    else:
        y = _FFNFn.apply(x2, w1, b1, w2, b2, act, (None, None, None, None), skip)
    return y.reshape(*lead, w2.shape[0])


class _LinearTCPackedFn(torch.autograd.Function):
    """One GEMM for several Linear layers that share their input (the Q, K, V projections): weights, biases and their
    gradient sinks are single views over adjacent arena spans (``optim.fused.PackedParams``), so forward, dgrad and wgrad
    are one tcgen05 GEMM each instead of three, and the dgrad needs no gradient accumulation."""

    @staticmethod
    def forward(ctx, x2, wpack, bpack, skip=None):
        ctx.skip = skip
        w = wpack.data
        y = gemm(x2, w, False, False, bias=bpack.data if bpack is not None else None)
        ctx.save_for_backward(x2, w)
        ctx.packs = (wpack, bpack)
        return y

    @staticmethod
    def backward(ctx, dy):
        from .functional import colsum

        x2, w = ctx.saved_tensors
        wpack, bpack = ctx.packs
        dy = dy.contiguous()
        dx = _dgrad_plus_skip(dy, w, ctx.skip) if ctx.needs_input_grad[0] else None
        if ctx.skip is not None:
            ctx.skip.dres = None
        if wpack.first_arrival():
            gemm(dy, x2, True, True, out=wpack.grad)
        else:                                    # the packed layers are applied more than once per forward pass
            wpack.grad.add_(gemm(dy, x2, True, True, out_dtype=wpack.grad.dtype))
        wpack.deliver()
        if bpack is not None:
            if bpack.first_arrival():
                colsum(dy, out=bpack.grad)
            else:
                bpack.grad.add_(colsum(dy).to(bpack.grad.dtype))
            bpack.deliver()
        return dx, None, None, None


def linear_packed(x, wpack, bpack=None, skip=None):
    """``x @ [W0; W1; ...]^T + [b0; b1; ...]`` for parameters packed by the FusedOptimizer; returns ``[..., sum(out_i)]``."""
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    if torch.is_grad_enabled():
        wpack.note_use()
        if bpack is not None:
            bpack.note_use()
    y = _LinearTCPackedFn.apply(x2, wpack, bpack, skip)
    return y.reshape(*lead, wpack.data.shape[0])


def linear_packed_ok(x, wpack, bpack) -> bool:
    return (wpack is not None and x.is_cuda and x.dtype == torch.bfloat16 and wpack.data.dtype == torch.bfloat16
            and wpack.grad.dtype == torch.bfloat16 and gemm_supported(x.reshape(-1, x.shape[-1]), wpack.data)
            and (bpack is None or bpack.data.dtype == torch.bfloat16))


def split3(x: torch.Tensor, stack_rows: bool, pattern: int) -> torch.Tensor:
    """fp32 [R, C] -> the tripled bf16 operand of the three-term split (see csrc/elementwise.cu::split3_kernel)."""
    R, C = x.shape
    out = torch.empty((3 * R, C) if stack_rows else (R, 3 * C), dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().mlt_split3_bf16(ptr(x), _c_ll(x.stride(0)), ptr(out), _c_ll(R), _c_ll(C), _c_int(int(stack_rows)), _c_int(pattern),
                                     stream_handle()), "mlt_split3_bf16")
    launches.bump("split3_bf16")
    return out


class _LinearF32SplitFn(torch.autograd.Function):
    """fp32 ``act(x W^T + b)`` on the bf16 tensor cores through the three-term split (~2^-16 relative, fp32 accumulation):
    forward, data gradient and weight gradient are one tcgen05 GEMM each over a tripled reduction dimension.  This is what a
    user's fp32 ``nn.Linear`` runs on instead of a cuBLAS call (reference contract: any ``nn.Module``, main.py:30)."""

    @staticmethod
    def forward(ctx, x2, weight, bias, act, wparam, bparam):
        a = split3(x2, False, 0)                     # [M, 3K]  (hi, hi, lo)
        b = split3(weight, False, 1)                 # [N, 3K]  (hi, lo, hi)
        need_pre = act not in (None, "none")
        y = gemm(a, b, False, False, out_dtype=torch.float32, bias=bias, act=None if need_pre else act, split_k=False)
        ctx.save_for_backward(x2, weight, y if need_pre else x2.new_empty(0))
        ctx.cfg = (act, bias is not None, need_pre)
        ctx.params = (wparam, bparam)
        if need_pre:                                 # the pre-activation is kept for backward; the activation itself is one pass
            from .functional import _torch_act

            return _torch_act(y, act)
        return y

    @staticmethod
    def backward(ctx, dy):
        from .functional import _deliver, _sink, act_backward, colsum

        x2, weight, pre = ctx.saved_tensors
        act, has_bias, need_pre = ctx.cfg
        wparam, bparam = ctx.params
        dy = dy.contiguous().float()
        if need_pre:
            dy = act_backward(dy, pre, act, from_pre=True)
        dx = None
        if ctx.needs_input_grad[0]:                  # dX = dY W: reduce over N
            dx = gemm(split3(dy, False, 0), split3(weight, True, 1), False, True, out_dtype=torch.float32, split_k=False)
        wsink = _sink(wparam)
        dw = gemm(split3(dy, True, 0), split3(x2, True, 1), True, True, out_dtype=torch.float32,     # dW = dY^T X: reduce over M
                  out=wsink if (wsink is not None and wsink.dtype == torch.float32 and wsink.is_contiguous()) else None, split_k=False)
        gw = _deliver(wparam, dw, wsink if (wsink is not None and wsink.dtype == torch.float32 and wsink.is_contiguous()) else None) \
            if wparam is not None else dw
        gb = None
        if has_bias:
            bsink = _sink(bparam)
            db = colsum(dy, out=bsink)
            gb = _deliver(bparam, db, bsink) if bparam is not None else db
        return dx, gw, gb, None, None, None


def f32_split_ok(x2: torch.Tensor, weight: torch.Tensor) -> bool:
    M, K = x2.shape
    N = weight.shape[0]
    return (x2.is_cuda and x2.dtype == torch.float32 and weight.dtype == torch.float32 and x2.stride(1) == 1 and weight.stride(1) == 1
            and K % 8 == 0 and N % 8 == 0 and M % 8 == 0)


def linear_act_f32(x, weight, bias=None, act: Optional[str] = None):
    """fp32 linear on the tensor cores (three-term bf16 split); shapes the GEMM cannot take go to the PyTorch statement."""
    from .functional import _has_sink, _torch_act

    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    if not f32_split_ok(x2, weight) or act not in (None, "none", "relu", "tanh", "gelu"):
        return _torch_act(TF.linear(x, weight, bias), act)
    if _has_sink(weight, bias):
        y = _LinearF32SplitFn.apply(x2, weight.detach(), None if bias is None else bias.detach(), act, weight, bias)
    else:
        y = _LinearF32SplitFn.apply(x2, weight, bias, act, None, None)
    return y.reshape(*lead, weight.shape[0])


def linear_act_tc(x, weight, bias=None, act: Optional[str] = None):
    """Tensor-core linear for bf16 activations/weights; anything else goes to the PyTorch statement."""
    from .functional import _has_sink

    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    if not (gemm_supported(x2, weight) and weight.shape[0] % 8 == 0):
        if x2.dtype == torch.float32 and weight.dtype == torch.float32:
            return linear_act_f32(x, weight, bias, act)         # fp32 layers: three-term bf16 split on the same tcgen05 GEMM
        y = TF.linear(x, weight, bias)
        from .functional import _torch_act

        return _torch_act(y, act)
    if _has_sink(weight, bias):
        y = _LinearTCFn.apply(x2, weight.detach(), None if bias is None else bias.detach(), act, weight, bias)
    else:
        y = _LinearTCFn.apply(x2, weight, bias, act, None, None)
    return y.reshape(*lead, weight.shape[0])
