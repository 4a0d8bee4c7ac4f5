"""Functional op layer: autograd wrappers over the sm_100a kernel library.

Every op has two statements of the same math: the native kernel (used whenever the tensors
are on a CUDA device - the library is then mandatory, see ``_lib.native_available``) and a
plain PyTorch one (CPU execution and the oracle the GPU tests compare against).

Weight gradients: when a parameter carries ``_mlt_grad_sink`` (set by the FusedOptimizer) the
backward kernel writes the gradient straight into the flat gradient arena and signals the
reducer through ``_mlt_ready()``; autograd then sees ``None`` for that input, so there is no
AccumulateGrad kernel and no flatten copy (SURVEY §7.3 "grads produced directly into bucket views").
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch
import torch.nn.functional as TF

from . import _lib
from ._lib import check, launches, native_available, ptr, stream_handle  # noqa: F401 (re-exported)

ACT_IDS = {None: 0, "none": 0, "relu": 1, "gelu": 2, "tanh": 3}
_c_int, _c_ll = ctypes.c_int, ctypes.c_longlong


def use_native(*tensors) -> bool:
    return bool(tensors) and all(t is None or t.is_cuda for t in tensors) and native_available()


# ----------------------------------------------------------------------------------------------
# per-device scratch (zero-initialised once; kernels re-zero their tickets themselves)
# ----------------------------------------------------------------------------------------------
_scratch: dict = {}


def scratch(device: torch.device, key: str, nbytes: int) -> torch.Tensor:
    k = (device.index, key)
    t = _scratch.get(k)
    if t is None or t.numel() < nbytes:
        t = torch.zeros(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _scratch[k] = t
    return t


# ----------------------------------------------------------------------------------------------
# optimizer
# ----------------------------------------------------------------------------------------------
_hyper_ring: dict = {}
_RING = 4096


def upload_hyper(dev_hyper: torch.Tensor, values) -> None:
    """Async H2D of the 8-float hyper vector through a ring of pinned slots (safe to run ahead of the GPU)."""
    key = dev_hyper.data_ptr()
    ent = _hyper_ring.get(key)
    if ent is None:
        ent = [torch.zeros(_RING, len(values), dtype=torch.float32).pin_memory(), 0]
        _hyper_ring[key] = ent
    ring, i = ent
    slot = ring[i % _RING]
    slot.copy_(torch.tensor(values, dtype=torch.float32))
    ent[1] = i + 1
    dev_hyper.copy_(slot, non_blocking=True)


def fused_update(opt, lo: int, hi: int) -> None:
    """Apply ``opt.kind`` to arena slice [lo, hi) in one launch (see csrc/optim.cu)."""
    from ..optim.rules import RULE_IDS

    n = hi - lo
    if n <= 0:
        return
    L = _lib.lib()
    shadow = None if opt.model is opt.master else opt.model[lo:hi]
    s2 = opt.s2[lo:hi] if opt.s2.numel() > 1 else None
    check(L.mlt_fused_update(_c_int(RULE_IDS[opt.kind]), ptr(opt.master[lo:hi]), ptr(opt.grad[lo:hi]),
                             _c_int(int(opt.grad.dtype == torch.bfloat16)), ptr(shadow), ptr(opt.s1[lo:hi]), ptr(s2),
                             ptr(opt.hyper), _c_ll(n), stream_handle()), "mlt_fused_update")
    launches.bump("fused_update")


# ----------------------------------------------------------------------------------------------
# loss / metric
# ----------------------------------------------------------------------------------------------
def softmax_xent(logits: torch.Tensor, targets: torch.Tensor, acc3: Optional[torch.Tensor] = None,
                 want_acc: bool = False, need_grad: bool = True):
    """Mean CE loss and d(loss)/d(logits) in one pass.  Returns ``(loss, dlogits)``.

    ``acc3`` (``[loss_sum, metric_sum, steps]``) is updated on the device when given.
    """
    B, C = logits.shape
    if use_native(logits, targets):
        logits = logits.contiguous()
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        dl = torch.empty_like(logits) if need_grad else None
        ws = scratch(logits.device, "xent", 16)
        check(_lib.lib().mlt_softmax_xent(ptr(logits), _c_int(int(logits.dtype == torch.bfloat16)), ptr(targets),
                                          ptr(dl), ptr(loss), ptr(acc3), ptr(ws), _c_int(B), _c_int(C),
                                          _c_int(int(want_acc)), stream_handle()), "mlt_softmax_xent")
        launches.bump("softmax_xent")
        return loss, dl
    lf = logits.float()
    logp = torch.log_softmax(lf, dim=-1)
    loss = TF.nll_loss(logp, targets)
    dl = None
    if need_grad:
        dl = torch.exp(logp)
        dl[torch.arange(B, device=logits.device), targets] -= 1.0
        dl = (dl / B).to(logits.dtype)
    if acc3 is not None:
        acc3[0] += loss.detach()
        if want_acc:
            acc3[1] += (lf.argmax(-1) == targets).float().mean()
        acc3[2] += 1.0
    return loss, dl


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets):
        loss, dl = softmax_xent(logits, targets, need_grad=True)
        ctx.save_for_backward(dl)
        return loss.to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g.to(dl.dtype), None


def cross_entropy(logits, targets):
    if logits.dim() != 2 or targets.dtype != torch.int64:
        return TF.cross_entropy(logits, targets)
    return _CrossEntropyFn.apply(logits, targets)


REGRESSION_KINDS = {"l1": 0, "l2": 1, "custom": 1}


def regression_loss(pred: torch.Tensor, target: torch.Tensor, kind: str = "l2", acc3: Optional[torch.Tensor] = None,
                    want_mcrmse: bool = False, need_grad: bool = True):
    """Mean L1 / L2 loss over all elements and d(loss)/d(pred) in one pass; returns ``(loss, dpred)``.

    ``kind``: ``l1`` (torch ``L1Loss``), ``l2`` (``MSELoss``) or ``custom`` (the reference's ``custom_loss_function``,
    mean((out - target)**2), src/utils/functions.py:15-17).  With ``want_mcrmse`` the reference's second metric (mean over
    columns of the column-wise RMSE, src/trainer.py:160-163) is accumulated into ``acc3[1]`` by the same kernel."""
    k = REGRESSION_KINDS[kind]
    if (use_native(pred, target) and pred.dtype in (torch.float32, torch.bfloat16) and target.dtype in (torch.float32, torch.bfloat16)
            and pred.shape == target.shape and pred.numel() > 0):
        pred, target = pred.contiguous(), target.contiguous()
        C = pred.shape[-1] if (want_mcrmse and pred.dim() >= 2) else 0
        if want_mcrmse and not (0 < C <= 1024):
            want_mcrmse, C = False, 0
        loss = torch.empty((), dtype=torch.float32, device=pred.device)
        dp = torch.empty_like(pred) if need_grad else None
        L = _lib.lib()
        L.mlt_regression_loss_ws_bytes.restype = _c_ll
        ws = scratch(pred.device, "regloss", max(int(L.mlt_regression_loss_ws_bytes(_c_int(1024))), 256))
        check(L.mlt_regression_loss(_c_int(k), ptr(pred), _c_int(int(pred.dtype == torch.bfloat16)), ptr(target),
                                    _c_int(int(target.dtype == torch.bfloat16)), ptr(dp), ptr(loss), ptr(acc3), ptr(ws),
                                    _c_ll(ws.numel()), _c_ll(pred.numel()), _c_int(C), _c_int(int(want_mcrmse)), stream_handle()),
              "mlt_regression_loss")
        launches.bump("regression_loss")
        return loss, dp
    d = pred.float() - target.float()
    loss = d.abs().mean() if k == 0 else (d * d).mean()
    dp = None
    if need_grad:
        dp = (torch.sign(d) if k == 0 else 2.0 * d) / d.numel()
        dp = dp.to(pred.dtype)
    if acc3 is not None:
        acc3[0] += loss.detach()
        if want_mcrmse:
            acc3[1] += torch.mean(torch.sqrt(torch.mean(d.reshape(-1, d.shape[-1]) ** 2, dim=0)))
        acc3[2] += 1.0
    return loss, dp


def nll_loss(logp: torch.Tensor, targets: torch.Tensor, acc3: Optional[torch.Tensor] = None, want_acc: bool = False,
             need_grad: bool = True):
    """Mean negative log-likelihood of log-probabilities ``logp`` [B, C] (torch ``NLLLoss``; the reference's ``neg-loss``,
    src/trainer.py:143-144) with its gradient and the arg-max accuracy from one kernel; returns ``(loss, dlogp)``."""
    B, C = logp.shape
    if use_native(logp, targets) and logp.dtype in (torch.float32, torch.bfloat16):
        logp = logp.contiguous()
        loss = torch.empty((), dtype=torch.float32, device=logp.device)
        dl = torch.empty_like(logp) if need_grad else None
        ws = scratch(logp.device, "xent", 16)
        check(_lib.lib().mlt_nll_loss(ptr(logp), _c_int(int(logp.dtype == torch.bfloat16)), ptr(targets), ptr(dl), ptr(loss),
                                      ptr(acc3), ptr(ws), _c_int(B), _c_int(C), _c_int(int(want_acc)), stream_handle()), "mlt_nll_loss")
        launches.bump("nll_loss")
        return loss, dl
    loss = TF.nll_loss(logp.float(), targets)
    dl = None
    if need_grad:
        dl = torch.zeros_like(logp)
        dl[torch.arange(B, device=logp.device), targets] = -1.0 / B
    if acc3 is not None:
        acc3[0] += loss.detach()
        if want_acc:
            acc3[1] += (logp.argmax(-1) == targets).float().mean()
        acc3[2] += 1.0
    return loss, dl


class _FusedLossFn(torch.autograd.Function):
    """Autograd face of the fused loss kernels for callers outside the engine: forward keeps the gradient the kernel produced."""

    @staticmethod
    def forward(ctx, x, target, which):
        if which == "nll":
            loss, d = nll_loss(x, target, need_grad=True)
        else:
            loss, d = regression_loss(x, target, which, need_grad=True)
        ctx.save_for_backward(d)
        return loss.to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return d * g.to(d.dtype), None, None


def argmax_accuracy(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    if use_native(logits, targets) and logits.dim() == 2:
        out = torch.empty((), dtype=torch.float32, device=logits.device)
        logits = logits.contiguous()
        check(_lib.lib().mlt_argmax_accuracy(ptr(logits), _c_int(int(logits.dtype == torch.bfloat16)), ptr(targets),
                                             ptr(out), _c_int(logits.shape[0]), _c_int(logits.shape[1]),
                                             stream_handle()), "mlt_argmax_accuracy")
        launches.bump("argmax_accuracy")
        return out
    return (logits.argmax(-1) == targets).float().mean()


# ----------------------------------------------------------------------------------------------
# gradient sinks
# ----------------------------------------------------------------------------------------------
# A parameter may be used more than once in one forward pass (a module applied twice, tied weights).  Gradient sinks are
# WRITTEN, not accumulated, so uses are counted per step: ``_has_sink`` (forward) counts them, and in backward the first
# arrival writes the sink, later arrivals write a temporary that is added into it, and the reducer hears "ready" only
# when the last use has delivered (its bucket must not be reduced - nor the weight updated - before that).
_STEP = [0]


def new_step() -> None:
    """Start a new step: use counters of every parameter restart lazily (called by ``FusedOptimizer.begin_step``)."""
    _STEP[0] += 1


def _note_use(p) -> None:
    if p is None:
        return
    if getattr(p, "_mlt_stamp", -1) != _STEP[0]:
        p._mlt_stamp, p._mlt_uses, p._mlt_done = _STEP[0], 0, 0
    p._mlt_uses += 1


def _sink(param) -> Optional[torch.Tensor]:
    """Where this backward call should write the parameter's gradient: the arena sink for the first arrival of the step,
    ``None`` (= allocate a temporary; ``_deliver`` adds it into the sink) for later arrivals of a reused parameter."""
    if param is None:
        return None
    s = getattr(param, "_mlt_grad_sink", None)
    if s is not None and getattr(param, "_mlt_uses", 1) > 1 and getattr(param, "_mlt_done", 0) > 0:
        return None
    return s


def _deliver(param, grad_written: Optional[torch.Tensor], sink: Optional[torch.Tensor]):
    """What autograd should see for a weight whose gradient ``grad_written`` was just produced.

    With a sink the kernel already wrote into the arena: tell the reducer and hand autograd ``None``.
    """
    real = getattr(param, "_mlt_grad_sink", None) if param is not None else None
    if real is None:
        return grad_written
    uses = getattr(param, "_mlt_uses", 1)
    if uses > 1:
        if sink is None and grad_written is not None:
            real.add_(grad_written.reshape(real.shape).to(real.dtype))
        param._mlt_done = getattr(param, "_mlt_done", 0) + 1
        if param._mlt_done < uses:
            return None
    ready = getattr(param, "_mlt_ready", None)
    if ready is not None:
        ready()
    return None


# ----------------------------------------------------------------------------------------------
# small-shape CNN path (direct conv + pool, small linear)
# ----------------------------------------------------------------------------------------------
DIRECT_CONV_MAX_TAPS = 1023


class _ConvReluPoolFn(torch.autograd.Function):
    """conv2d(stride 1, no pad) + bias [+ ReLU] [+ 2x2 max-pool] as one kernel; fp32 NCHW."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu, pool, wparam, bparam):
        B, Cin, H, W = x.shape
        Cout, _, KH, KW = weight.shape
        OH, OW = H - KH + 1, W - KW + 1
        PH, PW = (OH // 2, OW // 2) if pool else (OH, OW)
        x = x.contiguous()
        y = torch.empty((B, Cout, PH, PW), dtype=torch.float32, device=x.device)
        idx = torch.empty((B, Cout, PH, PW), dtype=torch.uint8, device=x.device) if pool else None
        check(_lib.lib().mlt_conv2d_direct_fwd(ptr(x), ptr(weight), ptr(bias), ptr(y), ptr(idx), _c_int(B), _c_int(Cin),
                                               _c_int(H), _c_int(W), _c_int(Cout), _c_int(KH), _c_int(KW),
                                               _c_int(int(relu)), _c_int(int(pool)), stream_handle()),
              "mlt_conv2d_direct_fwd")
        launches.bump("conv2d_direct_fwd")
        ctx.save_for_backward(x, weight, y, idx if idx is not None else torch.empty(0, device=x.device))
        ctx.cfg = (relu, pool, bias is not None)
        ctx.params = (wparam, bparam)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y, idx = ctx.saved_tensors
        relu, pool, has_bias = ctx.cfg
        wparam, bparam = ctx.params
        B, Cin, H, W = x.shape
        Cout, _, KH, KW = weight.shape
        dy = dy.contiguous()
        need_dx = ctx.needs_input_grad[0]
        dx = torch.empty_like(x) if need_dx else None
        wsink, bsink = _sink(wparam), _sink(bparam) if has_bias else None
        dw = wsink if wsink is not None else torch.empty_like(weight)
        db = (bsink if bsink is not None else torch.empty(Cout, dtype=torch.float32, device=x.device)) if has_bias else None
        L = _lib.lib()
        L.mlt_conv2d_direct_bwd_scratch_bytes.restype = _c_ll
        nbytes = L.mlt_conv2d_direct_bwd_scratch_bytes(_c_int(B), _c_int(Cin), _c_int(Cout), _c_int(KH), _c_int(KW))
        ws = scratch(x.device, f"convbwd{Cout}x{Cin}x{KH}x{KW}", nbytes)
        check(L.mlt_conv2d_direct_bwd(ptr(x), ptr(weight), ptr(dy), ptr(y), ptr(idx if pool else None), ptr(dx), ptr(dw),
                                      ptr(db), ptr(ws), _c_ll(ws.numel()), _c_int(B), _c_int(Cin), _c_int(H), _c_int(W),
                                      _c_int(Cout), _c_int(KH), _c_int(KW), _c_int(int(relu)), _c_int(int(pool)),
                                      _c_int(0), stream_handle()), "mlt_conv2d_direct_bwd")
        launches.bump("conv2d_direct_bwd", 2 if need_dx else 1)
        return dx, _deliver(wparam, dw, wsink), (_deliver(bparam, db, bsink) if has_bias else None), None, None, None, None


def conv_relu_pool(x, weight, bias=None, relu=True, pool=True):
    """Fused small-channel convolution block.  CPU / oracle: ``max_pool2d(relu(conv2d(x)))``."""
    taps = weight.shape[1] * weight.shape[2] * weight.shape[3]
    if use_native(x, weight) and x.dtype == torch.float32 and taps <= DIRECT_CONV_MAX_TAPS:
        return _ConvReluPoolFn.apply(x, weight.detach(), None if bias is None else bias.detach(), relu, pool, weight, bias) \
            if _has_sink(weight, bias) else _ConvReluPoolFn.apply(x, weight, bias, relu, pool, None, None)
    y = TF.conv2d(x, weight, bias)
    if relu:
        y = TF.relu(y)
    if pool:
        y = TF.max_pool2d(y, 2, 2)
    return y


def _has_sink(p, *also) -> bool:
    """True when ``p``'s gradient goes straight into the optimizer arena.  Called once per USE in the forward pass (it counts
    the use of ``p`` and of the parameters in ``also`` that ride along, e.g. the bias)."""
    ok = getattr(p, "_mlt_grad_sink", None) is not None and torch.is_grad_enabled() and p.requires_grad
    if ok:
        _note_use(p)
        for q in also:
            _note_use(q)
    return ok


class _LinearSmallFn(torch.autograd.Function):
    """CUDA-core linear for shapes far below a tensor-core tile (fp32 or bf16 tensors)."""

    @staticmethod
    def forward(ctx, x, weight, bias, act, wparam, bparam):
        B, K = x.shape
        N = weight.shape[0]
        x = x.contiguous()
        y = torch.empty((B, N), dtype=x.dtype, device=x.device)
        bf16 = int(x.dtype == torch.bfloat16)
        check(_lib.lib().mlt_linear_small_fwd(ptr(x), ptr(weight), ptr(bias), ptr(y), _c_int(B), _c_int(K), _c_int(N),
                                              _c_int(act), _c_int(bf16), stream_handle()), "mlt_linear_small_fwd")
        launches.bump("linear_small_fwd")
        ctx.save_for_backward(x, weight, y)
        ctx.cfg = (act, bias is not None)
        ctx.params = (wparam, bparam)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        act, has_bias = ctx.cfg
        wparam, bparam = ctx.params
        B, K = x.shape
        N = weight.shape[0]
        dy = dy.contiguous()
        need_dx = ctx.needs_input_grad[0]
        dx = torch.empty_like(x) if need_dx else None
        wsink, bsink = _sink(wparam), _sink(bparam) if has_bias else None
        gdt = wsink.dtype if wsink is not None else weight.dtype
        dw = wsink if wsink is not None else torch.empty(weight.shape, dtype=gdt, device=x.device)
        db = (bsink if bsink is not None else torch.empty(N, dtype=gdt, device=x.device)) if has_bias else None
        check(_lib.lib().mlt_linear_small_bwd(ptr(dy), ptr(y), ptr(x), ptr(weight), ptr(dx), ptr(dw), ptr(db), _c_int(B),
                                              _c_int(K), _c_int(N), _c_int(act), _c_int(0),
                                              _c_int(int(x.dtype == torch.bfloat16)), _c_int(int(gdt == torch.bfloat16)),
                                              stream_handle()), "mlt_linear_small_bwd")
        launches.bump("linear_small_bwd")
        return dx, _deliver(wparam, dw, wsink), (_deliver(bparam, db, bsink) if has_bias else None), None, None, None


SMALL_LINEAR_MAX = 1 << 22      # B*K*N below this stays on the CUDA-core path


def _torch_act(y, act):
    if act in (None, "none"):
        return y
    return {"relu": TF.relu, "gelu": TF.gelu, "tanh": torch.tanh}[act](y)


def linear_act(x, weight, bias=None, act: Optional[str] = None):
    """``act(x @ W^T + b)`` with the activation fused into the GEMM epilogue."""
    if use_native(x, weight):
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        B, K = x2.shape
        N = weight.shape[0]
        small_ok = x.dtype in (torch.float32, torch.bfloat16) and weight.dtype == x.dtype and \
            (bias is None or bias.dtype == x.dtype) and act in (None, "none", "relu", "tanh")
        if small_ok and (B * K * N <= SMALL_LINEAR_MAX or (x.dtype == torch.bfloat16 and (N % 8 != 0 or K % 8 != 0))):
            a = ACT_IDS[act]
            if _has_sink(weight, bias):
                y = _LinearSmallFn.apply(x2, weight.detach(), None if bias is None else bias.detach(), a, weight, bias)
            else:
                y = _LinearSmallFn.apply(x2, weight, bias, a, None, None)
            return y.reshape(*lead, N)
        from .gemm import linear_act_tc

        return linear_act_tc(x, weight, bias, act)
    return _torch_act(TF.linear(x, weight, bias), act)


# ----------------------------------------------------------------------------------------------
# helpers around the tensor-core GEMMs
# ----------------------------------------------------------------------------------------------
def act_backward(dy: torch.Tensor, saved: torch.Tensor, act: str, from_pre: bool) -> torch.Tensor:
    """``dy * act'(.)`` where ``saved`` is the pre-activation (``from_pre``) or the activation output."""
    if use_native(dy, saved) and dy.dtype in (torch.float32, torch.bfloat16) and dy.dtype == saved.dtype:
        dy, saved = dy.contiguous(), saved.contiguous()
        dz = torch.empty_like(dy)
        check(_lib.lib().mlt_act_backward(ptr(dy), ptr(saved), ptr(dz), _c_ll(dy.numel()), _c_int(ACT_IDS[act]),
                                          _c_int(int(from_pre)), _c_int(int(dy.dtype == torch.bfloat16)), stream_handle()),
              "mlt_act_backward")
        launches.bump("act_backward")
        return dz
    s = saved.float()
    if act == "relu":
        g = (s > 0).float()
    elif act == "tanh":
        t = torch.tanh(s) if from_pre else s
        g = 1 - t * t
    elif act == "gelu":
        g = 0.5 * (1 + torch.erf(s * 0.7071067811865476)) + s * torch.exp(-0.5 * s * s) * 0.3989422804014327
    else:
        g = torch.ones_like(s)
    return (dy.float() * g).to(dy.dtype)


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Column sums of a 2-D tensor (bias gradient)."""
    M, N = x.shape
    if use_native(x) and x.dtype in (torch.float32, torch.bfloat16) and x.stride(1) == 1:
        if out is None:
            out = torch.empty(N, dtype=x.dtype, device=x.device)
        L = _lib.lib()
        if (x.dtype == torch.bfloat16 and x.is_contiguous() and N % 8 == 0 and N <= 16384 and M >= 1024 and x.data_ptr() % 16 == 0
                and out.dtype in (torch.float32, torch.bfloat16)):
            ws = scratch(x.device, "colsum_flat", 256 + 16384 * 4)
            check(L.mlt_colsum_flat(ptr(x), ptr(out), _c_int(int(out.dtype == torch.bfloat16)), ptr(ws), _c_ll(ws.numel()), _c_int(M),
                                    _c_int(N), _c_int(0), stream_handle()), "mlt_colsum_flat")
            launches.bump("colsum_flat")
            return out
        L.mlt_colsum_scratch_bytes.restype = _c_ll
        nbytes = L.mlt_colsum_scratch_bytes(_c_int(M), _c_int(N))
        ws = scratch(x.device, "colsum", nbytes)
        check(L.mlt_colsum(ptr(x), _c_int(int(x.dtype == torch.bfloat16)), ptr(out), _c_int(int(out.dtype == torch.bfloat16)),
                           ptr(ws), _c_ll(ws.numel()), _c_int(M), _c_int(N), _c_ll(x.stride(0)), _c_int(0), stream_handle()),
              "mlt_colsum")
        launches.bump("colsum")
        return out
    r = x.float().sum(0)
    if out is not None:
        out.copy_(r)
        return out
    return r.to(x.dtype)


# ----------------------------------------------------------------------------------------------
# dropout: Philox mask = f(seed, device step counter, call site, element); nothing stored for backward
# ----------------------------------------------------------------------------------------------
_drop = {"seed": 0x5DEECE66D, "call": 0, "step": {}}


def dropout_seed(seed: int) -> None:
    _drop["seed"] = int(seed) & 0xFFFFFFFFFFFFFFFF


def _drop_step(device: torch.device) -> torch.Tensor:
    device = torch.device(device)
    idx = device.index if device.index is not None else torch.cuda.current_device()
    t = _drop["step"].get(idx)
    if t is None:
        t = torch.zeros(1, dtype=torch.int64, device=device)
        _drop["step"][idx] = t
    return t


def dropout_begin_step(device: torch.device) -> None:
    """Called once per training step (inside the captured graph): advances the device-side step counter the masks are
    keyed on and restarts the call-site numbering."""
    _drop["call"] = 0
    if device.type == "cuda":
        _drop_step(device).add_(1)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, call_id):
        x = x.contiguous()
        y = torch.empty_like(x)
        ctx.cfg = (p, call_id)
        _DropoutFn._run(x, y, p, call_id)
        return y

    @staticmethod
    def _run(x, y, p, call_id):
        check(_lib.lib().mlt_dropout(ptr(x), ptr(y), _c_ll(x.numel()), ctypes.c_float(p), ctypes.c_ulonglong(_drop["seed"]),
                                     ptr(_drop_step(x.device)), ctypes.c_uint(call_id), _c_int(int(x.dtype == torch.bfloat16)),
                                     stream_handle()), "mlt_dropout")
        launches.bump("dropout")

    @staticmethod
    def backward(ctx, dy):
        p, call_id = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        _DropoutFn._run(dy, dx, p, call_id)          # same mask: same (seed, step, call site)
        return dx, None, None


def dropout(x: torch.Tensor, p: float, training: bool = True) -> torch.Tensor:
    """Inverted dropout.  Native path: mask regenerated in backward from a counter-based generator (csrc/dropout.cu)."""
    if not training or p <= 0.0:
        return x
    if use_native(x) and x.dtype in (torch.float32, torch.bfloat16):
        _drop["call"] += 1
        return _DropoutFn.apply(x, float(p), _drop["call"])
    return torch.nn.functional.dropout(x, p, True)


# ----------------------------------------------------------------------------------------------
# transformer pieces: LayerNorm (+ residual), attention core, embeddings  (bf16 activations)
# ----------------------------------------------------------------------------------------------
SKIP_GRAD_FUSION = os.environ.get("MLT_SKIP_GRAD_FUSION", "1") != "0"


class SkipGrad:
    """Side channel for the gradient of an identity skip connection.  In ``y = relu(bn(conv_c(..conv_a(x)..)) + x)`` the input
    ``x`` receives two gradients: the data gradient of ``conv_a`` and the masked output gradient that BatchNorm's backward hands
    to its residual input.  Autograd would add them with one more elementwise kernel (three passes over the block input; 1.2 ms
    of a ResNet-50 step).  With a ``SkipGrad`` shared by the two calls, BatchNorm's backward parks the residual gradient here
    and returns none, and ``conv_a``'s data-gradient GEMM adds it in its epilogue (``aux_mode`` 5).  The same pattern serves the transformer
    block: ``LN(f(x) + x)`` parks LayerNorm's input gradient, the data-gradient GEMM of ``f``'s first layer (the packed Q/K/V
    projection, the first MLP layer) picks it up."""

    __slots__ = ("dres",)

    def __init__(self):
        self.dres = None


class _LayerNormFn(torch.autograd.Function):
    """``LN(x [+ residual]) * gamma + beta``; the residual add is fused, its gradient is the LN input gradient."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps, gparam, bparam, skip=None):
        H = x.shape[-1]
        x2 = x.reshape(-1, H).contiguous()
        r2 = residual.reshape(-1, H).contiguous() if residual is not None else None
        T = x2.shape[0]
        y = torch.empty_like(x2)
        hsum = torch.empty_like(x2) if r2 is not None else None
        ctx.skip = skip if residual is not None else None
        mean = torch.empty(T, dtype=torch.float32, device=x.device)
        rstd = torch.empty(T, dtype=torch.float32, device=x.device)
        check(_lib.lib().mlt_layernorm_fwd(ptr(x2), ptr(r2), ptr(gamma), ptr(beta), _c_int(int(gamma.dtype == torch.bfloat16)),
                                           ptr(y), ptr(hsum), ptr(mean), ptr(rstd), _c_int(T), _c_int(H),
                                           ctypes.c_float(eps), stream_handle()), "mlt_layernorm_fwd")
        launches.bump("layernorm_fwd")
        ctx.save_for_backward(hsum if hsum is not None else x2, mean, rstd, gamma)
        ctx.shape = x.shape
        ctx.has_res = residual is not None
        ctx.params = (gparam, bparam)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        h, mean, rstd, gamma = ctx.saved_tensors
        gparam, bparam = ctx.params
        T, H = h.shape
        dy2 = dy.reshape(T, H).contiguous()
        dx = torch.empty_like(h)
        gsink, bsink = _sink(gparam), _sink(bparam)
        gdt = gsink.dtype if gsink is not None else gamma.dtype
        dg = gsink if gsink is not None else torch.empty(H, dtype=gdt, device=h.device)
        db = bsink if bsink is not None else torch.empty(H, dtype=gdt, device=h.device)
        L = _lib.lib()
        L.mlt_layernorm_bwd_scratch_bytes.restype = _c_ll
        ws = scratch(h.device, "ln_bwd", L.mlt_layernorm_bwd_scratch_bytes(_c_int(T), _c_int(H)))
        check(L.mlt_layernorm_bwd(ptr(dy2), ptr(h), ptr(mean), ptr(rstd), ptr(gamma), _c_int(int(gamma.dtype == torch.bfloat16)),
                                  ptr(dx), ptr(dg), ptr(db), _c_int(int(gdt == torch.bfloat16)), ptr(ws), _c_ll(ws.numel()),
                                  _c_int(T), _c_int(H), stream_handle()), "mlt_layernorm_bwd")
        launches.bump("layernorm_bwd")
        dxv = dx.view(ctx.shape)
        gres = dxv if ctx.has_res else None
        if ctx.skip is not None and ctx.has_res:
            ctx.skip.dres, gres = dx, None            # [T, H]: added by the data-gradient GEMM that produces the other half
        return dxv, gres, _deliver(gparam, dg, gsink), _deliver(bparam, db, bsink), None, None, None, None


def layer_norm_native_ok(x, gamma) -> bool:
    H = x.shape[-1]
    return bool(use_native(x, gamma) and x.dtype == torch.bfloat16 and H % 8 == 0 and H <= 1024
                and gamma.dtype in (torch.bfloat16, torch.float32))


def layer_norm(x, gamma, beta, eps: float = 1e-12, residual=None, skip=None):
    """``skip`` (with ``residual``): see :class:`SkipGrad`; only pass one when :func:`layer_norm_native_ok` holds."""
    H = x.shape[-1]
    if layer_norm_native_ok(x, gamma):
        if _has_sink(gamma, beta):
            return _LayerNormFn.apply(x, residual, gamma.detach(), beta.detach(), eps, gamma, beta, skip)
        return _LayerNormFn.apply(x, residual, gamma, beta, eps, None, None, skip)
    assert skip is None, "layer_norm: a SkipGrad needs the native path"
    h = x if residual is None else x + residual
    return TF.layer_norm(h.float(), (H,), gamma.float(), beta.float(), eps).to(x.dtype)


class _RMSNormFn(torch.autograd.Function):
    """RMSNorm(x [+ residual]) * gamma (csrc/rmsnorm.cu); saves the pre-norm sum and rstd, one-pass backward."""

    @staticmethod
    def forward(ctx, x, residual, gamma, eps, gparam):
        shape = x.shape
        H = shape[-1]
        x2 = x.contiguous().view(-1, H)
        T = x2.shape[0]
        r2 = residual.contiguous().view(-1, H) if residual is not None else None
        y = torch.empty_like(x2)
        hsum = torch.empty_like(x2) if r2 is not None else None
        rstd = torch.empty(T, dtype=torch.float32, device=x.device)
        check(_lib.lib().mlt_rmsnorm_fwd(ptr(x2), ptr(r2), ptr(gamma), _c_int(int(gamma.dtype == torch.bfloat16)), ptr(y), ptr(hsum),
                                         ptr(rstd), _c_int(T), _c_int(H), ctypes.c_float(eps), stream_handle()), "mlt_rmsnorm_fwd")
        launches.bump("rmsnorm_fwd")
        ctx.save_for_backward(hsum if hsum is not None else x2, rstd, gamma)
        ctx.shape, ctx.has_res, ctx.gparam, ctx.gdtype = shape, r2 is not None, gparam, gamma.dtype
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy):
        h, rstd, gamma = ctx.saved_tensors
        gparam = ctx.gparam
        T, H = h.shape
        dy2 = dy.contiguous().view(T, H)
        dx = torch.empty_like(h)
        gsink = _sink(gparam)
        gdt = gsink.dtype if gsink is not None else ctx.gdtype
        dg = gsink if gsink is not None else torch.empty(H, dtype=gdt, device=h.device)
        L = _lib.lib()
        L.mlt_rmsnorm_bwd_scratch_bytes.restype = _c_ll
        ws = scratch(h.device, "rms_bwd", L.mlt_rmsnorm_bwd_scratch_bytes())
        check(L.mlt_rmsnorm_bwd(ptr(dy2), ptr(h), ptr(rstd), ptr(gamma), _c_int(int(gamma.dtype == torch.bfloat16)), ptr(dx), ptr(dg),
                                _c_int(int(gdt == torch.bfloat16)), ptr(ws), _c_ll(ws.numel()), _c_int(T), _c_int(H), stream_handle()),
              "mlt_rmsnorm_bwd")
        launches.bump("rmsnorm_bwd")
        dxv = dx.view(ctx.shape)
        return dxv, (dxv if ctx.has_res else None), _deliver(gparam, dg, gsink), None, None


def rms_norm(x, gamma, eps: float = 1e-6, residual=None):
    """``RMSNorm(x [+ residual]) * gamma`` over the last dimension (what ``torch.nn.RMSNorm`` computes)."""
    H = x.shape[-1]
    if (use_native(x, gamma) and x.dtype == torch.bfloat16 and H % 8 == 0 and 8 <= H <= 8192
            and gamma.dtype in (torch.bfloat16, torch.float32) and x.numel() > 0):
        if _has_sink(gamma):
            return _RMSNormFn.apply(x, residual, gamma.detach(), eps, gamma)
        return _RMSNormFn.apply(x, residual, gamma, eps, None)
    h = x if residual is None else x + residual
    hf = h.float()
    return (hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps) * gamma.float()).to(x.dtype)


class _AttentionFn(torch.autograd.Function):
    """softmax(Q K^T / sqrt(d)) V per (sample, head) as batched tcgen05 GEMMs + row-softmax kernels.

    q, k, v: [B, S, heads*d] bf16 (row stride = heads*d); output the same layout.  P is kept for backward.
    """

    @staticmethod
    def forward(ctx, q, k, v, heads):
        from .gemm import gemm_batched

        B, S, HD = q.shape
        d = HD // heads
        scale = 1.0 / (d ** 0.5)
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        P = torch.empty((B, heads, S, S), dtype=torch.bfloat16, device=q.device)
        # scores[b,h] = Q[b,:,h,:] K[b,:,h,:]^T          operand (z1 = head, z2 = sample)
        gemm_batched(q, HD, d, S * HD, False, k, HD, d, S * HD, False, P, S, S * S, heads * S * S, S, S, d, heads, B)
        check(_lib.lib().mlt_softmax_rows(ptr(P), _c_ll(B * heads * S), _c_int(S), ctypes.c_float(scale), stream_handle()),
              "mlt_softmax_rows")
        launches.bump("softmax_rows")
        ctxt = torch.empty_like(q)
        # out[b,:,h,:] = P[b,h] V[b,:,h,:]               B operand stored [keys][d]: MN-major
        gemm_batched(P, S, S * S, heads * S * S, False, v, HD, d, S * HD, True, ctxt, HD, d, S * HD, S, d, S, heads, B)
        ctx.save_for_backward(q, k, v, P)
        ctx.cfg = (heads, scale)
        return ctxt

    @staticmethod
    def backward(ctx, do):
        from .gemm import gemm_batched

        q, k, v, P = ctx.saved_tensors
        heads, scale = ctx.cfg
        B, S, HD = q.shape
        d = HD // heads
        do = do.contiguous()
        L = _lib.lib()
        dP = torch.empty_like(P)
        # dP = dO V^T
        gemm_batched(do, HD, d, S * HD, False, v, HD, d, S * HD, False, dP, S, S * S, heads * S * S, S, S, d, heads, B)
        # dV = P^T dO
        dv = torch.empty_like(v)
        gemm_batched(P, S, S * S, heads * S * S, True, do, HD, d, S * HD, True, dv, HD, d, S * HD, S, d, S, heads, B)
        check(L.mlt_softmax_rows_bwd(ptr(dP), ptr(P), _c_ll(B * heads * S), _c_int(S), ctypes.c_float(scale), stream_handle()),
              "mlt_softmax_rows_bwd")
        launches.bump("softmax_rows_bwd")
        # dQ = dS K ;  dK = dS^T Q
        dq, dk = torch.empty_like(q), torch.empty_like(k)
        gemm_batched(dP, S, S * S, heads * S * S, False, k, HD, d, S * HD, True, dq, HD, d, S * HD, S, d, S, heads, B)
        gemm_batched(dP, S, S * S, heads * S * S, True, q, HD, d, S * HD, True, dk, HD, d, S * HD, S, d, S, heads, B)
        return dq, dk, dv, None


def _qkv_ld(q, k, v):
    """Common row stride of q, k, v when they can be handed to the fused kernels without a copy, else None."""
    B, S, HD = q.shape
    ld = q.stride(1)
    for t in (q, k, v):
        if t.stride(2) != 1 or t.stride(1) != ld or t.stride(0) != S * ld or t.data_ptr() % 16:
            return None
    return ld if ld % 8 == 0 and ld >= HD else None


def _key_bias(key_bias, B, S, device):
    """``[B, S]`` fp32 contiguous additive key bias (HF's extended mask ``[B, 1, 1, S]`` is accepted), or None."""
    if key_bias is None:
        return None
    if getattr(key_bias, "_mlt_key_bias", False):        # prepared once per forward pass by the model (prepare_key_bias)
        return key_bias
    kb = key_bias.reshape(B, S) if key_bias.numel() == B * S else key_bias.expand(B, 1, 1, S).reshape(B, S)
    kb = kb.to(device=device, dtype=torch.float32).clamp_min(-30000.0)       # finfo.min would overflow the exp2-domain arithmetic
    return kb.contiguous()


def prepare_key_bias(attention_mask: torch.Tensor) -> torch.Tensor:
    """HuggingFace-style ``attention_mask`` [B, S] (1 = token, 0 = padding) -> the additive key bias the attention kernels take,
    prepared once per forward pass (every layer reuses it)."""
    kb = ((1.0 - attention_mask.float()) * -30000.0).contiguous()
    kb._mlt_key_bias = True
    return kb


def _drop_args(drop, device):
    """ctypes tail ``(p, seed, step_ptr, call_id)`` of the ``_ex`` attention entry points; ``drop`` is None or ``(p, call_id)``."""
    if drop is None:
        return ctypes.c_float(0.0), ctypes.c_ulonglong(0), None, ctypes.c_uint(0)
    return ctypes.c_float(drop[0]), ctypes.c_ulonglong(_drop["seed"]), ptr(_drop_step(device)), ctypes.c_uint(drop[1])


def _new_drop(p_drop: float, training: bool):
    """A fresh call site for in-kernel attention dropout, or None when dropout is off."""
    if not training or p_drop <= 0.0:
        return None
    _drop["call"] += 1
    return (float(p_drop), _drop["call"])


def _mix32(x: torch.Tensor) -> torch.Tensor:
    M = 0xFFFFFFFF
    x = ((x ^ (x >> 16)) * 0x7FEB352D) & M
    x = ((x ^ (x >> 15)) * 0x846CA68B) & M
    return x ^ (x >> 16)


def attention_keep_mask(B: int, heads: int, S: int, p_drop: float, call_id: int, step: int = None, seed: int = None,
                        device="cpu") -> tuple:
    """Host statement of the mask the fused attention kernels draw (csrc/attention_tcgen05.cu, ``FaDrop``): returns
    ``(keep [B, heads, S, S] bool, scale)`` with ``scale = 256 / thr``.  Used by the tests as the oracle's mask."""
    M = 0xFFFFFFFF
    seed = _drop["seed"] if seed is None else seed
    if step is None:
        step = int(_drop_step(device).item()) if torch.device(device).type == "cuda" else 0
    thr = min(max(int(round((1.0 - p_drop) * 256.0)), 1), 256)
    as_t = lambda v: torch.tensor(v & M, dtype=torch.int64)
    k0 = _mix32(as_t((seed & M) ^ ((call_id * 0x9E3779B9) & M) ^ (step & M)))
    k1 = _mix32(as_t(((seed >> 32) & M) ^ ((call_id * 0x85EBCA6B) & M) ^ ((step >> 32) & M) ^ 0x5BD1E995))
    S2 = S // 2
    bh = torch.arange(B * heads, dtype=torch.int64).view(-1, 1, 1)
    i = torch.arange(S, dtype=torch.int64).view(1, -1, 1)
    j = torch.arange(S, dtype=torch.int64).view(1, 1, -1)
    idx = (((bh * S2 + (i >> 1)) & M) * S2 + (j >> 1)) & M
    x = idx ^ k0
    x = ((x ^ (x >> 16)) * 0x7FEB352D) & M
    x = ((x ^ (x >> 15) ^ k1) * 0x846CA68B) & M
    x = x ^ (x >> 16)
    byte = (x >> (((i & 1) * 2 + (j & 1)) * 8)) & 0xFF
    return (byte < thr).view(B, heads, S, S).to(device), 256.0 / thr


def attention_fused_fwd(q, k, v, heads: int, want_lse: bool = False, key_bias=None, drop=None):
    """Fused tcgen05 attention forward (csrc/attention_tcgen05.cu).  Returns ``out`` or ``(out, lse)``.

    q, k, v may be column slices of one packed ``[B, S, 3*H*d]`` projection (no copy is made).  ``key_bias``: optional
    ``[B, S]`` additive bias per key (0 for tokens, a large negative number for padding), added after the 1/sqrt(d) scaling.
    ``drop``: None, or ``(p, call_id)`` for dropout on the probabilities drawn inside the kernel."""
    B, S, HD = q.shape
    ld = _qkv_ld(q, k, v)
    if ld is None:
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        ld = HD
    out = torch.empty((B, S, HD), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, heads, S), dtype=torch.float32, device=q.device) if want_lse else None
    check(_lib.lib().mlt_attention_fwd_ex(ptr(q), ptr(k), ptr(v), _c_ll(ld), ptr(key_bias), ptr(out), ptr(lse), _c_int(B), _c_int(S),
                                          _c_int(heads), _c_int(HD // heads), *_drop_args(drop, q.device), stream_handle()),
          "mlt_attention_fwd")
    launches.bump("attention_fwd_tcgen05")
    return (out, lse) if want_lse else out


def _attention_bwd(q, k, v, out, lse, do, heads, dq, dk, dv, key_bias=None, drop=None):
    B, S, HD = q.shape
    ld = _qkv_ld(q, k, v)
    ldd = _qkv_ld(dq, dk, dv)
    assert ld is not None and ldd is not None
    delta = torch.empty_like(lse)
    check(_lib.lib().mlt_attention_bwd_ex(ptr(q), ptr(k), ptr(v), _c_ll(ld), ptr(key_bias), ptr(out), ptr(do), ptr(lse), ptr(delta),
                                          ptr(dq), ptr(dk), ptr(dv), _c_ll(ldd), _c_int(B), _c_int(S), _c_int(heads),
                                          _c_int(HD // heads), *_drop_args(drop, q.device), stream_handle()), "mlt_attention_bwd")
    launches.bump("attention_bwd_tcgen05", 3)


class _FusedAttentionFn(torch.autograd.Function):
    """Flash attention on tcgen05: forward keeps only the per-row log-sum-exp, backward recomputes P on the tensor cores
    (csrc/attention_tcgen05.cu: delta, dQ and dK/dV kernels, no atomics)."""

    @staticmethod
    def forward(ctx, q, k, v, heads, key_bias=None, drop=None):
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        out, lse = attention_fused_fwd(q, k, v, heads, want_lse=True, key_bias=key_bias, drop=drop)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.heads, ctx.key_bias, ctx.drop = heads, key_bias, drop
        return out

    @staticmethod
    def backward(ctx, do):
        q, k, v, out, lse = ctx.saved_tensors
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _attention_bwd(q, k, v, out, lse, do.contiguous(), ctx.heads, dq, dk, dv, ctx.key_bias, ctx.drop)
        return dq, dk, dv, None, None, None


class _PackedAttentionFn(torch.autograd.Function):
    """Same kernels on a packed ``[B, S, 3*H*d]`` QKV projection: q, k, v are read as column slices and dq, dk, dv are
    written straight into one packed gradient, so neither direction needs a split or a concat."""

    @staticmethod
    def forward(ctx, qkv, heads, key_bias=None, drop=None):
        HD = qkv.shape[-1] // 3
        q, k, v = qkv[..., :HD], qkv[..., HD:2 * HD], qkv[..., 2 * HD:]
        out, lse = attention_fused_fwd(q, k, v, heads, want_lse=True, key_bias=key_bias, drop=drop)
        ctx.save_for_backward(qkv, out, lse)
        ctx.heads, ctx.key_bias, ctx.drop = heads, key_bias, drop
        return out

    @staticmethod
    def backward(ctx, do):
        qkv, out, lse = ctx.saved_tensors
        HD = qkv.shape[-1] // 3
        dqkv = torch.empty_like(qkv)
        _attention_bwd(qkv[..., :HD], qkv[..., HD:2 * HD], qkv[..., 2 * HD:], out, lse, do.contiguous(), ctx.heads,
                       dqkv[..., :HD], dqkv[..., HD:2 * HD], dqkv[..., 2 * HD:], ctx.key_bias, ctx.drop)
        return dqkv, None, None, None


def attention_packed(qkv, heads: int, key_bias=None, p_drop: float = 0.0, training: bool = False):
    """Self-attention core on a packed projection ``[B, S, 3*H*d]`` (q | k | v along the last dim).  ``p_drop``: dropout on the
    attention probabilities (active when ``training``), drawn inside the fused kernels."""
    HD = qkv.shape[-1] // 3
    q = qkv[..., :HD]
    if (use_native(qkv) and qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and fused_attention_ok(q, heads)
            and qkv.data_ptr() % 16 == 0):
        kb = _key_bias(key_bias, qkv.shape[0], qkv.shape[1], qkv.device)
        drop = _new_drop(p_drop, training)
        if torch.is_grad_enabled() and qkv.requires_grad:
            return _PackedAttentionFn.apply(qkv, heads, kb, drop)
        return attention_fused_fwd(q, qkv[..., HD:2 * HD], qkv[..., 2 * HD:], heads, key_bias=kb, drop=drop)
    return attention(q, qkv[..., HD:2 * HD], qkv[..., 2 * HD:], heads, key_bias, p_drop, training)


def fused_attention_ok(q, heads: int) -> bool:
    B, S, HD = q.shape
    return HD // heads == 64 and S % 64 == 0 and 64 <= S <= 512


def attention(q, k, v, heads: int, key_bias=None, p_drop: float = 0.0, training: bool = False):
    """Multi-head self-attention core.  q, k, v: [B, S, heads*d]; ``key_bias``: optional additive ``[B, S]`` (or HF's
    ``[B, 1, 1, S]``) key-padding bias; ``p_drop``: dropout on the probabilities while ``training``.  Head dim 64 and S a
    multiple of 64 up to 512 run as one fused tcgen05 kernel per direction (mask and dropout included); other bf16 shapes
    without mask or dropout use batched tcgen05 GEMMs + row softmax; the rest (and the CPU) is stated with torch ops."""
    B, S, HD = q.shape
    d = HD // heads
    dropping = training and p_drop > 0.0
    if use_native(q, k, v) and q.dtype == torch.bfloat16 and d % 8 == 0 and S % 8 == 0 and S <= 1024:
        needs_grad = torch.is_grad_enabled() and (q.requires_grad or k.requires_grad or v.requires_grad)
        if fused_attention_ok(q, heads):
            kb = _key_bias(key_bias, B, S, q.device)
            drop = _new_drop(p_drop, training)
            if needs_grad:
                return _FusedAttentionFn.apply(q, k, v, heads, kb, drop)
            return attention_fused_fwd(q, k, v, heads, key_bias=kb, drop=drop)
        if key_bias is None and not dropping:
            return _AttentionFn.apply(q, k, v, heads)
    qh, kh, vh = (t.view(B, S, heads, d).transpose(1, 2) for t in (q, k, v))
    mask = None if key_bias is None else key_bias.reshape(B, 1, 1, S).float()
    if dropping:
        scores = (qh.float() @ kh.float().transpose(-1, -2)) * (1.0 / d ** 0.5)
        if mask is not None:
            scores = scores + mask
        o = dropout(torch.softmax(scores, dim=-1), p_drop, True) @ vh.float()
    else:
        o = TF.scaled_dot_product_attention(qh.float(), kh.float(), vh.float(), attn_mask=mask)
    return o.transpose(1, 2).reshape(B, S, HD).to(q.dtype)


class _EmbeddingFn(torch.autograd.Function):
    """word[ids] + position[0..S) + token_type[0] -> [B, S, H] bf16 (BERT embeddings before their LayerNorm)."""

    @staticmethod
    def forward(ctx, ids, word, pos, typ, wparam, pparam, tparam):
        B, S = ids.shape
        H = word.shape[1]
        out = torch.empty((B, S, H), dtype=torch.bfloat16, device=ids.device)
        ids = ids.contiguous()
        check(_lib.lib().mlt_embedding_sum_fwd_v(ptr(ids), ptr(word), ptr(pos), ptr(typ), _c_int(int(word.dtype == torch.bfloat16)),
                                                 ptr(out), _c_int(B * S), _c_int(S), _c_int(H), _c_ll(word.shape[0]), stream_handle()),
              "mlt_embedding_sum_fwd")
        launches.bump("embedding_fwd")
        ctx.save_for_backward(ids)
        ctx.meta = (word.shape, pos.shape, typ.shape, word.dtype)
        ctx.params = (wparam, pparam, tparam)
        return out

    @staticmethod
    def backward(ctx, dh):
        (ids,) = ctx.saved_tensors
        wshape, pshape, tshape, wdtype = ctx.meta
        wparam, pparam, tparam = ctx.params
        B, S = ids.shape
        H = wshape[1]
        dh = dh.contiguous()
        ws, ps, ts = _sink(wparam), _sink(pparam), _sink(tparam)
        gdt = ws.dtype if ws is not None else wdtype
        dev = dh.device
        dword = ws if ws is not None else torch.empty(wshape, dtype=gdt, device=dev)
        dpos = ps if ps is not None else torch.zeros(pshape, dtype=gdt, device=dev)
        dtyp = ts if ts is not None else torch.zeros(tshape, dtype=gdt, device=dev)
        check(_lib.lib().mlt_embedding_bwd(ptr(ids), ptr(dh), ptr(dword), _c_ll(wshape[0]), ptr(dpos),
                                           _c_int(int(gdt == torch.bfloat16)), _c_int(B), _c_int(S), _c_int(H), stream_handle()),
              "mlt_embedding_bwd")
        launches.bump("embedding_bwd", 2)
        colsum(dh.view(B * S, H), out=dtyp[0])
        return None, _deliver(wparam, dword, ws), _deliver(pparam, dpos, ps), _deliver(tparam, dtyp, ts), None, None, None


def bert_embeddings(ids, word, pos, typ):
    if use_native(ids, word) and word.dtype in (torch.bfloat16, torch.float32) and word.shape[1] % 2 == 0:
        if _has_sink(word, pos, typ):
            return _EmbeddingFn.apply(ids, word.detach(), pos.detach(), typ.detach(), word, pos, typ)
        return _EmbeddingFn.apply(ids, word, pos, typ, None, None, None)
    S = ids.shape[1]
    return (word[ids] + pos[:S].unsqueeze(0) + typ[0]).to(torch.bfloat16 if word.is_cuda else word.dtype)
