"""Flat-arena fused optimizer (replaces ``torch.optim.*`` at reference src/trainer.py:123-138).

Layout
------
All trainable parameters live in ONE flat buffer per role:

* ``master``  fp32 weights (what checkpoints are written from),
* ``model``   compute-dtype weights the modules read (aliases ``master`` for fp32,
  a bf16 shadow otherwise),
* ``grad``    compute-dtype gradients; every ``param.grad`` destination is a view of it,
* ``s1/s2``   fp32 optimizer state.

Parameters are laid out in *reverse* registration order, so the gradients autograd
produces first (last layers) sit at the lowest offsets and a gradient bucket is a
contiguous ``[lo, hi)`` slice that can be reduced and updated by one kernel launch
without a flatten copy.  Offsets are padded to 128 elements so bf16 weight views stay
16-byte aligned for TMA descriptors.

The update itself is one launch over a slice (``step_range``) that fuses the
1/world gradient scale, the bf16->fp32 cast, the rule, and the compute-dtype
write-back.  Hyper-parameters (lr, bias corrections...) are read from a small device
vector refreshed from the host each step, so a captured CUDA graph replays with a new
learning rate.  On CPU the same rules run through :mod:`.rules`.

The class subclasses ``torch.optim.Optimizer`` with a single param group so that the
stock ``torch.optim.lr_scheduler`` classes the reference uses drive it unchanged.
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional

import torch

from . import rules
from .rules import DEFAULT_EPS, HYPER_LEN, RULE_IDS, RULE_STATES, Hyper

ALIGN = 128  # elements


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class PackedParams:
    """Several parameters that sit back to back in the arena, seen as one tensor (``data``) with one gradient sink."""

    def __init__(self, data, grad, members):
        self.data, self.grad, self.members = data, grad, list(members)

    def ready(self):
        for p in self.members:
            r = getattr(p, "_mlt_ready", None)
            if r is not None:
                r()

    # uses per step are counted on the first member (see ops.functional: sinks are written, not accumulated)
    def note_use(self):
        from ..ops.functional import _note_use

        _note_use(self.members[0])

    def first_arrival(self) -> bool:
        p = self.members[0]
        return not (getattr(p, "_mlt_uses", 1) > 1 and getattr(p, "_mlt_done", 0) > 0)

    def deliver(self):
        p = self.members[0]
        uses = getattr(p, "_mlt_uses", 1)
        if uses > 1:
            p._mlt_done = getattr(p, "_mlt_done", 0) + 1
            if p._mlt_done < uses:
                return
        self.ready()


class FusedOptimizer(torch.optim.Optimizer):
    def __init__(
        self,
        params: Iterable[torch.nn.Parameter],
        kind: str = "sgd",
        lr: float = 1e-3,
        momentum: float = 0.9,
        weight_decay: float = 0.0,
        betas=(0.9, 0.999),
        eps: Optional[float] = None,
        compute_dtype: Optional[torch.dtype] = None,
        allocator: Optional[Callable[[int, torch.dtype, torch.device], torch.Tensor]] = None,
    ):
        if kind not in RULE_IDS:
            raise ValueError(f"unknown optimizer {kind!r}; expected one of {tuple(RULE_IDS)}")
        plist = [p for p in params if p.requires_grad]
        if not plist:
            raise ValueError("optimizer got an empty parameter list")
        self.kind = kind
        defaults = dict(lr=lr, momentum=momentum, weight_decay=weight_decay,
                        betas=tuple(betas), eps=DEFAULT_EPS[kind] if eps is None else eps)
        super().__init__(plist, defaults)
        self.device = plist[0].device
        self.compute_dtype = compute_dtype or plist[0].dtype
        if self.compute_dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("compute dtype must be fp32 or bf16")
        self._alloc = allocator or (lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev))
        self.step_count = 0
        self._build_arena(plist)

    # ------------------------------------------------------------------ arena
    @staticmethod
    def _apply_packs(order):
        """Honour ``param._mlt_pack = (key, index, count)``: the members of a complete pack are laid out back to back
        (in index order) so that ONE tensor view covers all of them - e.g. the Q, K and V projection weights of an
        attention block become a single [3*H, H] GEMM operand and a single [3*H, H] gradient sink."""
        groups = {}
        for p in order:
            tag = getattr(p, "_mlt_pack", None)
            if tag is not None:
                groups.setdefault(tag[0], {})[tag[1]] = p
        complete = {}
        for key, members in groups.items():
            n = next(iter(members.values()))._mlt_pack[2]
            ps = [members.get(i) for i in range(n)]
            if any(q is None for q in ps):
                continue
            if any(q.numel() % ALIGN or q.shape[1:] != ps[0].shape[1:] or q.dtype != ps[0].dtype for q in ps):
                continue
            complete[key] = ps
        out, seen = [], set()
        for p in order:
            if id(p) in seen:
                continue
            tag = getattr(p, "_mlt_pack", None)
            members = complete.get(tag[0]) if tag is not None else None
            for q in (members or [p]):
                seen.add(id(q))
                out.append(q)
        return out, list(complete.values())

    def _build_arena(self, plist):
        order, packs = self._apply_packs(list(reversed(plist)))           # gradient-ready order
        self.params_in_order = order
        self.offsets, self.sizes = [], []
        off = 0
        for p in order:
            self.offsets.append(off)
            self.sizes.append(p.numel())
            off += _round_up(p.numel(), ALIGN)
        self.numel = off
        dev = self.device
        self.master = self._alloc(off, torch.float32, dev)
        self.grad = self._alloc(off, self.compute_dtype, dev)
        self.model = self.master if self.compute_dtype == torch.float32 else self._alloc(off, self.compute_dtype, dev)
        nstate = RULE_STATES[self.kind]
        # optimizer state comes from the same allocator as the weights: with the NVLink backend that is peer-addressable
        # arena memory, which lets a checkpoint pull the slices other ranks own (sharded update kernels)
        self.s1 = self._alloc(off, torch.float32, dev)
        self.s2 = self._alloc(off, torch.float32, dev) if nstate > 1 else torch.zeros(1, dtype=torch.float32, device=dev)
        self.hyper = torch.zeros(HYPER_LEN, dtype=torch.float32, device=dev)
        self.index_of = {}
        # conv weights marked channels-last live physically as [Cout, KH, KW, Cin] (the GEMM operand layout of the
        # NHWC convolution path) while keeping their logical [Cout, Cin, KH, KW] shape
        self._cl = [bool(getattr(p, "_mlt_channels_last", False)) and p.dim() == 4 for p in order]
        with torch.no_grad():
            for i, p in enumerate(order):
                src = p.detach().clone()
                self._view(self.master, i).copy_(src.float())
                if self.model is not self.master:
                    self._view(self.model, i).copy_(src)
                p.data = self._view(self.model, i)
                p.grad = None
                p._mlt_grad_sink = self._view(self.grad, i)   # where backward kernels write
                p._mlt_index = i
                self.index_of[id(p)] = i
            for members in packs:
                i0 = self.index_of[id(members[0])]
                lo, n = self.offsets[i0], sum(q.numel() for q in members)
                shape = (sum(q.shape[0] for q in members),) + tuple(members[0].shape[1:])
                members[0]._mlt_packed = PackedParams(self.model[lo:lo + n].view(shape), self.grad[lo:lo + n].view(shape), members)

    def _view(self, flat: torch.Tensor, i: int) -> torch.Tensor:
        o, n = self.offsets[i], self.sizes[i]
        shape = self.params_in_order[i].shape
        if self._cl[i]:
            Co, Ci, KH, KW = shape
            return flat[o:o + n].view(Co, KH, KW, Ci).permute(0, 3, 1, 2)
        return flat[o:o + n].view(shape)

    def grad_view(self, i: int) -> torch.Tensor:
        return self._view(self.grad, i)

    def master_view(self, i: int) -> torch.Tensor:
        return self._view(self.master, i)

    def span(self, i_lo: int, i_hi: int):
        """Element range covering params ``i_lo .. i_hi-1`` (arena order)."""
        lo = self.offsets[i_lo]
        hi = self.numel if i_hi >= len(self.offsets) else self.offsets[i_hi]
        return lo, hi

    # ------------------------------------------------------------------ hyper
    def hyper_for_next_step(self, grad_scale: float = 1.0) -> Hyper:
        g = self.param_groups[0]
        b1 = g["momentum"] if self.kind == "sgd" else g["betas"][0]
        return Hyper(lr=float(g["lr"]), wd=float(g["weight_decay"]), b1=float(b1), b2=float(g["betas"][1]),
                     eps=float(g["eps"]), step=self.step_count + 1, grad_scale=grad_scale)

    def hyper_rows(self, n: int, grad_scale: float = 1.0, lrs=None) -> torch.Tensor:
        """The next ``n`` steps' hyper vectors as an ``[n, 8]`` fp32 tensor (what the native executors upload per run):
        ``Hyper.pack()`` for steps ``step_count+1 .. step_count+n``, vectorised.  ``lrs``: optional per-step learning rates."""
        import numpy as np

        g = self.param_groups[0]
        b1 = float(g["momentum"] if self.kind == "sgd" else g["betas"][0])
        b2 = float(g["betas"][1])
        if lrs is None and self.kind in ("sgd", "adagrad"):
            # these rules read no bias corrections: every row of the epoch is the same eight numbers
            key = (n, float(g["lr"]), float(g["weight_decay"]), b1, b2, float(g["eps"]), float(grad_scale))
            if getattr(self, "_rows_key", None) != key:
                row = torch.tensor([key[1], key[2], b1, b2, key[5], 1.0 - b1, (1.0 - b2) ** 0.5, key[6]], dtype=torch.float32)
                self._rows_cache, self._rows_key = row.repeat(n, 1).contiguous(), key
            return self._rows_cache
        steps = np.arange(self.step_count + 1, self.step_count + n + 1, dtype=np.float64)
        rows = np.empty((n, 8), dtype=np.float64)
        rows[:, 0] = float(g["lr"]) if lrs is None else np.asarray(lrs, dtype=np.float64)
        rows[:, 1] = float(g["weight_decay"])
        rows[:, 2] = b1
        rows[:, 3] = b2
        rows[:, 4] = float(g["eps"])
        rows[:, 5] = 1.0 - np.power(b1, steps)
        rows[:, 6] = np.sqrt(1.0 - np.power(b2, steps))
        rows[:, 7] = grad_scale
        return torch.from_numpy(rows.astype(np.float32))

    # ------------------------------------------------------------------- step
    def begin_step(self, grad_scale: float = 1.0, upload: bool = True) -> Hyper:
        """Advance the step counter and publish this step's hyper-parameters."""
        from ..ops import functional as F

        F.new_step()                 # per-step use counters of parameters that appear more than once in the forward pass
        h = self.hyper_for_next_step(grad_scale)
        self.step_count += 1
        self._opt_called = True      # what torch's lr_scheduler looks at to order step() calls
        self._hyper_host = h
        if upload and self.device.type == "cuda":
            F.upload_hyper(self.hyper, h.pack())
        return h

    def step_range(self, lo: int, hi: int, hyper: Optional[Hyper] = None):
        """Apply the rule to arena elements ``[lo, hi)``; gradients must already be reduced."""
        h = hyper or self._hyper_host
        if self.device.type == "cuda":
            from ..ops import functional as F

            F.fused_update(self, lo, hi)
            return
        s2 = self.s2[lo:hi] if self.s2.numel() > 1 else self.s2
        p = self.master[lo:hi]
        rules.apply_rule(self.kind, p, self.grad[lo:hi], self.s1[lo:hi], s2, h)
        if self.model is not self.master:
            self.model[lo:hi].copy_(p)

    @torch.no_grad()
    def step(self, closure=None):
        """Whole-arena update (used when no reducer drives ``step_range`` per bucket)."""
        loss = closure() if closure is not None else None
        self.collect_stray_grads()
        self.begin_step()
        self.step_range(0, self.numel)
        return loss

    @torch.no_grad()
    def collect_stray_grads(self):
        """Move gradients autograd left in ``param.grad`` into the flat buffer."""
        for i, p in enumerate(self.params_in_order):
            if p.grad is not None:
                self.grad_view(i).copy_(p.grad)
                p.grad = None

    def zero_grad(self, set_to_none: bool = True):
        # Gradient slots are overwritten (never accumulated) each step, so there is
        # nothing to clear; only drop stray autograd-owned grads.
        for p in self.params_in_order:
            p.grad = None

    # ------------------------------------------------------------- checkpoint
    def master_state_dict(self, model: torch.nn.Module, prefix: str = ""):
        """fp32 ``state_dict`` of ``model`` with parameters taken from the master copy."""
        from collections import OrderedDict

        out = OrderedDict()
        for name, t in model.state_dict().items():
            out[prefix + name] = t.detach()
        for name, p in model.named_parameters():
            i = self.index_of.get(id(p))
            if i is not None:
                out[prefix + name] = self.master_view(i).detach()
        return out

    @torch.no_grad()
    def reload_from_params(self):
        """Re-read master/model copies after someone overwrote ``param.data`` contents in place."""
        for i, p in enumerate(self.params_in_order):
            self._view(self.master, i).copy_(p.detach().float())

    @torch.no_grad()
    def load_master(self, named_tensors: dict, model: torch.nn.Module):
        for name, p in model.named_parameters():
            i = self.index_of.get(id(p))
            if i is None or name not in named_tensors:
                continue
            src = named_tensors[name].to(self.device)
            self._view(self.master, i).copy_(src.float())
            if self.model is not self.master:
                self._view(self.model, i).copy_(src)

    def state_dict(self):
        return {
            "kind": self.kind, "step_count": self.step_count,
            "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
            "master": self.master.detach().cpu(), "s1": self.s1.detach().cpu(), "s2": self.s2.detach().cpu(),
        }

    @torch.no_grad()
    def load_state_dict(self, sd):
        if sd["kind"] != self.kind or sd["master"].numel() != self.numel:
            raise ValueError("optimizer state does not match this model/optimizer")
        self.step_count = int(sd["step_count"])
        for g, saved in zip(self.param_groups, sd["param_groups"]):
            g.update(saved)
        self.master.copy_(sd["master"])
        self.s1.copy_(sd["s1"])
        self.s2.copy_(sd["s2"])
        if self.model is not self.master:
            self.model.copy_(self.master)
