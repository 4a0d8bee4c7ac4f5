"""The reference's default model (reference src/model.py:7-24): a 62,006-parameter LeNet-style CNN.

conv(3->6,5x5) -> ReLU -> maxpool2 -> conv(6->16,5x5) -> ReLU -> maxpool2 -> flatten 400
-> fc120 -> ReLU -> fc84 -> ReLU -> fc10.

Same attribute names, parameter shapes, ``state_dict`` keys and initialisation order as the
reference (so a given seed yields identical initial weights and its checkpoints load both ways),
but ``forward`` runs five fused kernels instead of eleven library calls:
conv+bias+ReLU+pool (x2), linear+bias+ReLU (x2), linear+bias.
"""
from __future__ import annotations

import torch
from torch import nn

import ctypes

from ..ops import functional as F


class MLModel(nn.Module):
    _mlt_native = True

    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 6, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, num_classes)

    def forward(self, x):
        x = F.conv_relu_pool(x, self.conv1.weight, self.conv1.bias, relu=True, pool=True)
        x = F.conv_relu_pool(x, self.conv2.weight, self.conv2.bias, relu=True, pool=True)
        x = x.view(-1, 16 * 5 * 5)
        x = F.linear_act(x, self.fc1.weight, self.fc1.bias, "relu")
        x = F.linear_act(x, self.fc2.weight, self.fc2.bias, "relu")
        x = F.linear_act(x, self.fc3.weight, self.fc3.bias, None)
        return x


    # ------------------------------------------------------------------ fused single-kernel step
    def make_fused_step(self, optimizer, comm=None):
        """Return a callable running the whole train/eval step as ONE kernel (csrc/lenet_step.cu), or None.

        Conditions: CUDA, fp32 compute (model == master), default 10-class head, parameters owned by
        ``optimizer``'s arena, and a communicator that is local or the native P2P backend.
        """
        from ..ops import _lib

        if not (F.native_available() and optimizer.device.type == "cuda" and optimizer.model is optimizer.master):
            return None
        if self.fc3.out_features != 10:
            return None
        world = getattr(comm, "world_size", 1) if comm is not None else 1
        ctx = getattr(comm, "ctx", None) if world > 1 else None
        if world > 1 and ctx is None:
            return None                      # library collectives (gloo / nccl): use the layer kernels + reducer
        if world > 1 and optimizer.numel * 4 > comm.oneshot_max:
            return None
        params = [self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias, self.fc1.weight, self.fc1.bias,
                  self.fc2.weight, self.fc2.bias, self.fc3.weight, self.fc3.bias]
        if any(id(p) not in optimizer.index_of for p in params):
            return None
        return _FusedLeNetStep(params, optimizer, ctx)


class _FusedLeNetStep:
    def __init__(self, params, optimizer, ctx):
        from ..ops import _lib
        from ..optim.rules import RULE_IDS

        self.L = _lib.lib()
        self.L.mlt_lenet_scratch_bytes.restype = ctypes.c_longlong
        self.opt = optimizer
        self.ctx = ctx
        self.rule = RULE_IDS[optimizer.kind]
        self.off = (ctypes.c_longlong * 10)(*[optimizer.offsets[optimizer.index_of[id(p)]] for p in params])
        self._scratch = None
        self._scratch_B = 0

    def supports(self, x, y) -> bool:
        return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and tuple(x.shape[1:]) == (3, 32, 32)
                and x.is_contiguous() and y.dtype == torch.int64 and y.dim() == 1)

    def ensure_scratch(self, B: int, device):
        if self._scratch is None or B > self._scratch_B:
            nbytes = int(self.L.mlt_lenet_scratch_bytes(ctypes.c_int(B)))
            self._scratch = torch.zeros(nbytes, dtype=torch.uint8, device=device)
            self._scratch_B = B

    def __call__(self, x, y, train: bool, acc3=None, want_acc=False, loss_out=None):
        from ..ops._lib import check, launches, ptr, stream_handle

        B = x.shape[0]
        self.ensure_scratch(B, x.device)
        o = self.opt
        s2 = o.s2 if o.s2.numel() > 1 else None
        check(self.L.mlt_lenet_step(ptr(o.master), ptr(o.s1), ptr(s2), ptr(o.hyper), ctypes.c_longlong(o.numel), self.off,
                                    ctypes.c_int(self.rule), ptr(x), ptr(y), ctypes.c_int(B), ctypes.c_int(int(train)),
                                    ctypes.c_int(int(want_acc)), ptr(acc3), ptr(loss_out), ptr(self._scratch),
                                    ctypes.c_longlong(self._scratch.numel()),
                                    ctypes.byref(self.ctx) if self.ctx is not None else None, stream_handle()),
              "mlt_lenet_step")
        launches.bump("lenet_step")

    def run_pool(self, x_pool, y_pool, nsteps: int, train: bool = True, hyper_table=None, losses=None, acc3=None,
                 want_acc: bool = False):
        """``nsteps`` steps in ONE kernel launch over a device-resident pool of batches (step i uses batch i % P).

        x_pool: [P, B, 3, 32, 32] fp32, y_pool: [P, B] int64, both contiguous on the device.  hyper_table: optional
        [nsteps, 8] fp32 device tensor of per-step hyper rows (else the optimizer's current row is used for every step);
        losses: optional [nsteps] fp32 device tensor that receives the per-step losses."""
        from ..ops._lib import check, launches, ptr, stream_handle

        P, B = int(x_pool.shape[0]), int(x_pool.shape[1])
        assert x_pool.is_contiguous() and y_pool.is_contiguous() and x_pool.dtype == torch.float32 and y_pool.dtype == torch.int64
        self.ensure_scratch(B, x_pool.device)
        o = self.opt
        s2 = o.s2 if o.s2.numel() > 1 else None
        xs = (ctypes.c_void_p * 1)(x_pool.data_ptr())
        ys = (ctypes.c_void_p * 1)(y_pool.data_ptr())
        check(self.L.mlt_lenet_steps(ptr(o.master), ptr(o.s1), ptr(s2), ptr(o.hyper), ctypes.c_longlong(o.numel), self.off,
                                     ctypes.c_int(self.rule), xs, ys, ctypes.c_int(1), ctypes.c_longlong(P), ctypes.c_int(B), None,
                                     ptr(hyper_table), ptr(losses), None, None, ctypes.c_int(nsteps), ctypes.c_int(int(train)),
                                     ctypes.c_int(int(want_acc)), ptr(acc3), ptr(self._scratch), ctypes.c_longlong(self._scratch.numel()),
                                     ctypes.byref(self.ctx) if self.ctx is not None else None, stream_handle()), "mlt_lenet_steps")
        launches.bump("lenet_step", 1)
