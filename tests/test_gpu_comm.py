"""Fused all-reduce + update kernels, with the ranks emulated on ONE GPU (gridDim.y = rank, cooperative launch).

Kernels that wait on each other must not run as separate launches on one GPU, so the multi-rank protocol is
exercised as a single cooperative kernel over every rank's arena; real multi-GPU runs are in test_multigpu.py.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

C_INT, C_LL, C_VP = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p
DEV = "cuda:0"


def _align(x, a):
    return (x + a - 1) // a * a


class FakeWorld:
    """`world` arenas on one device with the production layout (flags | scalars | recv | heap)."""

    def __init__(self, world, n, gdtype, slot_bytes=1 << 20):
        from ml_trainer_b200.ops import _lib
        from ml_trainer_b200.parallel.p2p import P2PCtx

        self.L = _lib.lib()
        self.L.mlt_p2p_flag_block_bytes.restype = C_LL
        self.world, self.n, self.gdtype = world, n, gdtype
        flags = _align(int(self.L.mlt_p2p_flag_block_bytes()), 4096)
        self.scalar_off = flags
        self.recv_off = flags + 4096
        gsz = 2 if gdtype == torch.bfloat16 else 4
        self.grad_off = _align(self.recv_off + 2 * world * slot_bytes, 4096)
        self.model_off = _align(self.grad_off + n * gsz, 4096)
        total = self.model_off + n * 4 + 4096
        self.arenas = [torch.zeros(total, dtype=torch.uint8, device=DEV) for _ in range(world)]
        self.ctx = P2PCtx()
        for r in range(8):
            self.ctx.base[r] = self.arenas[r].data_ptr() if r < world else 0
        self.ctx.mc_base = None
        self.ctx.world, self.ctx.rank = world, 0
        self.ctx.flags_off, self.ctx.recv_off, self.ctx.recv_slot_bytes = 0, self.recv_off, slot_bytes

    def grad(self, r):
        gsz = 2 if self.gdtype == torch.bfloat16 else 4
        return self.arenas[r][self.grad_off:self.grad_off + self.n * gsz].view(self.gdtype)

    def model(self, r, dtype):
        sz = 2 if dtype == torch.bfloat16 else 4
        return self.arenas[r][self.model_off:self.model_off + self.n * sz].view(dtype)


def _ptrs(ts):
    return (C_VP * len(ts))(*[t.data_ptr() if t is not None else 0 for t in ts])


@pytest.mark.parametrize("algo", [0, 1])
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("kind", ["sgd", "adamw"])
def test_emulated_allreduce_update(algo, world, gdtype, kind):
    from ml_trainer_b200.optim import rules
    from ml_trainer_b200.optim.rules import Hyper, RULE_IDS

    torch.manual_seed(0)
    n = 128 * 37                                   # not a multiple of world * vector width on purpose
    fw = FakeWorld(world, n, gdtype)
    model_is_master = gdtype == torch.float32
    p0 = torch.randn(n)
    masters, s1s, s2s, hypers = [], [], [], []
    for r in range(world):
        m = fw.model(r, torch.float32) if model_is_master else torch.empty(n, device=DEV)
        m.copy_(p0)
        masters.append(m)
        s1s.append(torch.zeros(n, device=DEV))
        s2s.append(torch.zeros(n, device=DEV))
        hypers.append(torch.zeros(8, device=DEV))
    ref_p, ref_s1, ref_s2 = p0.clone(), torch.zeros(n), torch.zeros(n)
    for step in range(1, 4):                       # several steps: tickets and receive-slot parity must keep working
        grads = [torch.randn(n).to(gdtype) for _ in range(world)]
        for r in range(world):
            fw.grad(r).copy_(grads[r])
        h = Hyper(lr=0.05, wd=0.01, b1=0.9, b2=0.999, eps=1e-8 if kind != "sgd" else 0.0, step=step, grad_scale=1.0 / world)
        for r in range(world):
            hypers[r].copy_(torch.tensor(h.pack()))
        rc = fw.L.mlt_p2p_allreduce_update(ctypes.byref(fw.ctx), C_INT(algo), C_INT(RULE_IDS[kind]),
                                           C_INT(int(gdtype == torch.bfloat16)), C_LL(fw.grad_off), C_LL(fw.model_off),
                                           C_INT(int(model_is_master)), _ptrs(masters), _ptrs(s1s), _ptrs(s2s), _ptrs(hypers),
                                           C_LL(n), C_INT(3), C_INT(world), C_VP(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, fw.L.mlt_last_error()
        torch.cuda.synchronize()
        gsum = sum(g.float() for g in grads)
        rules.apply_rule(kind, ref_p, gsum, ref_s1, ref_s2, h)
        for r in range(world):
            w = fw.model(r, torch.float32 if model_is_master else torch.bfloat16).float().cpu()
            tol = 1e-5 if model_is_master else 1e-2
            assert torch.allclose(w, ref_p, rtol=tol, atol=tol), (step, r, (w - ref_p).abs().max())
        if algo == 0:                              # one-shot: every rank holds the full fp32 master
            for r in range(world):
                assert torch.allclose(masters[r].cpu(), ref_p, rtol=1e-5, atol=1e-6)
        # every replica bit-identical
        w0 = fw.model(0, torch.float32 if model_is_master else torch.bfloat16)
        for r in range(1, world):
            assert torch.equal(fw.model(r, w0.dtype), w0)
    out = (ctypes.c_uint * 6)()
    assert fw.L.mlt_p2p_check_error(ctypes.byref(fw.ctx), out) == 0


@pytest.mark.parametrize("world", [2, 8])
def test_emulated_broadcast_and_scalar_reduce(world):
    fw = FakeWorld(world, 128 * 8, torch.float32)
    for r in range(world):
        fw.grad(r).fill_(float(r + 1))
    st = C_VP(torch.cuda.current_stream().cuda_stream)
    root = world - 1
    assert fw.L.mlt_p2p_broadcast(ctypes.byref(fw.ctx), C_LL(fw.grad_off), C_LL(128 * 8 * 4), C_INT(root), C_INT(world), st) == 0
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.all(fw.grad(r) == float(root + 1))
    vals = [torch.tensor([float(r), 1.0, 10.0 * r], device=DEV) for r in range(world)]
    table = torch.tensor([v.data_ptr() for v in vals], dtype=torch.int64, device=DEV)
    for rep in range(3):
        for r in range(world):
            vals[r].copy_(torch.tensor([float(r), 1.0, 10.0 * r]))
        assert fw.L.mlt_p2p_reduce_scalars(ctypes.byref(fw.ctx), C_LL(fw.scalar_off), C_VP(table.data_ptr()), C_INT(3),
                                           C_INT(world), st) == 0
        torch.cuda.synchronize()
        expect = torch.tensor([sum(range(world)), float(world), 10.0 * sum(range(world))])
        for r in range(world):
            assert torch.allclose(vals[r].cpu(), expect), (rep, r, vals[r])


@pytest.mark.parametrize("algo", [0, 1])
@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
def test_emulated_mean_only_allreduce(algo, gdtype):
    """Rule 5 ("mean only"): the fused kernels act as a plain in-place mean all-reduce over the gradient buffers."""
    world, n = 4, 128 * 9
    fw = FakeWorld(world, n, gdtype)
    grads = [torch.randn(n).to(gdtype) for _ in range(world)]
    for r in range(world):
        fw.grad(r).copy_(grads[r])
    tmp = [torch.empty(n, device=DEV) for _ in range(2 * world)]
    hyp = [torch.tensor([0, 0, 0, 0, 0, 1, 1, 1.0 / world], device=DEV) for _ in range(world)]
    rc = fw.L.mlt_p2p_allreduce_update(ctypes.byref(fw.ctx), C_INT(algo), C_INT(5), C_INT(int(gdtype == torch.bfloat16)),
                                       C_LL(fw.grad_off), C_LL(fw.grad_off), C_INT(0), _ptrs(tmp[:world]), _ptrs(tmp[world:]),
                                       _ptrs([None] * world), _ptrs(hyp), C_LL(n), C_INT(2), C_INT(world),
                                       C_VP(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, fw.L.mlt_last_error()
    torch.cuda.synchronize()
    ref = sum(g.float() for g in grads) / world
    for r in range(world):
        got = fw.grad(r).float().cpu()
        assert torch.allclose(got, ref, rtol=1e-2 if gdtype == torch.bfloat16 else 1e-6, atol=1e-2 if gdtype == torch.bfloat16 else 1e-6)
