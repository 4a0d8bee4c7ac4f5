"""Analysis helper: run one GEMM shape a few times (for ncu).  usage: gemm_one.py M N K [mode]  mode: fwd|dgrad|wgrad"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ml_trainer_b200.ops.gemm import gemm

M, N, K = (int(v) for v in sys.argv[1:4])
mode = sys.argv[4] if len(sys.argv) > 4 else "fwd"
x = torch.randn(M, K, device="cuda").bfloat16()
w = torch.randn(N, K, device="cuda").bfloat16()
dy = torch.randn(M, N, device="cuda").bfloat16()
fn = {"fwd": lambda: gemm(x, w), "dgrad": lambda: gemm(dy, w, False, True), "wgrad": lambda: gemm(dy, x, True, True)}[mode]
for _ in range(6):
    fn()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(10):
    fn()
e.record()
torch.cuda.synchronize()
print("us", s.elapsed_time(e) * 100)
