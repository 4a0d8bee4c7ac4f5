"""Analysis helper: one implicit-GEMM convolution (forward + backward) a few times (for ncu / timing).
usage: conv_one.py B H Cin Cout K stride pad"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ml_trainer_b200.ops import conv as C

B, H, Ci, Co, K, stride, pad = (int(v) for v in sys.argv[1:8])
x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16).requires_grad_()
w = (torch.randn(Co, Ci, K, K, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
y = C.conv2d_nhwc(x, w, stride, pad)
dy = torch.randn_like(y)


def step():
    yy = C.conv2d_nhwc(x, w, stride, pad)
    torch.autograd.grad(yy, (x, w), dy)


for _ in range(3):
    step()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(5):
    step()
e.record()
torch.cuda.synchronize()
OH = y.shape[1]
fl = 3 * 2.0 * B * OH * OH * Co * Ci * K * K
print("fwd+bwd us", s.elapsed_time(e) * 200, "TFLOP/s", fl / (s.elapsed_time(e) / 5 * 1e-3) / 1e12)
