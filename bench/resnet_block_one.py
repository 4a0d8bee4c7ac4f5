"""Analysis helper: one 3x3 convolution + BatchNorm(+ReLU) forward/backward at a ResNet-50 layer-2 shape (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ml_trainer_b200.ops import conv as C

B, H, Ci, Co = 256, 28, 128, 128
x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16).requires_grad_()
w = (torch.randn(Co, Ci, 3, 3, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
bn = torch.nn.BatchNorm2d(Co).to("cuda").to(torch.bfloat16).train()
for _ in range(4):
    y = C.batch_norm_nhwc(C.conv2d_nhwc(x, w, 1, 1), bn, relu=True)
    torch.autograd.grad(y, (x, w, bn.weight, bn.bias), torch.ones_like(y))
torch.cuda.synchronize()
print("ok")
