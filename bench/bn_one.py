"""One BatchNorm forward + backward on an [M, C] bf16 tensor (for ncu captures).  Usage: python bench/bn_one.py M C [relu] [res]"""
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200.ops import conv as C  # noqa: E402

M, ch = int(sys.argv[1]), int(sys.argv[2])
relu, res = "relu" in sys.argv[3:], "res" in sys.argv[3:]
bn = torch.nn.BatchNorm2d(ch).cuda().to(torch.bfloat16).train()
x = torch.randn(M, ch, device="cuda").to(torch.bfloat16).view(M, 1, 1, ch).requires_grad_()
r = torch.randn(M, ch, device="cuda").to(torch.bfloat16).view(M, 1, 1, ch).requires_grad_() if res else None
go = torch.randn(M, ch, device="cuda").to(torch.bfloat16).view(M, 1, 1, ch)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.zero_()
    y = C.batch_norm_nhwc(x, bn, relu=relu, residual=r)
    y.backward(go)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
