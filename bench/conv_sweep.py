"""Analysis helper: forward+backward time of every distinct ResNet-50 convolution shape at batch 256 (x multiplicity)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ml_trainer_b200.ops import conv as C

B = 256
# (H, Cin, Cout, K, stride, pad, count)
SHAPES = [(56, 64, 64, 1, 1, 0, 1), (56, 64, 64, 3, 1, 1, 3), (56, 64, 256, 1, 1, 0, 4), (56, 256, 64, 1, 1, 0, 2),
          (56, 256, 128, 1, 1, 0, 1), (56, 128, 128, 3, 2, 1, 1), (28, 128, 512, 1, 1, 0, 4), (28, 512, 128, 1, 1, 0, 3),
          (28, 128, 128, 3, 1, 1, 3), (56, 256, 512, 1, 2, 0, 1),
          (28, 512, 256, 1, 1, 0, 1), (28, 256, 256, 3, 2, 1, 1), (14, 256, 1024, 1, 1, 0, 6), (14, 1024, 256, 1, 1, 0, 5),
          (14, 256, 256, 3, 1, 1, 5), (28, 512, 1024, 1, 2, 0, 1),
          (14, 1024, 512, 1, 1, 0, 1), (14, 512, 512, 3, 2, 1, 1), (7, 512, 2048, 1, 1, 0, 3), (7, 2048, 512, 1, 1, 0, 2),
          (7, 512, 512, 3, 1, 1, 2), (14, 1024, 2048, 1, 2, 0, 1)]
tot = 0.0
for (H, Ci, Co, K, s, p, cnt) in SHAPES:
    x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16).requires_grad_()
    w = (torch.randn(Co, Ci, K, K, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    y = C.conv2d_nhwc(x, w, s, p)
    dy = torch.randn_like(y)
    OH = y.shape[1]

    def fwd():
        return C.conv2d_nhwc(x, w, s, p)

    def both():
        torch.autograd.grad(C.conv2d_nhwc(x, w, s, p), (x, w), dy)

    res = []
    for fn in (fwd, both):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            fn()
        b.record()
        torch.cuda.synchronize()
        res.append(a.elapsed_time(b) / 5 * 1e3)
    fl = 2.0 * B * OH * OH * Co * Ci * K * K
    tf, tb = res[0], res[1] - res[0]
    tot += cnt * res[1]
    print(f"{H:3d} {Ci:4d}->{Co:4d} k{K} s{s} x{cnt}: fwd {tf:7.1f} us {fl / tf / 1e6:7.0f} TF/s | bwd {tb:7.1f} us {2 * fl / tb / 1e6:7.0f} TF/s | x{cnt} = {cnt * res[1] / 1e3:6.2f} ms")
print("total conv ms per step", tot / 1e3)
