"""Analysis helper: per-kernel durations of the fused attention kernels at equal CTA count and varying step count."""
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from ml_trainer_b200.ops import functional as F  # noqa: E402

H, d = 12, 64
for B, S in ((128, 128), (64, 256), (32, 512)):
    q, k, v = ((torch.randn(B, S, H * d, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_() for _ in range(3))
    go = torch.randn(B, S, H * d, device="cuda").to(torch.bfloat16)
    for _ in range(3):
        F.attention(q, k, v, H).backward(go)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(10):
            F.attention(q, k, v, H).backward(go)
        torch.cuda.synchronize()
    print(f"B={B} S={S} steps={S // 64}")
    for e in prof.key_averages():
        if "attention" in e.key:
            print(f"   {e.key[:48]:50s} {e.device_time_total / e.count:8.1f} us  x{e.count}")
