"""Analysis helper: per-CTA timeline of the dQ kernel (globaltimer stamps)."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200.ops import _lib, functional as F  # noqa: E402

H, d = 12, 64
B, S = int(sys.argv[1]), int(sys.argv[2])
q, k, v = ((torch.randn(B, S, H * d, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_() for _ in range(3))
go = torch.randn(B, S, H * d, device="cuda").to(torch.bfloat16)
for _ in range(3):
    F.attention(q, k, v, H).backward(go)
n = B * H * (S // 128)
buf = torch.zeros(n * 8, dtype=torch.int64, device="cuda")
L = _lib.lib()
L.mlt_attention_set_debug(ctypes.c_void_p(buf.data_ptr()))
F.attention(q, k, v, H).backward(go)
torch.cuda.synchronize()
L.mlt_attention_set_debug(ctypes.c_void_p(0))
t = buf.view(n, 8).cpu().double()
t0 = t[:, 0].min()
names = ["start", "setup_done", "q_landed", "scores0_ready", "loop_end", "mma_done", "stores_issued", "exit"]
t = t[:, [0, 1, 2, 4, 5, 6, 3, 7]]
print("kernel span us", (t[:, 7].max() - t0).item() / 1e3)
dt = (t[:, 1:] - t[:, :-1]) / 1e3
for i in range(7):
    print(f"  {names[i]:>16s} -> {names[i + 1]:<16s} mean {dt[:, i].mean():7.2f} us  p10 {dt[:, i].quantile(0.1):7.2f}  p90 {dt[:, i].quantile(0.9):7.2f}")
life = (t[:, 7] - t[:, 0]) / 1e3
print("  CTA lifetime mean", life.mean().item(), "p10", life.quantile(0.1).item(), "p90", life.quantile(0.9).item())
order = t[:, 0].argsort()
print("  start times of CTAs (us) at ranks 0,295,296,600,900,1200,1535:", [round((t[order[i], 0] - t0).item() / 1e3, 2) for i in (0, 295, 296, 600, 900, 1200, min(n - 1, 1535))])
