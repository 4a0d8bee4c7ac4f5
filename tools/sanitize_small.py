"""Small single-GPU exercise of this round's new kernels, meant to run under `compute-sanitizer --tool memcheck`
(tools/gpu_sanitize.sh): loss kernels, RMSNorm, uint8 augmentation, the fp32 three-term split, masked attention, clamped
embedding ids.  Prints OK when every result is finite."""
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200 import DeviceAugment  # noqa: E402
from ml_trainer_b200.ops import functional as F  # noqa: E402
from ml_trainer_b200.ops.gemm import split3  # noqa: E402

dev = "cuda"
torch.manual_seed(0)
acc = torch.zeros(3, device=dev)
p, t = torch.randn(33, 6, device=dev), torch.randn(33, 6, device=dev)
for kind in ("l1", "l2"):
    loss, dp = F.regression_loss(p, t, kind, acc, True)
lp = torch.log_softmax(torch.randn(17, 10, device=dev), -1)
loss, dl = F.nll_loss(lp, torch.randint(0, 10, (17,), device=dev), acc, True)
x = torch.randn(37, 1032, device=dev).bfloat16().requires_grad_()
g = torch.ones(1032, device=dev, requires_grad=True)
F.rms_norm(x, g, 1e-6).sum().backward()
aug = DeviceAugment((3, 32, 32), padding=4, seed=1)
img = torch.randint(0, 256, (9, 3072), dtype=torch.uint8, device=dev)
out = aug.apply_cuda(img, pos0=5)
s = split3(torch.randn(40, 24, device=dev), False, 0)
q = (torch.randn(2, 192, 128, device=dev) * 0.5).bfloat16().requires_grad_()
kb = F.prepare_key_bias((torch.arange(192, device=dev)[None, :] < torch.tensor([[192], [100]], device=dev)).long())
F.attention(q, q, q, 2, kb).sum().backward()
word = torch.randn(50, 64, device=dev).bfloat16()
emb = F.bert_embeddings(torch.tensor([[1, 49, 70, -3]], device=dev), word, torch.randn(8, 64, device=dev).bfloat16(),
                        torch.randn(2, 64, device=dev).bfloat16())
torch.cuda.synchronize()
ok = all(torch.isfinite(v.float()).all().item() for v in (acc, dp, dl, x.grad, out, s, q.grad, emb))
print("OK" if ok else "NON-FINITE")
