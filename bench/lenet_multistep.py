"""Analysis helper: time K LeNet steps as K launches vs ONE multi-step launch over a pool larger than L2."""
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200 import SyntheticDataset, Trainer, build_model  # noqa: E402

ds = SyntheticDataset((3, 32, 32), 10, length=64, seed=1)
tr = Trainer(build_model("mlmodel"), datasets=(ds, ds), epochs=1, batch_size=32, optimizer="sgd", lr=1e-3, progress=False,
             model_dir="/tmp/mlt_ms", metric=None, native_loop=False)
f = tr.engine.fused
P, B, K = 512, 32, 2000                      # 512 x 393 KB = 201 MB of inputs: every step reads a batch that is not in L2
xp = torch.randn(P, B, 3, 32, 32, device="cuda")
yp = torch.randint(0, 10, (P, B), device="cuda")
losses = torch.zeros(K, device="cuda")
tr.optimizer.begin_step(1.0)
for _ in range(2):
    f.run_pool(xp, yp, 50, True, losses=losses)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); f.run_pool(xp, yp, K, True, losses=losses); b.record()
torch.cuda.synchronize()
t1 = a.elapsed_time(b) * 1e3 / K
a.record()
for i in range(K):
    f(xp[i % P], yp[i % P], True, None, False, losses[i:i + 1])
b.record()
torch.cuda.synchronize()
t2 = a.elapsed_time(b) * 1e3 / K
print(f"multi-step launch: {t1:.2f} us/step ({B / t1 * 1e6:,.0f} samples/s);  one launch per step: {t2:.2f} us/step ({B / t2 * 1e6:,.0f} samples/s)")
