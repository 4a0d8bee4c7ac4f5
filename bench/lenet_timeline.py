"""Analysis helper: per-CTA stage timeline of the single-kernel LeNet step (%globaltimer stamps)."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200 import SyntheticDataset, Trainer, build_model  # noqa: E402
from ml_trainer_b200.ops import _lib  # noqa: E402

ds = SyntheticDataset((3, 32, 32), 10, length=32 * 8, seed=1)
tr = Trainer(build_model("mlmodel"), datasets=(ds, ds), epochs=1, batch_size=32, optimizer="sgd", lr=1e-3, progress=False,
             model_dir="/tmp/mlt_tl", metric=None, native_loop=False)
tr.model.train()
feed = iter(tr._feed(tr.train_loader))
x, y = next(feed)
for _ in range(5):
    tr.engine.train_step(x, y)
torch.cuda.synchronize()
nblk = 148
buf = torch.zeros(nblk * 16, dtype=torch.int64, device="cuda")
L = _lib.lib()
L.mlt_lenet_set_debug(ctypes.c_void_p(buf.data_ptr()))
tr.engine.train_step(x, y)
torch.cuda.synchronize()
L.mlt_lenet_set_debug(ctypes.c_void_p(0))
t = buf.view(nblk, 16).cpu().double()
used = t[:, 0] > 0
t = t[used]
t0 = t[:, 0].min()
names = ["start", "x+w loaded", "conv1", "conv2", "fc1", "fc2", "fc3..fc1bwd", "conv2 bwd", "conv1 wgrad(end A)", "grid barrier", "phase B"]
act = t[:, 1] > 0          # CTAs that carried a sample
print("CTAs", int(used.sum()), "with sample", int(act.sum()), "kernel span us", (t[:, 10].max() - t0).item() / 1e3)
for i in range(1, 11):
    if i <= 8:
        d = (t[act, i] - t[act, i - 1]) / 1e3
    else:
        d = (t[:, i] - t[:, i - 1].clamp(min=t0)) / 1e3 if i == 9 else (t[:, i] - t[:, i - 1]) / 1e3
    print(f"  {names[i]:>22s}: mean {d.mean():6.2f} us  max {d.max():6.2f}")
print("  phase A end (max over CTAs) at", ((t[act, 8].max() - t0) / 1e3).item(), "us; barrier release at", ((t[:, 9].min() - t0) / 1e3).item(),
      "..", ((t[:, 9].max() - t0) / 1e3).item())
