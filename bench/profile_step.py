"""Analysis helper: kernel-level breakdown of one training step (eager, under torch.profiler; numbers are for ranking
kernels, never reported as bench values).  usage: python bench/profile_step.py bert_base|resnet50 [ours|reference]"""
import sys
from collections import defaultdict

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
model_name = sys.argv[1] if len(sys.argv) > 1 else "bert_base"
impl = sys.argv[2] if len(sys.argv) > 2 else "ours"
dev = torch.device("cuda", 0)

if impl == "ours":
    from ml_trainer_b200 import Trainer, build_model
    from ml_trainer_b200.data.loader import SyntheticDataset

    if model_name == "bert_base":
        B = 32
        ds = SyntheticDataset((512,), 2, length=B * 4, seed=1, kind="tokens", pool=256)
        opt, lr = "adamw", 2e-5
    else:
        B = 256
        ds = SyntheticDataset((3, 224, 224), 1000, length=B * 4, seed=1, pool=512)
        opt, lr = "sgd", 0.1
    tr = Trainer(build_model(model_name), datasets=(ds, ds), epochs=1, batch_size=B, optimizer=opt, lr=lr, momentum=0.9,
                 dtype="bf16", graph=False, native_loop=False, progress=False, model_dir="/tmp/mlt_prof", seed=1)
    tr.model.train()
    feed = iter(tr._feed(tr.train_loader))
    x, y = next(feed)
    step = lambda: tr.engine.train_step(x, y)
else:
    from bench.ref_models import build_reference_model

    m = build_reference_model(model_name, "bf16").to(dev).train()
    if model_name == "bert_base":
        x = torch.randint(0, 30000, (32, 512), device=dev)
        y = torch.randint(0, 2, (32,), device=dev)
        o = torch.optim.AdamW(m.parameters(), lr=2e-5)
    else:
        x = torch.randn(256, 3, 224, 224, device=dev)
        y = torch.randint(0, 1000, (256,), device=dev)
        o = torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9)
    crit = torch.nn.CrossEntropyLoss()

    def step():
        o.zero_grad()
        loss = crit(m(x), y)
        loss.backward()
        o.step()

for _ in range(3):
    step()
torch.cuda.synchronize()
N = 3
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(N):
        step()
    torch.cuda.synchronize()
tot = defaultdict(lambda: [0.0, 0])
for e in prof.key_averages():
    tot[e.key][0] += e.device_time_total / N
    tot[e.key][1] += e.count // N
allt = sum(v[0] for v in tot.values())
print(f"{model_name} {impl}: total device time per step {allt / 1e3:.2f} ms, {sum(v[1] for v in tot.values())} launches")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0])[:32]:
    print(f"  {v[0] / 1e3:8.3f} ms {100 * v[0] / allt:5.1f}%  x{v[1]:4d}  {k[:110]}")
