import sys, time, torch
sys.path.insert(0, ".")
from ml_trainer_b200 import SyntheticDataset, Trainer, build_model, Loader
import ml_trainer_b200.runtime as rt
K=200
warm = SyntheticDataset((3,32,32),10,length=32*24,seed=7,pool=8192)
timed = SyntheticDataset((3,32,32),10,length=32*K,seed=32,pool=min(32*K,8192))
tr = Trainer(build_model("mlmodel"), datasets=(warm, warm), epochs=1, batch_size=32, optimizer="sgd", lr=1e-3, momentum=0.9, metric=None, progress=False, model_dir="/tmp/ep", seed=32)
tr.train_epoch(1)
tr.train_loader = Loader(timed, batch_size=32, shuffle=True, pin_memory=True, num_workers=0)
orig_ei = rt.epoch_indices
def ei(l):
    t=time.perf_counter(); r=orig_ei(l); print("epoch_indices ms", (time.perf_counter()-t)*1e3); return r
rt.epoch_indices = ei
orig_run = rt.NativeLoop.run_lenet
def run(self,*a,**k):
    t=time.perf_counter(); r=orig_run(self,*a,**k); print("run_lenet ms", (time.perf_counter()-t)*1e3, "steps", len(r)); return r
rt.NativeLoop.run_lenet = run
for ep in (2,3):
    torch.cuda.synchronize(); t=time.perf_counter(); tr.train_epoch(ep); torch.cuda.synchronize(); print("train_epoch ms", (time.perf_counter()-t)*1e3)
