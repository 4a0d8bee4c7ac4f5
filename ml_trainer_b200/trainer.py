"""The training runtime (reference src/trainer.py, class Trainer).

Same constructor signature, keyword whitelist, public methods (``fit``, ``test``,
``save_model``, ``save_history_``, ``clear``) and attributes (``train_loader``,
``val_loader``, ``history``, ``train_losses`` ...) as the reference, plus the
``evaluate`` / ``save`` / ``load`` aliases.  What changed underneath:

* no ``DistributedDataParallel``: a :class:`~ml_trainer_b200.parallel.reducer.Reducer`
  launches one fused all-reduce + optimizer kernel per gradient bucket on a side stream;
* no ``torch.optim``: a flat-arena :class:`~ml_trainer_b200.optim.fused.FusedOptimizer`
  (fp32 master weights, optional bf16 compute copy) that the stock LR schedulers drive;
* loss and metric accumulate on the device and are read back once per epoch (the
  reference calls ``loss.item()`` and two ``.cpu()`` per step, src/trainer.py:186,166);
* inputs arrive through a pinned, double-buffered copy-stream prefetcher;
* the whole step (forward, loss, backward, reduce, update) can be captured in a CUDA graph;
* checkpoints are snapshotted asynchronously from the master weights without moving
  the live model, in the reference's exact file format.

Deliberate behavioural fixes, each documented in DESIGN.md: ``neg-loss``/``l2``/``custom``
criteria work, ``ReduceLROnPlateau`` is stepped with the validation loss, the sampler's
``set_epoch`` is called, validation is sharded and reduced across ranks, ``history.pkl``
is written by rank 0 only, ``test()`` runs under ``eval()``/``no_grad``.
"""
from __future__ import annotations

import gc
import os
import pickle
from collections import OrderedDict
from typing import Optional

import torch
import torch.distributed as dist
from torch.utils.data import distributed as tud

from .config import CRITERIA, OPTIMIZERS, SCHEDULERS, TrainerConfig, validate_kwargs
from .data.loader import Loader, PinnedPrefetcher
from .engine import StepEngine
from .optim.fused import FusedOptimizer
from .parallel.comm import Comm, LocalComm, make_comm
from .parallel.reducer import Reducer
from .utils import metrics as M
from .utils.checkpoint import SnapshotWriter, atomic_torch_save, save_state_dict
from .utils.functions import custom_loss_function
from .utils.logging import get_logger
from .utils.utils import strip_module_prefix

logger = get_logger("ml_trainer_b200")

try:  # tqdm is a reference dependency but keep the trainer importable without it
    from tqdm import tqdm
except Exception:  # pragma: no cover
    def tqdm(it, **kw):
        class _Bar:
            def __init__(self, it): self.it = it
            def __enter__(self): return self
            def __exit__(self, *a): return False
            def __iter__(self): return iter(self.it)
            def set_postfix(self, **kw): pass
        return _Bar(it)


class Trainer:
    allowed_kwargs = TrainerConfig.allowed()

    def __init__(self, model, datasets=None, epochs=None, batch_size=None,
                 is_parallel=False, save_history=False, **config):
        logger.info("Config inputs.", config=config)
        self.validate_kwargs(config, self.allowed_kwargs)
        cfg = self.cfg = TrainerConfig(**config)
        # Unpack kwargs (attribute names kept from the reference)
        self.epochs = epochs
        self.scheduler_type = cfg.scheduler
        self.optimizer_type = cfg.optimizer
        self.momentum = cfg.momentum
        self.weight_decay = cfg.weight_decay
        self.lr = cfg.lr
        self.criterion_type = cfg.criterion
        self.metric = cfg.metric
        self.pred_function_type = cfg.pred_function
        self.model_dir = cfg.model_dir
        self.is_parallel = is_parallel
        self.save_history = save_history
        self.train_losses, self.val_losses = [], []
        self.train_metrics, self.val_metrics = [], []
        self.history = {}
        self.start_epoch = 1
        torch.manual_seed(cfg.seed)

        # ---- device first, then the process group (the reference sets the device after wrapping)
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0")) if is_parallel else 0
        if torch.cuda.is_available():
            if is_parallel:
                torch.cuda.set_device(self.local_rank % torch.cuda.device_count())
            self.device = torch.device("cuda", torch.cuda.current_device())
        else:
            self.device = torch.device("cpu")
        logger.info(f"Training on device: {self.device}.")
        self.comm: Comm = make_comm(cfg.backend, self.device, is_parallel, algo=cfg.comm_algo) if is_parallel else LocalComm()
        self.rank, self.world_size = self.comm.rank, self.comm.world_size

        # ---- data
        self.train_loader = self.val_loader = None
        self.train_sampler = self.val_sampler = None
        self.global_batch_size = batch_size
        if datasets:
            train_set, val_set = datasets
            if batch_size is None:
                raise ValueError("batch_size is required when datasets are given")
            if is_parallel and self.world_size > 1:
                self.train_sampler = tud.DistributedSampler(train_set, num_replicas=self.world_size, rank=self.rank)
                if cfg.shard_val:
                    self.val_sampler = tud.DistributedSampler(val_set, num_replicas=self.world_size,
                                                              rank=self.rank, shuffle=False)
                # global-batch semantics, as the reference: scale batch size by world size
                batch_size = max(batch_size // self.world_size, 1)
            self.batch_size = batch_size
            pin = cfg.pin_memory if cfg.pin_memory is not None else self.device.type == "cuda"
            lkw = dict(num_workers=cfg.num_workers, pin_memory=pin, drop_last=cfg.drop_last)
            logger.info("Preparing the data.")
            self.train_loader = Loader(train_set, batch_size=batch_size, shuffle=self.train_sampler is None,
                                       sampler=self.train_sampler, **lkw)
            self.val_loader = Loader(val_set, batch_size=batch_size, shuffle=self.val_sampler is None,
                                     sampler=self.val_sampler, num_workers=cfg.num_workers, pin_memory=pin)
            for name, ld in (("train", self.train_loader), ("validation", self.val_loader)):
                logger.debug("Processes {}/{} ({:.0f}%) of {} data".format(
                    len(ld.sampler), len(ld.dataset), 100.0 * len(ld.sampler) / max(len(ld.dataset), 1), name))
        else:
            self.batch_size = batch_size
            logger.warning("Testing only available. No datasets in arguments.")

        # ---- model, criterion, optimizer, reducer
        logger.info("Loading the model.")
        self.model = model.to(self.device)
        if cfg.convert and self.device.type == "cuda":
            from .ops.convert import convert_model

            self.model = convert_model(self.model)
        self.compute_dtype = {"fp32": torch.float32, "bf16": torch.bfloat16}[cfg.dtype]
        self.criterion = self._get_criterion()
        self.pred_function = self._get_prediction_function()
        padded = sum((p.numel() + 127) // 128 * 128 for p in self.model.parameters() if p.requires_grad)
        from .optim.rules import RULE_STATES

        self.comm.prepare(padded, self.compute_dtype, n_states=RULE_STATES.get(self.optimizer_type, 2))
        self.optimizer = self._get_optimizer()
        if self.world_size > 1:
            self._sync_initial_state()
        self.reducer = Reducer(self.optimizer, self.comm, bucket_mb=cfg.bucket_mb)
        self.scheduler = self._get_scheduler()
        self.phase_timer = None
        if self.cfg.trace:
            from .utils import tracing

            self.phase_timer = tracing.PhaseTimer()
            tracing.enable(True, self.phase_timer)
        self.engine = StepEngine(self.model, self.criterion, self.criterion_type, self.metric, self.reducer,
                                 self.device, graph=cfg.graph, compute_dtype=self.compute_dtype,
                                 fused_step=cfg.fused_step)
        self._snap = SnapshotWriter(self.device)
        if cfg.resume:
            self.load(cfg.resume)

    # ----------------------------------------------------------------- factories
    def _get_prediction_function(self):
        return M.get_prediction_function(self.pred_function_type)

    def _get_optimizer(self):
        if self.optimizer_type not in OPTIMIZERS:
            raise ValueError(f"unknown optimizer {self.optimizer_type!r}; expected one of {OPTIMIZERS}")
        return FusedOptimizer(self.model.parameters(), kind=self.optimizer_type, lr=self.lr,
                              momentum=self.momentum, weight_decay=self.weight_decay,
                              betas=self.cfg.betas, eps=self.cfg.eps, compute_dtype=self.compute_dtype,
                              allocator=self.comm.alloc)

    def _get_criterion(self):
        kind = self.criterion_type
        from .ops import modules as M_

        table = {"cross_entropy": M_.CrossEntropyLoss, "neg-loss": M_.NLLLoss, "l1": M_.L1Loss, "l2": M_.MSELoss,
                 "custom": M_.CustomLoss}      # every criterion: loss + gradient + metric from one kernel (csrc/loss.cu)
        if kind in table:
            return table[kind]()
        raise ValueError(f"unknown criterion {kind!r}; expected one of {CRITERIA}")

    def _get_scheduler(self):
        """Only the selected scheduler is built (the reference builds all three eagerly)."""
        kind = self.scheduler_type
        if not kind:
            return None
        sched = torch.optim.lr_scheduler
        table = {
            "CosineAnnealingWarmRestarts": lambda: sched.CosineAnnealingWarmRestarts(self.optimizer, T_0=5, eta_min=1e-7),
            "ReduceLROnPlateau": lambda: sched.ReduceLROnPlateau(self.optimizer, "min", min_lr=1e-7),
            "StepLR": lambda: sched.StepLR(self.optimizer, step_size=2),
        }
        return table[kind]()        # KeyError on an unknown name, as the reference

    @property
    def scheduler_options(self):
        return SCHEDULERS

    # -------------------------------------------------------------- distributed
    def _sync_initial_state(self):
        """Rank 0's parameters and buffers become everyone's (DDP ctor broadcast)."""
        self.comm.broadcast(self.optimizer.master, src=0)
        if self.optimizer.model is not self.optimizer.master:
            self.optimizer.model.copy_(self.optimizer.master)
        for b in self.model.buffers():
            self.comm.broadcast(b, src=0)

    def _sync_buffers(self):
        """Rank 0's module buffers (BatchNorm running statistics) become everyone's.

        DDP broadcasts them before EVERY forward.  Training-mode BatchNorm never reads them, so doing it
        once before each evaluation gives the same numbers for one collective per epoch instead of one per step."""
        if self.world_size > 1:
            for b in self.model.buffers():
                if b.is_floating_point():
                    self.comm.broadcast(b, src=0)

    def _average_gradients(self):
        """Reference API (src/trainer.py:152-158): SUM all-reduce then divide by world size."""
        self.optimizer.collect_stray_grads()
        self.comm.all_reduce_mean(self.optimizer.grad)

    # ------------------------------------------------------------------ metrics
    def _evaluate(self, outputs, targets):
        return M.evaluate(self.metric, outputs, targets)

    def _get_predictions(self, outputs):
        return M.get_predictions(outputs, self.pred_function)

    # -------------------------------------------------------------------- loops
    def _feed(self, loader):
        aug = getattr(loader, "_augment", None)
        if aug is None or self.device.type != "cuda":
            return PinnedPrefetcher(loader, self.device, slots=max(self.cfg.prefetch, 1))   # dtype casts happen in the step
        # uint8 image batches: upload as uint8, then crop / flip / normalise on the device (csrc/augment.cu)
        loader._augment_on_device = True
        feed = PinnedPrefetcher(loader, self.device, slots=max(self.cfg.prefetch, 1))
        train = loader._augment_train

        class _Augmented:
            h2d_bytes = property(lambda s_: feed.h2d_bytes)

            def __len__(s_):
                return len(feed)

            def __iter__(s_):
                # a small ring of output buffers: stable addresses (captured CUDA graphs are keyed by them), and the
                # previous batch may still be in use by the step that was launched with it
                ring, k = [None] * (feed.slots + 1), 0
                for x, y in feed:
                    buf = ring[k]
                    if buf is None or buf.shape[0] != x.shape[0]:
                        buf = torch.empty((x.shape[0], *aug.out_shape), dtype=torch.float32, device=x.device)
                        if x.shape[0] == loader.batch_size:
                            ring[k] = buf
                    yield aug.apply_cuda(x, out=buf, train=train), y
                    k = (k + 1) % len(ring)

        return _Augmented()

    def _bar(self, it, total):
        disable = (not self.cfg.progress) or self.rank != 0
        return tqdm(it, unit="batch", total=total, disable=disable)

    def _reduce_epoch(self, sums):
        """``sums`` = [loss_sum, metric_sum, batches] (device tensor, or host floats) -> cross-rank mean loss / metric."""
        if isinstance(sums, torch.Tensor) and sums.device.type == "cpu":
            sums = sums.tolist()
        if isinstance(sums, (list, tuple)):
            # host values (the native epoch loop hands back per-step losses in pinned memory): sum over the ranks and read the
            # peers-stopped-answering word in ONE round trip to the device
            loss_sum, metric_sum, n = self.comm.reduce_scalars_host(sums) if self.world_size > 1 else sums
            if self.world_size == 1:
                self.comm.check_errors()
        else:
            if self.world_size > 1:
                sums = self.comm.reduce_scalars(sums)
            loss_sum, metric_sum, n = sums.tolist()
            self.comm.check_errors()            # a peer that stopped answering: raise here, not a checkpoint later
        n = max(n, 1.0)
        return loss_sum / n, metric_sum / n

    # ----------------------------------------------------------- native epochs
    def _native_ok(self, loader, model) -> bool:
        """Whole-epoch C++ execution: single-kernel step + in-memory tensor dataset of the model's input shape."""
        f = self.engine.fused
        if f is None or model is not self.model or not self.cfg.native_loop or not getattr(loader, "_fast", False):
            return False
        x, y = loader._backing[0], loader._backing[1]
        if not (y.dtype == torch.int64 and y.dim() == 1 and y.is_contiguous() and x.is_contiguous()):
            return False
        aug = getattr(loader, "_augment", None)
        if aug is not None:                     # uint8 rows, augmented on the device into the step's [B, 3, 32, 32] input
            return x.dtype == torch.uint8 and tuple(aug.out_shape) == (3, 32, 32)
        return x.dtype == torch.float32 and x.dim() == 4 and tuple(x.shape[1:]) == (3, 32, 32)

    def _native_graph_ok(self, loader, model) -> bool:
        """Whole-epoch C++ execution of CUDA-graph steps: any model, in-memory tensor dataset, full batches."""
        eng = self.engine
        return (eng.fused is None and eng.use_graph and model is self.model and self.cfg.native_loop
                and getattr(loader, "_fast", False) and self.device.type == "cuda" and getattr(loader, "_augment", None) is None
                and all(t.is_contiguous() for t in loader._backing[:2]) and len(loader._backing) == 2)

    def _train_epoch_native_graph(self, epoch):
        """Python drives the first steps (eager warm-up, then one graph capture per ring slot); the C++ executor
        replays the captured graphs for the rest of the epoch; a trailing partial batch runs eagerly."""
        from .runtime import epoch_indices

        eng, opt, loader = self.engine, self.optimizer, self.train_loader
        loader.draw_base_seed()                 # what iter(DataLoader) would draw before the sampler (same random stream)
        idx, sizes = epoch_indices(loader)
        B = loader.batch_size
        nb = len(sizes)
        full = sum(1 for s in sizes if s == B)
        loop = self._native_loop_for(loader)
        x_be, y_be = loader._backing
        if not hasattr(self, "_slot_views") or self._slot_views_key != id(loop):
            self._slot_views = [loop.slot_tensors(s, (B, *x_be.shape[1:]), x_be.dtype, (B, *y_be.shape[1:]), y_be.dtype)
                                for s in range(loop.slots)]
            self._slot_views_key = id(loop)
        # large samples (images): page-lock the dataset once, samples then go device-ward by DMA straight from it
        loop.set_direct_rows(self._page_lock(x_be))
        pin_losses = torch.zeros(nb, dtype=torch.float32).pin_memory()
        step = 0

        def python_step(i):
            sel = idx[i * B:i * B + sizes[i]]
            if sizes[i] == B:
                xs, ys = self._slot_views[i % loop.slots]
                xs.copy_(x_be[sel]); ys.copy_(y_be[sel])
            else:
                xs, ys = x_be[sel].to(self.device), y_be[sel].to(self.device)
            eng.train_step(xs, ys)
            pin_losses[i].copy_(eng.last_loss, non_blocking=True)
            if self.scheduler_type == "CosineAnnealingWarmRestarts":
                self.scheduler.step(epoch - 1 + i / nb)

        while step < full and any(eng.graph_exec_handle(*self._slot_views[s]) is None for s in range(loop.slots)):
            python_step(step)
            step += 1
        if step < full:
            # ring alignment: the executor's step j uses slot j % S, so start it on a multiple of S
            while step % loop.slots != 0 and step < full:
                python_step(step)
                step += 1
        if step < full:
            rows = []
            for i in range(step, full):
                rows.append(opt.hyper_for_next_step(self.reducer.grad_scale).pack())
                opt.step_count += 1
                opt._opt_called = True
                if self.scheduler_type == "CosineAnnealingWarmRestarts":
                    self.scheduler.step(epoch - 1 + i / nb)
            execs = [eng.graph_exec_handle(*self._slot_views[s]) for s in range(loop.slots)]
            torch.cuda.current_stream().synchronize()
            losses = loop.run_graphs(execs, opt.hyper, x_be, y_be, idx[step * B:full * B], B, torch.tensor(rows, dtype=torch.float32),
                                     eng.loss_static, eng.launches_per_step)
            pin_losses[step:full] = losses
            eng.graph_replays += full - step
            self.last_native_stats = dict(loop.last_stats)
            step = full
        for i in range(step, nb):
            python_step(i)
        self.last_step_losses = pin_losses
        return nb

    def _page_lock(self, x: torch.Tensor) -> bool:
        """Page-lock an in-memory dataset tensor whose rows are large enough for per-sample DMA (>= 32 KB) so the executor can
        copy samples device-ward without a host gather.  Registrations are remembered per tensor and undone in ``close()``:
        memory that is freed while still registered poisons later CUDA calls."""
        row_bytes = x[0].numel() * x.element_size()
        nbytes = x.numel() * x.element_size()
        if row_bytes < 32768 or nbytes > (8 << 30) or self.device.type != "cuda":
            return False
        if x.is_pinned():
            return True
        reg = self.__dict__.setdefault("_registered", {})
        key = x.data_ptr()
        if key in reg:
            return True
        try:
            rc = torch.cuda.cudart().cudaHostRegister(key, nbytes, 0)
            if int(rc) != 0:
                raise RuntimeError(f"cudaHostRegister -> {rc}")
        except Exception:                              # registration refused (ulimit -l): keep the host gather
            return False
        reg[key] = x                                   # keeps the tensor alive for as long as it is registered
        return True

    def _page_unlock_all(self):
        reg = self.__dict__.get("_registered") or {}
        if reg and torch.cuda.is_available():
            torch.cuda.synchronize(self.device)
            for ptr_ in list(reg):
                try:
                    torch.cuda.cudart().cudaHostUnregister(ptr_)
                except Exception:
                    pass
        reg.clear()

    def _native_loop_for(self, loader, deep: bool = False):
        from .runtime import NativeLoop

        x, y = loader._backing[0], loader._backing[1]
        key = (loader.batch_size, x[0].numel() * x.element_size(), y.element_size(), deep)
        loops = self.__dict__.setdefault("_nloops", {})
        if key in loops:
            self._nloop, self._nloop_key = loops[key], key
        if getattr(self, "_nloop_key", None) != key:
            # Ring depth.  A deep ring lets the host enqueue a dozen single-kernel steps ahead of the GPU.  Measured (K = 20, end to
            # end): one GPU 44.7 us/step with 12 slots against 47.7 with 3 (profiles/r2/bench_ours_n1_ring{12,3}.json); at eight
            # GPUs it buys nothing (83.9 against 82.6 us/step, bench_ours_n8_with_step_stamps.json / _headline.json; with the
            # old spinning waits it was slower, 118 against 100) - so deep only without peers.
            # Graph-replay steps keep the shallow ring (one captured graph per slot; ResNet batches are 38 MB each).
            slots = max(self.cfg.prefetch + 1, 3)
            if deep and key[0] * (key[1] + key[2]) <= (2 << 20):
                slots = max(slots, int(os.environ.get("MLT_RING_SLOTS", "12" if self.world_size == 1 else "3")))
                if os.environ.get("MLT_LENET_PERSISTENT", "auto") == "1":
                    slots = min(slots, 8)          # the persistent step kernel takes at most 8 ring slots
            self._nloop = loops[key] = NativeLoop(self.device, key[0], key[1], key[2], slots=slots)
            self._nloop_key = key
        return self._nloop

    def _run_native_epoch(self, loader, epoch, train: bool, use_acc: bool = True):
        from .runtime import epoch_indices

        eng, opt = self.engine, self.optimizer
        loader.draw_base_seed()                 # what iter(DataLoader) would draw before the sampler (same random stream)
        idx, sizes = epoch_indices(loader)
        nb = len(sizes)
        hyper = None
        if train:
            # the LR schedule is evaluated up front, one row of hyper-parameters per step
            lrs = None
            if self.scheduler_type == "CosineAnnealingWarmRestarts":        # stepped per batch (as per pytorch docs)
                lrs = []
                for i in range(nb):
                    lrs.append(float(opt.param_groups[0]["lr"]))
                    opt._opt_called = True
                    self.scheduler.step(epoch - 1 + i / nb)
            hyper = opt.hyper_rows(nb, self.reducer.grad_scale, lrs)
            opt.step_count += nb
            opt._opt_called = True
        loop = self._native_loop_for(loader, deep=True)
        x, y = loader._backing[0], loader._backing[1]
        B = loader.batch_size
        aug = getattr(loader, "_augment", None)
        if aug is not None:
            if getattr(self, "_aug_out", None) is None or self._aug_out.shape[0] < B:
                self._aug_out = torch.empty((B, *aug.out_shape), dtype=torch.float32, device=self.device)
            is_train = train and loader._augment_train
            loop.set_augment(aug, self._aug_out, aug.position, is_train)
            if is_train:
                aug.position += int(idx.numel())
        else:
            loop.set_augment(None)
        uniform = sizes[-1] == B and (len(sizes) < 2 or (min(sizes) == B and max(sizes) == B))
        losses = loop.run_lenet(eng.fused, x, y, idx, B, None if uniform else sizes, hyper, train, eng.acc if use_acc else None,
                                self.metric == "accuracy", eng._loss_buf)
        if train:
            self.last_step_losses = losses
            self.last_native_stats = dict(loop.last_stats)
        return nb

    def _train_one_epoch(self, epoch):
        if not self.model.training:
            self.model.train()
        if self.train_sampler is not None:
            self.train_sampler.set_epoch(epoch)
        eng = self.engine
        nb = len(self.train_loader)
        if self._native_ok(self.train_loader, self.model):
            # every step's loss lands in pinned host memory anyway (the kernel stores it there), so without a metric the
            # epoch mean needs no device accumulator: no zeroing kernel before the epoch, no read-back after it
            use_acc = bool(self.metric)
            if use_acc:
                eng.reset_accumulators()
            self._run_native_epoch(self.train_loader, epoch, train=True, use_acc=use_acc)
            if self.scheduler_type == "StepLR":
                self.scheduler.step()
            if use_acc:
                sums = eng.accumulators()
            else:
                sums = [float(self.last_step_losses.sum(dtype=torch.float64)), 0.0, float(nb)]
            loss, metric = self._reduce_epoch(sums)
            self.train_losses.append(loss)
            if self.metric:
                self.train_metrics.append(metric)
            return
        eng.reset_accumulators()
        if self._native_graph_ok(self.train_loader, self.model):
            self._train_epoch_native_graph(epoch)
            if self.scheduler_type == "StepLR":
                self.scheduler.step()
            loss, metric = self._reduce_epoch(eng.accumulators())
            self.train_losses.append(loss)
            if self.metric:
                self.train_metrics.append(metric)
            return
        every = max(int(self.cfg.progress_every), 1)
        feed = self._feed(self.train_loader)
        pin_losses = torch.zeros(nb, dtype=torch.float32).pin_memory() if self.device.type == "cuda" else None
        with self._bar(feed, nb) as tepoch:
            for i, (inputs, targets) in enumerate(tepoch):
                eng.train_step(inputs, targets)
                if pin_losses is not None and eng.last_loss is not None:
                    pin_losses[i].copy_(eng.last_loss, non_blocking=True)      # per-step loss without a sync
                if self.scheduler_type == "CosineAnnealingWarmRestarts":
                    self.scheduler.step(epoch - 1 + i / nb)  # as per pytorch docs
                if self.cfg.progress and self.rank == 0 and (i + 1) % every == 0:
                    ls, ms, n = eng.read_accumulators()
                    post = {"loss": ls / max(n, 1)}
                    if self.metric:
                        post["metric"] = ms / max(n, 1)
                    tepoch.set_postfix(**post)
        if self.scheduler_type == "StepLR":
            self.scheduler.step()
        loss, metric = self._reduce_epoch(eng.accumulators())
        self.last_step_losses = pin_losses
        self.last_native_stats = {"h2d_bytes": getattr(feed, "h2d_bytes", 0), "d2h_bytes": 4 * nb, "steps": nb}
        self.train_losses.append(loss)
        if self.metric:
            self.train_metrics.append(metric)

    @torch.no_grad()
    def _validate_one_epoch(self):
        self._sync_buffers()
        loss, metric = self._eval_loop(self.model, self.val_loader, reduce=self.val_sampler is not None)
        self.val_losses.append(loss)
        if self.metric:
            self.val_metrics.append(metric)
        if self.scheduler_type == "ReduceLROnPlateau":
            self.scheduler.step(loss)       # the reference never steps it (a no-op scheduler)

    @torch.no_grad()
    def _eval_loop(self, model, loader, reduce=False):
        was_training = model.training
        model.eval()
        eng = self.engine
        eng.reset_accumulators()
        if self._native_ok(loader, model):
            self._run_native_epoch(loader, 0, train=False)
        else:
            with self._bar(self._feed(loader), len(loader)) as tepoch:
                for inputs, targets in tepoch:
                    eng.eval_step(model, inputs, targets)
        sums = eng.accumulators()
        if reduce and self.world_size > 1:
            sums = self.comm.reduce_scalars(sums)
        loss_sum, metric_sum, n = sums.tolist()
        model.train(was_training)
        n = max(n, 1.0)
        return loss_sum / n, metric_sum / n

    # -------------------------------------------------------------- checkpoints
    def _state_dict_for_save(self) -> "OrderedDict[str, torch.Tensor]":
        prefix = "module." if self.is_parallel else ""
        return self.optimizer.master_state_dict(self.model, prefix=prefix)

    def consolidate(self):
        """Collective: make this rank's fp32 master weights and optimizer state complete and check the communicator.

        The sharded all-reduce + update kernels (two-shot / NVLS) keep master and state current only on the rank that
        owns a slice; every rank must call this before any rank writes a checkpoint."""
        self.comm.check_errors()
        if self.world_size > 1 and self.engine.fused is None:
            self.comm.consolidate(self.optimizer, self.reducer.buckets)

    def save_model(self, model_dir):
        logger.info("Saving the model.")
        return save_state_dict(self._state_dict_for_save(), model_dir, "model.pth", writer=self._snap)

    def save_history_(self, model_dir):
        logger.info("Saving the training history.")
        os.makedirs(model_dir, exist_ok=True)
        path = os.path.join(model_dir, "history.pkl")
        with open(path, "wb") as fp:
            pickle.dump(self.history, fp)
        return path

    def save(self, model_dir: Optional[str] = None, epoch: Optional[int] = None):
        """``model.pth`` (reference format) + ``trainer_state.pth`` sidecar for resume."""
        model_dir = model_dir or self.model_dir
        path = self.save_model(model_dir)
        state = {
            "epoch": epoch if epoch is not None else len(self.train_losses),
            "optimizer": self.optimizer.state_dict(),
            "scheduler": self.scheduler.state_dict() if self.scheduler is not None else None,
            "train_losses": self.train_losses, "val_losses": self.val_losses,
            "train_metrics": self.train_metrics, "val_metrics": self.val_metrics,
            "rng": torch.get_rng_state(),
        }
        atomic_torch_save(state, os.path.join(model_dir, "trainer_state.pth"))
        return path

    def load(self, path: str):
        """Load ``model.pth`` (either key style) and, if present next to it, the resume sidecar."""
        model_dir = path if os.path.isdir(path) else os.path.dirname(path)
        model_path = os.path.join(model_dir, "model.pth") if os.path.isdir(path) else path
        if os.path.basename(model_path) == "trainer_state.pth":
            model_path = os.path.join(model_dir, "model.pth")
        state = strip_module_prefix(torch.load(model_path, map_location="cpu"))
        own = self.model.state_dict()
        missing = [k for k in own if k not in state]
        if missing:
            raise KeyError(f"checkpoint {model_path} lacks keys {missing[:4]}...")
        self.optimizer.load_master(state, self.model)
        with torch.no_grad():
            for name, buf in self.model.named_buffers():
                if name in state:
                    buf.copy_(state[name])
        side = os.path.join(model_dir, "trainer_state.pth")
        if os.path.exists(side):
            st = torch.load(side, map_location="cpu", weights_only=False)
            self.optimizer.load_state_dict(st["optimizer"])
            if self.scheduler is not None and st.get("scheduler") is not None:
                self.scheduler.load_state_dict(st["scheduler"])
            self.train_losses, self.val_losses = list(st["train_losses"]), list(st["val_losses"])
            self.train_metrics, self.val_metrics = list(st["train_metrics"]), list(st["val_metrics"])
            self.start_epoch = int(st["epoch"]) + 1
            torch.set_rng_state(st["rng"])
        self.engine.invalidate()
        return self

    # ---------------------------------------------------------------------- fit
    def fit(self):
        if self.train_loader is None:
            raise RuntimeError("fit() needs datasets; this Trainer was built for testing only")
        logger.info("Start training..")
        for epoch in range(self.start_epoch, self.epochs + 1):
            logger.info(f"{'-' * 30} EPOCH {epoch} / {self.epochs} {'-' * 30}")
            self._train_one_epoch(epoch)
            self.clear()
            self._validate_one_epoch()
            self.clear()
            # Save model on master node; everyone waits so no rank races ahead of the file.
            if epoch % max(self.cfg.save_every, 1) == 0 or epoch == self.epochs:
                self.consolidate()          # collective: owners hand their master / state slices to everyone
                if self.rank == 0:
                    self.save(self.model_dir, epoch=epoch)
                self.comm.barrier()
            if self.phase_timer is not None:
                self.phase_ms = self.phase_timer.summary()       # mean device ms per phase over the eager steps of this epoch
                if self.phase_ms and self.rank == 0:
                    logger.info("step phases (ms): " + ", ".join(f"{k} {v:.3f}" for k, v in sorted(self.phase_ms.items())))
            if self.metric:
                logger.info(f"train loss: {self.train_losses[-1]} - train {self.metric}: {self.train_metrics[-1]}")
                logger.info(f"valid loss: {self.val_losses[-1]} - valid {self.metric}: {self.val_metrics[-1]}\n\n")
            else:
                logger.info(f"train loss: {self.train_losses[-1]}")
                logger.info(f"valid loss: {self.val_losses[-1]}\n\n")
        self.history = {
            "epochs": [*range(1, self.epochs + 1)],
            "train_loss": self.train_losses,
            "val_loss": self.val_losses,
            "train_metric": self.train_metrics,
            "val_metric": self.val_metrics,
            "metric_type": self.metric,
        }
        if self.save_history and self.rank == 0:
            self.save_history_(self.model_dir)
        self.comm.barrier()
        logger.info("Training Complete.")
        return self.history

    train_epoch = _train_one_epoch      # public name for one pass over the training loader

    # --------------------------------------------------------------------- test
    @torch.no_grad()
    def test(self, model, test_loader):
        """Evaluate an external ``model`` on ``test_loader``; returns ``(loss, metric)`` or ``loss``."""
        logger.info("Testing..")
        model = model.to(self.device)
        loss, metric = self._eval_loop(model, test_loader, reduce=False)
        if self.metric:
            return loss, metric
        return loss

    def evaluate(self, loader=None):
        """North-star alias: evaluate the trainer's own model (default: the validation loader)."""
        loader = loader if loader is not None else self.val_loader
        if loader is None:
            raise RuntimeError("evaluate() needs a loader")
        return self.test(self.model, loader)

    def clear(self):
        gc.collect()        # no empty_cache(): it is an implicit device sync and frees graph pools

    def close(self):
        self.reducer.remove()
        for lp in (self.__dict__.get("_nloops") or {}).values():
            lp.close()
        self.__dict__["_nloops"] = {}
        self._nloop = None
        self._nloop_key = None
        self._page_unlock_all()
        self.comm.close()

    def __del__(self):  # best effort: never leave a freed dataset page-locked
        try:
            self._page_unlock_all()
        except Exception:
            pass

    def validate_kwargs(self, kwargs, allowed_kwargs, error_message="Keyword argument not understood:"):
        """Checks that all keyword arguments are in the set of allowed keys."""
        validate_kwargs(kwargs, allowed_kwargs, error_message)
