"""One training / evaluation step, eager or replayed from a CUDA graph.

Hot loop of the reference (src/trainer.py:180-197): ``zero_grad`` -> H2D -> forward ->
loss -> ``loss.item()`` -> backward -> ``optimizer.step()`` -> metric on the host.
Here a step is: forward -> fused loss -> backward, during which the reducer launches the
fused all-reduce + update per bucket on a side stream -> join.  Loss and metric are
added into a 3-float device accumulator ``[loss_sum, metric_sum, steps]``; nothing
touches the host.

Graph mode: after ``WARMUP`` eager steps the whole sequence is captured once per
distinct (input buffer, target buffer) pair: the prefetcher cycles through a fixed set
of device slots, so there are as many graphs as slots.  Per-step host work then shrinks
to uploading eight floats of hyper-parameters and one ``cudaGraphLaunch``.  Learning
rate and bias corrections live in device memory so a replay sees new values.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

WARMUP = 3


class StepEngine:
    def __init__(self, model, criterion, criterion_type, metric, reducer, device,
                 graph="auto", compute_dtype=torch.float32, fused_step=True):
        self.model = model
        self.criterion = criterion
        self.criterion_type = criterion_type
        self.metric = metric
        self.reducer = reducer
        self.device = device
        self.compute_dtype = compute_dtype
        self.acc = torch.zeros(3, dtype=torch.float32, device=device)
        self.loss_static = torch.zeros((), dtype=torch.float32, device=device)   # last step's loss at a fixed address
        self._one = torch.ones((), dtype=torch.float32, device=device)
        want = graph if isinstance(graph, bool) else (str(graph).lower() in ("1", "true", "auto", "yes"))
        self.use_graph = bool(want) and device.type == "cuda"
        self._graph_forced = graph is True        # graph=True: a failed capture is an error; "auto": fall back to eager steps
        self._graphs: Dict[Tuple, torch.cuda.CUDAGraph] = {}
        self._eager_steps = 0
        self._pool = None
        self.graph_replays = 0
        self.launches_per_step = 0
        self.kernels_per_step: Dict[str, int] = {}      # per-kernel launch counts of one captured step
        self.last_loss: Optional[torch.Tensor] = None
        # single-kernel step (forward + loss + backward + all-reduce + update) when the model provides one
        self.fused = None
        if fused_step and getattr(criterion, "is_cross_entropy", False) and hasattr(model, "make_fused_step") \
                and device.type == "cuda" and metric in (None, "accuracy"):
            self.fused = model.make_fused_step(reducer.opt, reducer.comm)
        if self.fused is not None:
            self._loss_buf = torch.zeros((), dtype=torch.float32, device=device)

    # ------------------------------------------------------------ accumulators
    def reset_accumulators(self):
        self.acc.zero_()

    def accumulators(self) -> torch.Tensor:
        return self.acc.clone()

    def read_accumulators(self):
        return self.acc.tolist()

    def graph_replays_since(self, last_n_steps: int) -> int:
        """How many of the last ``last_n_steps`` train steps were graph replays (all of them once captured)."""
        return min(self.graph_replays, last_n_steps) if self.use_graph else 0

    def graph_exec_handle(self, inputs, targets):
        """Raw ``cudaGraphExec_t`` of the step captured on these input buffers (None until it has been captured)."""
        key = (inputs.data_ptr(), targets.data_ptr(), tuple(inputs.shape), tuple(targets.shape))
        g = self._graphs.get(key)
        return None if g is None else int(g.raw_cuda_graph_exec())

    def invalidate(self):
        """Drop captured graphs (after a checkpoint load or any out-of-band state change)."""
        self._graphs.clear()
        self._eager_steps = 0

    # -------------------------------------------------------------------- loss
    def _forward_loss(self, model, inputs, targets, train: bool):
        """Forward + loss + device-side accumulation.  Returns ``(outputs, loss, dlogits_or_None)``.

        With the native cross-entropy criterion the loss, its gradient w.r.t. the logits, the accuracy
        and the accumulator update come out of ONE kernel and no autograd node is built for the loss.
        """
        from .utils import metrics as M

        if self.compute_dtype != torch.float32 and inputs.is_floating_point() and inputs.dtype != self.compute_dtype \
                and not getattr(model, "_mlt_accepts_fp32_input", False):
            inputs = inputs.to(self.compute_dtype)       # inside the captured graph; library models convert themselves
        outputs = model(inputs)
        crit = self.criterion
        if getattr(crit, "native_fused", False) and crit.accepts(outputs, targets):
            in_kernel = self.metric is not None and self.metric in crit.metrics     # metric accumulated by the loss kernel
            loss, dl = crit.fused_step(outputs.detach(), targets, self.acc, in_kernel, need_grad=train)
            if self.metric and not in_kernel:
                with torch.no_grad():
                    self.acc[1] += M.evaluate(self.metric, outputs.detach(), targets).float()
            return outputs, loss, dl
        if self.criterion_type in ("l1", "l2", "custom") and targets.is_floating_point() \
                and targets.dtype != outputs.dtype:
            targets = targets.to(outputs.dtype)
        loss = self.criterion(outputs, targets)
        with torch.no_grad():
            if self.metric:
                self.acc[1] += M.evaluate(self.metric, outputs.detach(), targets).float()
            self.acc[0] += loss.detach().float()
            self.acc[2] += self._one
        return outputs, loss, None

    # ------------------------------------------------------------------- train
    def _train_body(self, inputs, targets):
        from .utils.tracing import phase

        self.reducer.arm()
        if getattr(self.model, "_mlt_uses_dropout", False):
            from .ops.functional import dropout_begin_step

            dropout_begin_step(self.device)
        with phase("forward+loss"):
            outputs, loss, dl = self._forward_loss(self.model, inputs, targets, train=True)
        with phase("backward"):
            if dl is not None:
                outputs.backward(dl)
            else:
                loss.backward()
        with phase("reduce+update"):
            self.reducer.finish_step()
        self.loss_static.copy_(loss.detach().float())
        self.last_loss = self.loss_static

    def train_step(self, inputs, targets):
        opt = self.reducer.opt
        if self.fused is not None and self.fused.supports(inputs, targets):
            opt.begin_step(self.reducer.grad_scale)
            self.fused(inputs, targets, True, self.acc, self.metric == "accuracy", self._loss_buf)
            self.last_loss = self._loss_buf
            self.launches_per_step = 1
            return
        if not self.use_graph:
            opt.begin_step(self.reducer.grad_scale)
            self._train_body(inputs, targets)
            return
        key = (inputs.data_ptr(), targets.data_ptr(), tuple(inputs.shape), tuple(targets.shape))
        g = self._graphs.get(key)
        if g is not None:
            opt.begin_step(self.reducer.grad_scale)
            g.replay()
            self.graph_replays += 1
            return
        if self._eager_steps < WARMUP:
            self._eager_steps += 1
            opt.begin_step(self.reducer.grad_scale)
            if self._graph_forced or getattr(self.model, "_mlt_native", False):
                self._train_body(inputs, targets)
            else:
                self._eager_step_watching_for_syncs(inputs, targets)
            return
        # capture: hyper-parameters are uploaded outside the graph, everything else inside
        opt.begin_step(self.reducer.grad_scale)
        torch.cuda.synchronize(self.device)
        from .ops._lib import launches

        n0, by0 = launches.n, dict(launches.by_name)
        g = torch.cuda.CUDAGraph()
        prev_stream = torch.cuda.current_stream(self.device)
        try:
            with torch.cuda.graph(g, pool=self._pool):
                self._train_body(inputs, targets)
        except Exception as ex:  # noqa: BLE001 - a user module that syncs (.item()), branches on data, allocates outside the pool
            if self._graph_forced:
                raise
            import logging

            from .ops import functional as F

            logging.getLogger("ml_trainer_b200").warning(
                "CUDA-graph capture of the training step failed (%s: %s); continuing with eager steps", type(ex).__name__, ex)
            self.use_graph = False
            self._graphs.clear()
            self._heal_after_failed_capture(prev_stream)
            self.model.zero_grad(set_to_none=True)
            F.new_step()                      # forget the uses the aborted forward pass registered
            self._train_body(inputs, targets)
            return
        self.launches_per_step = launches.n - n0      # native kernels recorded into (and replayed by) the graph
        self.kernels_per_step = {k: v - by0.get(k, 0) for k, v in launches.by_name.items() if v - by0.get(k, 0) > 0}
        if self._pool is None:
            self._pool = g.pool()
        if len(self._graphs) >= 32:           # input buffers keep changing (no stable ring of slots): do not grow without bound
            self._graphs.pop(next(iter(self._graphs)))
        self._graphs[key] = g
        g.replay()          # capture does not execute; run the step that was just recorded
        self.graph_replays += 1

    def _eager_step_watching_for_syncs(self, inputs, targets):
        """The reference accepts any ``nn.Module`` (/root/reference/main.py:30, src/trainer.py:41-47).  A warm-up step of a
        user-supplied module with torch's sync detector on: a forward / loss that reads values back to the
        host (``.item()``, ``bool(tensor)``, ``.cpu()``: what data-dependent control flow looks like from here) cannot be
        captured, so graphs are switched off BEFORE a doomed capture is ever started.  The step itself runs normally."""
        import logging
        import warnings

        mode = torch.cuda.get_sync_debug_mode()
        with warnings.catch_warnings(record=True) as caught:
            warnings.simplefilter("always")
            torch.cuda.set_sync_debug_mode("warn")        # (announces itself as a prototype feature: recorded, not shown)
            try:
                self._train_body(inputs, targets)
            finally:
                torch.cuda.set_sync_debug_mode(mode)
        if any("synchroniz" in str(w.message) and "prototype" not in str(w.message) for w in caught):
            logging.getLogger("ml_trainer_b200").warning(
                "the training step synchronises with the host (e.g. .item() in forward): CUDA graphs are off for this model")
            self.use_graph = False

    def _heal_after_failed_capture(self, prev_stream):
        """torch leaves three things behind when ``capture_end`` raises: the capture side stream stays current, and the default
        CUDA generator stays in 'capturing' mode (every later torch.randn on the device would assert).  Restore the stream and
        run one trivial successful capture, whose epilogue resets the generator."""
        torch.cuda.set_stream(prev_stream)
        torch.cuda.synchronize(self.device)
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._one.mul_(1.0)
            del g
        except Exception:  # noqa: BLE001
            pass
        torch.cuda.set_stream(prev_stream)
        torch.cuda.synchronize(self.device)

    # -------------------------------------------------------------------- eval
    @torch.no_grad()
    def eval_step(self, model, inputs, targets):
        if self.fused is not None and model is self.model and self.fused.supports(inputs, targets):
            self.fused(inputs, targets, False, self.acc, self.metric == "accuracy", self._loss_buf)
            return None
        from .utils.tracing import phase

        with phase("eval"):
            outputs, _, _ = self._forward_loss(model, inputs, targets, train=False)
        return outputs
