"""Trainer keyword whitelist and defaults (reference src/trainer.py:26-41,307-311).

The eleven reference keys keep their names and defaults.  New keys are added to the
whitelist rather than bypassing it, so an unknown key still raises
``TypeError("Keyword argument not understood:", key)`` exactly as the reference does.
"""
from __future__ import annotations

from dataclasses import dataclass, field, fields
from typing import Any, Optional

REFERENCE_KEYS = (
    "seed", "scheduler", "optimizer", "momentum", "weight_decay", "lr",
    "criterion", "metric", "pred_function", "model_dir", "backend",
)

OPTIMIZERS = ("sgd", "adam", "adagrad", "adamax", "adamw")
CRITERIA = ("cross_entropy", "neg-loss", "l1", "l2", "custom")
SCHEDULERS = ("CosineAnnealingWarmRestarts", "ReduceLROnPlateau", "StepLR")


@dataclass
class TrainerConfig:
    # ---- reference keys (same defaults) --------------------------------------
    seed: int = 32
    scheduler: Optional[str] = None
    optimizer: str = "sgd"
    momentum: float = 0.9
    weight_decay: float = 0.0
    lr: float = 0.001
    criterion: str = "cross_entropy"
    metric: Optional[str] = "accuracy"
    pred_function: Optional[str] = "softmax"
    model_dir: str = "model_output"
    backend: str = "smddp"
    # ---- B200 additions ------------------------------------------------------
    dtype: str = "fp32"            # compute dtype: fp32 | bf16 (fp32 master weights either way)
    graph: Any = "auto"            # capture the whole train step in a CUDA graph: auto | True | False
    fused_step: bool = True        # models that provide a single-kernel train step (MLModel) use it
    native_loop: bool = True       # run whole epochs in the C++ executor when data and step allow it
    convert: bool = True           # swap torch.nn layers for the sm_100a op library on CUDA
    prefetch: int = 2              # pinned staging slots of the H2D prefetcher (0 = synchronous copies)
    num_workers: int = 0
    pin_memory: Optional[bool] = None   # None -> pinned whenever the device is CUDA
    progress_every: int = 50       # steps between progress-bar read-backs (the reference syncs every step)
    progress: bool = True          # tqdm bars on/off
    bucket_mb: float = 25.0        # gradient bucket cap (DDP default is 25 MiB)
    comm_algo: str = "auto"        # auto | oneshot | twoshot | nvls
    shard_val: bool = True         # shard + reduce validation across ranks (the reference replicates it)
    betas: tuple = (0.9, 0.999)
    eps: Optional[float] = None    # None -> the torch default of the chosen optimizer
    resume: Optional[str] = None   # path to a trainer_state.pth sidecar
    save_every: int = 1            # epochs between model.pth snapshots
    drop_last: bool = False
    trace: bool = False            # NVTX ranges per step phase + a per-epoch device-time summary (utils/tracing.py)

    extra: dict = field(default_factory=dict, repr=False)

    @classmethod
    def allowed(cls) -> set:
        return {f.name for f in fields(cls) if f.name != "extra"}

    @classmethod
    def from_kwargs(cls, kwargs: dict) -> "TrainerConfig":
        validate_kwargs(kwargs, cls.allowed())
        return cls(**kwargs)

    def reference_view(self) -> dict:
        return {k: getattr(self, k) for k in REFERENCE_KEYS}


def validate_kwargs(kwargs, allowed_kwargs, error_message="Keyword argument not understood:"):
    """Raise ``TypeError(error_message, key)`` on the first key outside the whitelist."""
    for kwarg in kwargs:
        if kwarg not in allowed_kwargs:
            raise TypeError(error_message, kwarg)
