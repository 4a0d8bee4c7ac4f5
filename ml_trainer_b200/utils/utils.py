"""Checkpoint / history helpers (reference src/utils/utils.py:9-68).

File formats are kept byte-compatible with the reference:
``model.pth`` is a ``torch.save``d ``OrderedDict[str, fp32 Tensor]`` whose keys carry
a ``module.`` prefix when written from parallel mode, and ``history.pkl`` is a pickled
dict with the six keys ``epochs, train_loss, val_loss, train_metric, val_metric,
metric_type``.  matplotlib is imported only inside :func:`plot_history` so that the
loaders work on boxes without a plotting stack (the reference imports it at module
scope, utils.py:5).
"""
from __future__ import annotations

import os
import pickle
from collections import OrderedDict

import torch

HISTORY_KEYS = ("epochs", "train_loss", "val_loss", "train_metric", "val_metric", "metric_type")
_DDP_PREFIX = "module."


def load_history(file_dir: str) -> dict:
    with open(os.path.join(file_dir, "history.pkl"), "rb") as fp:
        return pickle.load(fp)


def strip_module_prefix(state_dict) -> "OrderedDict[str, torch.Tensor]":
    """Drop a leading ``module.`` from every key that has one (and only those)."""
    out = OrderedDict()
    for k, v in state_dict.items():
        out[k[len(_DDP_PREFIX):] if k.startswith(_DDP_PREFIX) else k] = v
    return out


def load_model(model: torch.nn.Module, PATH: str, map_location="cpu") -> torch.nn.Module:  # noqa: N803
    """Load ``model.pth`` written by either local or parallel training.

    Unlike the reference (which blindly slices 7 characters off every key and relies
    on an exception to fall back) the prefix is removed only where present, so mixed
    or already-clean checkpoints load without the double read.
    """
    state = torch.load(PATH, map_location=map_location)
    model.load_state_dict(strip_module_prefix(state))
    return model


def plot_history(history: dict, show: bool = True, save_path: str | None = None):
    """Loss (and metric) curves.  Headless-safe: uses Agg when there is no display."""
    import matplotlib

    if not os.environ.get("DISPLAY"):
        matplotlib.use("Agg")
    import numpy as np
    from matplotlib import pyplot as plt

    x = history["epochs"]
    has_metric = history.get("metric_type") is not None
    fig, axes = plt.subplots(2 if has_metric else 1, 1, figsize=(10, 10 if has_metric else 5))
    axes = list(axes) if has_metric else [axes]
    panels = [("Loss", "train_loss", "val_loss", "Training Loss vs. Validation Loss")]
    if has_metric:
        name = history["metric_type"]
        panels.append((name, "train_metric", "val_metric", f"{name} - Training vs. Validation"))
    for ax, (ylabel, tr, va, title) in zip(axes, panels):
        ax.plot(x, history[tr], c="C0", label="train")
        ax.plot(x, history[va], c="C1", label="validation")
        if len(x) > 25:
            ticks = np.arange(0, len(x) + 1, 5)
            ax.set_xticks(ticks)
            ax.set_xticklabels(ticks, rotation=45)
        else:
            ax.set_xticks(x)
        ax.set_ylabel(ylabel)
        ax.set_title(title)
        ax.legend()
    axes[0].set_xlabel("Epochs")
    fig.tight_layout()
    if save_path:
        fig.savefig(save_path)
    if show:
        plt.show()
    return fig
