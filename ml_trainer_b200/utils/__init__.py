from .functions import custom_loss_function, custom_pre_process_function
from .utils import load_history, load_model, plot_history

__all__ = ["custom_loss_function", "custom_pre_process_function", "load_history", "load_model", "plot_history"]
