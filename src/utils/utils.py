"""``src.utils.utils`` of the reference -> :mod:`ml_trainer_b200.utils.utils`."""
from ml_trainer_b200.utils.utils import load_history, load_model, plot_history  # noqa: F401
