"""``src.model`` of the reference (src/model.py) -> :mod:`ml_trainer_b200.models`."""
from ml_trainer_b200.models import MLModel, build_model  # noqa: F401
