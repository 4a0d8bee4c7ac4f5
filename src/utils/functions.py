"""``src.utils.functions`` of the reference -> :mod:`ml_trainer_b200.utils.functions`."""
from ml_trainer_b200.utils.functions import (TensorAugment, custom_loss_function,  # noqa: F401
                                             custom_pre_process_function)
