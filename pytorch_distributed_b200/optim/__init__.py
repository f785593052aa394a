from .sgd import SGD
from .lr_scheduler import StepLR

__all__ = ["SGD", "StepLR"]
