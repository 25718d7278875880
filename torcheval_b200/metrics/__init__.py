"""``torcheval_b200.metrics`` — the ranking & counting metrics of ``torcheval.metrics`` (same class
names, constructors and update / compute / merge_state / reset contract; reference
torcheval/metrics/__init__.py:70-133), computed by sm_100a kernels.

    from torcheval_b200.metrics import BinaryAUROC, MulticlassConfusionMatrix
    from torcheval_b200.metrics import functional as F
    from torcheval_b200.metrics.toolkit import sync_and_compute
"""
from . import functional
from .classification import *  # noqa: F401,F403
from .classification import __all__ as _classification_all
from .metric import Metric
from .window import WindowedBinaryAUROC

__all__ = ["Metric", "functional", "WindowedBinaryAUROC", *_classification_all]
