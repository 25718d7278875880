from .auroc import WindowedBinaryAUROC

__all__ = ["WindowedBinaryAUROC"]
