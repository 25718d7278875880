from .accuracy import BinaryAccuracy, MulticlassAccuracy
from .auprc import BinaryAUPRC, MulticlassAUPRC, MultilabelAUPRC
from .auroc import BinaryAUROC, MulticlassAUROC
from .binned_auprc import BinaryBinnedAUPRC, MulticlassBinnedAUPRC, MultilabelBinnedAUPRC
from .binned_auroc import BinaryBinnedAUROC, MulticlassBinnedAUROC
from .binned_precision_recall_curve import (
    BinaryBinnedPrecisionRecallCurve,
    MulticlassBinnedPrecisionRecallCurve,
    MultilabelBinnedPrecisionRecallCurve,
)
from .confusion_matrix import BinaryConfusionMatrix, MulticlassConfusionMatrix
from .f1_score import BinaryF1Score, MulticlassF1Score
from .precision import BinaryPrecision, MulticlassPrecision
from .precision_recall_curve import (
    BinaryPrecisionRecallCurve,
    MulticlassPrecisionRecallCurve,
    MultilabelPrecisionRecallCurve,
)
from .recall import BinaryRecall, MulticlassRecall
from .recall_at_fixed_precision import BinaryRecallAtFixedPrecision, MultilabelRecallAtFixedPrecision

__all__ = [
    "BinaryAccuracy",
    "BinaryAUPRC",
    "BinaryAUROC",
    "BinaryBinnedAUPRC",
    "BinaryBinnedAUROC",
    "BinaryBinnedPrecisionRecallCurve",
    "BinaryConfusionMatrix",
    "BinaryF1Score",
    "BinaryPrecision",
    "BinaryPrecisionRecallCurve",
    "BinaryRecall",
    "BinaryRecallAtFixedPrecision",
    "MulticlassAccuracy",
    "MulticlassAUPRC",
    "MulticlassAUROC",
    "MulticlassBinnedAUPRC",
    "MulticlassBinnedAUROC",
    "MulticlassBinnedPrecisionRecallCurve",
    "MulticlassConfusionMatrix",
    "MulticlassF1Score",
    "MulticlassPrecision",
    "MulticlassPrecisionRecallCurve",
    "MulticlassRecall",
    "MultilabelAUPRC",
    "MultilabelBinnedAUPRC",
    "MultilabelBinnedPrecisionRecallCurve",
    "MultilabelPrecisionRecallCurve",
    "MultilabelRecallAtFixedPrecision",
]
