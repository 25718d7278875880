"""Functional metrics of the ranking & counting hot path — same names and signatures as
``torcheval.metrics.functional`` (reference functional/__init__.py:60-111) for the rows in scope
(SURVEY.md §8)."""
from .classification.accuracy import binary_accuracy, multiclass_accuracy
from .classification.auprc import binary_auprc, multiclass_auprc, multilabel_auprc
from .classification.auroc import binary_auroc, multiclass_auroc
from .classification.binned_auprc import binary_binned_auprc, multiclass_binned_auprc, multilabel_binned_auprc
from .classification.binned_auroc import binary_binned_auroc, multiclass_binned_auroc
from .classification.binned_precision_recall_curve import (
    binary_binned_precision_recall_curve,
    multiclass_binned_precision_recall_curve,
    multilabel_binned_precision_recall_curve,
)
from .classification.confusion_matrix import binary_confusion_matrix, multiclass_confusion_matrix
from .classification.f1_score import binary_f1_score, multiclass_f1_score
from .classification.precision import binary_precision, multiclass_precision
from .classification.precision_recall_curve import (
    binary_precision_recall_curve,
    multiclass_precision_recall_curve,
    multilabel_precision_recall_curve,
)
from .classification.recall import binary_recall, multiclass_recall
from .classification.recall_at_fixed_precision import (
    binary_recall_at_fixed_precision,
    multilabel_recall_at_fixed_precision,
)

__all__ = [
    "binary_accuracy",
    "binary_auprc",
    "binary_auroc",
    "binary_binned_auprc",
    "binary_binned_auroc",
    "binary_binned_precision_recall_curve",
    "binary_confusion_matrix",
    "binary_f1_score",
    "binary_precision",
    "binary_precision_recall_curve",
    "binary_recall",
    "binary_recall_at_fixed_precision",
    "multiclass_accuracy",
    "multiclass_auprc",
    "multiclass_auroc",
    "multiclass_binned_auprc",
    "multiclass_binned_auroc",
    "multiclass_binned_precision_recall_curve",
    "multiclass_confusion_matrix",
    "multiclass_f1_score",
    "multiclass_precision",
    "multiclass_precision_recall_curve",
    "multiclass_recall",
    "multilabel_auprc",
    "multilabel_binned_auprc",
    "multilabel_binned_precision_recall_curve",
    "multilabel_precision_recall_curve",
    "multilabel_recall_at_fixed_precision",
]
