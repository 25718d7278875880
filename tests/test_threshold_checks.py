"""The threshold checks of the binned metrics (reference: `_binned_precision_recall_curve_param_check`,
`_binary_binned_auprc_param_check`, `_binary_binned_auroc_param_check`): same errors, same order, but
thresholds built from an int or a list are judged on the host when they are made (no device read),
and a tensor handed in costs one read instead of five."""
import pytest
import torch

from torcheval_b200.metrics.functional.classification.binned_auprc import _binary_binned_auprc_param_check
from torcheval_b200.metrics.functional.classification.binned_auroc import _binary_binned_auroc_param_check
from torcheval_b200.metrics.functional.classification.binned_precision_recall_curve import (
    _binned_precision_recall_curve_param_check,
)
from torcheval_b200.metrics.functional.tensor_utils import (
    THR_FIRST, THR_LAST, THR_RANGE, THR_UNSORTED, _FAULTS_ATTR, _create_threshold_tensor, _threshold_faults,
)

CPU = torch.device("cpu")
SORTED = "The `threshold` should be a sorted tensor."
RANGE = r"The values in `threshold` should be in the range of \[0, 1\]."
FIRST = "First value in `threshold` should be 0."
LAST = "Last value in `threshold` should be 1."


def test_faults_known_at_creation():
    assert getattr(_create_threshold_tensor(100, CPU), _FAULTS_ATTR) == 0
    assert getattr(_create_threshold_tensor(2, CPU), _FAULTS_ATTR) == 0
    assert getattr(_create_threshold_tensor(1, CPU), _FAULTS_ATTR) == THR_LAST  # linspace(0, 1, 1) = [0.]
    assert getattr(_create_threshold_tensor([0.0, 0.5, 1.0], CPU), _FAULTS_ATTR) == 0
    assert getattr(_create_threshold_tensor([0.5, 0.2, 1.0], CPU), _FAULTS_ATTR) == THR_UNSORTED | THR_FIRST
    assert getattr(_create_threshold_tensor([-0.5, 0.2, 1.5], CPU), _FAULTS_ATTR) == THR_RANGE | THR_FIRST | THR_LAST
    t = torch.tensor([0.0, 1.0])
    assert _create_threshold_tensor(t, CPU) is t and not hasattr(t, _FAULTS_ATTR)  # a caller's tensor: judged when checked
    assert _threshold_faults(t) == 0
    assert _threshold_faults(torch.tensor([0.2, 0.1])) == THR_UNSORTED | THR_FIRST | THR_LAST
    assert _threshold_faults(torch.empty(0)) is None and _threshold_faults(torch.zeros(2, 2)) is None


@pytest.mark.parametrize("make", [lambda v: _create_threshold_tensor(v, CPU), lambda v: torch.tensor(v)], ids=["list", "tensor"])
def test_errors_and_their_order(make):
    ok = make([0.0, 0.25, 1.0])
    _binned_precision_recall_curve_param_check(ok)
    _binary_binned_auroc_param_check(1, ok)
    _binary_binned_auprc_param_check(1, ok)
    for check in (_binned_precision_recall_curve_param_check, lambda t: _binary_binned_auroc_param_check(1, t),
                  lambda t: _binary_binned_auprc_param_check(1, t)):
        with pytest.raises(ValueError, match=SORTED):
            check(make([0.0, 0.7, 0.3, 1.0]))
        with pytest.raises(ValueError, match=SORTED):  # unsorted is reported before out of range
            check(make([0.0, 1.7, 0.3]))
        with pytest.raises(ValueError, match=RANGE):
            check(make([0.0, 0.5, 1.5]))
    # the curve and the AUROC accept thresholds that do not span [0, 1]; the AUPRC does not
    part = make([0.1, 0.5, 0.9])
    _binned_precision_recall_curve_param_check(part)
    _binary_binned_auroc_param_check(1, part)
    with pytest.raises(ValueError, match=FIRST):
        _binary_binned_auprc_param_check(1, part)
    with pytest.raises(ValueError, match=LAST):
        _binary_binned_auprc_param_check(1, make([0.0, 0.5, 0.9]))
    with pytest.raises(ValueError, match="1-dimensional"):
        _binary_binned_auprc_param_check(1, torch.zeros(2, 2))
    with pytest.raises(ValueError, match="num_tasks"):
        _binary_binned_auprc_param_check(0, ok)
    with pytest.raises(ValueError, match=LAST):
        _binary_binned_auprc_param_check(1, _create_threshold_tensor(1, CPU))
