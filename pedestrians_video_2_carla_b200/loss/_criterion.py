"""The kernels behind the loss mirrors compute the mean (or sum) of squared errors, i.e. torch.nn.MSELoss -- the only
criterion the reference wires in (loss/__init__.py:28-35).  Anything else is refused instead of silently computing MSE."""
from torch import nn


def require_mse(criterion, name, reductions=('mean',)):
    """None is accepted (the dependent modes of the reference pass no criterion); returns the reduction."""
    if criterion is None:
        return reductions[0]
    if not isinstance(criterion, nn.MSELoss):
        raise NotImplementedError('%s: the accelerated path implements nn.MSELoss only, got %s'
                                  % (name, type(criterion).__name__))
    if criterion.reduction not in reductions:
        raise NotImplementedError('%s: nn.MSELoss(reduction=%r) is not implemented (supported: %s)'
                                  % (name, criterion.reduction, ', '.join(reductions)))
    return criterion.reduction
