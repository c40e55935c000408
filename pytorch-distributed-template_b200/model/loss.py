"""`loss` plug-ins (reference: model/loss.py:4-5)."""
from pdt import ops


def nll_loss(output, target):
    """Mean negative log-likelihood over rows whose target is not -100, as F.nll_loss.
    One fused kernel also counts top-1 / top-3 hits; they ride along as `loss._pdt_stats`."""
    return ops.nll_loss_with_stats(output, target)[0]


# Read by Trainer._step: this criterion can be evaluated by the model's head kernel (MnistModel.expect_target).
nll_loss.pdt_fused_head = True
