"""`metrics` plug-ins (reference: model/metric.py:4-20).  Both run the fused loss+metric
kernel, which produces integer correct-counts; the divide by len(target) happens here."""
from pdt import ops


def accuracy(output, target):
    assert output.shape[0] == len(target)
    top1, _, n = ops.metric_counts(output, target, 1)
    return top1 / n


def top_k_acc(output, target, k=3):
    assert output.shape[0] == len(target)
    _, topk, n = ops.metric_counts(output, target, k)
    return topk / n


# Tags read by Trainer._valid_epoch / test.py: a metric that is a ratio of one of the fused kernel's integer
# counts can be evaluated per rank and combined with a 16-byte all-reduce instead of gathering every output.
accuracy.pdt_count = ("top1", 1)
top_k_acc.pdt_count = ("topk", 3)
