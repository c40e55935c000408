"""Per-tensor gradient error of the CUDA path against the fp32 oracle and the fp64 truth, by batch size and by
kernel path (general kernels / fused kernels).  Diagnostic; prints one line per (n, path)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT, os.path.join(ROOT, "tests")]
import torch

from oracle import ref_step as R
from util_gpu import make_model, max_err


def main():
    from model.loss import nll_loss
    from pdt import ops
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    for n in (1024, 2048, 4096, 8192):
        params = R.init_params(0)
        x, t, k2, k1 = R.synthetic_batch(n, 800 + n)
        _, _, g32 = R.loss_and_grads(params, x, t, k2, k1)
        p64 = {k: v.double() for k, v in params.items()}
        _, _, g64 = R.loss_and_grads(p64, x.double(), t, k2, k1)
        for path, rows in (("general", 0), ("fused", 1 << 30)):
            old = ops.FUSED_ROWS
            ops.FUSED_ROWS = rows
            try:
                model = make_model(params)
                model.train()
                model.inject_dropout_masks([(k2, k1)])
                td = t.cuda()
                if rows:
                    model.expect_target(td)
                loss = nll_loss(model(x.cuda()), td)
                loss.backward()
                torch.cuda.synchronize()
            finally:
                ops.FUSED_ROWS = old
            line = []
            for (k, _), p in zip(R.PARAM_SHAPES, model._ordered):
                line.append("%s ours-ref32 %.1e ours-f64 %.1e ref32-f64 %.1e" % (
                    k, max_err(p.grad, g32[k]), max_err(p.grad, g64[k]), max_err(g32[k], g64[k])))
            print("n=%d %s | " % (n, path) + " | ".join(line[:4]))


if __name__ == "__main__":
    main()
