"""Stage-by-stage comparison of the CUDA path with the oracle; prints relative max errors.
Run on the GPU box:  python scripts/gpu_debug.py [n]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT, os.path.join(ROOT, "tests")]
import torch
import torch.nn.functional as F

from oracle import ref_step as R
from pdt import native, ops
from util_gpu import lane_to_rows, lane4_to_rows, flat_from_params, max_err

n = int(sys.argv[1]) if len(sys.argv) > 1 else 37
dev = torch.device("cuda")
params = R.init_params(0)
x, t, k2, k1 = R.synthetic_batch(n, 1234)
flat = flat_from_params(params, dev)
xg, tg, k2g, k1g = x.to(dev), t.to(dev), k2.to(dev), k1.to(dev)

for mode in ("eval", "train"):
    kk2, kk1 = (None, None) if mode == "eval" else (k2g, k1g)
    logp, saved = ops.mnist_forward_raw(xg, flat, kk2, kk1, True)
    torch.cuda.synchronize()
    p1, idx1, p2, idx2, hd = saved[:5]
    p = params
    r1 = F.relu(F.max_pool2d(F.conv2d(x, p["conv1.weight"], p["conv1.bias"]), 2))
    c2 = F.conv2d(r1, p["conv2.weight"], p["conv2.bias"])
    if mode == "train":
        c2 = c2 * (k2.float() * 2)[:, :, None, None]
    r2 = F.relu(F.max_pool2d(c2, 2)).view(-1, 320)
    h = F.relu(F.linear(r2, p["fc1.weight"], p["fc1.bias"]))
    if mode == "train":
        h = h * k1.float() * 2
    ref = R.mnist_forward(p, x, None if mode == "eval" else k2, None if mode == "eval" else k1)
    print(mode, "p1", max_err(lane4_to_rows(p1, n, 1440), r1.reshape(n, -1)),
          "p2", max_err(lane_to_rows(p2, n, 320), r2),
          "hd", max_err(lane_to_rows(hd, n, 50), h),
          "logp", max_err(logp, ref))

# loss + metrics
ref_out = R.mnist_forward(params, x, k2, k1)
loss, stats = ops.nll_loss_with_stats(ref_out.to(dev), tg)
st = ops.unpack_stats(stats)
print("loss", st, "ref", float(R.nll_loss(ref_out, t)), R.accuracy_count(ref_out, t), R.top_k_count(ref_out, t))

# backward
loss_ref, out_ref, g_ref = R.loss_and_grads(params, x, t, k2, k1)
logp, saved = ops.mnist_forward_raw(xg, flat, k2g, k1g, True)
loss, stats = ops.nll_loss_with_stats(logp.detach().requires_grad_(True), tg)
dlogp = torch.zeros(n, 10, device=dev)
native.call("pdt_nll_bwd", tg.data_ptr(), stats.data_ptr(), None, dlogp.data_ptr(), n, 10, -100, native.stream())
ref_dlogp = torch.zeros(n, 10)
ref_dlogp[torch.arange(n), t] = -1.0 / n
print("dlogp", max_err(dlogp, ref_dlogp))
grad = torch.zeros(21840 + 4, device=dev)
ops.mnist_backward_raw(xg, flat, saved, logp, dlogp, True, grad, 0.0, 1.0, stats)
torch.cuda.synchronize()
off = 0
for k, shape in R.PARAM_SHAPES:
    cnt = g_ref[k].numel()
    print("grad", k, max_err(grad[off:off + cnt].view(shape), g_ref[k]))
    off += cnt
print("tail", grad[21840:].tolist(), float(loss_ref))
