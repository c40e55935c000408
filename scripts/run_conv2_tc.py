"""Runs conv2's tensor-core forward at batch 1024 a few times (the command the ncu capture in profiles/ wraps)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch
from pdt import ops
from pdt.native import call, ptr
from model.model import MnistModel

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device("cuda")
torch.manual_seed(0)
m = MnistModel().to(dev)
x = torch.randn(n, 1, 28, 28, device=dev)
k2 = (torch.rand(n, 20, device=dev) < 0.5).to(torch.uint8)
k1 = (torch.rand(n, 50, device=dev) < 0.5).to(torch.uint8)
_, saved = ops.mnist_forward_raw(x, m.flat.data, k2, k1, True)
p1, idx1, p2, idx2 = saved[:4]
base = m.flat.data.data_ptr()
for _ in range(6):
    call("pdt_conv2_drop_pool_relu_fwd_tc", ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0, ptr(p2), ptr(idx2), n,
         torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("ok")
