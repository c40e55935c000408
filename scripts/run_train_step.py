"""A few eager training steps through the trainer-style fused path (for ncu captures of single kernels)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch
from model.model import MnistModel
from model.loss import nll_loss
from pdt import optim as pdt_optim

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.manual_seed(0)
m = MnistModel().cuda()
m.train()
m.flat.defer_reduce = True      # as the Trainer does: the gradient reduction is fused into the optimizer's kernel
opt = pdt_optim.Adam(m.parameters(), lr=1e-3, amsgrad=True)
g = torch.Generator().manual_seed(1)
x = torch.randn(n, 1, 28, 28, generator=g).cuda()
t = torch.randint(0, 10, (n,), generator=g).cuda()
one = torch.ones((), device="cuda")
for _ in range(steps):
    opt.zero_grad()
    m.expect_target(t)
    loss = nll_loss(m(x), t)
    loss.backward(one)
    opt.step()
torch.cuda.synchronize()
print("loss", float(loss))
