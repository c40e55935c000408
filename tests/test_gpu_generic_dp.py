"""FlatDataParallel around a module that is NOT the native MnistModel (reference train.py:46-52 wraps whatever
`arch` names): gradients are flattened, scaled by 1/world with pdt_flat_scale_cast and all-reduced once, fired by
the last parameter's gradient hook.  One process, world size 1, on one GPU (the scale is 1 and the all-reduce is the
identity, so plain torch autograd on an unwrapped copy is the reference)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.nn as nn

pytestmark = pytest.mark.gpu


@pytest.fixture
def world1():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29571")
    torch.cuda.set_device(0)
    if not dist.is_initialized():
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    yield
    dist.destroy_process_group()


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Linear(12, 7)
        self.b = nn.Linear(7, 3)
        self.unused = nn.Linear(7, 3)

    def forward(self, x, use_all=True):
        h = torch.relu(self.a(x))
        out = self.b(h)
        return out + self.unused(h) if use_all else out


@pytest.mark.parametrize("register", [False, True])
def test_generic_module_matches_plain_autograd_and_rejects_unused_parameters(world1, register):
    """register=True: the flat gradient buffer lives in ncclMemAlloc memory registered on the communicator."""
    from pdt import optim as O
    from pdt.parallel import FlatDataParallel
    torch.manual_seed(3)
    plain = Net().cuda()
    wrapped_core = Net().cuda()
    wrapped_core.load_state_dict(plain.state_dict())
    dp = FlatDataParallel(wrapped_core, device_ids=[0], comm="nccl", register_buffer=register)
    assert (dp._registered is not None) == register
    assert not dp.native and list(dp.state_dict())[0].startswith("module.")
    opt = O.SGD(wrapped_core.parameters(), lr=0.1, momentum=0.9)
    ref_opt = torch.optim.SGD(plain.parameters(), lr=0.1, momentum=0.9)
    x = torch.randn(16, 12, device="cuda")
    for _ in range(3):
        opt.zero_grad()
        ref_opt.zero_grad()
        dp(x).square().mean().backward()
        plain(x).square().mean().backward()
        for p, q in zip(wrapped_core.parameters(), plain.parameters()):
            assert torch.allclose(p.grad, q.grad, rtol=1e-6, atol=1e-7)
        opt.step()
        ref_opt.step()
    for p, q in zip(wrapped_core.parameters(), plain.parameters()):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)
    # a parameter without a gradient: the all-reduce cannot have fired; the next forward says so
    opt.zero_grad()
    dp(x, use_all=False).square().mean().backward()
    with pytest.raises(RuntimeError, match="gradients for 4 of 6 parameters"):
        dp(x)
    # and the counter starts afresh afterwards
    opt.zero_grad()
    dp(x).square().mean().backward()
    dp(x)
    dp.release_registered_buffer()


def test_model_moved_after_optimizer_is_built_raises():
    from model.model import MnistModel
    from pdt import native, optim as O
    model = MnistModel().cuda()
    opt = O.Adam(model.parameters(), lr=1e-3)
    fb = model.flat
    model.cuda()                    # no move: the buffer (and the optimizer's view of it) stays
    model.to(torch.float32)
    assert model.flat is fb
    model.cpu()
    model.cuda()                    # a real move: the optimizer's buffer is orphaned and says so
    assert model.flat is not fb
    with pytest.raises(native.PdtError, match="moved or cast"):
        opt.step()
    with pytest.raises(TypeError):
        O.Adam(model.parameters(), lr=1e-3, nonsense=1)
    O.Adam(model.parameters(), lr=1e-3, foreach=None, fused=None, capturable=False, maximize=False)
    with pytest.raises(ValueError):
        O.SGD(model.parameters(), lr=1e-3, maximize=True)
