"""Raw uint8 pixels as model input: ToTensor + Normalize((0.1307,), (0.3081,)) of the reference's
loader (data_loader/data_loaders.py:13-16) applied while conv1 loads the batch.  The product on uint8 input must equal
the product on the loader's normalised floats BIT FOR BIT (output, loss, every gradient, on every kernel path), and
match the oracle fed those floats."""
import pytest
import torch

from oracle import ref_step as R
from util_gpu import grad_matches, make_model, max_err, close

pytestmark = pytest.mark.gpu


def reference_loader_floats(xb):
    """torchvision's ToTensor (v / 255) and Normalize ((t - mean) / std), in fp32, on the CPU"""
    return xb.to(torch.float32).div(255).sub(0.1307).div(0.3081)


def _step(x, t, masks, path):
    from model.loss import nll_loss
    from pdt import ops
    model = make_model(R.init_params(0))
    model.train()
    model.inject_dropout_masks([masks])
    old = ops.FUSED_ROWS
    ops.FUSED_ROWS = 0 if path == "general" else 1 << 30
    try:
        td = t.cuda()
        if path == "trainer":
            model.expect_target(td)
        out = model(x.cuda())
        loss = nll_loss(out, td)
        loss.backward()
        torch.cuda.synchronize()
    finally:
        ops.FUSED_ROWS = old
    return out.detach().clone(), loss.detach().clone(), [p.grad.detach().clone() for p in model._ordered]


@pytest.mark.parametrize("n", [1, 37, 128, 1024])
@pytest.mark.parametrize("path", ["general", "fused", "trainer"])
def test_uint8_input_equals_normalised_float_input(n, path):
    g = torch.Generator().manual_seed(70 + n)
    xb = torch.randint(0, 256, (n, 1, 28, 28), generator=g, dtype=torch.uint8)
    t = torch.randint(0, 10, (n,), generator=g)
    k2 = (torch.rand(n, 20, generator=g) < 0.5).to(torch.uint8)
    k1 = (torch.rand(n, 50, generator=g) < 0.5).to(torch.uint8)
    xf = reference_loader_floats(xb)
    out_b, loss_b, g_b = _step(xb, t, (k2, k1), path)
    out_f, loss_f, g_f = _step(xf, t, (k2, k1), path)
    assert torch.equal(out_b, out_f) and torch.equal(loss_b, loss_f)
    for a, b in zip(g_b, g_f):
        assert torch.equal(a, b)


def test_uint8_input_matches_oracle_on_the_loaders_floats():
    n = 256
    params = R.init_params(0)
    g = torch.Generator().manual_seed(5)
    xb = torch.randint(0, 256, (n, 1, 28, 28), generator=g, dtype=torch.uint8)
    t = torch.randint(0, 10, (n,), generator=g)
    k2 = (torch.rand(n, 20, generator=g) < 0.5).to(torch.uint8)
    k1 = (torch.rand(n, 50, generator=g) < 0.5).to(torch.uint8)
    for attempt in range(64):        # raw pixels whose normalised floats are well-conditioned (see ref_step.robust_batch)
        bad = (R.decision_margins(params, reference_loader_floats(xb), k2, k1) < 2e-6).nonzero()[:, 0]
        if bad.numel() == 0:
            break
        xb[bad] = torch.randint(0, 256, (int(bad.numel()), 1, 28, 28), generator=g, dtype=torch.uint8)
    else:
        pytest.fail("no well-conditioned draw")
    xf = reference_loader_floats(xb)
    out, loss, grads = _step(xb, t, (k2, k1), "fused")
    ref_loss, ref_out, ref_g = R.loss_and_grads(params, xf, t, k2, k1)
    p64 = {k: v.double() for k, v in params.items()}
    _, _, g64 = R.loss_and_grads(p64, xf.double(), t, k2, k1)
    assert close(out, ref_out), max_err(out, ref_out)
    assert abs(loss.item() - float(ref_loss)) <= 1e-5 * abs(float(ref_loss))
    for (k, _), gr in zip(R.PARAM_SHAPES, grads):
        ok, msg = grad_matches(gr, ref_g[k], g64[k])
        assert ok, (k, msg)


def test_trainer_epoch_on_uint8_batches(tmp_path):
    """The Trainer stages uint8 batches as they are (a quarter of the bytes) and trains to the same weights as on
    the loader's floats (same dropout stream)."""
    from pdt import ops
    from test_gpu_trainer import ListLoader, make_trainer
    finals = []
    for as_bytes in (True, False):
        model = make_model(R.init_params(0))
        ops.seed_dropout(11)
        loader = ListLoader()
        loader.batch_size = 64
        for i in range(4):
            g = torch.Generator().manual_seed(200 + i)
            xb = torch.randint(0, 256, (64, 1, 28, 28), generator=g, dtype=torch.uint8)
            t = torch.randint(0, 10, (64,), generator=g)
            loader.append(((xb if as_bytes else reference_loader_floats(xb)).pin_memory(), t.pin_memory()))
        tr = make_trainer(tmp_path / str(as_bytes), model, loader)
        log = tr._train_epoch(1)
        finals.append((log["loss"], model.flat.data[:21840].clone()))
    assert finals[0][0] == finals[1][0]
    assert torch.equal(finals[0][1], finals[1][1])
