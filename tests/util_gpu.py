"""Helpers shared by the GPU parity tests."""
import torch

from oracle import ref_step as R


def lane_to_rows(t, n, feats):
    """[T][feats][32] lane layout -> [n][feats]."""
    T = (n + 31) // 32
    return t.view(T, feats, 32).permute(0, 2, 1).reshape(T * 32, feats)[:n]


def lane4_to_rows(t, n, feats):
    """[T][feats/4][32][4] (p1's lane4 layout) -> [n][feats]."""
    T = (n + 31) // 32
    return t.view(T, feats // 4, 32, 4).permute(0, 2, 1, 3).reshape(T * 32, feats)[:n]


def flat_from_params(params, device):
    return torch.cat([params[k].reshape(-1) for k, _ in R.PARAM_SHAPES]).to(device=device, dtype=torch.float32)


def params_from_flat(flat):
    out, off = {}, 0
    for k, shape in R.PARAM_SHAPES:
        n = 1
        for s in shape:
            n *= s
        out[k] = flat[off:off + n].view(shape).detach().cpu().clone()
        off += n
    return out


def close(a, b, rtol=1e-5):
    """rtol 1e-5 with atol = rtol * max|ref| per tensor (stated openly)."""
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    atol = rtol * float(b.abs().max())
    return torch.allclose(a, b, rtol=rtol, atol=atol)


def max_err(a, b):
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))


def grad_matches(ours, ref32, ref64, rtol=1e-5):
    """Gradient parity against the oracle: within rtol (of max|ref|) of the fp64 truth, and no further from the fp32
    oracle than rtol plus the fp32 oracle's OWN distance from the truth.  (At n = 4096 torch's fp32 CPU convolution
    backward is itself up to 2.5e-5 from the truth, depending on how many host threads split the sum -- measured on
    the GPU box, profiles/r02_grad_err.txt -- so "within 1e-5 of the fp32 oracle" alone would test the oracle's
    rounding, not this implementation.)  Returns (ok, message)."""
    e_truth, e_ref, e_o = max_err(ours, ref64), max_err(ref32, ref64), max_err(ours, ref32)
    ok = e_truth <= rtol and e_o <= rtol + e_ref
    return ok, "ours-fp64 %.2e, ours-oracle %.2e, oracle-fp64 %.2e" % (e_truth, e_o, e_ref)


def make_model(params=None, seed=0):
    from model.model import MnistModel
    torch.manual_seed(seed)
    m = MnistModel().to("cuda")
    if params is not None:
        m.load_state_dict({k: v for k, v in params.items()})
    return m
