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


def make_model(params=None, seed=0):
    from model.model import MnistModel
    torch.manual_seed(seed)
    m = MnistModel().to("cuda")
    if params is not None:
        m.load_state_dict({k: v for k, v in params.items()})
    return m
