"""conv2 + Dropout2d + max-pool + relu on the tensor cores (pdt_conv2_drop_pool_relu_fwd_tc: tcgen05.mma in
split-TF32 arithmetic) against the oracle's fp64 evaluation of model/model.py:17 and against the CUDA-core kernel
it stands beside.  Tolerance: 1e-5 of max|ref| (north_star's fp32 bound); arg-max codes exact on inputs whose pooling
windows have no near-tie (oracle robust_batch with a margin well above the arithmetic's 1e-6)."""
import pytest
import torch
import torch.nn.functional as F

from oracle import ref_step as R
from util_gpu import flat_from_params, lane_to_rows, max_err

pytestmark = pytest.mark.gpu


def _run(n, seed, train):
    from pdt import ops
    from pdt.native import call, ptr, stream
    params = R.init_params(0)
    x, t, k2, k1 = R.robust_batch(n, seed, margin=1e-4, modes=("train" if train else "eval",))
    dev = torch.device("cuda")
    flat = flat_from_params(params, dev)
    keep = k2.cuda() if train else None
    _, saved = ops.mnist_forward_raw(x.cuda(), flat, keep, k1.cuda() if train else None, True)
    p1 = saved[0]
    T = (n + 31) // 32
    base = flat.data_ptr()
    outs = []
    for entry in ("pdt_conv2_drop_pool_relu_fwd_tc", "pdt_conv2_drop_pool_relu_fwd"):      # tensor cores, CUDA cores
        p2 = torch.full((T * 320 * 32,), float("nan"), device=dev)
        idx2 = torch.full((T * 320 * 32,), 255, dtype=torch.uint8, device=dev)
        call(entry, ptr(p1), base + 4 * 260, base + 4 * 5260, ptr(keep), 2.0, ptr(p2), ptr(idx2), n, stream())
        outs.append((p2, idx2))
    (p2, idx2), (p2_core, idx2_core) = outs
    torch.cuda.synchronize()
    p64 = {k: v.double() for k, v in params.items()}
    h = F.relu(F.max_pool2d(F.conv2d(x.double(), p64["conv1.weight"], p64["conv1.bias"]), 2))
    c2 = F.conv2d(h, p64["conv2.weight"], p64["conv2.bias"])
    if train:
        c2 = c2 * (k2.double() * 2.0)[:, :, None, None]
    ref = F.relu(F.max_pool2d(c2, 2)).view(n, 320)
    return (lane_to_rows(p2, n, 320), lane_to_rows(idx2, n, 320), lane_to_rows(p2_core, n, 320),
            lane_to_rows(idx2_core, n, 320), ref)


@pytest.mark.parametrize("n", [1, 7, 37, 128, 1024])
@pytest.mark.parametrize("train", [False, True])
def test_tensor_core_conv2_matches_oracle_and_cuda_core_kernel(n, train):
    p2, idx2, p2_core, idx2_core, ref = _run(n, 4000 + n, train)
    assert not torch.isnan(p2).any()
    e_tc, e_core = max_err(p2, ref), max_err(p2_core, ref)
    assert e_tc <= 1e-5, (e_tc, e_core)
    assert e_tc <= 4 * e_core + 1e-6, (e_tc, e_core)       # no worse than a few fp32 roundings of the CUDA-core sum
    assert torch.equal(idx2.cpu(), idx2_core.cpu())
