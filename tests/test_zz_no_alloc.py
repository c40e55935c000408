"""pdt_nll_metrics_fwd_static: the loss/metric kernel for callers that may not allocate.

A host that has given all device memory to its own pool (or that must not hit cudaMalloc's
device-wide synchronisation in its loop) cannot supply the scratch buffer pdt_nll_metrics_fwd
asks for; this entry point uses a small array built into the library instead.  The test calls it
through the C ABI twice in a row (the second call sees whatever the first left in the shared
array) and requires the same record as the regular entry point AND as the oracle's independent
CPU computation of the same loss and counts: counts exact, loss within the suite's rtol 1e-5.
It lives in a file pytest collects last, and must pass both on its own in a fresh process and
after every other test has run.
"""
import struct

import pytest
import torch

pytestmark = pytest.mark.gpu


def _record(stats):
    raw = stats.cpu().numpy().tobytes()
    return struct.unpack("ffiiii", raw[:24])


def _inputs(n, seed, classes=10):
    g = torch.Generator().manual_seed(seed)
    logp = torch.log_softmax(torch.randn(n, classes, generator=g), dim=1).cuda()
    target = torch.randint(0, classes, (n,), generator=g)
    target[::11] = -100
    return logp, target.cuda()


def test_static_scratch_entry_matches_regular_entry_twice():
    from pdt import native
    lib = native.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = native.stream()
    for call_no, (n, seed) in enumerate([(1000, 1), (16384, 2)]):       # two calls in a row, second at the limit
        logp, target = _inputs(n, seed)
        ref = torch.zeros(8, dtype=torch.int32, device=dev)
        scratch = torch.zeros(lib.pdt_nll_scratch_bytes(n), dtype=torch.uint8, device=dev)
        native.call("pdt_nll_metrics_fwd", logp.data_ptr(), target.data_ptr(), n, 10, 3, -100, ref.data_ptr(),
                    scratch.data_ptr(), stream)
        out = torch.zeros(8, dtype=torch.int32, device=dev)
        native.call("pdt_nll_metrics_fwd_static", logp.data_ptr(), target.data_ptr(), n, 10, 3, -100,
                    out.data_ptr(), stream)
        torch.cuda.synchronize()
        a, b = _record(out), _record(ref)
        assert a[2:] == b[2:], (call_no, a, b)                       # top1, topk, n_valid, n_rows: exact
        assert abs(a[0] - b[0]) <= 1e-5 * abs(b[0]) and abs(a[1] - b[1]) <= 1e-5 * abs(b[1]), (call_no, a, b)
        assert a[4] == int((target != -100).sum())
        # independent of the kernel both entry points share: the oracle's loss and counts on the CPU
        from oracle import ref_step as R
        lc, tc = logp.cpu(), target.cpu()
        assert R.rows_are_tie_free(lc, 3)
        assert a[2] == R.accuracy_count(lc, tc) and a[3] == R.top_k_count(lc, tc, 3), (call_no, a)
        ref_loss = float(R.nll_loss(lc, tc))
        assert abs(a[0] - ref_loss) <= 1e-5 * abs(ref_loss), (call_no, a[0], ref_loss)


def test_static_scratch_entry_rejects_oversize():
    from pdt import native
    lib = native.load()
    logp, target = _inputs(16385, 3)
    out = torch.zeros(8, dtype=torch.int32, device="cuda")
    rc = lib.pdt_nll_metrics_fwd_static(logp.data_ptr(), target.data_ptr(), 16385, 10, 3, -100, out.data_ptr(),
                                        native.stream())
    assert rc == -1
    msg = lib.pdt_last_error().decode()
    assert "16384" in msg and "pdt_nll_metrics_fwd" in msg
