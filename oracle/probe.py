"""TEST INFRASTRUCTURE (checker only, never on the product path): a data-parallel parity probe.

Runs K training steps of the PRODUCT (MnistModel under FlatDataParallel, one rank per GPU) on
slices of seeded global batches and compares, on rank 0, against
  (a) the oracle's W-rank restatement of the reference step (oracle/ref_step.py: forward, nll_loss,
      backward per rank, DDP's divide-by-W-and-sum of train.py:46-52, reduce_loss of
      base_trainer.py:165-174, torch Adam+amsgrad), in fp32 and in fp64, and
  (b) the product itself on ONE GPU over the concatenated global batch (SURVEY.md section 4:
      "W-GPU run == 1-GPU run on the global batch").
Used by scripts/ddp_check.py (tests/test_gpu_multi.py) and by bench.py's multi-GPU line, which
prints the result as "parity" so the driver's scaling record carries it.

Tolerances (north_star): metric counts bit-exact; loss, gradients and weights within 1e-5 of max|ref|
per tensor.  That is well-posed for every quantity of an SGD run (the update is linear in the gradient),
and both optimizers are run.  Under Adam the update is lr * m / (sqrt(v) + eps): wherever |g| is not
large against the rounding noise of a gradient sum, a difference in the last bits of g moves the weight
by a visible fraction of lr, and from the second step on every gradient is taken at those weights.  Two
fp32 runs that sum in different orders -- the reference's own fp32 run and its fp64 run included -- then
part by far more than 1e-5 (measured: 2e-5 of max|w| after 3 steps at batch 64, 1e-4 at batch 1024).
So for Adam the probe holds to rtol 1e-5 what is well-posed (every step's loss, the first step's
gradients, taken at identical weights) and requires of the weights after K steps that their mean
distance from the fp64 truth be no more than twice the fp32 reference's own, per tensor: as close to
the truth as the reference itself is.
"""
import torch

from . import ref_step as R


def _rel(a, b):
    a, b = a.detach().cpu().double().reshape(-1), b.detach().cpu().double().reshape(-1)
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))


def _make_opt(kind):
    return R.Adam(lr=1e-3, weight_decay=0.0, amsgrad=True) if kind == "adam" else R.SGD(lr=0.01)


def _oracle_run(params, batches, world, dtype, kind):
    """The reference step restated, keeping everything the comparison needs: final weights, the reduced loss of
    every step, the averaged gradients of every step, the metric counts of the last step."""
    p = {k: v.to(dtype) for k, v in params.items()}
    opt = _make_opt(kind)
    losses, grads, top1, top3 = [], [], 0, 0
    for step in batches:
        per_rank, local, top1, top3 = [], [], 0, 0
        for (x, t, k2, k1) in step:
            loss, out, g = R.loss_and_grads(p, x.to(dtype), t, k2, k1)
            per_rank.append(g)
            local.append(loss)
            top1 += R.accuracy_count(out, t)
            top3 += R.top_k_count(out, t, 3)
        grads.append(R.ddp_average(per_rank))
        losses.append(float(R.reduce_loss(local)))
        p = opt.step(dict(p), grads[-1])
    return p, losses, grads, (top1, top3)


def _split(flat):
    out, off = {}, 0
    for k, shape in R.PARAM_SHAPES:
        cnt = 1
        for s in shape:
            cnt *= s
        out[k] = flat[off:off + cnt].view(shape).detach().cpu().clone()
        off += cnt
    return out


def _product_run(batches_for_me, device, wrap, trainer_path, kind):
    """K steps of the product through the plug-in calls.  batches_for_me: list of (x, t, k2, k1).
    Returns (model, wrapped model, the flat gradient of every step, last tail or stats, the reduced loss of
    every step)."""
    from model.model import MnistModel
    from model.loss import nll_loss
    from pdt import optim as O, ops
    torch.manual_seed(0)
    model = MnistModel().to(device)
    opt = O.Adam(model.parameters(), lr=1e-3, weight_decay=0, amsgrad=True) if kind == "adam" \
        else O.SGD(model.parameters(), lr=0.01)
    net = wrap(model) if wrap is not None else model
    net.train()
    if trainer_path:
        model.flat.defer_reduce = True
    one = torch.ones((), device=device)
    losses, grads, stats = [], [], None
    for (x, t, k2, k1) in batches_for_me:
        td = t.to(device)
        model.inject_dropout_masks([(k2, k1)])
        opt.zero_grad()
        if trainer_path:
            model.expect_target(td)
        out = net(x.to(device))
        loss = nll_loss(out, td)
        loss.backward(one)
        opt.step()
        grads.append(model.flat.grad[:21840].clone())
        if wrap is not None:
            losses.append(float(net.reduced_tail()[0]))
            stats = net.reduced_tail().clone().cpu()
        else:
            losses.append(float(loss.detach()))
            st = ops.unpack_stats(loss._pdt_stats)
            stats = torch.tensor([losses[-1], st["top1"], st["topk"], st["n_valid"]], dtype=torch.float64)
    torch.cuda.synchronize()
    return model, net, grads, stats, losses


def _mean_err_vs_truth(w, truth):
    """per tensor: mean |w - truth| / max|truth|"""
    return {k: float((w[k].double() - truth[k]).abs().mean() / truth[k].abs().max()) for k, _ in R.PARAM_SHAPES}


def _compare(kind, ours_w, ours_g, ours_l, ours_counts, ref_w, ref_g, ref_l, ref_counts, truth_w, truth_g):
    """SGD: everything is well-posed -> rtol 1e-5 (relative to max|ref| per tensor) on the losses of every step,
    the gradients of every step and the final weights.  Adam: losses of every step and the gradients of the
    FIRST step at rtol 1e-5; the weights after K steps by their mean distance from the fp64 truth, which must
    not exceed twice the fp32 reference's own (see the module docstring)."""
    out = {"loss_max_err": max(abs(a - b) / abs(b) for a, b in zip(ours_l, ref_l)),
           "counts_exact": tuple(ours_counts) == tuple(ref_counts)}
    steps = range(len(ref_g)) if kind == "sgd" else range(1)
    # gradients: within 1e-5 of the fp64 truth, and no further from the fp32 reference than 1e-5 + that reference's
    # own distance from the truth (torch's fp32 CPU sums carry up to 2.5e-5 of their own at batch 4096)
    out["grad_max_err"] = max(_rel(_split(ours_g[s])[k], truth_g[s][k]) for s in steps for k, _ in R.PARAM_SHAPES)
    out["grad_max_err_vs_fp32"] = max(_rel(_split(ours_g[s])[k], ref_g[s][k]) for s in steps for k, _ in R.PARAM_SHAPES)
    out["fp32_reference_grad_err"] = max(_rel(ref_g[s][k], truth_g[s][k]) for s in steps for k, _ in R.PARAM_SHAPES)
    ok = (out["loss_max_err"] <= 1e-5 and out["grad_max_err"] <= 1e-5 and out["counts_exact"]
          and out["grad_max_err_vs_fp32"] <= 1e-5 + out["fp32_reference_grad_err"])
    if kind == "sgd":
        out["weights_max_err"] = max(_rel(ours_w[k], ref_w[k]) for k, _ in R.PARAM_SHAPES)
        ok = ok and out["weights_max_err"] <= 1e-5
    else:
        mine, theirs = _mean_err_vs_truth(ours_w, truth_w), _mean_err_vs_truth(ref_w, truth_w)
        out["weights_mean_err_vs_fp64"] = max(mine.values())
        out["reference_fp32_mean_err_vs_fp64"] = max(theirs.values())
        out["weights_max_err_vs_reference_fp32"] = max(_rel(ours_w[k], ref_w[k]) for k, _ in R.PARAM_SHAPES)
        ok = ok and all(mine[k] <= 2 * theirs[k] + 1e-6 for k in mine)
    out["ok"] = bool(ok)
    return out


def _robust_batches(n, world, steps, seed, kind, rank, device):
    """Global batches of n * world samples, one per step, well-conditioned (ref_step.robust_batch) at the weights
    the fp64 run of the same optimizer holds at that step: no relu / arg-max decision of the step is within
    rounding noise of flipping.  Built on rank 0, shipped to every rank, returned as per-rank slices on the CPU."""
    import torch.distributed as dist
    total = n * world
    built = []
    if rank == 0:
        p64 = {k: v.double() for k, v in R.init_params(0).items()}
        opt = _make_opt(kind)
        for s in range(steps):
            b = R.robust_batch(total, seed + s, params=p64)
            built.append(b)
            if s + 1 < steps:
                _, _, g = R.loss_and_grads(p64, b[0].double(), b[1], b[2], b[3])
                p64 = opt.step(dict(p64), g)
    out = []
    for s in range(steps):
        if world > 1:
            if rank == 0:
                ts = [t.to(device) for t in built[s]]
            else:
                ts = [torch.empty((total, 1, 28, 28), dtype=torch.float32, device=device),
                      torch.empty((total,), dtype=torch.int64, device=device),
                      torch.empty((total, 20), dtype=torch.uint8, device=device),
                      torch.empty((total, 50), dtype=torch.uint8, device=device)]
            for t in ts:
                dist.broadcast(t, src=0)
            xg, tg, k2g, k1g = [t.cpu() for t in ts]
        else:
            xg, tg, k2g, k1g = built[s]
        out.append([(xg[r * n:(r + 1) * n], tg[r * n:(r + 1) * n], k2g[r * n:(r + 1) * n], k1g[r * n:(r + 1) * n])
                    for r in range(world)])
    return out


def dp_parity_probe(rank, world, device, n=1024, steps=3, seed=9100, trainer_path=True, single_gpu=True,
                    kinds=("sgd", "adam")):
    """See the module docstring.  Every rank must call it; rank 0 returns the report dict (others None)."""
    import torch.distributed as dist
    from pdt.parallel import FlatDataParallel
    wrap = (lambda m: FlatDataParallel(m, device_ids=[device.index], output_device=device.index)) if world > 1 else None
    report = {"world": world, "per_rank_batch": n, "steps": steps,
              "path": "trainer (fused)" if trainer_path else "plug-in calls",
              "against": "oracle W-rank restatement of the reference step", "ok": True, "max_err": 0.0,
              "counts_exact": True, "ranks_bit_identical": True} if rank == 0 else None
    for kind in kinds:
        batches = _robust_batches(n, world, steps, seed, kind, rank, device)
        # the fused tail (reduce + all-reduce + update in one kernel) exists for Adam; SGD takes the general calls
        model, net, grads, tail, reduced = _product_run([b[rank] for b in batches], device, wrap,
                                                        trainer_path and kind == "adam", kind)
        if world > 1 and net.p2p is not None:
            net.p2p.check()
        identical = True
        if world > 1:
            mine = model.flat.data[:21840].clone()
            ref0 = mine.clone()
            dist.broadcast(ref0, src=0)
            flag = torch.tensor([0 if torch.equal(mine, ref0) else 1], device=device)
            dist.all_reduce(flag)
            identical = int(flag.item()) == 0
        if rank == 0:
            params0 = R.init_params(0)
            f32, l32, g32, counts = _oracle_run(params0, batches, world, torch.float32, kind)
            f64, _, g64, _ = _oracle_run(params0, batches, world, torch.float64, kind)
            ours_w = _split(model.flat.data[:21840])
            ours_counts = (int(tail[1]), int(tail[2]), int(tail[3]))
            rep = _compare(kind, ours_w, grads, reduced, ours_counts, f32, g32, l32, counts + (n * world,), f64, g64)
            rep["ranks_bit_identical"] = bool(identical)
            rep["ok"] = bool(rep["ok"] and identical)
            if single_gpu and world > 1:
                glob = [tuple(torch.cat([b[r][i] for r in range(world)]) for i in range(4)) for b in batches]
                m1, _, g1, st1, l1 = _product_run(glob, device, None, False, kind)
                one_w = _split(m1.flat.data[:21840])
                one_g = [_split(g) for g in g1]
                sg = _compare(kind, ours_w, grads, reduced, ours_counts, one_w, one_g, l1,
                              (int(st1[1]), int(st1[2]), int(st1[3])), f64, g64)
                rep["vs_single_gpu_on_global_batch"] = sg
                rep["ok"] = bool(rep["ok"] and sg["ok"])
            report[kind] = rep
            report["ok"] = bool(report["ok"] and rep["ok"])
            report["counts_exact"] = bool(report["counts_exact"] and rep["counts_exact"])
            report["ranks_bit_identical"] = bool(report["ranks_bit_identical"] and identical)
            strict = [rep["loss_max_err"], rep["grad_max_err"]] + ([rep["weights_max_err"]] if kind == "sgd" else [])
            report["max_err"] = max([report["max_err"]] + strict)
        if world > 1:
            dist.barrier()
    return report
