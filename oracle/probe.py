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

Tolerances (north_star): metric counts bit-exact; loss, gradients and weights rtol 1e-5 with
atol = 1e-5 * max|ref| per tensor.  Adam weights are split by conditioning exactly as
tests/test_gpu_parity.py::check_adam_weights does (an element whose gradient is ~1e-7 in some step
amplifies rounding noise through 1/(sqrt(v)+eps); the reference's own fp32 run is ~2e-5 from the fp64
truth there): well-conditioned elements at rtol 1e-5, the others no further from the fp64 truth than
4x the oracle's fp32 run is (per tensor).
"""
import torch

from . import ref_step as R


def _rel(a, b):
    a, b = a.detach().cpu().double().reshape(-1), b.detach().cpu().double().reshape(-1)
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))


def _oracle_run(params, batches, world, dtype):
    """The reference step restated, keeping everything the comparison needs."""
    p = {k: v.to(dtype) for k, v in params.items()}
    opt = R.Adam(lr=1e-3, weight_decay=0.0, amsgrad=True)
    well = {k: torch.ones_like(v, dtype=torch.bool) for k, v in p.items()}
    losses, grads, top1, top3 = [], None, 0, 0
    for step in batches:
        per_rank, local, top1, top3 = [], [], 0, 0
        for (x, t, k2, k1) in step:
            loss, out, g = R.loss_and_grads(p, x.to(dtype), t, k2, k1)
            per_rank.append(g)
            local.append(loss)
            top1 += R.accuracy_count(out, t)
            top3 += R.top_k_count(out, t, 3)
        grads = R.ddp_average(per_rank)
        for k in well:
            well[k] &= (grads[k].abs() > 1e-6) | (grads[k] == 0)
        losses.append(float(R.reduce_loss(local)))
        p = opt.step(dict(p), grads)
    return p, losses, grads, (top1, top3), well


def _weights_report(ours, ref32, ref64, well):
    """ours/ref32/ref64: dict name -> tensor.  Returns (ok, worst well-conditioned relative error,
    worst ill-conditioned ratio ours64 / (4 ref64 + floor))."""
    ok, worst_well, worst_ill = True, 0.0, 0.0
    for k, _ in R.PARAM_SHAPES:
        o, a, b = ours[k].detach().cpu().double(), ref32[k].double(), ref64[k].double()
        scale = float(a.abs().max())
        m = well[k]
        if m.any():
            e = float((o - a).abs()[m].max()) / scale
            worst_well = max(worst_well, e)
            ok &= e <= 2e-5        # rtol 1e-5 * |ref| + atol 1e-5 * max|ref|
        if (~m).any():
            ours64 = float((o - b).abs()[~m].max())
            ref64e = float((a - b).abs()[~m].max())
            ratio = ours64 / (4 * ref64e + 1e-5 * scale)
            worst_ill = max(worst_ill, ratio)
            ok &= ratio <= 1.0
    return bool(ok), worst_well, worst_ill


def _split(flat):
    out, off = {}, 0
    for k, shape in R.PARAM_SHAPES:
        cnt = 1
        for s in shape:
            cnt *= s
        out[k] = flat[off:off + cnt].view(shape).detach().cpu().clone()
        off += cnt
    return out


def _product_run(batches_for_me, device, wrap, trainer_path):
    """K steps of the product through the plug-in calls.  batches_for_me: list of (x, t, k2, k1).
    Returns (model, wrapped model, last flat gradient, last tail or stats, the reduced loss of every step)."""
    from model.model import MnistModel
    from model.loss import nll_loss
    from pdt import optim as O, ops
    torch.manual_seed(0)
    model = MnistModel().to(device)
    opt = O.Adam(model.parameters(), lr=1e-3, weight_decay=0, amsgrad=True)
    net = wrap(model) if wrap is not None else model
    net.train()
    if trainer_path:
        model.flat.defer_reduce = True
    one = torch.ones((), device=device)
    losses, stats = [], None
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
        if wrap is not None:
            losses.append(float(net.reduced_tail()[0]))
            stats = net.reduced_tail().clone().cpu()
        else:
            losses.append(float(loss))
            st = ops.unpack_stats(loss._pdt_stats)
            stats = torch.tensor([losses[-1], st["top1"], st["topk"], st["n_valid"]], dtype=torch.float64)
    torch.cuda.synchronize()
    return model, net, model.flat.grad[:21840].clone(), stats, losses


def dp_parity_probe(rank, world, device, n=1024, steps=3, seed=9100, trainer_path=True, single_gpu=True):
    """See the module docstring.  Every rank must call it; rank 0 returns the report dict (others None)."""
    import torch.distributed as dist
    from pdt.parallel import FlatDataParallel
    batches = []
    for s in range(steps):
        xg, tg, k2g, k1g = R.synthetic_batch(n * world, seed + s)
        batches.append([(xg[r * n:(r + 1) * n], tg[r * n:(r + 1) * n], k2g[r * n:(r + 1) * n], k1g[r * n:(r + 1) * n])
                        for r in range(world)])
    wrap = (lambda m: FlatDataParallel(m, device_ids=[device.index], output_device=device.index)) if world > 1 else None
    model, net, grad, tail, reduced = _product_run([b[rank] for b in batches], device, wrap, trainer_path)
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
    report = None
    if rank == 0:
        params0 = R.init_params(0)
        f32, losses32, g32, counts, _ = _oracle_run(params0, batches, world, torch.float32)
        f64, _, _, _, well = _oracle_run(params0, batches, world, torch.float64)
        ours_w, ours_g = _split(model.flat.data[:21840]), _split(grad)
        grad_err = max(_rel(ours_g[k], g32[k]) for k, _ in R.PARAM_SHAPES)
        loss_err = max(abs(a - b) / abs(b) for a, b in zip(reduced, losses32))
        w_ok, w_well, w_ill = _weights_report(ours_w, f32, f64, well)
        counts_exact = (int(tail[1]), int(tail[2]), int(tail[3])) == (counts[0], counts[1], n * world)
        report = {
            "world": world, "per_rank_batch": n, "steps": steps, "path": "trainer (fused)" if trainer_path else "plug-in calls",
            "against": "oracle W-rank restatement of the reference step (fp32; fp64 for conditioning)",
            "max_err": max(grad_err, loss_err, w_well), "grad_max_err": grad_err, "loss_max_err": loss_err,
            "weights_well_conditioned_max_err": w_well, "weights_ill_conditioned_ratio": w_ill,
            "counts_exact": bool(counts_exact), "counts": [int(tail[1]), int(tail[2]), int(tail[3])],
            "ranks_bit_identical": bool(identical),
            "ok": bool(grad_err <= 1e-5 and loss_err <= 1e-5 and w_ok and counts_exact and identical),
        }
        if single_gpu and world > 1:
            glob = [(torch.cat([b[r][0] for r in range(world)]), torch.cat([b[r][1] for r in range(world)]),
                     torch.cat([b[r][2] for r in range(world)]), torch.cat([b[r][3] for r in range(world)])) for b in batches]
            m1, _, g1, st1, l1 = _product_run(glob, device, None, False)
            w1 = _split(m1.flat.data[:21840])
            sg_grad = max(_rel(ours_g[k], _split(g1)[k]) for k, _ in R.PARAM_SHAPES)
            sg_loss = max(abs(a - b) / abs(b) for a, b in zip(reduced, l1))
            # weights: the single-GPU product run takes the oracle's place as "ref32"
            sg_ok, sg_well, sg_ill = _weights_report(ours_w, w1, f64, well)
            sg_counts = (int(tail[1]), int(tail[2]), int(tail[3])) == (int(st1[1]), int(st1[2]), int(st1[3]))
            report["vs_single_gpu_on_global_batch"] = {
                "grad_max_err": sg_grad, "loss_max_err": sg_loss, "weights_well_conditioned_max_err": sg_well,
                "weights_ill_conditioned_ratio": sg_ill, "counts_exact": bool(sg_counts),
                "ok": bool(sg_grad <= 1e-5 and sg_loss <= 1e-5 and sg_ok and sg_counts)}
            report["ok"] = bool(report["ok"] and report["vs_single_gpu_on_global_batch"]["ok"])
    if world > 1:
        dist.barrier()
    return report
