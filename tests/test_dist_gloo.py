"""N>1 path on CPU (gloo, world_size 2): shard reference batches over ranks, each rank fills its rows
of the per-batch confusion tensor (here through the oracle, standing in for the GPU engine), one
all-reduce, and every rank must hold the single-process result."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import metrics_oracle as mo


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, ref_batch, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from image_segmentation_b200.evaluator import RobustnessEvaluator, batch_range
    g = torch.Generator().manual_seed(0)
    pred = torch.randint(0, 3, (n, 16, 16), generator=g)
    tgt = torch.randint(0, 3, (n, 16, 16), generator=g)
    nb = -(-n // ref_batch)
    b0, b1 = batch_range(nb, rank, world)
    cm = torch.zeros((nb, 3, 3), dtype=torch.int64)
    for b in range(b0, b1):
        sl = slice(b * ref_batch, min((b + 1) * ref_batch, n))
        cm[b] = torch.from_numpy(mo.confusion_matrix(pred[sl].numpy(), tgt[sl].numpy()))
    dist.all_reduce(cm, op=dist.ReduceOp.SUM)
    full = torch.stack([torch.from_numpy(mo.confusion_matrix(pred[b * ref_batch:(b + 1) * ref_batch].numpy(),
                                                             tgt[b * ref_batch:(b + 1) * ref_batch].numpy()))
                        for b in range(nb)])
    ok = torch.equal(cm, full)
    scores = RobustnessEvaluator.wrapper_scores(cm)
    out.put((rank, ok, scores))
    dist.destroy_process_group()


def test_sharded_counts_equal_single_process():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    n, ref_batch, world = 37, 8, 2  # ragged last batch
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, ref_batch, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res)
    assert res[0][2] == res[1][2]


# ---------------------------------------------------------------------------------------------------------------------
class _StubEngine:
    """Stands in for engine.Engine on the CPU: fills the confusion tensor of the images it is given from a deterministic
    'prediction' (a function of the pixel values and of the perturbation's iparam), through the oracle's counting."""
    arch, max_microbatch = "unet", 148

    def eval_batches(self, images, targets, perturbation, ref_batch, want_dice=False, out_cm=None, out_dice=None, **kw):
        n = images.shape[0]
        shift = 0 if perturbation is None else int(perturbation.struct.iparam)
        pred = ((images[..., 0].long() + shift) // 86).clamp(0, 2)
        for b in range(-(-n // ref_batch)):
            sl = slice(b * ref_batch, min((b + 1) * ref_batch, n))
            out_cm[b] = torch.from_numpy(mo.confusion_matrix(pred[sl].numpy(), targets[sl].numpy()))
            if want_dice:
                out_dice[b] = out_cm[b].double()
        return dict(cm=out_cm, dice=out_dice)


class _StubModel:
    _n_classes = 3

    def engine(self, device, max_microbatch=None):
        return _StubEngine()


def _evaluator_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from image_segmentation_b200.engine import Perturbation
    from image_segmentation_b200.evaluator import RobustnessEvaluator
    g = torch.Generator().manual_seed(1)
    n = 45                                        # 6 reference batches of 8, the last one of 5
    imgs = torch.randint(0, 256, (n, 6, 7, 3), generator=g, dtype=torch.uint8)
    masks = torch.randint(0, 3, (n, 6, 7), generator=g, dtype=torch.uint8)
    ev = RobustnessEvaluator(_StubModel(), imgs, masks, 8, device="cpu", rank=rank, world_size=world)
    one = RobustnessEvaluator(_StubModel(), imgs, masks, 8, device="cpu", rank=0, world_size=1)
    specs = [None] + [Perturbation(iparam=k) for k in (20, 40, 60, 80)]
    want = []
    for s in specs:
        c, d = one.run_level(s, want_dice=True, reduce=False)
        want.append((c.clone(), d.clone()))
    ok = True
    for s, (c, d) in zip(specs, want):           # synchronous levels
        cm, dice = ev.run_level(s, want_dice=True)
        ok = ok and torch.equal(cm, c) and torch.equal(dice, d)
    pend = [ev.submit_level(s, want_dice=True) for s in specs]   # more levels in flight than ring slots (3)
    for p, (c, d) in list(zip(pend, want))[2:]:                  # slots 0 and 1 were recycled: their results are gone
        cm, dice = p.wait()
        ok = ok and torch.equal(cm, c) and torch.equal(dice, d)
    rates = ev.rebalance(imgs, masks, 1.0 if rank == 0 else 2.0)  # rank 1 twice as fast: 2 + 4 batches
    cm, _ = ev.run_level(specs[2])
    ok = ok and torch.equal(cm, want[2][0]) and rates == [1.0, 2.0]
    out.put((rank, ok, (ev.b0, ev.b1)))
    dist.destroy_process_group()


def test_sharded_evaluator_ring_and_rebalance_on_gloo():
    """RobustnessEvaluator itself (sharding by reference batch, run_level, the async submit_level ring, weighted
    re-balancing) over gloo with world_size 2, the engine replaced by a CPU stub: every rank ends with the single-rank
    counts.  (The same with the real engine under NCCL: tests/test_gpu_multi.py.)"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_evaluator_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    assert [r[2] for r in res] == [(0, 2), (2, 6)]
