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
