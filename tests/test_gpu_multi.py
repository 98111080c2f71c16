"""N > 1 on real GPUs (skipped on a 1-GPU box): the sharded evaluator under NCCL must reproduce the single-GPU counts
bit for bit — per-rank slices, micro-batch splits, parity-mode aux slicing, the async all-reduce ring, weighted
re-balancing — and two contexts on two devices of ONE process must both work (the shared-memory opt-in is per device)."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _need_two():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on the box (gpurun --gpus 2)")


def _model(seed=42):
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.UNet import UNet
    torch.manual_seed(seed)
    m = UNet().eval()
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    m.load_state_dict(sd)
    return m


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from image_segmentation_b200 import synthetic
        from image_segmentation_b200.engine import Perturbation as P
        from image_segmentation_b200.evaluator import RobustnessEvaluator
        from image_segmentation_b200.scripts.robustness_evaluation import parity_spec
        import random
        m = _model().to(dev)
        imgs, masks = synthetic.make_dataset(37, 64, 128, seed=13)   # 5 reference batches of 8, the last one ragged
        ev = RobustnessEvaluator(m, imgs, masks, 8, dev, rank=rank, world_size=world, max_microbatch=6)
        one = RobustnessEvaluator(m, imgs, masks, 8, dev, rank=0, world_size=1, max_microbatch=148)
        out = {}
        random.seed(3); torch.manual_seed(3)
        specs = [("none", None), ("lut", P.u8_contrast(0.6)), ("blur", P.u8_blur(3)),
                 ("noise", parity_spec("gaussian_noise", 10, 37, 64, 128)),          # aux for the WHOLE set: local_aux slices
                 ("occlusion", parity_spec("occlusion", 25, 37, 64, 128)),
                 ("salt", parity_spec("salt_pepper_noise", 0.06, 37, 64, 128))]
        ok = True
        for name, spec in specs:
            cm, dice = ev.run_level(spec, want_dice=True)
            cm1, dice1 = one.run_level(spec, want_dice=True, reduce=False)
            same = torch.equal(cm, cm1) and torch.allclose(dice, dice1, rtol=1e-5, atol=1e-3)  # fp32 partial sums regroup
            ok = ok and same and int(cm.sum()) == 37 * 64 * 128
            out[name] = bool(same)
        # the async ring: several levels in flight, consumed late, in any order
        want = [one.run_level(s, reduce=False)[0].clone() for _, s in specs[:3]]
        pend = [ev.submit_level(s) for _, s in specs[:3]]
        got = [p.wait()[0].clone() for p in reversed(pend)][::-1]
        out["async_ring"] = all(torch.equal(a, b) for a, b in zip(got, want))
        # weighted re-balancing: rank 1 pretends to be 3x faster; the counts must not change
        rates = ev.rebalance(imgs, masks, 1.0 if rank == 0 else 3.0)
        out["rebalance_rows"] = (ev.b0, ev.b1)
        cm, _ = ev.run_level(specs[1][1])
        out["rebalanced"] = bool(torch.equal(cm, one.run_level(specs[1][1], reduce=False)[0]))
        ok = ok and out["async_ring"] and out["rebalanced"]
        q.put((rank, ok, out))
    finally:
        dist.destroy_process_group()


def test_sharded_evaluator_matches_single_gpu_under_nccl():
    _need_two()
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
    for rank, ok, out in sorted(res):
        print(f"rank {rank}: {out}")
        assert ok, (rank, out)
    rows = dict((r, o["rebalance_rows"]) for r, _, o in res)
    assert rows[0] == (0, 1) and rows[1] == (1, 5)   # 5 batches split 1 : 3 -> edges round(1.25) = 1


def test_two_contexts_on_two_devices_of_one_process():
    """cudaFuncSetAttribute(MaxDynamicSharedMemorySize) belongs to a (function, device) pair: the second context of a
    process, on another GPU, must make its own opt-in (a per-process flag would leave its kernels at the 48 KB
    default and every launch would fail)."""
    _need_two()
    from image_segmentation_b200 import synthetic
    m = _model()
    imgs, masks = synthetic.make_dataset(5, 64, 64, seed=2)
    outs = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            eng = m.to(dev).engine(dev, 4)
            r = eng.eval_batches(imgs.to(dev), masks.to(dev), None, 4, want_logits=True)
            torch.cuda.synchronize(dev)
            outs.append((r["cm"].cpu(), r["logits"].cpu()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert int(outs[0][0].sum()) == 5 * 64 * 64
