"""CPU-only checks of the host side: C-ABI library loads and exports what include/unet_b200.h declares,
LUT construction, metric formulas, sharding arithmetic, drop-in module surface, loud failure on CPU."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import metrics_oracle as mo
from oracle import perturb_oracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build()
    from image_segmentation_b200 import _lib
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "unet_b200.h")).read()
    declared = set(re.findall(r"\b(ub_[a-z0-9_]+)\s*\(", header))
    assert {"ub_create", "ub_eval_batches", "ub_forward", "ub_load_weights"} <= declared
    from image_segmentation_b200 import _lib
    assert declared == set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ub_abi_version() == 1


def test_create_fails_loudly_without_gpu(lib):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = ctypes.c_void_p()
    rc = lib.ub_create(ctypes.byref(h), 0, 0, 3, 3, 8)
    assert rc < 0 and not h.value and len(lib.ub_last_error(None)) > 0


@pytest.mark.parametrize("kind,param,increase", [(0, 1.25, None), (0, 0.1, None), (0, 1.01, None), (1, 45, True),
                                                 (2, 45, False), (1, 0, True), (2, 5, False)])
def test_lut_matches_oracle(lib, kind, param, increase):
    from image_segmentation_b200 import _lib
    p = _lib.ub_perturb()
    assert lib.ub_make_lut(ctypes.byref(p), kind, float(param)) == 0
    want = po.contrast_lut(param) if kind == 0 else po.brightness_lut(param, increase)
    assert np.array_equal(np.array(list(p.lut), dtype=np.uint8), want)
    assert p.kind == _lib.UB_P_U8_LUT


def test_all_script_luts(lib):
    from image_segmentation_b200 import _lib
    from oracle.make_golden import SCRIPT_LEVELS
    for fam in ("contrast_increase", "contrast_decrease"):
        for s in SCRIPT_LEVELS[fam][1]:
            p = _lib.ub_perturb()
            lib.ub_make_lut(ctypes.byref(p), 0, float(s))
            assert np.array_equal(np.array(list(p.lut), dtype=np.uint8), po.contrast_lut(s)), s


def test_metric_formulas_match_oracle():
    from image_segmentation_b200 import metrics
    rng = np.random.default_rng(0)
    cms = rng.integers(0, 70000, (64, 3, 3))
    cms[5, 1, :] = 0
    cms[6, :, :] = 0
    cms[6, 0, 0] = 10
    t = torch.from_numpy(cms)
    iou = metrics.iou_from_cm(t).numpy()
    pa = metrics.pixel_accuracy_from_cm(t).numpy()
    for i in range(64):
        assert iou[i] == mo.iou_from_cm(cms[i])
        assert pa[i] == mo.pixel_accuracy_from_cm(cms[i])
        assert metrics.wrapper_dice(torch.tensor(iou[i])).numpy() == mo.wrapper_dice_from_iou(iou[i])
    sums = torch.rand((7, 3, 3), dtype=torch.float64) * 100
    sums[2, 1, 2] = 0
    d = metrics.dice_from_sums(sums)
    for i in range(7):
        assert abs(float(d[i]) - mo.dice_from_sums(sums[i].numpy())) < 1e-12


@pytest.mark.parametrize("nb,world", [(460, 8), (368, 8), (7, 4), (3, 8), (1, 2), (460, 1)])
def test_batch_range_partitions(nb, world):
    from image_segmentation_b200.evaluator import batch_range
    spans = [batch_range(nb, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == nb
    for a, b in zip(spans, spans[1:]):
        assert a[1] == b[0]
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1


def test_dropin_surface_and_cpu_refusal():
    from image_segmentation_b200.models import UNet as U
    from image_segmentation_b200.models import losses, processing_blocks as pb
    m = U.UNet().eval()
    assert len(m.state_dict()) == 124 and sum(p.numel() for p in m.parameters()) == 7755907
    assert sum(p.numel() for p in U.LargeUNet().parameters()) == 31096451
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 3, 64, 64))  # no CPU fallback
    with pytest.raises(RuntimeError):
        pb.ContrastChange(1.1)(torch.zeros(1, 3, 8, 8))
    with pytest.raises(RuntimeError):
        losses.IoU()(torch.zeros(1, 3, 8, 8), torch.zeros(1, 8, 8, dtype=torch.long))
    from image_segmentation_b200.customDatasets.datasets import CustomImageDataset
    with pytest.raises(ValueError):
        CustomImageDataset(split="bogus")
    ds = CustomImageDataset(split="test", n=2, h=32, w=32)
    x, t = ds[0]
    assert x.dtype == torch.float32 and x.shape == (3, 32, 32) and t.dtype == torch.int64
    assert torch.equal((x * 255).round() / 255.0, x)


def test_rng_replay_helpers_follow_reference_order():
    """Host draws of the drop-in wrappers == oracle replay (same `random` stream consumption)."""
    import random
    from image_segmentation_b200.customDatasets import perturbations as P
    random.seed(42)
    a = [P.OcclusionIncrease(None, 15).draw_corner(40, 56) for _ in range(5)]
    random.seed(42)
    b = [po.replay_occlusion_corner(15, 40, 56) for _ in range(5)]
    assert a == b
    random.seed(7)
    c = P.SaltPepperNoise(None, 0.1).draw_code(24, 24)
    random.seed(7)
    d = po.replay_salt_pepper_code(0.1, 24, 24)
    assert np.array_equal(c, d)


def test_weighted_batch_range_partitions_every_batch_once():
    """evaluator.batch_range with per-rank weights (re-balancing by measured rate): contiguous, complete, monotone,
    proportional within one batch, and identical to the balanced split for equal weights up to rounding."""
    from image_segmentation_b200.evaluator import batch_range
    for n, w in [(460, [1.0] * 8), (460, [1, 1, 1, 1, 1, 1, 1, 0.9]), (5, [1.0, 3.0]), (3, [1.0] * 8), (0, [1.0, 2.0]),
                 (368, [25.1, 24.2, 25.9, 23.8])]:
        world = len(w)
        edges = [batch_range(n, r, world, w) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        for a, b in zip(edges, edges[1:]):
            assert a[1] == b[0] and a[0] <= a[1]
        for r, (a, b) in enumerate(edges):
            assert abs((b - a) - n * w[r] / sum(w)) <= 1.0
    assert [batch_range(460, r, 8) for r in range(8)][0] == (0, 58)
    import pytest
    with pytest.raises(ValueError):
        batch_range(10, 0, 2, [1.0, 0.0])


@pytest.mark.parametrize("n_in,n_out", [(256, 224), (224, 224), (200, 224), (300, 224), (512, 224), (100, 224), (1000, 224),
                                        (97, 64), (131, 160)])
def test_clip_resize_taps_reproduce_antialiased_bilinear(lib, n_in, n_out):
    """`ub_clip_resize_taps` (host code; the tables `clip_patch_kernel` resamples with): applied separably in numpy they
    reproduce torchvision's Resize on a tensor = F.interpolate(bilinear, antialias=True, align_corners=False)
    (CustomClipPreprocessor, models/processing_blocks.py:156-170)."""
    from image_segmentation_b200 import _lib
    T = _lib.UB_CLIP_MAX_TAPS
    start = np.zeros(n_out, dtype=np.int32)
    count = np.zeros(n_out, dtype=np.int32)
    w = np.zeros((n_out, T), dtype=np.float32)
    assert lib.ub_clip_resize_taps(n_in, n_out, start.ctypes.data, count.ctypes.data, w.ctypes.data) == 0
    assert (count >= 1).all() and (count <= T).all() and (start >= 0).all() and (start + count <= n_in).all()
    assert np.allclose(w.sum(1), 1.0, atol=1e-6) and (w >= 0).all()
    m = np.zeros((n_out, n_in), dtype=np.float64)            # the resampling matrix of one axis
    for i in range(n_out):
        m[i, start[i]:start[i] + count[i]] = w[i, :count[i]]
    g = torch.Generator().manual_seed(n_in * 1000 + n_out)
    x = torch.rand((1, 2, n_in, 57), generator=g)
    want = torch.nn.functional.interpolate(x, size=(n_out, 57), mode="bilinear", antialias=True, align_corners=False)
    got = np.einsum("oi,bciw->bcow", m, x.numpy().astype(np.float64))
    assert np.abs(got - want.numpy()).max() < 2e-6
    want_w = torch.nn.functional.interpolate(x.transpose(2, 3), size=(57, n_out), mode="bilinear", antialias=True, align_corners=False)
    assert np.abs(np.einsum("oi,bcwi->bcwo", m, x.transpose(2, 3).numpy().astype(np.float64)) - want_w.numpy()).max() < 2e-6


def test_clip_resize_taps_refuse_more_than_the_tap_limit(lib):
    from image_segmentation_b200 import _lib
    n_out = 224
    start = np.zeros(n_out, dtype=np.int32); count = np.zeros(n_out, dtype=np.int32)
    w = np.zeros((n_out, _lib.UB_CLIP_MAX_TAPS), dtype=np.float32)
    assert lib.ub_clip_resize_taps(1400, n_out, start.ctypes.data, count.ctypes.data, w.ctypes.data) == -3   # UB_ERR_UNSUPPORTED
    assert b"taps per axis" in lib.ub_last_error(None)
    assert lib.ub_clip_resize_taps(0, n_out, start.ctypes.data, count.ctypes.data, w.ctypes.data) == -1


def test_build_recipe_depends_on_every_source_file():
    """The one translation unit includes every file of csrc/: the rebuild check must watch all of them (a fixed list once
    missed the newer headers and left a stale library in place)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ub_build", os.path.join(ROOT, "image-segmentation_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    deps = set(b.deps())
    csrc = os.path.join(ROOT, "image-segmentation_b200", "csrc")
    for f in os.listdir(csrc):
        if f.endswith((".cu", ".cuh", ".inl")):
            assert f in deps, f
    assert any(d.endswith("unet_b200.h") for d in deps)
    src = open(os.path.join(csrc, "ub_api.cu")).read()
    for inc in re.findall(r'#include "([^"]+)"', src):
        assert os.path.basename(inc) in {os.path.basename(d) for d in deps}, inc


def _convt_resample_weights(o, n):
    """numpy fp32 restatement of csrc/ub_conv_tc.cuh: convt_resample_weights (the arithmetic of resample_bilinear_kernel,
    i.e. of F.interpolate(align_corners=True) from 2n to n samples); returns (w_even, w_odd, taps_are_children_of_o)."""
    f32 = np.float32
    ni = 2 * n
    s = f32(ni - 1) / f32(n - 1) if n > 1 else f32(0)
    f = f32(s * f32(o))
    i0 = int(f)
    i1 = i0 + (1 if i0 < ni - 1 else 0)
    l = f32(f - f32(i0))
    h = f32(f32(1) - l)
    w0 = (f32(0) if i0 & 1 else h) + (f32(0) if i1 & 1 else l)
    w1 = (h if i0 & 1 else f32(0)) + (l if i1 & 1 else f32(0))
    return w0, w1, (i0 >> 1) == o and (i1 >> 1) == o


def test_convtranspose_then_resize_reads_only_the_children_of_one_pixel():
    """The premise of the EPI_CONVT_BLEND epilogue: output sample o of the 2n -> n align_corners resize reads the two
    children of ConvTranspose input sample o and nothing else.  True for every n up to 2048; beyond that fp32 rounding of
    o * (2n-1)/(n-1) can land sample n-2 exactly on an odd source index (first at n = 2082), which is why build_plan
    evaluates the predicate per size (convt_resize_blends) and keeps the two-kernel path when it fails."""
    for n in range(1, 2049):
        for o in (range(n) if n <= 96 else sorted({0, 1, n // 2, n - 3, n - 2, n - 1})):
            w0, w1, ok = _convt_resample_weights(o, n)
            assert ok, (n, o)
            assert abs(float(w0 + w1) - 1.0) < 1e-6
    assert not _convt_resample_weights(2080, 2082)[2]


@pytest.mark.parametrize("h,w", [(1, 1), (2, 3), (7, 5), (32, 32), (33, 16)])
def test_convtranspose_plus_resize_is_a_blend_of_four_parity_products(h, w):
    """models/processing_blocks.py:106-107 on the first decoder stage: ConvTranspose2d(k2,s2) then F.interpolate(size=skip,
    bilinear, align_corners=True) with skip = the ConvTranspose INPUT size.  The library evaluates the pair as
    sum_q c_q(y,x) * (W[:,:,a,b]^T x[y,x]) + bias with q = 2a+b (column 4*co + q of the re-packed weights): same result."""
    g = torch.Generator().manual_seed(5)
    cin, cout = 6, 4
    x = torch.randn(2, cin, h, w, generator=g)
    wt = torch.randn(cin, cout, 2, 2, generator=g)
    bias = torch.randn(cout, generator=g)
    ref = torch.nn.functional.interpolate(torch.nn.functional.conv_transpose2d(x, wt, bias, stride=2), size=(h, w),
                                          mode="bilinear", align_corners=True)
    wy = np.array([_convt_resample_weights(o, h)[:2] for o in range(h)], dtype=np.float32)  # [h][2]
    wx = np.array([_convt_resample_weights(o, w)[:2] for o in range(w)], dtype=np.float32)
    out = torch.zeros_like(ref)
    for a in range(2):
        for b in range(2):
            v = torch.einsum("nihw,io->nohw", x, wt[:, :, a, b])
            c = torch.from_numpy(np.outer(wy[:, a], wx[:, b]))
            out += c[None, None] * v
    out += bias[None, :, None, None]
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)


def _simulate_row_window(W, n_tiles, spare):
    """Bookkeeping model of the sliding TMEM window of csrc/ub_conv_rs2.cuh (the issuing thread's loop + rs_issue_chunk's
    CASE table + the epilogue's reads), one k-step per tile.  A block holds a set of (input row, ky) contributions; a
    non-accumulating MMA replaces the set.  Returns (what the epilogue read per output row, MMAs issued per tile)."""
    blocks = {b: {"garbage"} for b in range(W + 2)}
    read, mmas = {}, []

    def mma(first_block, n_blocks, tau, ky_of_first, accumulate):
        for i in range(n_blocks):
            contrib = (tau, ky_of_first - i)          # B rows are ordered ky = 2, 1, 0 = output rows tau-1, tau, tau+1
            b = first_block + i
            blocks[b] = (blocks[b] | {contrib}) if accumulate else {contrib}

    blk = W - 1                                        # home of row tau-1
    for tau in range(n_tiles):
        issued = 0
        if spare:                                      # linear window: blocks blk..blk+2 physically contiguous (W, W+1: overflow)
            if blk == 0:
                mma(0, 3, tau, 2, False); issued += 1                       # CASE 3: all three are first touches
            else:
                mma(blk, 2, tau, 2, True); mma(blk + 2, 1, tau, 0, False); issued += 2   # CASE 0, first k-step
        elif blk <= W - 3:
            mma(blk, 2, tau, 2, True); mma(blk + 2, 1, tau, 0, False); issued += 2
        elif blk == W - 2:                             # CASE 1: row tau+1 wraps to block 0
            mma(blk, 2, tau, 2, True); mma(0, 1, tau, 0, False); issued += 2
        else:                                          # CASE 2: rows tau, tau+1 wrap to blocks 0, 1
            mma(blk, 1, tau, 2, True); mma(0, 1, tau, 1, True); mma(1, 1, tau, 0, False); issued += 3
        mmas.append(issued)
        if tau >= 1:                                   # row tau-1 is complete: its epilogue reads home (+ overflow)
            sigma = tau - 1
            home = sigma % W
            got = set(blocks[home])
            if spare and home < 2:
                got |= blocks[W + home]
            read[sigma] = got
        blk = (blk + 1) % W
    return read, mmas


@pytest.mark.parametrize("W,spare", [(4, False), (6, False), (8, False), (6, True), (4, True)])
def test_sliding_tmem_window_gives_every_row_its_three_contributions(W, spare):
    """Every output row sigma >= 1 must be read as exactly {(sigma-1, ky=0), (sigma, ky=1), (sigma+1, ky=2)} — in the circular
    window (split MMAs where the span wraps) and in the linear one with two overflow blocks (Rs2Cfg::SPARE: the first
    contributions of the rows whose home is block 0 / 1 sit in blocks W / W+1 and the epilogue adds them)."""
    read, mmas = _simulate_row_window(W, 5 * W + 3, spare)
    for sigma, got in read.items():
        if sigma == 0:
            continue                                   # the first row of a CTA has no tile -1: masked in the kernel (rho = 0)
        assert got == {(sigma - 1, 0), (sigma, 1), (sigma + 1, 2)}, (sigma, got)
    if spare:                                          # never more than the first k-step's split; one MMA when all blocks are fresh
        assert set(mmas) == {1, 2} and mmas.count(1) == len([t for t in range(len(mmas)) if (W - 1 + t) % W == 0])
    else:
        assert 3 in mmas                               # the wrap costs extra MMAs on every lap
