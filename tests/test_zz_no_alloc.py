"""`ub_blur_images_static` (include/unet_b200.h): the blur families through the C-ABI with the library's file-scope
static scratch planes instead of per-call stream-ordered allocations.  The entry point has no counterpart in the
reference; its yardstick is the existing entry point `ub_perturb_images`, which the parity tests pin bit-exactly to the
oracle — so the bar here is bit-equality too (the tolerance the blur tests use), on two calls in a row that share the
planes, with a different image set, resolution and pass count in between so stale plane contents would show.

Collected last on purpose.  Passes alone in a fresh process and after the other tests.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng(dev):
    from image_segmentation_b200.engine import Engine
    e = Engine(dev, "unet", 3, 3, 4)    # perturbation entry points need no weights
    yield e, dev
    e.close()


def _images(n, h, w, seed, dev):
    g = torch.Generator().manual_seed(seed)
    u8 = torch.randint(0, 256, (n, h, w, 3), generator=g, dtype=torch.uint8)
    xf = (u8.permute(0, 3, 1, 2).float() / 255.0).contiguous()
    return u8.to(dev), xf.to(dev)


def _check_pair(e, dev, n, h, w, passes, seed):
    from image_segmentation_b200.engine import Perturbation as P
    u8, xf = _images(n, h, w, seed, dev)
    out = torch.full((n, 3, h, w), float("nan"), device=dev)       # allocated by the CALLER, before the call
    for src, spec in ((u8, P.u8_blur(passes)), (xf, P.f_blur(passes)), (xf, P.u8_blur(passes))):
        want = e.perturb(src, spec)                                 # existing entry point: ub_perturb_images
        out.fill_(float("nan"))
        got = e.blur_static(src, spec, out)                         # new entry point, static planes
        torch.cuda.synchronize(dev)
        assert torch.equal(got, want), (n, h, w, passes, src.dtype)


def test_static_blur_equals_perturb_images_twice_in_a_row(eng):
    e, dev = eng
    # call 1 and call 2 back to back on the same planes; then a smaller, ragged shape with other pass counts (0, 1 and
    # beyond the fused kernel's limit) and the first shape again: the planes' old contents must never leak
    _check_pair(e, dev, 5, 256, 256, 3, seed=1)
    _check_pair(e, dev, 5, 256, 256, 3, seed=2)
    for passes in (0, 1, 12, 13, 20):
        _check_pair(e, dev, 2, 100, 72, passes, seed=10 + passes)
    _check_pair(e, dev, 1, 512, 512, 5, seed=3)                     # exactly at the limit's scale (786,432 elements)
    _check_pair(e, dev, 5, 256, 256, 3, seed=1)


def test_static_blur_performs_no_device_allocation(eng):
    """Neither the driver's free-memory figure nor torch's allocator moves across calls (the static planes were mapped
    when the library was loaded)."""
    from image_segmentation_b200.engine import Perturbation as P
    e, dev = eng
    u8, _ = _images(4, 256, 256, 5, dev)
    out = torch.empty((4, 3, 256, 256), device=dev)
    e.blur_static(u8, P.u8_blur(20), out)                           # set-up call: symbol address looked up once
    torch.cuda.synchronize(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    reserved0 = torch.cuda.memory_reserved(dev)
    for _ in range(3):
        e.blur_static(u8, P.u8_blur(20), out)
    torch.cuda.synchronize(dev)
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0 and torch.cuda.memory_reserved(dev) == reserved0   # nothing new was taken from the device


def test_static_blur_rejects_sizes_beyond_the_limit_and_other_families(eng):
    from image_segmentation_b200 import _lib as L
    from image_segmentation_b200.engine import Perturbation as P
    e, dev = eng
    u8, _ = _images(6, 256, 256, 6, dev)                            # 6*3*256*256 = 1,179,648 > 1,048,576
    out = torch.empty((6, 3, 256, 256), device=dev)
    with pytest.raises(RuntimeError, match="UB_STATIC_BLUR_MAX_ELEMS"):
        e.blur_static(u8, P.u8_blur(2), out)
    with pytest.raises(RuntimeError, match="blur families only"):
        e.blur_static(u8[:1].contiguous(), P.u8_contrast(1.1), out[:1].contiguous())
    assert L.UB_STATIC_BLUR_MAX_ELEMS == 1 << 20
    # the context is still usable after the refusals
    _check_pair(e, dev, 1, 64, 64, 2, seed=9)
