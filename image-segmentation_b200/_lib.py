"""ctypes binding of csrc/libunetb200.so — the C ABI declared in include/unet_b200.h.

Fails loudly: a missing library or a missing symbol is an ImportError/RuntimeError, never a fallback.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libunetb200.so")

c_void_p, c_int, c_char_p, c_int64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int64

(UB_ARCH_UNET, UB_ARCH_LARGE_UNET, UB_ARCH_CLIP_UNET, UB_ARCH_CLIP_UNET_PROMPT,
 UB_ARCH_AUTOENCODER, UB_ARCH_ENCODER, UB_ARCH_SKIP_DECODER, UB_ARCH_CLIP_VIT) = 0, 1, 2, 3, 4, 5, 6, 7
UB_IMG_F32_NCHW, UB_IMG_U8_NHWC = 0, 1
UB_TGT_I64, UB_TGT_U8 = 0, 1
(UB_P_NONE, UB_P_U8_NOISE, UB_P_U8_BLUR, UB_P_U8_LUT, UB_P_U8_OCCLUSION, UB_P_U8_SALT_PEPPER, UB_P_F_NOISE,
 UB_P_F_BLUR, UB_P_F_CONTRAST, UB_P_F_BRIGHTNESS, UB_P_F_OCCLUSION, UB_P_F_SALT_PEPPER) = range(12)
UB_LUT_CONTRAST, UB_LUT_BRIGHTNESS_INCREASE, UB_LUT_BRIGHTNESS_DECREASE = 0, 1, 2
UB_CLIP_MAX_TAPS = 12  # ub_clip_resize_taps: weights per output coordinate
UB_STATIC_BLUR_MAX_ELEMS = 1 << 20  # ub_blur_images_static: n*3*h*w limit of the static scratch planes


class ub_perturb(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int32), ("iparam", ctypes.c_int32), ("fparam", ctypes.c_float),
                ("fparam2", ctypes.c_float), ("seed", ctypes.c_uint64), ("lut", ctypes.c_uint8 * 256)]


# name -> (restype, argtypes); every symbol include/unet_b200.h declares
SIGNATURES = {
    "ub_abi_version": (c_int, []),
    "ub_make_lut": (c_int, [ctypes.POINTER(ub_perturb), c_int, ctypes.c_double]),
    "ub_create": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int]),
    "ub_destroy": (None, [c_void_p]),
    "ub_last_error": (c_char_p, [c_void_p]),
    "ub_load_weights": (c_int, [c_void_p, c_int, ctypes.POINTER(c_char_p), ctypes.POINTER(c_void_p),
                                ctypes.POINTER(ctypes.POINTER(c_int64)), ctypes.POINTER(c_int)]),
    "ub_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ub_eval_batches": (c_int, [c_void_p, ctypes.POINTER(ub_perturb), c_void_p, c_int, c_void_p, c_int, c_void_p,
                                c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p]),
    "ub_eval_batches_clip": (c_int, [c_void_p, ctypes.POINTER(ub_perturb), c_void_p, c_int, c_void_p, c_int, c_void_p,
                                     c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p]),
    "ub_forward_prompt": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ub_encoder_forward": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, ctypes.POINTER(c_void_p), c_int, c_void_p]),
    "ub_skip_decoder_forward": (c_int, [c_void_p, ctypes.POINTER(c_void_p), c_int, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ub_perturb_images": (c_int, [c_void_p, ctypes.POINTER(ub_perturb), c_void_p, c_int, c_void_p, c_int, c_int,
                                  c_int, c_void_p, c_void_p]),
    "ub_clip_image_features": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ub_clip_resize_taps": (c_int, [c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ub_clip_dims": (c_int, [c_void_p, ctypes.POINTER(c_int), ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "ub_debug_clip_preprocess": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ub_blur_images_static": (c_int, [c_void_p, ctypes.POINTER(ub_perturb), c_void_p, c_int, c_int, c_int, c_int,
                                      c_void_p, c_void_p]),
    "ub_decode_masks": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "ub_pack_dataset": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ub_metrics": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                           c_void_p, c_void_p]),
    "ub_debug_conv3x3": (c_int, [c_void_p, c_char_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "ub_debug_convt2x2": (c_int, [c_void_p, c_char_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ub_debug_resample": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "ub_launch_count": (c_int64, [c_void_p]),
    "ub_profile_begin": (c_int, [c_void_p]),
    "ub_profile_end": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_int64), c_char_p, c_int]),
}

_lib = None


def load():
    """Load the shared library and bind every declared entry point (raises if anything is missing)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU/PyTorch fallback for this path)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.ub_abi_version() != 1:
        raise ImportError("libunetb200.so ABI version mismatch")
    _lib = lib
    return lib
