"""ctypes binding of libvo2w.so (include/vo2w.h).  No fallback: a missing library is an error."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VO2W_LIB") or os.path.join(HERE, "libvo2w.so")   # VO2W_LIB: A/B builds in experiments

OK, E_INVALID, E_CUDA, E_WORKSPACE, E_NODEVICE = 0, -1, -2, -3, -4
DUR_I32, DUR_I64, DUR_F32, DUR_F64 = 0, 1, 2, 3
PREC_FP32, PREC_BF16 = 0, 1


class GenConfig(C.Structure):
    _fields_ = [
        ("num_upsamples", C.c_int32),
        ("upsample_rates", C.c_int32 * 4),
        ("upsample_kernel_sizes", C.c_int32 * 4),
        ("upsample_initial_channel", C.c_int32),
        ("num_kernels", C.c_int32),
        ("resblock_kernel_sizes", C.c_int32 * 3),
        ("resblock_dilation_sizes", (C.c_int32 * 3) * 3),
        ("num_mels", C.c_int32),
    ]


# every symbol include/vo2w.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "vo2w_last_error": (C.c_char_p, []),
    "vo2w_version": (C.c_int, []),
    "vo2w_device_check": (C.c_int, []),
    "vo2w_generator_create": (C.c_int, [C.POINTER(GenConfig), C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)]),
    "vo2w_generator_destroy": (C.c_int, [C.c_void_p]),
    "vo2w_generator_workspace_bytes": (C.c_size_t, [C.c_void_p, C.c_int, C.c_int]),
    "vo2w_generator_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_size_t, C.c_int, C.c_void_p]),
    "vo2w_generator_forward_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_float, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_void_p]),
    "vo2w_generator_forward_host_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.c_float, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.c_void_p]),
    "vo2w_generator_debug_tap": (C.c_longlong, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_size_t]),
    "vo2w_generator_has_tap": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vo2w_debug_layer": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vo2w_debug_pack": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "vo2w_debug_pair": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                  C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                  C.c_void_p]),
    "vo2w_debug_pair_post": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_float, C.c_void_p,
                                       C.c_void_p, C.c_float, C.c_void_p]),
    "vo2w_debug_stream_pack": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "vo2w_generator_last_launches": (C.c_int, [C.c_void_p]),
    "vo2w_generator_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "vo2w_generator_get_timing": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "vo2w_postnet_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "vo2w_postnet_destroy": (C.c_int, [C.c_void_p]),
    "vo2w_postnet_workspace_bytes": (C.c_size_t, [C.c_void_p, C.c_int, C.c_int]),
    "vo2w_postnet_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                       C.c_size_t, C.c_void_p]),
    "vo2w_linear_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_void_p]),
    "vo2w_decoder_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
                                      C.POINTER(C.c_void_p)]),
    "vo2w_decoder_destroy": (C.c_int, [C.c_void_p]),
    "vo2w_decoder_workspace_bytes": (C.c_size_t, [C.c_void_p, C.c_int, C.c_int]),
    "vo2w_decoder_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                       C.c_void_p, C.c_size_t, C.c_void_p]),
    "vo2w_varpred_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "vo2w_varpred_destroy": (C.c_int, [C.c_void_p]),
    "vo2w_varpred_workspace_bytes": (C.c_size_t, [C.c_void_p, C.c_int, C.c_int]),
    "vo2w_varpred_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_size_t, C.c_void_p]),
    "vo2w_varpred_forward_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "vo2w_variance_embed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float,
                                      C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]),
    "vo2w_duration_round": (C.c_int, [C.c_void_p, C.c_float, C.c_void_p, C.c_longlong, C.c_void_p]),
    "vo2w_debug_tf32_conv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "vo2w_debug_tf32_conv_ex": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                          C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                          C.c_void_p]),
    "vo2w_lr_lengths": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vo2w_length_regulate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_void_p,
                                       C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vo2w_mask_from_lengths": (C.c_int, [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libvo2w.so (built in-tree by build.py / __graft_entry__.build()).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m visual_onoma_to_wave_b200.build` "
                "(there is no CPU or PyTorch fallback for this path)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)          # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int, what: str = "libvo2w") -> None:
    """Reference error convention: Python exceptions only."""
    if rc < 0:
        msg = lib().vo2w_last_error()
        raise RuntimeError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
