"""ctypes binding of libisg.so (include/isg.h).  Fails loudly when the library is missing:
there is no Python/CPU implementation of the hot path to fall back to."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# ISG_LIB: experiments only (an A/B build made by scripts/build_variant.sh); still a libisg build, never a fallback
LIB_PATH = os.environ.get("ISG_LIB") or os.path.join(_HERE, "libisg.so")
_lib = None

MODE_FP32, MODE_BF16, MODE_BF16X3 = 0, 1, 2

# every symbol include/isg.h declares: name -> (restype, argtypes)
_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float
SIGNATURES = {
    "isg_last_error": (C.c_char_p, []),
    "isg_version": (_i, []),
    "isg_state_count": (_i, [_i]),
    "isg_state_name": (C.c_char_p, [_i, _i]),
    "isg_state_numel": (_i64, [_i, _i]),
    "isg_state_is_int64": (_i, [_i, _i]),
    "isg_plan_create": (_i, [_i, _i, _i, _i, _i, _i, C.POINTER(_vp)]),
    "isg_plan_dry_run": (_i, [_i, _i, _i, _i, _i]),
    "isg_plan_destroy": (_i, [_vp]),
    "isg_plan_workspace_bytes": (C.c_size_t, [_vp]),
    "isg_plan_load_weights": (_i, [_vp, C.POINTER(_vp), C.POINTER(_i64), _i]),
    "isg_forward": (_i, [_vp, _vp, _i, _vp, _vp]),
    "isg_forward_u8": (_i, [_vp, _vp, _i, C.POINTER(_f), C.POINTER(_f), _vp, _vp]),
    "isg_plan_launches_per_forward": (_i, [_vp]),
    "isg_plan_macs_per_image": (_i64, [_vp]),
    "isg_plan_tap_count": (_i, [_vp]),
    "isg_plan_tap_name": (C.c_char_p, [_vp, _i]),
    "isg_plan_tap_shape": (_i, [_vp, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)]),
    "isg_plan_tap_read": (_i, [_vp, _i, _i, _vp]),
    "isg_decode": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _i, _i, _f, _vp, _vp, _vp, _vp, _vp]),
    "isg_decode_tile_rows": (_i, []),
    "isg_decode_workspace_bytes": (_i64, [_i]),
    "isg_decode_bind_workspace": (_i, [_vp, _i]),
    "isg_decode_tiled": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _i, _f, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "isg_idmap": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _f, _vp, _vp]),
    "isg_pack_mask_bits": (_i, [_vp, _i64, _vp, _vp]),
    "isg_mask_iou_counts": (_i, [_vp, _vp, _i64, _vp, _vp]),
    "isg_mask_iou_counts_thr": (_i, [_vp, _vp, _i64, _i, _i, _vp, _vp]),
    "isg_keypoint_heatmaps": (_i, [_vp, _i, _i, _i, _i, _f, _f, _vp, _i64, _vp]),
    "isg_crop_resize_u8": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "isg_adam_step": (_i, [_vp, _vp, _vp, _vp, _i64, _f, _f, _f, _f, _f, _i, _vp]),
    "isg_bn_max_split": (_i, []),
    "isg_bn_train_forward": (_i, [_vp, _i, _i, _i, _i64, _vp, _vp, _vp, _vp, _vp, _f, _f, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "isg_bn_train_backward": (_i, [_vp, _vp, _i, _i, _i, _i64, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "isg_dwconv_forward": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "isg_dwconv_wgrad": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "isg_dwconv_wgrad_workspace_bytes": (_i64, [_i]),
    "isg_convt_k2s_forward": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "isg_bce_workspace_bytes": (_i64, []),
    "isg_bce_sigmoid": (_i, [_vp, _i, _vp, _i64, _vp, _vp, _vp, _vp]),
    "isg_bce_sigmoid_static_max_n": (_i64, []),
    "isg_bce_sigmoid_static": (_i, [_vp, _i, _vp, _i64, _vp, _vp, _vp]),
}


class IsgError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise IsgError(
                f"{LIB_PATH} not found: build it with `python -m instancesegmentation_b200._build` "
                "(nvcc, sm_100a).  There is no fallback implementation.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


def check(status):
    if status != 0:
        raise IsgError(f"libisg error {status}: {lib().isg_last_error().decode()}")
