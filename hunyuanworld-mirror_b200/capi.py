"""ctypes binding of libwm_b200.so (include/wm_b200.h).  There is no fallback: if the library has not been
built (python hunyuanworld-mirror_b200/build.py) importing the symbols raises, and every compute entry point
returns WM_ERR_ARCH on anything that is not a B200."""
from __future__ import annotations

import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "lib", "libwm_b200.so")

PROF_CLASSES = ("layernorm", "gemm_qkv", "attn_local", "attn_global", "gemm_proj", "gemm_fc1", "gemm_fc2", "embed", "allgather")
WM_OK, WM_ERR_INVALID, WM_ERR_ARCH, WM_ERR_CUDA, WM_ERR_NCCL, WM_ERR_STATE = range(6)


class WmConfig(C.Structure):
    _fields_ = [("depth", C.c_int), ("dino_depth", C.c_int), ("intermediate", C.c_int * 4), ("n_intermediate", C.c_int)]


class WmError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"libwm_b200 error {status}: {message}")
        self.status = status


_vp, _i, _f, _u = C.c_void_p, C.c_int, C.c_float, C.c_uint
_fp = C.c_void_p  # float* passed as raw addresses (tensor.data_ptr())

# name -> (restype, argtypes); every symbol include/wm_b200.h declares
SIGNATURES = {
    "wm_version": (C.c_int, []),
    "wm_last_error": (C.c_char_p, []),
    "wm_create": (C.c_int, [C.POINTER(WmConfig), C.POINTER(_vp)]),
    "wm_destroy": (None, [_vp]),
    "wm_load_weight": (C.c_int, [_vp, C.c_char_p, _fp, C.POINTER(C.c_int64), _i]),
    "wm_finalize_weights": (C.c_int, [_vp]),
    "wm_set_dino_pos_table": (C.c_int, [_vp, _i, _i, _fp]),
    "wm_workspace_bytes": (C.c_size_t, [_vp, _i, _i, _i, _i]),
    "wm_forward": (C.c_int, [_vp, _vp, _fp, _fp, _fp, _fp, C.POINTER(C.c_int), _i, _i, _i, _i, C.POINTER(_vp), _vp,
                             C.c_size_t]),
    "wm_forward_tokens": (C.c_int, [_vp, _vp, _fp, _i, _i, _i, C.POINTER(_vp), _vp, C.c_size_t]),
    "wm_embed": (C.c_int, [_vp, _vp, _fp, _fp, _fp, _fp, C.POINTER(C.c_int), _i, _i, _i, _i, _fp, _vp, C.c_size_t]),
    "wm_profile": (C.c_int, [_vp, _u]),
    "wm_profile_read": (C.c_int, [_vp, _i, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "wm_launch_count": (C.c_longlong, [_vp]),
    "wm_nccl_unique_id": (C.c_int, [C.c_char_p, _vp]),
    "wm_nccl_init": (C.c_int, [_vp, C.c_char_p, _vp, _i, _i]),
    "wm_nccl_allgather": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "wm_p2p_alloc": (C.c_int, [_vp, _i, _i, _i, _vp]),
    "wm_p2p_attach": (C.c_int, [_vp, _vp]),
    "wm_p2p_status": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "wm_flash_attn_gathered": (C.c_int, [_vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _f, _i, _vp, C.c_size_t, _i, _i,
                                         _vp, _u, _vp]),
    "wm_p2p_raise_flag": (C.c_int, [_vp, _vp, _u]),
    "wm_layernorm": (C.c_int, [_vp, _fp, _fp, _fp, _f, _i, _vp, _i]),
    "wm_gemm_bias_act": (C.c_int, [_vp, _vp, _i, _vp, _i, _fp, _i, _i, _i, _vp, _i, _i]),
    "wm_gemm_qkv": (C.c_int, [_vp, _vp, _vp, _fp, _fp, _fp, _fp, _fp, _f, _fp, _fp, _i, _i, _i, _f, _i, _vp, _vp]),
    "wm_gemm_qkv_guard": (C.c_int, [_vp, _vp, _vp, _fp, _fp, _fp, _fp, _fp, _f, _fp, _fp, _i, _i, _i, _f, _i, _vp, _vp, _vp]),
    "wm_flash_attn_guarded": (C.c_int, [_vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, C.c_size_t, _vp, _f]),
    "wm_guard_words": (C.c_int, [_vp, _i, C.POINTER(C.c_float)]),
    "wm_set_int": (C.c_int, [_vp, C.c_char_p, _i]),
    "wm_get_int": (C.c_int, [_vp, C.c_char_p]),
    "wm_fp16_saturation": (C.c_int, [_vp, _vp, C.c_int]),
    "wm_logit_bound": (C.c_float, [_vp, _i, _i]),
    "wm_gemm_ls_residual": (C.c_int, [_vp, _vp, _i, _vp, _i, _fp, _fp, _i, _i, _i, _fp, _i]),
    "wm_gemm_ls_residual_ws": (C.c_int, [_vp, _vp, _i, _vp, _i, _fp, _fp, _i, _i, _i, _fp, _i, _vp, C.c_size_t]),
    "wm_gemm_workspace_bytes": (C.c_size_t, []),
    "wm_copy_cols": (C.c_int, [_vp, _fp, C.c_longlong, _fp, _i, _i]),
    "wm_gemm_table": (C.c_int, [_vp, _vp, _i, _vp, _i, _fp, _i, _i, _i, _i, _fp, _i]),
    "wm_flash_attn": (C.c_int, [_vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _i]),
    "wm_flash_attn_ws": (C.c_int, [_vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _i, _vp, C.c_size_t]),
    "wm_attn_workspace_bytes": (C.c_size_t, []),
    "wm_umma_probe": (C.c_int, [_vp, _vp, _vp, _fp, _i, _u, _u]),
    "wm_patch_gather": (C.c_int, [_vp, _fp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "wm_rope_tables": (None, [_i, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "wm_conv2d_nhwc": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _vp, _i, _i, _fp, _vp, _i, _vp, _i, _vp]),
    "wm_conv2d_final": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _vp, _fp, _fp, _fp, _i, _i, _fp, _fp]),
    "wm_layernorm_rows": (C.c_int, [_vp, _fp, _fp, _fp, _f, C.c_longlong, _i, _i, _i, _i, _vp]),
    "wm_bilinear_nhwc": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _i, _fp, _fp, _vp]),
    "wm_pixel_shuffle": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "wm_im2col_s2": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "wm_skinny_linear": (C.c_int, [_vp, _fp, _i, _fp, _fp, _i, _i, _i, _i, _i, _fp, _fp, _fp, _i]),
    "wm_cam_layernorm": (C.c_int, [_vp, _fp, C.c_longlong, _fp, _fp, _f, _i, _fp, _fp]),
    "wm_cam_attention": (C.c_int, [_vp, _fp, _i, _fp]),
    "wm_cam_embed": (C.c_int, [_vp, _fp, _i, _i, _fp, _fp, _i, _i, _fp]),
    "wm_cam_predict": (C.c_int, [_vp, _fp, _i, _fp, _fp, _i, _i, _fp, _fp]),
    "wm_normalize_depth": (C.c_int, [_vp, _fp, _i, _i, C.c_double, _i, _i, _fp]),
    "wm_pose_prior": (C.c_int, [_vp, _fp, _i, _fp, _i, _i, _i, C.c_double, _fp, _fp]),
    "wm_im2row7": (C.c_int, [_vp, _fp, _i, _i, _i, _vp]),
    "wm_conv2d_project": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _vp, _fp, _fp, _fp, _i, _fp]),
    "wm_splat_cameras": (C.c_int, [_vp, _fp, _i, _i, _i, _fp]),
    "wm_splat_prepare": (C.c_int, [_vp, _fp, _fp, _fp, _fp, _i, _i, _i, _fp, _fp, _fp, _fp, _fp, _fp]),
    "wm_prune_workspace_bytes": (C.c_int, [C.c_longlong, C.POINTER(C.c_size_t)]),
    "wm_prune_splats": (C.c_int, [_vp, C.c_longlong, _fp, _fp, _fp, _fp, _fp, _fp, C.c_double, _vp, C.c_size_t, _fp, _fp, _fp,
                                  _fp, _fp, _vp]),
    "wm_prune_splats_bounds": (C.c_int, [_vp, C.c_longlong, _fp, _fp, _fp, _fp, _fp, _fp, C.c_double, _vp, C.c_size_t, _fp, _fp, _fp,
                                         _fp, _fp, _vp, _vp]),
    "wm_voxel_bounds": (C.c_int, [_vp, C.c_longlong, _fp, C.c_double, _vp]),
    "wm_voxel_keys": (C.c_int, [_vp, C.c_longlong, _fp, C.c_double, _vp, _vp]),
    "wm_conf_filter_workspace_bytes": (C.c_int, [C.c_longlong, C.POINTER(C.c_size_t)]),
    "wm_conf_filter_splats": (C.c_int, [_vp, C.c_longlong, C.c_longlong, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _vp, C.c_size_t, _fp, _fp,
                                        _fp, _fp, _fp, _fp]),
    "wm_raster_project": (C.c_int, [_vp, _fp, _fp, _fp, _fp, _fp, _fp, _i, C.c_longlong, _i, _i, _f, _f, _f, _fp, _fp, _fp, _vp, _vp,
                                    _vp, _vp, C.c_size_t]),
    "wm_raster_workspace_bytes": (C.c_int, [C.c_longlong, C.c_longlong, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "wm_raster_render": (C.c_int, [_vp, _fp, _fp, _fp, _vp, _vp, _fp, _fp, _i, C.c_longlong, _i, _i, C.c_longlong, _vp, C.c_size_t,
                                   _vp, _fp, _fp, _fp]),
    "wm_load_image_workspace_bytes": (C.c_int, [_i, _i, _i, _i, C.POINTER(C.c_size_t)]),
    "wm_load_image": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _fp, _vp, C.c_size_t]),
    "wm_load_image_static_bytes": (C.c_size_t, []),
    "wm_load_image_static": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _fp]),
    "wm_rope2d": (C.c_int, [_vp, _fp, _vp, _i, _i, _i, _i, _fp, _fp, _fp]),
    "wm_pack_bf16": (C.c_int, [_vp, _fp, C.c_longlong, _i, _i, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libwm_b200.so once and attach the prototypes.  Raises if it is missing (no CPU fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python hunyuanworld-mirror_b200/build.py` "
                "(nvcc, sm_100a).  This package has no CPU or PyTorch fallback.")
        handle = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the header and the library ever disagree
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def last_error() -> str:
    msg = lib().wm_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(status: int) -> None:
    if status != WM_OK:
        raise WmError(status, last_error())


class nvtx_range:
    """`with nvtx_range("name"):` pushes an NVTX range when WM_NVTX=1 (for nsys / ncu --nvtx timelines); free otherwise."""
    enabled = os.environ.get("WM_NVTX", "0") == "1"

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if self.enabled:
            import torch

            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *a):
        if self.enabled:
            import torch

            torch.cuda.nvtx.range_pop()
        return False
