"""Image loader behind the reference's interface (src/utils/inference_utils.py:187-248 prepare_images_to_tensor), SURVEY.md
section 8f row 4.  Files are decoded on the host with PIL exactly as the reference does (`Image.open`, RGBA composited over
white, `convert("RGB")`); the decoded bytes go to the GPU once and everything else - Pillow's bicubic resize, ToTensor,
centre crop, constant-1.0 padding, batching - is csrc/loader.cu, bit-equal to the reference.  There is no CPU path."""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np
import torch

from . import capi

PATCH = 14


def _resize_dims(orig_w: int, orig_h: int, max_dim: int, strategy: str):
    """inference_utils.py:159-171."""
    if strategy == "pad":
        if orig_w >= orig_h:
            new_w = max_dim
            new_h = round(orig_h * (new_w / orig_w) / PATCH) * PATCH
        else:
            new_h = max_dim
            new_w = round(orig_w * (new_h / orig_h) / PATCH) * PATCH
    else:
        new_w = max_dim
        new_h = round(orig_h * (new_w / orig_w) / PATCH) * PATCH
    return new_w, new_h


def prepare_arrays_to_tensor(arrays: Sequence[np.ndarray], resize_strategy: str = "crop", target_size: int = 518, device=None) -> torch.Tensor:
    """Decoded RGB uint8 arrays [H, W, 3] -> fp32 CUDA tensor [1, N, 3, H', W'] in [0, 1], as prepare_images_to_tensor."""
    if len(arrays) == 0:
        raise ValueError("At least 1 image is required")
    if resize_strategy not in ("crop", "pad"):
        raise ValueError("Strategy must be 'crop' or 'pad'")
    lib = capi.lib()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    plans = []
    for a in arrays:
        if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3:
            raise ValueError(f"expected a decoded RGB uint8 array [H, W, 3], got {a.dtype} {a.shape}")
        h, w, _ = a.shape
        new_w, new_h = _resize_dims(w, h, target_size, resize_strategy)
        crop_top, crop_h = ((new_h - target_size) // 2, target_size) if (resize_strategy == "crop" and new_h > target_size) else (0, new_h)
        out_h, out_w, top, left = crop_h, new_w, 0, 0
        if resize_strategy == "pad":  # _apply_padding(tensor, target_size)
            h_pad, w_pad = target_size - out_h, target_size - out_w
            if h_pad > 0 or w_pad > 0:
                top, left, out_h, out_w = h_pad // 2, w_pad // 2, out_h + h_pad, out_w + w_pad
        plans.append(dict(a=a, new_w=new_w, new_h=new_h, crop_top=crop_top, crop_h=crop_h, out_h=out_h, out_w=out_w, top=top, left=left))
    dims = {(p["out_h"], p["out_w"]) for p in plans}
    if len(dims) > 1:  # inference_utils.py:234-240: the images that are not of the maximal size are padded to a square
        print(f"Warning: Mixed image dimensions found: {dims}")
        mh, mw = max(d[0] for d in dims), max(d[1] for d in dims)
        side = max(mh, mw)
        for p in plans:
            if p["out_h"] != mh or p["out_w"] != mw:
                h_pad, w_pad = side - p["out_h"], side - p["out_w"]
                if h_pad > 0 or w_pad > 0:
                    p["top"] += h_pad // 2
                    p["left"] += w_pad // 2
                    p["out_h"] += h_pad
                    p["out_w"] += w_pad
        sizes = [(p["out_h"], p["out_w"]) for p in plans]
        if len(set(sizes)) > 1:  # the reference fails here in torch.stack
            raise RuntimeError(f"stack expects each tensor to be equal size, but got {list(sizes[0])} at entry 0 and "
                               f"{list(next(s for s in sizes if s != sizes[0]))} (mixed image sizes whose maximum is not square)")
    out_h, out_w = plans[0]["out_h"], plans[0]["out_w"]
    batch = torch.empty(1, len(plans), 3, out_h, out_w, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        for i, p in enumerate(plans):
            a = np.ascontiguousarray(p["a"])
            if not a.flags.writeable:
                a = a.copy()
            h, w, _ = a.shape
            rgb = torch.from_numpy(a).to(dev, non_blocking=False)
            need = C.c_size_t(0)
            capi.check(lib.wm_load_image_workspace_bytes(h, w, p["new_w"], p["new_h"], C.byref(need)))
            ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
            capi.check(lib.wm_load_image(stream, rgb.data_ptr(), h, w, p["new_w"], p["new_h"], p["crop_top"], p["crop_h"], out_h, out_w,
                                         p["top"], p["left"], batch[0, i].data_ptr(), ws.data_ptr(), ws.numel()))
    return batch


def prepare_images_to_tensor(file_paths: List[str], resize_strategy: str = "crop", target_size: int = 518, device=None) -> torch.Tensor:
    """prepare_images_to_tensor (inference_utils.py:187-248): same arguments, the batch comes back on the GPU."""
    if not file_paths:
        raise ValueError("At least 1 image is required")
    if resize_strategy not in ("crop", "pad"):
        raise ValueError("Strategy must be 'crop' or 'pad'")
    from PIL import Image

    arrays = []
    for path in file_paths:
        img = Image.open(path)
        if img.mode == "RGBA":  # _handle_alpha_channel (:148-156): blend with a white background
            img = Image.alpha_composite(Image.new("RGBA", img.size, (255, 255, 255, 255)), img)
        arrays.append(np.array(img.convert("RGB")))  # a writable copy
    return prepare_arrays_to_tensor(arrays, resize_strategy, target_size, device)
