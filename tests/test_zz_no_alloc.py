"""wm_load_image_static: the image loader with its scratch in a static device array inside the library (no caller
workspace, no allocation; include/wm_b200.h).  It has no counterpart in the reference.
Checked through the C ABI against the existing entry point wm_load_image (bit-equal: integer / byte
arithmetic) on calls in a row over images of different geometry, so that what one call leaves in the shared array would
show in the next.  This file is collected last on purpose."""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

from oracle import loader_oracle as L


def _plan(a, strat, target=518):
    h, w, _ = a.shape
    new_w, new_h = L.resize_dims(w, h, target, strat)
    crop_top, crop_h = ((new_h - target) // 2, target) if (strat == "crop" and new_h > target) else (0, new_h)
    out_h, out_w, top, left = crop_h, new_w, 0, 0
    if strat == "pad":
        h_pad, w_pad = target - out_h, target - out_w
        if h_pad > 0 or w_pad > 0:
            top, left, out_h, out_w = h_pad // 2, w_pad // 2, out_h + h_pad, out_w + w_pad
    return new_w, new_h, crop_top, crop_h, out_h, out_w, top, left


@pytest.mark.gpu
def test_static_scratch_loader_matches_workspace_loader():
    capi = importlib.import_module("hunyuanworld-mirror_b200.capi")
    lib = capi.lib()
    dev = torch.device("cuda", 0)
    cap = lib.wm_load_image_static_bytes()
    assert cap == 8 << 20
    photos = L.synthetic_photos()
    # (photo, strategy): different source sizes and both strategies back to back, the first one again at the end
    cases = [(0, "crop"), (2, "pad"), (6, "crop"), (5, "pad"), (0, "crop")]
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        prepared = []
        for idx, strat in cases:  # everything the two entry points need is allocated BEFORE the calls under test
            a = np.ascontiguousarray(photos[idx])
            h, w, _ = a.shape
            new_w, new_h, crop_top, crop_h, out_h, out_w, top, left = _plan(a, strat)
            need = C.c_size_t(0)
            capi.check(lib.wm_load_image_workspace_bytes(h, w, new_w, new_h, C.byref(need)))
            assert need.value <= cap
            prepared.append(dict(rgb=torch.from_numpy(a.copy()).to(dev), geom=(h, w, new_w, new_h, crop_top, crop_h, out_h, out_w, top, left),
                                 ws=torch.empty(need.value, dtype=torch.uint8, device=dev),
                                 ref=torch.full((3, out_h, out_w), -1.0, device=dev), got=torch.full((3, out_h, out_w), -2.0, device=dev),
                                 host=a, strat=strat))
        for p in prepared:
            capi.check(lib.wm_load_image(stream, p["rgb"].data_ptr(), *p["geom"], p["ref"].data_ptr(), p["ws"].data_ptr(), p["ws"].numel()))
        torch.cuda.synchronize(dev)
        free_before = None
        for i, p in enumerate(prepared):  # calls in a row on one stream, no synchronisation in between
            if i == 1:  # after the first call (module load, address look-up): device memory must not move any more
                torch.cuda.synchronize(dev)
                free_before = torch.cuda.mem_get_info(dev)[0]
            capi.check(lib.wm_load_image_static(stream, p["rgb"].data_ptr(), *p["geom"], p["got"].data_ptr()))
        torch.cuda.synchronize(dev)
        assert torch.cuda.mem_get_info(dev)[0] == free_before, "wm_load_image_static changed the device's free memory"
        for p in prepared:
            assert torch.equal(p["got"], p["ref"]), (p["geom"], p["strat"])
        # and against the CPU oracle for one of them (the existing entry point is itself tested in test_loader.py)
        p = prepared[1]
        assert np.array_equal(p["got"].cpu().numpy(), L.prepare_arrays([p["host"]], p["strat"], 518)[0, 0])
        # a second round over the same array in the opposite order
        for p in reversed(prepared):
            p["got"].fill_(-3.0)
            capi.check(lib.wm_load_image_static(stream, p["rgb"].data_ptr(), *p["geom"], p["got"].data_ptr()))
        torch.cuda.synchronize(dev)
        for p in prepared:
            assert torch.equal(p["got"], p["ref"]), (p["geom"], p["strat"])

        # beyond the array: rejected with a clear error before anything is launched
        h, w, new_w, new_h = 6000, 4000, 518, 777
        need = C.c_size_t(0)
        capi.check(lib.wm_load_image_workspace_bytes(h, w, new_w, new_h, C.byref(need)))
        assert need.value > cap
        dummy = torch.empty(16, dtype=torch.uint8, device=dev)
        out = torch.zeros(3, 518, 518, device=dev)
        rc = lib.wm_load_image_static(stream, dummy.data_ptr(), h, w, new_w, new_h, 129, 518, 518, 518, 0, 0, out.data_ptr())
        assert rc == capi.WM_ERR_INVALID
        msg = capi.last_error()
        assert "static array holds" in msg and str(need.value) in msg, msg
        torch.cuda.synchronize(dev)
        assert float(out.abs().max()) == 0.0
