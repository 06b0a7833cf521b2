"""CPU checks of the drop-in boundary: libwm_b200.so loads without a GPU, exports every symbol that
include/wm_b200.h declares, and refuses to compute (no CPU fallback).  No kernel is launched here."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT

pkg = importlib.import_module("hunyuanworld-mirror_b200")
capi = pkg.capi
from oracle import aggregator_oracle as O  # noqa: E402


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "wm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(wm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = capi.lib()
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/wm_b200.h but not exported"
    assert set(names) == set(capi.SIGNATURES), "ctypes table and header disagree"
    assert lib.wm_version() == 100


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_no_cpu_fallback():
    lib = capi.lib()
    cfg = capi.WmConfig()
    cfg.depth, cfg.dino_depth, cfg.n_intermediate = 1, 1, 0
    h = C.c_void_p()
    assert lib.wm_create(C.byref(cfg), C.byref(h)) == capi.WM_ERR_ARCH
    assert "no CPU path" in capi.last_error()
    buf = (C.c_float * 2048)()
    assert lib.wm_layernorm(None, buf, buf, buf, 1e-5, 2, buf, 1) == capi.WM_ERR_ARCH
    m = pkg.B200VisualGeometryTransformer(depth=1, dino_depth=1)
    with pytest.raises(capi.WmError):
        m.load_state_dict(pkg.weights.synthetic_state_dict(depth=1, dino_depth=1))


def test_argument_validation_needs_no_gpu():
    lib = capi.lib()
    cfg = capi.WmConfig()
    cfg.depth, cfg.dino_depth, cfg.n_intermediate = 2, 1, 2
    cfg.intermediate[0], cfg.intermediate[1] = 1, 1  # not increasing
    h = C.c_void_p()
    assert lib.wm_create(C.byref(cfg), C.byref(h)) == capi.WM_ERR_INVALID
    assert lib.wm_workspace_bytes(None, 2, 252, 250, 1) == 0  # W not a multiple of 14
    n8 = lib.wm_workspace_bytes(None, 8, 518, 518, 1)
    n8w = lib.wm_workspace_bytes(None, 8, 518, 518, 8)
    T = 8 * 1376
    assert n8 > T * 1024 * (4 + 4 + 2 + 2 + 4 + 2 + 8) * 0.99 and n8w - n8 == 7 * T * 2048 * 2


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_round2_entry_points_refuse_without_a_gpu():
    """The entry points added in round 2 (the no-allocation image loader, the device-side attention choice) load, report their
    constants and refuse to compute on a machine without a B200 -- no silent fallback."""
    lib = capi.lib()
    assert lib.wm_load_image_static_bytes() == 8 << 20
    buf = (C.c_float * 64)()
    assert lib.wm_load_image_static(None, buf, 4, 4, 2, 2, 0, 2, 2, 2, 0, 0, buf) == capi.WM_ERR_ARCH
    assert lib.wm_flash_attn_guarded(None, buf, 1024, buf, buf, 2048, buf, 1024, 1, 128, 128, 16, 10, 4, None, 0, buf, 80.0) == capi.WM_ERR_ARCH
    assert lib.wm_flash_attn_guarded(None, buf, 1024, buf, buf, 2048, buf, 1024, 1, 128, 128, 8, 10, 4, None, 0, buf, 80.0) == capi.WM_ERR_INVALID
    assert lib.wm_gemm_qkv_guard(None, buf, buf, buf, None, None, None, None, 1e-5, None, None, 1, 0, 1, 1.0, 128, buf, buf, None) == capi.WM_ERR_INVALID
    assert lib.wm_guard_words(None, 0, buf) == capi.WM_ERR_INVALID


def test_rope_tables_match_oracle():
    lib = capi.lib()
    n = 38
    cos = (C.c_float * ((n + 1) * 16))()
    sin = (C.c_float * ((n + 1) * 16))()
    lib.wm_rope_tables(n, cos, sin)
    c_ref, s_ref = O.rope_tables(n + 1, 32)
    np.testing.assert_allclose(np.array(cos).reshape(n + 1, 16), c_ref.numpy(), atol=3e-6)
    np.testing.assert_allclose(np.array(sin).reshape(n + 1, 16), s_ref.numpy(), atol=3e-6)


def test_module_interface_errors_on_cpu():
    m = pkg.B200VisualGeometryTransformer(depth=1, dino_depth=1)
    with pytest.raises(ValueError):
        m(torch.rand(1, 2, 4, 28, 28))  # channels != 3 (visual_transformer.py:272-273)
    with pytest.raises(AssertionError):
        m(torch.rand(1, 2, 3, 29, 28))  # patch_embed.py:67-68
    with pytest.raises(RuntimeError):
        m(torch.rand(1, 2, 3, 28, 28))  # weights not loaded
    with pytest.raises(NotImplementedError):
        pkg.B200VisualGeometryTransformer(embed_dim=768)
    with pytest.raises(NotImplementedError):
        pkg.B200VisualGeometryTransformer(enable_cond=False)


def test_schema_matches_reference_counts():
    s = pkg.weights.state_dict_schema()
    assert len(s) == 1222
    total = sum(int(np.prod(v)) for v in s.values())
    assert abs(total - 916.2e6) < 0.2e6
