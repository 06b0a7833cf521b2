"""`install_into()` on the real (constructor-shimmed) reference `WorldMirror`, build container only: the hand-off of keys
and shapes between the reference module and the B200 aggregator, and the loud failure of the product path on a box
without a B200.  The GPU half of the seam (a batch of two scenes through `B200VisualGeometryTransformer.forward`) is in
tests/test_parity_gpu.py::test_batch_of_two_scenes."""
import importlib
import os
import sys

import pytest
import torch

from conftest import REFERENCE

pkg = importlib.import_module("hunyuanworld-mirror_b200")


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="the reference tree exists in the build container only")
def test_install_into_shimmed_reference_worldmirror():
    sys.path[:0] = [REFERENCE, os.path.join(REFERENCE, "submodules", "gsplat")]
    try:
        import src.models.models.worldmirror as wm  # reference
    finally:
        del sys.path[:2]
    V = wm.VisualGeometryTransformer
    # WorldMirror passes two keywords its aggregator class does not take
    wm.VisualGeometryTransformer = lambda *a, enable_interpolation=False, max_resolution=2044, **k: V(*a, **k)
    try:
        model = wm.WorldMirror(enable_gs=False)  # CPU construction with the reference's own random init
    finally:
        wm.VisualGeometryTransformer = V
    ref = model.visual_geometry_transformer
    ref_sd = ref.state_dict()
    schema = pkg.weights.state_dict_schema(len(ref.frame_blocks), 24)
    assert set(ref_sd.keys()) == set(schema.keys())
    for k, shape in schema.items():
        assert tuple(ref_sd[k].shape) == tuple(shape), k
    assert list(ref.intermediate_idxs) == [4, 11, 17, 23] and ref.patch_start_idx == pkg.aggregator.PATCH_START
    sd = pkg.weights.synthetic_state_dict(seed=0)
    if torch.cuda.is_available():
        agg = pkg.install_into(model, sd)
        assert model.visual_geometry_transformer is agg and agg.depth == 24 and agg.intermediate_idxs == [4, 11, 17, 23]
    else:
        # no B200 here: the swap must fail loudly (no CPU fallback), and leave the reference module in place
        with pytest.raises(RuntimeError):
            pkg.install_into(model, sd)
        assert model.visual_geometry_transformer is ref
