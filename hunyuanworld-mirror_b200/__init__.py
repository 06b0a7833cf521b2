"""B200-native drop-in for the WorldMirror aggregator hot path.

`B200VisualGeometryTransformer` mirrors the reference `VisualGeometryTransformer`
(src/models/models/visual_transformer.py:20) over libwm_b200.so (include/wm_b200.h)."""
from .aggregator import B200VisualGeometryTransformer, install_into, nccl_library_path  # noqa: F401
from . import blocks, capi, heads, loader, parallel, priors, raster, weights  # noqa: F401
from .blocks import B200Attention, B200Block, B200RotaryPositionEmbedding2D  # noqa: F401
from .heads import B200CameraHead, B200DPTHead, B200GaussianSplatRenderer, B200WorldMirror  # noqa: F401
