"""Generate the image-loader fixture from the REAL reference (run in the build container only).

    PYTHONPATH=/root/reference python tests/golden/make_golden_loader.py

Writes the seeded photos of oracle/loader_oracle.py `synthetic_photos` as PNG files (lossless) into a scratch directory
inside the repo, runs the reference `prepare_images_to_tensor` (src/utils/inference_utils.py:187-248; PIL + torchvision)
on several groups of them with both strategies, and stores the resulting batches as uint8 (every value is k / 255, so
round(x * 255) is lossless) as shape + CRC32 + per-row sums in tests/golden/loader_cases.npz.  Pins oracle/loader_oracle.py."""
import os
import shutil
import sys
import types

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import loader_oracle as L  # noqa: E402  (only for the seeded inputs)

GROUPS = {"same_aspect": [0, 3], "landscape_mixed": [0, 1], "portrait_crop": [2], "portrait_mixed": [2, 5, 0], "square": [4],
          "upscale": [6, 3], "all": [0, 1, 2, 3, 4, 5, 6]}


def main():
    # the module imports its video helper (cv2 / moviepy, absent here) at import time; the loader does not use it
    stub = types.ModuleType("src.utils.video_utils")
    stub.video_to_image_frames = lambda *a, **k: []
    sys.modules["src.utils.video_utils"] = stub
    from src.utils.inference_utils import prepare_images_to_tensor  # reference

    tmp = os.path.join(HERE, "_loader_tmp")
    os.makedirs(tmp, exist_ok=True)
    photos = L.synthetic_photos()
    paths = []
    for i, img in enumerate(photos):
        p = os.path.join(tmp, f"photo{i}.png")
        Image.fromarray(img).save(p)
        paths.append(p)
    fx = {}
    for name, idx in GROUPS.items():
        for strat in ("crop", "pad"):
            try:
                out = prepare_images_to_tensor([paths[i] for i in idx], resize_strategy=strat, target_size=518)
            except RuntimeError as e:  # mixed sizes whose maximum is not square: the reference pads only the others and
                fx[f"{name}:{strat}"] = np.zeros(0, np.uint8)  # torch.stack fails (inference_utils.py:234-242); recorded as such
                print(name, strat, "RuntimeError:", str(e)[:60])
                continue
            q = (out * 255).round()
            assert (q / 255 - out).abs().max().item() < 1e-7
            u8 = q.to(dtype=__import__("torch").uint8).numpy()
            # the whole batch enters a CRC32 and per-row sums (a failure can be localised); the bytes themselves would be 19 MB
            fx[f"{name}:{strat}"] = np.array(u8.shape, np.int64)
            fx[f"{name}:{strat}:crc"] = np.array([__import__("zlib").crc32(u8.tobytes())], np.int64)
            fx[f"{name}:{strat}:rowsum"] = u8.astype(np.int64).sum(axis=-1)
            print(name, strat, tuple(out.shape))
    shutil.rmtree(tmp)
    path = os.path.join(HERE, "loader_cases.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()
