"""Image loader: the numpy oracle against Pillow itself and against the fixture the
reference's `prepare_images_to_tensor` produced (bit-exact: integer arithmetic), and the CUDA loader against the oracle."""
import importlib
import os
import zlib

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import loader_oracle as L

GROUPS = {"same_aspect": [0, 3], "landscape_mixed": [0, 1], "portrait_crop": [2], "portrait_mixed": [2, 5, 0], "square": [4],
          "upscale": [6, 3], "all": [0, 1, 2, 3, 4, 5, 6]}


@pytest.fixture(scope="module")
def photos():
    return L.synthetic_photos()


def test_resize_bit_equal_to_pillow(photos):
    from PIL import Image

    for img in photos:
        h, w, _ = img.shape
        for strat in ("crop", "pad"):
            nw, nh = L.resize_dims(w, h, 518, strat)
            ref = np.asarray(Image.fromarray(img).resize((nw, nh), Image.Resampling.BICUBIC))
            assert np.array_equal(L.resize_bicubic(img, nw, nh), ref), (img.shape, strat)


def _check_against_fixture(fx, name, strat, fn):
    key = f"{name}:{strat}"
    imgs = [L.synthetic_photos()[i] for i in GROUPS[name]]
    if fx[key].size == 0:  # the reference fails on this group (mixed sizes whose maximum is not square): so must we
        with pytest.raises((RuntimeError, ValueError)):
            fn(imgs, strat, 518)
        return
    out = fn(imgs, strat, 518)
    assert tuple(out.shape) == tuple(fx[key])
    u8 = np.rint(out * 255).astype(np.uint8)
    assert np.array_equal(u8.astype(np.float32) / np.float32(255), out)  # every value is exactly k / 255 in fp32
    assert np.array_equal(u8.astype(np.int64).sum(axis=-1), fx[key + ":rowsum"]), key
    assert zlib.crc32(u8.tobytes()) == int(fx[key + ":crc"][0]), key


@pytest.mark.parametrize("name", sorted(GROUPS))
@pytest.mark.parametrize("strat", ["crop", "pad"])
def test_oracle_bit_equal_to_reference_loader(name, strat):
    _check_against_fixture(np.load(os.path.join(GOLDEN, "loader_cases.npz")), name, strat, L.prepare_arrays)


def test_oracle_argument_errors(photos):
    with pytest.raises(ValueError):
        L.prepare_arrays([], "crop")
    with pytest.raises(ValueError):
        L.prepare_arrays(photos[:1], "stretch")


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GROUPS))
@pytest.mark.parametrize("strat", ["crop", "pad"])
def test_cuda_loader_bit_equal_to_reference_loader(name, strat):
    loader = importlib.import_module("hunyuanworld-mirror_b200.loader")
    fn = lambda imgs, s, t: loader.prepare_arrays_to_tensor(imgs, resize_strategy=s, target_size=t).cpu().numpy()
    _check_against_fixture(np.load(os.path.join(GOLDEN, "loader_cases.npz")), name, strat, fn)


@pytest.mark.gpu
def test_cuda_loader_files_and_errors(tmp_path, photos):
    from PIL import Image

    loader = importlib.import_module("hunyuanworld-mirror_b200.loader")
    paths = []
    for i in (0, 3):
        p = tmp_path / f"im{i}.png"
        Image.fromarray(photos[i]).save(p)
        paths.append(str(p))
    rgba = np.concatenate([photos[3], np.full(photos[3].shape[:2] + (1,), 128, np.uint8)], axis=-1)
    Image.fromarray(rgba, "RGBA").save(tmp_path / "alpha.png")
    out = loader.prepare_images_to_tensor(paths, resize_strategy="crop", target_size=518)
    assert out.is_cuda and tuple(out.shape) == (1, 2, 3, 392, 518)
    assert np.array_equal(out.cpu().numpy(), L.prepare_arrays([photos[0], photos[3]], "crop", 518))
    # RGBA is composited over white on the host exactly as the reference does (_handle_alpha_channel)
    a = loader.prepare_images_to_tensor([str(tmp_path / "alpha.png")], resize_strategy="pad", target_size=518)
    white = Image.alpha_composite(Image.new("RGBA", rgba.shape[1::-1], (255, 255, 255, 255)), Image.fromarray(rgba, "RGBA")).convert("RGB")
    assert np.array_equal(a.cpu().numpy(), L.prepare_arrays([np.asarray(white)], "pad", 518))
    with pytest.raises(ValueError):
        loader.prepare_images_to_tensor([], "crop")
    with pytest.raises(ValueError):
        loader.prepare_images_to_tensor(paths, "stretch")
