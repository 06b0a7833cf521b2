"""Forward rasterisation: time of one rasterize_splats call (projection, binning, streamed compositing) and a checksum of its
outputs.  (Session 31 ran this against the first, gather-per-batch compositing kernel, since removed: same checksum,
profiles/r02_s31_raster.log.)
    python tools/bench_raster.py [splats] [size]"""
import importlib
import os
import sys
import zlib

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
raster = importlib.import_module("hunyuanworld-mirror_b200.raster")
from oracle import raster_oracle as R  # noqa: E402  (scene generator only)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    size = int(sys.argv[2]) if len(sys.argv) > 2 else 518
    g = np.random.default_rng(2)
    sc = R.synthetic_scene(16, 4, size, size, 1)
    sc.update({"means": np.concatenate([g.uniform(-1.5, 1.5, size=(n, 2)), g.uniform(0.5, 4.0, size=(n, 1))], axis=1).astype(np.float32),
               "quats": g.normal(size=(n, 4)).astype(np.float32), "scales": np.exp(g.uniform(np.log(2e-3), np.log(2e-2), size=(n, 3))).astype(np.float32),
               "opacities": g.uniform(0.0, 1.0, size=n).astype(np.float32), "sh": g.normal(size=(n, 3)).astype(np.float32)})
    t = {k: torch.from_numpy(sc[k]).cuda() for k in ("means", "quats", "scales", "opacities", "sh", "camtoworlds", "Ks")}
    rz = raster.B200Rasterizer()
    run = lambda: rz.rasterize_splats(t["means"], t["quats"], t["scales"], t["opacities"], t["sh"][:, None, :], t["camtoworlds"], t["Ks"],  # noqa: E731
                                      size, size, sh_degree=0)
    out = run()
    torch.cuda.synchronize()
    crc = 0
    for o in out:
        crc = zlib.crc32(o.contiguous().cpu().numpy().tobytes(), crc)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    print(f"streamed compositing: {n} splats, 4 views {size}x{size}, "
          f"{rz.last['n_isects']} intersections, {e0.elapsed_time(e1) / 5:.3f} ms per rasterize_splats, crc32 {crc:08x}", flush=True)


if __name__ == "__main__":
    main()
