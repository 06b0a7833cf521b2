"""Golden fixture for the depth-prior normalisation from the REAL reference (build container only):
    PYTHONPATH=/root/reference python tests/golden/make_golden_priors.py
runs `normalize_depth` (src/models/utils/priors.py:107) on oracle/priors_oracle.py's seeded cases."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import priors_oracle as PO  # noqa: E402


def main():
    from src.models.utils.priors import normalize_depth  # reference

    x = PO.synthetic_depth_cases(seed=0)
    with torch.no_grad():
        y = normalize_depth(torch.from_numpy(x.copy())).numpy()
    np.savez_compressed(os.path.join(HERE, "priors_depth.npz"), out=y, seed=0, torch=torch.__version__)
    print("wrote priors_depth.npz", y.shape, float(y.mean()))


if __name__ == "__main__":
    main()
