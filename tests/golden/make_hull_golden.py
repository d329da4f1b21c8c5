"""Writes tests/golden/ref_hull.npz: outputs of the REFERENCE's own get_convexhull_volume (src/mapper/__init__.py:8-90,
with OpenCV, scikit-learn and SciPy as installed in this image) on the seeded panoramas of oracle/hull_oracle.panorama.
Runs only where /root/reference exists (this container)."""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import hull_oracle as ho  # noqa: E402

spec = importlib.util.spec_from_file_location("ref_mapper", "/root/reference/src/mapper/__init__.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

SEEDS = list(range(12))

if __name__ == "__main__":
    out = {"seeds": np.array(SEEDS)}
    for s in SEEDS:
        depth, inv = ho.panorama(s)
        last_invisibility, last_volume = ref.get_convexhull_volume(depth, inv, None)  # 3rd positional = hfov, unused
        out[f"invisibility_{s}"] = np.float64(last_invisibility)
        out[f"volume_{s}"] = np.float64(last_volume)
        print(s, float(last_invisibility), float(last_volume))
    np.savez(os.path.join(HERE, "ref_hull.npz"), **out)
