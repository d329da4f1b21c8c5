"""Writes tests/golden/ref_clusters.npz: outputs of the REFERENCE's own get_invisibility_clusters
(src/mapper/__init__.py:92-117) on the seeded images of oracle/scoring_oracle.blob_image.  Runs only where
/root/reference exists (this container).  The reference module imports cv2 at the top for get_convexhull_volume;
OpenCV is not installed here and get_invisibility_clusters does not use it, so an empty stand-in module satisfies
the import."""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import scoring_oracle as so  # noqa: E402

if "cv2" not in sys.modules:
    sys.modules["cv2"] = types.ModuleType("cv2")
spec = importlib.util.spec_from_file_location("ref_mapper", "/root/reference/src/mapper/__init__.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

SEEDS = [0, 1, 2, 3, 4, 5]

if __name__ == "__main__":
    out = {}
    for s in SEEDS:
        img = so.blob_image(s)
        centers, sums = ref.get_invisibility_clusters(img, 30)
        out[f"centers_{s}"] = np.array(centers, dtype=np.float64).reshape(-1, 2)
        out[f"sums_{s}"] = np.array(sums, dtype=np.float32)
        print(s, len(sums), [round(float(x), 2) for x in sums])
    out["seeds"] = np.array(SEEDS)
    np.savez(os.path.join(HERE, "ref_clusters.npz"), **out)
