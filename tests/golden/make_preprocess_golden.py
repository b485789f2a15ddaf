"""Generates tests/golden/preprocess_*.npz with OpenCV ITSELF (run in the build container, cv2 4.13.0): seeded 8-bit pictures of the
sizes found under data/circle(340x345) (345x340, 345x337, 344x340 - SURVEY F1) and what the reference's own three calls
(models/CvT(Par).py:421-423) make of them.  These are outputs of the dependency that holds the algorithm, not of the oracle."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = {"345x340_bgr": (345, 340, 3, 11), "345x337_bgr": (345, 337, 3, 12), "344x340_gray3": (344, 340, 3, 13), "200x131_bgr": (200, 131, 3, 14)}


def picture(h, w, c, seed, gray3=False):
    """smooth blobs + noise, so that both interpolation and rounding paths are exercised; gray3: B == G == R like the real JPEGs"""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = 120 + 100 * np.sin(xx / 17.0 + seed) * np.cos(yy / 23.0) + rng.normal(0, 25, (h, w))
    if gray3:
        img = np.repeat(base[..., None], 3, axis=2)
    else:
        img = np.stack([base + rng.normal(0, 30, (h, w)) for _ in range(c)], axis=2)
    return np.clip(img, 0, 255).astype(np.uint8)


def main():
    import cv2
    for name, (h, w, c, seed) in CASES.items():
        src = picture(h, w, c, seed, gray3="gray3" in name)
        r = cv2.resize(src, (128, 128))
        g = cv2.cvtColor(r, cv2.COLOR_BGR2GRAY)
        out = g / 255.0
        np.savez_compressed(os.path.join(HERE, "preprocess_%s.npz" % name), src=src, resized=r, gray=g, out=out.astype(np.float64),
                            cv2_version=np.array(cv2.__version__))
        print(name, src.shape, out.mean())


if __name__ == "__main__":
    main()
