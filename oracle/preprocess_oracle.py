"""CPU restatement of the image preparation inside preprocess_data (models/CvT(Par).py:420-423):

    image = cv2.imread(filename)                       # uint8 [h, w, 3] BGR (JPEG decode: stays on the host, not restated)
    image = cv2.resize(image, (image_width, image_height))   # INTER_LINEAR, 8-bit fixed point
    image = cv2.cvtColor(image, cv2.COLOR_BGR2GRAY)          # 8-bit fixed point
    image = image / 255.0                                    # float64; Keras casts to float32 at the model input

TEST INFRASTRUCTURE ONLY (tests/ and smoke may import it; the product may not).

The arithmetic lives in OpenCV (an un-vendored dependency of the reference; no version pinned).  PARITY PINNED: this file was
checked bit for bit against cv2 4.13.0 (the version in the build image) by tests/golden/make_preprocess_golden.py, whose outputs
(cv2's own results on seeded inputs) are committed as tests/golden/preprocess_*.npz and re-checked by tests/test_preprocess.py.

Algorithm (opencv/modules/imgproc/src/resize.cpp, resizeGeneric_ with HResizeLinear / VResizeLinear for CV_8U):
  per axis: fx = (float)((d + 0.5) * scale - 0.5), s = floor(fx), fx -= s, clamped at both borders (fx = 0);
            weights = cvRound((1 - fx) * 2048), cvRound(fx * 2048)            (INTER_RESIZE_COEF_BITS = 11, round half to even)
  horizontal: h = S[s] * a0 + S[s+1] * a1                                       (int, no shift)
  vertical:   out = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2
grey (color_rgb.simd.hpp, RGB2Gray<uchar>): (B * 3735 + G * 19235 + R * 9798 + (1 << 14)) >> 15
"""
import numpy as np


def axis_tables(src: int, dst: int, clamp_weights: bool = True):
    """x axis (clamp_weights): out-of-range source columns get weight (2048, 0) on the border column.  y axis: resize.cpp keeps the
    fractional weights and clips the two ROW INDICES instead (resizeGeneric_Invoker: sy = clip(sy0 + k, 0, ssize.height)), which rounds
    differently at the first / last rows when upscaling."""
    scale = 1.0 / (float(dst) / float(src))
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    f = (f - s.astype(np.float32)).astype(np.float32)
    if clamp_weights:
        lo = s < 0
        f[lo] = 0
        s[lo] = 0
        hi = s >= src - 1
        f[hi] = 0
        s[hi] = src - 1
    w0 = np.rint((np.float32(1.0) - f) * np.float32(2048)).astype(np.int32)
    w1 = np.rint(f * np.float32(2048)).astype(np.int32)
    return s, w0, w1


def resize_linear_u8(img: np.ndarray, dh: int, dw: int) -> np.ndarray:
    """img uint8 [h, w] or [h, w, c] -> uint8 [dh, dw(, c)], cv2.resize(img, (dw, dh)) with the default INTER_LINEAR."""
    squeeze = img.ndim == 2
    if squeeze:
        img = img[..., None]
    sh, sw = img.shape[:2]
    sy, b0, b1 = axis_tables(sh, dh, clamp_weights=False)
    sx, a0, a1 = axis_tables(sw, dw)
    x1 = np.minimum(sx + 1, sw - 1)
    y0 = np.clip(sy, 0, sh - 1)
    y1 = np.clip(sy + 1, 0, sh - 1)
    I = img.astype(np.int32)

    def hrow(rows):
        r = I[rows]
        return r[:, sx] * a0[None, :, None] + r[:, x1] * a1[None, :, None]

    r0, r1 = hrow(y0), hrow(y1)
    out = (((b0[:, None, None] * (r0 >> 4)) >> 16) + ((b1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2
    out = out.astype(np.uint8)
    return out[..., 0] if squeeze else out


def bgr2gray_u8(img: np.ndarray) -> np.ndarray:
    I = img.astype(np.int32)
    return ((I[..., 0] * 3735 + I[..., 1] * 19235 + I[..., 2] * 9798 + (1 << 14)) >> 15).astype(np.uint8)


def preprocess(img: np.ndarray, dh: int, dw: int) -> np.ndarray:
    """uint8 [h, w, 3] BGR (or [h, w] / [h, w, 1] grey) -> float32 [dh, dw] in [0, 1]: models/CvT(Par).py:421-423 + the Keras input cast."""
    r = resize_linear_u8(img, dh, dw)
    g = bgr2gray_u8(r) if (r.ndim == 3 and r.shape[2] == 3) else r.reshape(dh, dw)
    return (g.astype(np.float64) / 255.0).astype(np.float32)
