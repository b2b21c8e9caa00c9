"""ORACLE (test infrastructure, never on the product path): numpy restatement of the reference's two
colour -> gray conversions, both done inside OpenCV / its bundled libpng (cv2 4.13, binary wheel).

  mode 0  cv2.cvtColor(img, COLOR_BGR2GRAY) after cv2.imread   (reference: src/utils/image_loader.py:23-28)
          = (3735 B + 19235 G + 9798 R + 2^14) >> 15            [OpenCV color_rgb: BY15, GY15, RY15, shift 15]
  mode 1  cv2.imread(path, IMREAD_GRAYSCALE) of a colour PNG     (reference: src/run_single_pair.py:31-32)
          = (9797 R + 19234 G + 3737 B) >> 15                   [libpng png_set_rgb_to_gray(.299, .587): weights
                                                                  truncated to 15 bits, blue = 32768 - r - g, no rounding]

Pinned bit-exact against live cv2 in tests/test_oracle_gray.py (random images, all 256^3 colours on a
lattice, a PNG round trip)."""
import numpy as np


def bgr_to_gray(bgr: np.ndarray, mode: int = 0) -> np.ndarray:
    b, g, r = (bgr[..., i].astype(np.uint32) for i in range(3))
    if mode == 0:
        v = (3735 * b + 19235 * g + 9798 * r + 16384) >> 15
    elif mode == 1:
        v = (3737 * b + 19234 * g + 9797 * r) >> 15
    else:
        raise ValueError("mode must be 0 (cvtColor) or 1 (IMREAD_GRAYSCALE of a PNG)")
    return v.astype(np.uint8)
