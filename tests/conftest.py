import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(ROOT, "tests", "golden")
IMAGES = os.path.join(GOLDEN, "images")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_gray(name):
    import cv2
    img = cv2.imread(os.path.join(IMAGES, name), cv2.IMREAD_GRAYSCALE)
    assert img is not None, name
    return img


# K of the reference's default single pair: CameraCalibration().get_matrix(640, 480)
# (reference: src/core/camera_calibration.py:65-87; base intrinsics @960x720 scaled)
def default_K(w=640, h=480):
    sx, sy = w / 960.0, h / 720.0
    return np.array([[924.829 * sx, 0, 468.249 * sx], [0, 920.477 * sy, 353.659 * sy], [0, 0, 1.0]])


@pytest.fixture(scope="session")
def sim_pair():
    return load_gray("sim_000000.png"), load_gray("sim_000015.png")


@pytest.fixture(scope="session")
def gpu_available():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def dump(name, **arrays):
    """Save diagnostics where gpurun brings them back."""
    d = os.path.join(ROOT, "gpurun_out")
    os.makedirs(d, exist_ok=True)
    np.savez_compressed(os.path.join(d, name + ".npz"), **arrays)
