#!/usr/bin/env python
"""Generate tests/golden/golden.json from the UNMODIFIED reference (imported from
/root/reference, which exists only in the build container) and live cv2.

    python tests/golden/make_golden.py        # rewrites tests/golden/golden.json

What is frozen (everything the reference itself can produce for this path; it has no tests):
  * per fixture pair: R, t, num_matches, inliers of ``PoseEstimator.estimate_with_debug``
    (reference: src/core/pose_estimator.py:571-688) with VP refinement off, K from the
    reference's ``CameraCalibration().get_matrix(w, h)`` (src/core/camera_calibration.py:65-87);
  * per fixture image: cv2.ORB (as the reference constructs it, pose_estimator.py:85-91)
    keypoint count and SHA-256 of the canonically ordered keypoint fields / descriptors;
  * per fixture pair on canonically ordered features: match list hash, E, RANSAC mask hash,
    recoverPose R / t / count (cv2 calls exactly as pose_estimator.py:144-151, :522-527, :533);
  * build-independent integer known answers: OpenCV RNG subsets, quotas, umax,
    rBRIEF table hash, the reference's Euler conversions on fixed rotations.
The fixtures in tests/golden/images are byte copies of reference frames:
  sim_0000NN  <- evaluation-runs/simulator-data/data/images/0000NN.png
  phone_/salah_ likewise from phone-data / vo_dataset_salah.
"""
import hashlib
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def import_reference():
    """The reference package needs plotly only for its visualiser: stub it."""
    for name in ("plotly", "plotly.graph_objects"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, REF)
    from src.core.camera_calibration import CameraCalibration
    from src.core.pose_estimator import PoseEstimator
    from src.utils import geometry
    return PoseEstimator, CameraCalibration, geometry


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


PAIRS = [("sim_000000.png", "sim_000015.png"), ("sim_000015.png", "sim_000030.png"),
         ("phone_000000.png", "phone_000005.png"), ("salah_000000.png", "salah_000005.png")]


def main():
    import cv2

    from oracle import geometry as og
    from oracle import orb as oo
    from oracle.orb_pattern import PATTERN
    from oracle.reference_path import PathConfig, ReferencePath

    PoseEstimator, CameraCalibration, rgeom = import_reference()
    out = {"cv2_version": cv2.__version__, "images": {}, "pairs": [], "known": {}}

    def load(name):
        return cv2.imread(os.path.join(HERE, "images", name), cv2.IMREAD_GRAYSCALE)

    names = sorted({n for p in PAIRS for n in p})
    p = oo.OrbParams()
    for n in names:
        img = load(n)
        orb = cv2.ORB_create(nfeatures=4000, scaleFactor=1.1, nlevels=12, fastThreshold=15,
                             scoreType=cv2.ORB_HARRIS_SCORE)
        kps, desc = orb.detectAndCompute(img, None)
        c = oo.canonicalize_cv2(kps, desc, oo.level_geometry(img.shape[1], img.shape[0], p))
        out["images"][n] = {"shape": list(img.shape), "n_keypoints": len(c), "xy": sha(np.stack([c.x, c.y])),
                            "angle": sha(c.angle), "response": sha(c.response), "octave": sha(c.octave),
                            "size": sha(c.size), "desc": sha(c.desc),
                            "per_level": np.bincount(c.octave, minlength=12).tolist()}

    for a, b in PAIRS:
        i1, i2 = load(a), load(b)
        K = CameraCalibration().get_matrix(i1.shape[1], i1.shape[0])
        est = PoseEstimator(K, "ORB", nfeatures=4000, use_vp_refinement=False)
        d = est.estimate_with_debug(i1, i2)
        rec = {"img1": a, "img2": b, "K": K.tolist(),
               "reference": {"R": d["R"].tolist(), "t": d["t"].ravel().tolist(), "num_matches": int(d["num_matches"]),
                             "inliers": int(d["inliers"]),
                             "euler_yup": list(rgeom.rotation_to_euler(d["R"], "yup")),
                             "euler_zyx": list(rgeom.rotation_to_euler(d["R"], "zyx"))}}
        can = ReferencePath(K, PathConfig(), canonical=True).estimate_with_debug(i1, i2)
        rec["canonical"] = {"R": can["R"].tolist(), "t": can["t"].ravel().tolist(), "num_matches": can["num_matches"],
                            "inliers": int(can["inliers"]), "matches": sha(can["matches"]),
                            "first_matches": can["matches"][:8].tolist(), "E": can["E"].tolist(),
                            "ransac_inliers": int(can["ransac_mask"].sum()), "ransac_mask": sha(can["ransac_mask"]),
                            "pts1": sha(can["pts1"]), "pts2": sha(can["pts2"])}
        out["pairs"].append(rec)

    rng = og.CvRNG()
    out["known"]["rng_subsets_500"] = [og.draw_subset(rng, 500) for _ in range(8)]
    out["known"]["quotas_4000_12"] = [lv.quota for lv in oo.level_geometry(640, 480, p)]
    out["known"]["quotas_5000_12"] = [lv.quota for lv in oo.level_geometry(1280, 720, oo.OrbParams(nfeatures=5000))]
    out["known"]["quotas_20000_8"] = [lv.quota for lv in oo.level_geometry(3840, 2160, oo.OrbParams(20000, 1.1, 8))]
    out["known"]["sizes_640x480"] = [[lv.w, lv.h] for lv in oo.level_geometry(640, 480, p)]
    out["known"]["umax"] = oo.UMAX
    out["known"]["pattern_sha256"] = sha(PATTERN.astype(np.int8))
    # the reference's Euler helpers on fixed rotations (src/utils/geometry.py)
    eul = []
    for ypr in [(10.0, 20.0, 30.0), (-170.0, 5.0, 0.5), (0.0, 89.9995, 12.0), (45.0, -60.0, -120.0)]:
        for conv in ("yup", "zyx"):
            R = rgeom.euler_to_rotation(*ypr, conv)
            eul.append({"convention": conv, "ypr_in": list(ypr), "R": R.tolist(),
                        "ypr_out": list(rgeom.rotation_to_euler(R, conv))})
    out["known"]["euler"] = eul
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote golden.json:", {k: v["n_keypoints"] for k, v in out["images"].items()},
          [(r["reference"]["num_matches"], r["reference"]["inliers"]) for r in out["pairs"]])


if __name__ == "__main__":
    main()
