#!/usr/bin/env python
"""Stage BASELINE config 2 (the reference's simulator sequence at step 15: 59 frames -> 58 consecutive pairs) where
the GPU box can read it: byte copies of the reference's DATA files (no source code) into the git-ignored
``baseline/_ref/simulator-data/`` (which travels with gpurun snapshots).

    python tests/golden/stage_cfg2.py          # in the build container, where /root/reference exists

``tests/test_gpu_configs.py::test_cfg2_simulator_sequence`` uses the staged frames and skips when they are absent.
"""
import os
import shutil

SRC = "/root/reference/evaluation-runs/simulator-data/data"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
DST = os.path.join(ROOT, "baseline", "_ref", "simulator-data")


def main():
    os.makedirs(os.path.join(DST, "images"), exist_ok=True)
    names = sorted(os.listdir(os.path.join(SRC, "images")))
    frames = [n for n in names if int(os.path.splitext(n)[0]) % 15 == 0]      # get_frame_indices(step=15)
    for n in frames:
        shutil.copyfile(os.path.join(SRC, "images", n), os.path.join(DST, "images", n))
    shutil.copyfile(os.path.join(SRC, "camera_poses.txt"), os.path.join(DST, "camera_poses.txt"))
    print(f"staged {len(frames)} frames ({frames[0]} .. {frames[-1]}) + camera_poses.txt into {DST}")


if __name__ == "__main__":
    main()
