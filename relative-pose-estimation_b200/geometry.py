"""Host mirror of the reference's Euler helpers (reference: src/utils/geometry.py:14-27, 48-80,
181-205).  The engine also computes these on the device per pair (rpe_estimate_batch `euler`
output); these functions exist so callers of the reference's utils keep working, e.g. for the
world-frame composition R_prev @ R_rel in BatchProcessor (src/core/batch_processor.py:88-101)."""
from __future__ import annotations

import math

import numpy as np

CONVENTION_YUP = "yup"
CONVENTION_ZYX = "zyx"


def rotation_to_euler_yup(R):
    # np.arcsin as the reference (geometry.py:60): rounding in R_prev @ R_rel can leave |R[2,1]| a hair above 1 near
    # gimbal lock; numpy then yields nan with a RuntimeWarning and the sequence goes on, math.asin would raise
    pitch = float(np.arcsin(R[2, 1]))
    if abs(R[2, 1]) > 0.9999:
        roll, yaw = math.atan2(-R[1, 2], R[1, 1]), 0.0
    else:
        yaw, roll = math.atan2(-R[2, 0], R[0, 0]), math.atan2(R[1, 0], R[1, 1])
    return math.degrees(yaw), math.degrees(pitch), math.degrees(roll)


def rotation_to_euler_zyx(R):
    sy = math.sqrt(R[0, 0] ** 2 + R[1, 0] ** 2)
    if sy < 1e-6:
        roll, pitch, yaw = math.atan2(-R[1, 2], R[1, 1]), math.atan2(-R[2, 0], sy), 0.0
    else:
        roll, pitch, yaw = math.atan2(R[2, 1], R[2, 2]), math.atan2(-R[2, 0], sy), math.atan2(R[1, 0], R[0, 0])
    return math.degrees(yaw), math.degrees(pitch), math.degrees(roll)


def rotation_to_euler(R, convention=CONVENTION_YUP):
    """(yaw, pitch, roll) in degrees."""
    return rotation_to_euler_zyx(R) if convention == CONVENTION_ZYX else rotation_to_euler_yup(R)


def euler_to_rotation(yaw_deg, pitch_deg, roll_deg, convention=CONVENTION_YUP):
    y, p, r = math.radians(yaw_deg), math.radians(pitch_deg), math.radians(roll_deg)
    cy, sy, cp, sp, cr, sr = math.cos(y), math.sin(y), math.cos(p), math.sin(p), math.cos(r), math.sin(r)
    if convention == CONVENTION_ZYX:      # R = Rz(yaw) Ry(pitch) Rx(roll)
        return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                         [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                         [-sp, cp * sr, cp * cr]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])          # R = Ry(yaw) Rx(pitch) Rz(roll)
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    Rz = np.array([[cr, -sr, 0], [sr, cr, 0], [0, 0, 1]])
    return Ry @ Rx @ Rz
