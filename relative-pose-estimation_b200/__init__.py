"""B200-native two-view relative-pose engine: drop-in for the ORB -> BFMatcher ->
findEssentialMat -> recoverPose path of ofekm5/relative-pose-estimation
(reference: src/core/pose_estimator.py).  Import as ``rpe_b200`` (the directory name carries
a hyphen; ``rpe_b200/`` at the repo root is the importable alias of this package)."""
from .engine import Engine, EngineConfig
from .pose_estimator import PoseEstimator
from .geometry import rotation_to_euler, euler_to_rotation, rotation_to_euler_yup, rotation_to_euler_zyx
from .batch_processor import BatchProcessor
from .pose_evaluator import PoseEvaluator
from . import _lib, sharding

__all__ = ["Engine", "EngineConfig", "PoseEstimator", "BatchProcessor", "PoseEvaluator", "sharding", "rotation_to_euler", "euler_to_rotation",
           "rotation_to_euler_yup", "rotation_to_euler_zyx"]
