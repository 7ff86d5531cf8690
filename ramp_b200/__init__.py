"""ramp_b200 -- B200-native state/motion validity checking behind RAMP's Python API.

The public names mirror ``ramp`` (reference src/ramp/__init__.py:24-40).  Importing the
package does not load CUDA; the first GPU query loads ``libramp_b200.so`` and fails loudly if
it is missing (there is no CPU fallback).
"""

from ramp_b200.motion_planner import GpuMotionValidator, MotionPlanner
from ramp_b200.robot_model import RobotModel, create_geometry_object, load_robot_model
from ramp_b200.robot_state import RobotState
from ramp_b200.se3 import SE3, Quaternion

__all__ = ["SE3", "GpuMotionValidator", "MotionPlanner", "Quaternion", "RobotModel", "RobotState",
           "create_geometry_object", "load_robot_model"]
