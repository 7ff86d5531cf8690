"""Error types with the reference's names (reference src/ramp/exceptions.py:4-39) plus the
errors of the GPU boundary."""


class MissingBaseLinkError(Exception):
    """The configured base link is not a frame of the model."""


class MissingJointError(Exception):
    """A group names a joint the model does not have."""


class RobotDescriptionNotFoundError(Exception):
    """``robot-descriptions::<name>`` has neither an upstream asset nor a stand-in here."""

    def __init__(self, name):
        super().__init__(
            f"No description available for '{name}': the upstream robot_descriptions assets are fetched "
            "from the network and this build only carries stand-ins for panda, fr3, ur5 and h1.")


class RobotDescriptionError(Exception):
    """A robot description cannot be turned into a model faithfully (an unreadable collision
    mesh, an unsupported collision geometry ...).  Raised instead of skipping the element: a
    link that silently loses its collision body makes every answer about it wrong."""


class MissingAccelerationLimitError(Exception):
    """Acceleration limits are given for some but not all group joints."""

    def __init__(self, joint_names, defined_joint_names):
        missing = set(joint_names) - set(defined_joint_names)
        super().__init__(f"Acceleration limits only defined for {list(defined_joint_names)} - Missing {missing}")


class FrameNotFound(Exception):  # noqa: N818  (pink.exceptions.FrameNotFound spelling)
    """``get_frame_pose`` was asked for a frame the model does not have."""

    def __init__(self, name, frames):
        super().__init__(f"Frame '{name}' not found; frames: {[f.name for f in frames]}")


class RampGpuError(RuntimeError):
    """A call into libramp_b200.so failed (status code and ramp_last_error text)."""


class RampLibraryMissing(ImportError):  # noqa: N818
    """libramp_b200.so is not built / not loadable.  There is no CPU fallback."""
