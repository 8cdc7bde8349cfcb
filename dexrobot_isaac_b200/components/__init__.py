"""Reference-named component classes (dexhand_env/components/*) for the fused step pipeline.

The reference builds these in ``DexHandBase._init_components`` (tasks/dexhand_base.py:453-537) and
calls them from ``pre_physics_step`` / ``StepProcessor.process_physics_step`` / ``reset_idx``.  Here
they are thin facades over one :class:`dexrobot_isaac_b200.pipeline.StepPipeline`: same
constructors, method names, argument meaning and error behaviour (RuntimeError, fail fast), but all
arithmetic runs in the two fused CUDA kernels behind the C ABI (include/dexstep.h).  Consequences
of the fusion that a caller can observe are spelled out in each class docstring; there is no
eager/CPU fallback anywhere.
"""
from .action_processor import ActionProcessor
from .observation_encoder import ObservationEncoder
from .reset_manager import ResetManager
from .reward_calculator import RewardCalculator
from .step_processor import StepProcessor
from .termination_manager import TerminationManager

__all__ = ["ActionProcessor", "ObservationEncoder", "RewardCalculator", "TerminationManager", "ResetManager",
           "StepProcessor"]
