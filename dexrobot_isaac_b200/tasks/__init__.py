"""Task-hook facades (reference: dexhand_env/tasks/).  ``DexTask`` is the plugin interface of the reference
(tasks/task_interface.py:14-405); ``BaseTask`` and ``BlindGraspingTask`` are the two shipped tasks whose per-step
arithmetic runs inside the fused kernels - their hooks hand out the kernel's results under the reference's names."""
from .base_task import BaseTask
from .blind_grasping_task import BlindGraspingTask
from .task_interface import DexTask

__all__ = ["DexTask", "BaseTask", "BlindGraspingTask"]
