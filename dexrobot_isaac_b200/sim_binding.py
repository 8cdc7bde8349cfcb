"""Graph-safe application of the fused resets to the simulator.

``dex_post_physics`` resets the finished envs in place: it rewrites their rows of ``dof_state`` and ``actor_root_state`` (the
gymtorch views, zero copy).  PhysX only sees those rows when the caller pushes them with Isaac Gym's *indexed* setters.  The
reference builds the index tensors on the host from ``torch.nonzero(reset_buf)`` (components/reset/reset_manager.py:139-168,
components/physics/physics_manager.py:121-160): one device->host sync per step and tensors whose length changes every step,
which rules out CUDA-graph capture.

``IndexedResetApplier`` keeps FIXED-CAPACITY int32 index buffers on the device, fills them with ``dex_reset_actor_indices``
(one small launch, no host read; padding = a harmless id, see include/dexstep.h) and always calls the setters with the same
counts.  It needs the ``gym`` object, the ``sim`` handle and ``gymtorch.unwrap_tensor`` of the host application - nothing here
imports Isaac Gym, so the class is exercised against a fake gym in tests/test_cuda_sim_binding.py."""
from __future__ import annotations

import ctypes
from typing import Callable, Optional

import torch

from . import _lib as L


class IndexedResetApplier:
    def __init__(self, pipeline, gym, sim, unwrap_tensor: Callable, dof_actor: int = 0, capacity: Optional[int] = None):
        """``dof_actor``: local actor index of the hand (reset_manager.py: hand_local_actor_index).  ``capacity``: how many
        resets per step the index buffers hold; the default (num_envs) can never overflow."""
        self.p, self.gym, self.sim, self.unwrap = pipeline, gym, sim, unwrap_tensor
        self.num_actors = int(pipeline.layout.num_actors)
        self.dof_actor = int(dof_actor)
        self.capacity = int(capacity) if capacity is not None else pipeline.n
        if not 0 < self.capacity <= pipeline.n:
            raise ValueError("capacity must be in [1, num_envs]")
        dev = pipeline.device
        self.root_actor_ids = torch.zeros(self.capacity * self.num_actors, dtype=torch.int32, device=dev)
        self.dof_actor_ids = torch.zeros(self.capacity, dtype=torch.int32, device=dev)
        self.overflow = torch.zeros(1, dtype=torch.int32, device=dev)  # resets that did not fit (capacity < num_envs only)
        st = pipeline.sim_tensors
        self._root = st["actor_root_state"].view(-1, 13)   # the FULL tensors, as the gym API expects (reset_manager.py:150-156)
        self._dof = st["dof_state"].view(-1, 2)

    def apply(self):
        """Call after ``process_physics_step`` / ``pipeline.post_step()`` of the same step.  No host sync, constant arguments."""
        p = self.p
        L.check(p.lib.dex_reset_actor_indices(ctypes.byref(p._epi), self.capacity, self.num_actors, self.dof_actor,
                                              self.root_actor_ids.data_ptr(), self.dof_actor_ids.data_ptr(),
                                              self.overflow.data_ptr(), p._stream()), "dex_reset_actor_indices")
        p.launches += 1
        self.gym.set_actor_root_state_tensor_indexed(self.sim, self.unwrap(self._root), self.unwrap(self.root_actor_ids),
                                                     self.capacity * self.num_actors)
        self.gym.set_dof_state_tensor_indexed(self.sim, self.unwrap(self._dof), self.unwrap(self.dof_actor_ids), self.capacity)
