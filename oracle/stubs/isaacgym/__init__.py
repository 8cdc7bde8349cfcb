"""Import-only stand-in for NVIDIA Isaac Gym Preview 4 (binary, not installable here).

TEST INFRASTRUCTURE ONLY.  It exists so that the *reference's own* component
modules (which do ``from isaacgym import gymapi, gymtorch`` at import time) can be
imported by ``oracle/ref_harness.py`` in the build container.  Nothing in the
product package imports it.
"""
from . import gymapi, gymtorch, torch_utils  # noqa: F401
