"""Task configuration for the step pipeline.

The reference composes its configs with Hydra (cfg/config.yaml:7-11, cfg/task/BlindGrasping.yaml:3-5)
and hands the components a plain nested dict (train.py:246, dexhand_base.py:116-121).  The
components here take the same dict; this module only provides the two shipped task configs
without Hydra: ``yaml.safe_load`` + a deep merge that honours ``_delete_: true``
(BlindGrasping.yaml:129).
"""
from __future__ import annotations

import copy
import os
from typing import Optional

import yaml

_CFG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfg", "task")


def deep_merge(base: dict, over: dict) -> dict:
    out = copy.deepcopy(base)
    for key, val in over.items():
        if isinstance(val, dict) and val.get("_delete_", False):
            out[key] = {k: copy.deepcopy(v) for k, v in val.items() if k != "_delete_"}
        elif isinstance(val, dict) and isinstance(out.get(key), dict):
            out[key] = deep_merge(out[key], val)
        else:
            out[key] = copy.deepcopy(val)
    return out


def load_task_config(task: str = "BaseTask", num_envs: int = 1024, control_mode: Optional[str] = None,
                     policy_controls_hand_base: Optional[bool] = None,
                     policy_controls_fingers: Optional[bool] = None, cfg_dir: Optional[str] = None) -> dict:
    """Return ``{"sim": ..., "env": ..., "task": ...}`` for ``BaseTask`` or ``BlindGrasping``."""
    cfg_dir = cfg_dir or _CFG_DIR
    with open(os.path.join(cfg_dir, "BaseTask.yaml")) as f:
        cfg = yaml.safe_load(f)
    cfg.pop("defaults", None)
    if task != "BaseTask":
        path = os.path.join(cfg_dir, f"{task}.yaml")
        if not os.path.exists(path):
            raise ValueError(f"unknown task '{task}' (no {path})")
        with open(path) as f:
            over = yaml.safe_load(f)
        over.pop("defaults", None)
        cfg = deep_merge(cfg, over)
    cfg.setdefault("env", {})["numEnvs"] = int(num_envs)
    if control_mode is not None:
        if control_mode not in ("position", "position_delta"):
            raise ValueError(f"Invalid control mode: {control_mode}")
        cfg["task"]["controlMode"] = control_mode
    if policy_controls_hand_base is not None:
        cfg["task"]["policy_controls_hand_base"] = bool(policy_controls_hand_base)
    if policy_controls_fingers is not None:
        cfg["task"]["policy_controls_fingers"] = bool(policy_controls_fingers)
    return cfg
