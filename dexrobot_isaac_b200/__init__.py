"""dexrobot_isaac_b200 — B200-native per-step tensor pipeline of dexhand_env."""
__version__ = "0.1.0"
