"""Import-only stand-in for ``gym==0.23.1`` (absent here): only ``gym.spaces.Box`` is
touched by the reference's hot-path modules (observation_encoder.py:995)."""


class _Box:
    def __init__(self, low=None, high=None, shape=None, dtype=None):
        self.low, self.high, self.shape, self.dtype = low, high, shape, dtype


class spaces:  # noqa: N801 - mimics the module attribute
    Box = _Box
