"""gymapi stand-in: attribute-tolerant dummies, only what import/construct time touches."""


class _Dummy:
    def __init__(self, *args, **kwargs):
        self._args = args
        for k, v in kwargs.items():
            setattr(self, k, v)


class Vec3(_Dummy):
    pass


class Quat(_Dummy):
    pass


class Transform(_Dummy):
    pass


class AssetOptions(_Dummy):
    pass


class SimParams(_Dummy):
    pass


class PlaneParams(_Dummy):
    pass


DOMAIN_SIM = 0
DOMAIN_ENV = 1
DOMAIN_ACTOR = 2
MESH_VISUAL = 1
MESH_COLLISION = 2
MESH_VISUAL_AND_COLLISION = 3
DOF_MODE_POS = 1
DOF_MODE_NONE = 0
SIM_PHYSX = 1
UP_AXIS_Z = 1
