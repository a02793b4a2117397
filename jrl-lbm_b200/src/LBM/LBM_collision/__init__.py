from .LBM_collision_2d import LBMCollision2d  # noqa: F401
from .LBM_collision_MRT_2d import LBMCollisionMRT2d  # noqa: F401
from .LBM_collision_KBC_2d import LBMCollisionKBC2d  # noqa: F401
