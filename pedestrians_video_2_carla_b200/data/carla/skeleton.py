"""CARLA_SKELETON: bone -> tensor index.  The reference takes the enum from pedestrians_scenarios
(data/carla/skeleton.py:1); its values are the DFS pre-order indices of structure.yaml, built here from
the same file so the two cannot drift apart."""
import enum

from ...walker_control.pose import bone_order
from .utils import load

_NAMES, _PARENTS = bone_order(load('structure')['structure'])
CARLA_SKELETON = enum.Enum('CARLA_SKELETON', [(n, i) for i, n in enumerate(_NAMES)])
CARLA_PARENTS = tuple(_PARENTS)


def get_root_point():
    return CARLA_SKELETON.crl_hips__C.value


def get_neck_point():
    return CARLA_SKELETON.crl_neck__C.value
