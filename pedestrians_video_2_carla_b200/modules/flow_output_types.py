"""Output-type enums the projection layer switches on (reference modules/flow/output_types.py)."""
from enum import Enum


class MovementsModelOutputType(Enum):
    pose_changes = 0
    absolute_loc_rot = 1
    absolute_loc = 2
    relative_rot = 3
    pose_2d = 4


class TrajectoryModelOutputType(Enum):
    changes = 0
    loc_rot = 1
