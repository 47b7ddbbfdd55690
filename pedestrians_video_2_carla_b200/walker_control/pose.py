"""Bone-tree bookkeeping shared by the tensor classes (reference walker_control/pose.py:20-105): the structure YAML
walked depth-first gives the bone order of every tensor; here the walk also yields the parent table the kernels use."""
from collections import OrderedDict

from ..data.carla.utils import load


def bone_order(structure):
    """(names, parents): DFS pre-order of the tree, parent index per bone (-1 for the root)."""
    names, parents = [], []

    def add(node, parent):
        (name, children), = node.items()
        idx = len(names)
        names.append(name)
        parents.append(parent)
        for child in (children or []):
            add(child, idx)
    add(structure[0], -1)
    return names, parents


class Pose(object):
    """Holds the structure and the empty {bone: None} dict (pose.py:99-105); the CARLA-typed scalar FK of the
    reference's Pose (`carla.Transform.transform`, mul_carla_rotations) is CARLA actor control and out of scope."""

    def __init__(self, structure=None, *args, **kwargs):
        super().__init__()
        self._structure = structure if structure is not None else load('structure')['structure']
        self._bone_names, self._parents = bone_order(self._structure)

    @property
    def structure(self):
        return self._structure

    @property
    def empty(self):
        return OrderedDict((name, None) for name in self._bone_names)

    @property
    def bone_names(self):
        return list(self._bone_names)

    @property
    def parents(self):
        return list(self._parents)
