"""Reference-skeleton loading (reference data/carla/utils.py:16-77, data/base/utils.py:14-25): the UE4 dumps under
data/carla/files/*.yaml -> {bone: Transform} in metres, same file names and (age, gender) -> file mapping."""
import os
from collections import OrderedDict
from functools import lru_cache

import yaml

from ...carla_utils.transforms import Location, Rotation, Transform

FILES_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'files')

# data/carla/utils.py:27-33
_TYPE_TO_FILE = {
    ('adult', 'female'): 'sk_female_relative',
    ('adult', 'male'): 'sk_male_relative',
    ('child', 'female'): 'sk_girl_relative',
    ('child', 'male'): 'sk_kid_relative',
}


@lru_cache(maxsize=16)
def load(type_or_name, files_dir=FILES_DIR):
    """`load('structure')`, `load('sk_female_absolute')`, or `load('adult_female')` (-> relative skeleton)."""
    name = type_or_name
    parts = tuple(type_or_name.split('_'))
    if parts in _TYPE_TO_FILE:
        name = _TYPE_TO_FILE[parts]
    with open(os.path.join(files_dir, name + '.yaml'), 'r') as f:
        return yaml.safe_load(f)


def yaml_to_pose_dict(unreal_transforms, is_abs=False):
    """UE4 centimetres -> metres; relative: hips location zeroed; absolute: every bone relative to the hips
    (data/carla/utils.py:40-77)."""
    pose = OrderedDict()
    for bone, t in unreal_transforms.items():
        l, r = t['location'], t['rotation']
        pose[bone] = Transform(Location(l['x'] / 100.0, l['y'] / 100.0, l['z'] / 100.0),
                               Rotation(pitch=r['pitch'], yaw=r['yaw'], roll=r['roll']))
    hips = pose['crl_hips__C'].location
    hx, hy, hz = hips.x, hips.y, hips.z
    if is_abs:
        for t in pose.values():
            t.location = Location(t.location.x - hx, t.location.y - hy, t.location.z - hz)
    else:
        pose['crl_hips__C'].location = Location(0.0, 0.0, 0.0)
    return pose
