"""Re-export the UE4 reference-skeleton DATA (not code) that the hot path loads.

The reference ships five UE4-exported skeleton dumps plus the bone tree under
`src/pedestrians_video_2_carla/data/carla/files/*.yaml` (loaded by
`data/carla/utils.py:16-37`). Those numbers are inputs of the path (bone offsets
in cm, rotations in degrees) and the female relative/absolute pair is the only
golden vector the reference's tests hold (`tests/walker_control/test_pose.py:31-52`).
This script re-serialises them one bone per line (YAML flow style, same keys,
same values) into the package's `data/carla/files/` so the loader keeps the
reference's file contract while the GPU box (which has no /root/reference)
can run.  Run it in the build container only.
"""
import os
import sys

import yaml

SRC = '/root/reference/src/pedestrians_video_2_carla/data/carla/files'
SMPL_SRC = '/root/reference/src/pedestrians_video_2_carla/data/smpl/files'
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, '..', 'pedestrians_video_2_carla_b200', 'data', 'carla', 'files')
SMPL_DST = os.path.join(HERE, '..', 'pedestrians_video_2_carla_b200', 'data', 'smpl', 'files')


def fmt(v):
    s = '%.6f' % float(v)
    assert float(s) == float(v), v
    return s


def export_transforms(name):
    d = yaml.safe_load(open(os.path.join(SRC, name)))['transforms']
    lines = ['# UE4 skeleton dump (cm / degrees); one bone per line. Source data: reference %s' % name,
             'transforms:']
    for bone, t in d.items():
        l, r = t['location'], t['rotation']
        lines.append('  %s: {location: {x: %s, y: %s, z: %s}, rotation: {pitch: %s, yaw: %s, roll: %s}}' % (
            bone, fmt(l['x']), fmt(l['y']), fmt(l['z']), fmt(r['pitch']), fmt(r['yaw']), fmt(r['roll'])))
    open(os.path.join(DST, name), 'w').write('\n'.join(lines) + '\n')
    back = yaml.safe_load(open(os.path.join(DST, name)))['transforms']
    assert back == d, name


def flow(node):
    (name, children), = node.items()
    if children is None:
        return '{%s: null}' % name
    return '{%s: [%s]}' % (name, ', '.join(flow(c) for c in children))


def export_structure(src_dir, dst_dir):
    d = yaml.safe_load(open(os.path.join(src_dir, 'structure.yaml')))
    os.makedirs(dst_dir, exist_ok=True)
    out = '# bone tree; DFS order of this tree is the bone axis of every tensor\nstructure: [%s]\n' % (
        ', '.join(flow(n) for n in d['structure']))
    open(os.path.join(dst_dir, 'structure.yaml'), 'w').write(out)
    assert yaml.safe_load(out) == d


if __name__ == '__main__':
    os.makedirs(DST, exist_ok=True)
    for f in ('sk_female_relative.yaml', 'sk_male_relative.yaml', 'sk_girl_relative.yaml',
              'sk_kid_relative.yaml', 'sk_female_absolute.yaml'):
        export_transforms(f)
    export_structure(SRC, DST)
    if os.path.exists(os.path.join(SMPL_SRC, 'structure.yaml')):
        export_structure(SMPL_SRC, SMPL_DST)
    print('ok', sorted(os.listdir(DST)))
