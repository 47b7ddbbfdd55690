"""loc_2d_loc_rot_3d (reference loss/loc_2d_loc_rot_3d.py:6-21): loc_2d + loc_3d + rot_3d; None when one is
missing."""


def calculate_loss_loc_2d_loc_rot_3d(requirements=None, **kwargs):
    try:
        return requirements['loc_2d'] + requirements['loc_3d'] + requirements['rot_3d']
    except KeyError:
        return None
