"""loc_rot_3d (reference loss/loc_rot_3d.py:6-20): loc_3d + rot_3d; None when either is missing."""


def calculate_loss_loc_rot_3d(requirements=None, **kwargs):
    try:
        return requirements['loc_3d'] + requirements['rot_3d']
    except KeyError:
        return None
