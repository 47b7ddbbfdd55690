"""loc_2d_3d (reference loss/loc_2d_3d.py:6-17): the plain sum of the two requirements."""


def calculate_loss_loc_2d_3d(requirements=None, **kwargs):
    return requirements['loc_2d'] + requirements['loc_3d']
