"""weighted_loc_2d_loc_rot_3d (reference loss/weighted_loc_2d_loc_rot_3d.py:6-25): the weighted sum of loc_2d, loc_3d
and rot_3d (missing weights are 1); None when one requirement is missing.  The weights are Python scalars here: the
reference's per-call `torch.tensor(w, device=...)` is a host->device copy per term and step."""


def calculate_loss_weighted_loc_2d_loc_rot_3d(requirements=None, loss_weights=None, **kwargs):
    w = loss_weights or {}
    try:
        return (float(w.get('loc_2d', 1.0)) * requirements['loc_2d'] +
                float(w.get('loc_3d', 1.0)) * requirements['loc_3d'] +
                float(w.get('rot_3d', 1.0)) * requirements['rot_3d'])
    except KeyError:
        return None
