"""ReferenceSkeletonsDeNormalizer (reference transforms/pose/normalization/reference_skeletons_denormalizer.py:35-91):
rescales (optionally self-normalised) 2-D projections or 3-D absolute locations to the hips position and hips-neck
distance of the clip's CARLA reference skeleton.

The reference normalises, stacks one reference tensor per clip and denormalises in ~15 tensor ops per call; here the
(shift, scale) of the four reference skeletons is a 4-row device table built once per device, the clips carry a row
index, and normalise + denormalise is ONE kernel (ops.renormalize), forward and backward."""
import torch

from .... import ops
from ....data.carla import reference as R
from .hips_neck_extractor import HipsNeckExtractor
from .denormalizer import DeNormalizer, get_shift_scale

AGE_MAPPINGS = {
    'adult': 'adult', 'child': 'child',
    'senior': 'adult', 'young': 'child', float('nan'): 'adult', 'nan': 'adult',
}
GENDER_MAPPINGS = {
    'female': 'female', 'male': 'male',
    'neutral': 'female', float('nan'): 'female', 'nan': 'female',
}


def _lookup(table, key, default):
    if isinstance(key, float) and key != key:   # NaN keys never compare equal
        return default
    return table[key]


class ReferenceSkeletonsDeNormalizer(DeNormalizer):
    def __init__(self, extractor=None):
        self._extractor = extractor if extractor is not None else HipsNeckExtractor()
        self._tables = {}

    def _table(self, device, dim):
        key = (str(device), dim)
        if key not in self._tables:
            ref = R.get_absolute_tensors(device)[0] if dim == 3 else R.get_projections(device)[..., :2]
            self._tables[key] = get_shift_scale(self._extractor, ref)
        return self._tables[key]

    @staticmethod
    def skeleton_rows(meta, n, device):
        ages = meta.get('age', ['adult'] * n)
        genders = meta.get('gender', ['female'] * n)
        rows = [R.skeleton_index(_lookup(AGE_MAPPINGS, a, 'adult'), _lookup(GENDER_MAPPINGS, g, 'female'))
                for a, g in zip(ages, genders)]
        return torch.tensor(rows, dtype=torch.int32).to(device, non_blocking=True)

    def _run(self, frames, meta, autonormalize, dim, rows=None):
        if frames.dim() != 4:
            raise NotImplementedError('ReferenceSkeletonsDeNormalizer takes (N, L, J, C) frames')
        shift, scale = self._table(frames.device, dim)
        if rows is None:
            rows = self.skeleton_rows(meta, len(frames), frames.device)
        return ops.renormalize(frames, dim, self._extractor.hips, self._extractor.neck, autonormalize,
                               self._extractor.near_zero, shift, scale, rows)

    def from_projection(self, frames, meta, autonormalize=False, rows=None):
        """(N, L, J, 2|3) 2-D poses -> reference-skeleton pixel coordinates (:41-66)."""
        return self._run(frames, meta, autonormalize, 2, rows)

    def from_abs(self, frames, meta, autonormalize=False, rows=None):
        """(N, L, J, 3) absolute 3-D locations -> reference-skeleton metres (:68-91)."""
        return self._run(frames, meta, autonormalize, 3, rows)
