"""Normalizer (reference transforms/pose/normalization/normalizer.py:20-41): hips -> (0,0), |neck - hips| -> 1,
NaN/inf -> 0, xy zeroed where channel 2 < near_zero.  Forward and backward are one kernel each."""
from .... import ops
from .hips_neck_extractor import HipsNeckExtractor


class Normalizer(object):
    def __init__(self, extractor=None, near_zero=1e-5):
        self.extractor = extractor if extractor is not None else HipsNeckExtractor()
        self._near_zero = near_zero

    def __repr__(self):
        return '%s(extractor=%s)' % (self.__class__.__name__, self.extractor.__class__.__name__)

    def __call__(self, sample, dim=2, *args, **kwargs):
        kind = getattr(self.extractor, 'kind', 0)
        if kind:   # bbox / hips_neck_bbox: dataset-input transforms, forward only
            if dim != 2:
                raise NotImplementedError('the bounding-box extractors are 2-D (bbox_extractor.py:12-18)')
            return ops.normalize_extractor(sample, self.extractor.hips, self.extractor.neck, kind, self._near_zero)
        if dim == 2 and sample.shape[-1] in (2, 3):
            return ops.hips_neck_normalize(sample, self.extractor.hips, self.extractor.neck, self._near_zero)
        if dim == 3 and sample.shape[-1] == 3 and sample.dim() >= 3:   # absolute locations (denormaliser entry)
            return ops.renormalize(sample, 3, self.extractor.hips, self.extractor.neck, True, self._near_zero)
        raise NotImplementedError('Normalizer takes (..., J, 2|3) with dim=2 or (..., L, J, 3) with dim=3')
