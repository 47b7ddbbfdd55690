"""Normalizer (reference transforms/pose/normalization/normalizer.py:20-49): hips -> (0,0), |neck - hips| -> 1,
NaN/inf -> 0, xy zeroed where channel 2 < near_zero.  Forward and backward are one kernel each; `.scale` / `.shift`
(normalizer.py:43-49, read by data/base/mixins/dataset/projection_2d_mixin.py:229-230) report the extractor's values
of the LAST call, computed on demand by a small kernel from that call's input."""
from .... import ops
from .hips_neck_extractor import HipsNeckExtractor


class Normalizer(object):
    def __init__(self, extractor=None, near_zero=1e-5):
        self.extractor = extractor if extractor is not None else HipsNeckExtractor()
        self._near_zero = near_zero
        self._last = None          # (sample, dim) of the last call
        self._last_values = None   # (shift, scale), filled when first asked for

    def __repr__(self):
        return '%s(extractor=%s)' % (self.__class__.__name__, self.extractor.__class__.__name__)

    def __call__(self, sample, dim=2, *args, **kwargs):
        kind = getattr(self.extractor, 'kind', 0)
        self._last, self._last_values = (sample.detach(), dim), None
        if kind:   # bbox / hips_neck_bbox: dataset-input transforms, forward only
            if dim != 2:
                raise NotImplementedError('the bounding-box extractors are 2-D (bbox_extractor.py:12-18)')
            return ops.normalize_extractor(sample, self.extractor.hips, self.extractor.neck, kind, self._near_zero)
        if dim == 2 and sample.shape[-1] in (2, 3):
            return ops.hips_neck_normalize(sample, self.extractor.hips, self.extractor.neck, self._near_zero)
        if dim == 3 and sample.shape[-1] == 3 and sample.dim() >= 3:   # absolute locations (denormaliser entry)
            return ops.renormalize(sample, 3, self.extractor.hips, self.extractor.neck, True, self._near_zero)
        self._last = None
        raise NotImplementedError('Normalizer takes (..., J, 2|3) with dim=2 or (..., L, J, 3) with dim=3')

    def _shift_scale(self):
        if self._last is None:
            raise RuntimeError('Normalizer.shift / .scale: the normalizer has not been called yet')
        if self._last_values is None:
            sample, dim = self._last
            self._last_values = ops.extractor_shift_scale(sample, dim, self.extractor.hips, self.extractor.neck,
                                                          getattr(self.extractor, 'kind', 0), self._near_zero)
        return self._last_values

    @property
    def scale(self):
        """(...,) per-pose scale of the last call (a fresh tensor, like the reference's `.clone()`)."""
        return self._shift_scale()[1].clone()

    @property
    def shift(self):
        """(..., dim) per-pose shift point of the last call."""
        return self._shift_scale()[0].clone()
