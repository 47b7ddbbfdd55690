"""DeNormalizer (reference transforms/pose/normalization/denormalizer.py:7-37): sample[..., :dim] * scale + shift with
one (shift, scale) per clip; channel 2 of a 2-D sample is carried through.  One streaming kernel (ops.renormalize with
autonormalize off)."""
from .... import ops
from .hips_neck_extractor import HipsNeckExtractor


def get_shift_scale(extractor, reference):
    """Extractor.get_shift_scale (extractor.py:23-36) of a small reference table (rows, J, dim): init-time tensor
    arithmetic on a handful of skeletons, not part of the per-step path."""
    shift = reference[..., extractor.hips, :]
    scale = (reference[..., extractor.neck, :] - shift).norm(dim=-1)
    return shift.contiguous(), scale.contiguous()


class DeNormalizer(object):
    def __call__(self, sample, scale, shift, dim=2, *args, **kwargs):
        """sample (N, L, J, C); scale (N,), shift (N, dim)."""
        if sample.dim() != 4:
            raise NotImplementedError('DeNormalizer takes (N, L, J, C) samples')
        return ops.renormalize(sample, dim, 0, 0, False, 0.0, shift[..., :dim].contiguous(), scale.contiguous())

    @staticmethod
    def from_reference(extractor, reference):
        shift, scale = get_shift_scale(extractor if extractor is not None else HipsNeckExtractor(), reference)
        instance = DeNormalizer()
        return lambda sample, dim=2: instance(sample, scale, shift, dim)
