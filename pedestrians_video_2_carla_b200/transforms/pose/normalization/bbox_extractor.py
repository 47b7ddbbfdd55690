"""BBoxExtractor (reference transforms/pose/normalization/bbox_extractor.py:6-18): shift = centre of the bounding box
of the detected joints, scale point = its top centre.  The arithmetic lives in the kernel (ops.normalize_extractor)."""
from .hips_neck_extractor import HipsNeckExtractor


class BBoxExtractor(HipsNeckExtractor):
    kind = 1
