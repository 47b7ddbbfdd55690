"""HipsNeckBBoxFallbackExtractor (reference hips_neck_bbox_fallback_extractor.py:10-46; the default `hips_neck_bbox`
transform, data/base/base_datamodule.py:61,276): hips / neck normalisation whose scale falls back to 0.5748 x the
bounding-box scale when hips or neck are undetected.  The reference's fallback shift is assigned into a temporary and
has no effect (:27-31); reproduced."""
from .hips_neck_extractor import HipsNeckExtractor


class HipsNeckBBoxFallbackExtractor(HipsNeckExtractor):
    kind = 2
