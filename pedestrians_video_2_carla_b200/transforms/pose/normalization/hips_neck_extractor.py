"""HipsNeckExtractor (reference transforms/pose/normalization/hips_neck_extractor.py:6-13, extractor.py:23-36):
names the shift joint (hips) and the scale joint (neck) of a skeleton; the arithmetic lives in the kernel."""
from ....data.carla.skeleton import CARLA_SKELETON, get_neck_point, get_root_point


class HipsNeckExtractor(object):
    def __init__(self, input_nodes=CARLA_SKELETON, near_zero=1e-5):
        self.input_nodes, self.near_zero = input_nodes, near_zero
        if input_nodes is CARLA_SKELETON:
            self.hips, self.neck = get_root_point(), get_neck_point()
        else:  # any Skeleton-like enum exposing get_hips_point / get_neck_point (single joints)
            self.hips = input_nodes.get_hips_point().value
            self.neck = input_nodes.get_neck_point().value
