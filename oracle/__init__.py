"""CPU oracle for the FK + projection + loc_2d_3d hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under `pedestrians_video_2_carla_b200/` may import
this package; only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline`
/ `--impl reference` legs do, and there only as the checker / the timed CPU baseline.

Parity status (see DESIGN.md "Oracle"):
  * FK, Euler->matrix, YAML loading: PINNED to the reference's golden pair
    sk_female_relative.yaml -> sk_female_absolute.yaml
    (reference tests/walker_control/test_pose.py:31-52).
  * FK recursion, frame loop, world scan, normaliser, mask, loc_2d / loc_3d / loc_2d_3d:
    PINNED to outputs of the reference's own Python code executed in the build
    container with third-party stubs (tests/golden/make_golden.py ->
    tests/golden/reference_run.npz).
  * pytorch3d 0.6.0 arithmetic (rotation_6d_to_matrix, euler_angles_to_matrix,
    look_at_view_transform, PerspectiveCameras.transform_points_screen): pytorch3d is
    not installed anywhere reachable, so these are restated from its published
    formulas: "parity unpinned" for the camera closed form, cross-checked against
    scipy (Euler) and float64 finite differences (gradients).
"""
from .fk_oracle import *  # noqa: F401,F403
from .generator_oracle import *  # noqa: F401,F403
