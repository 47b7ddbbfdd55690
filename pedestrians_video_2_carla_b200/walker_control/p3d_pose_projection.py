"""P3dPoseProjection: pinhole projection of p3d-space joints to pixels on the GPU, drop-in for reference
walker_control/p3d_pose_projection.py (constructor, forward, update_camera, image_size, camera).

The reference builds a pytorch3d PerspectiveCameras(in_ndc=False) and calls transform_points_screen; for that camera
the chain reduces to  c = w R + T ; u = px - f c.x / c.z ; v = py - f c.y / c.z ; out.z = 1 / c.z  (pytorch3d 0.6.0,
SURVEY 8a row a10), which is what the kernel evaluates.  Camera constants follow p3d_pose_projection.py:37-69,
pose_projection.py:18-39,67-75 and carla_utils/setup.py:37-62."""
import numpy as np
import torch

from .. import _lib, ops


def look_at_view_transform(eye, at, up):
    """pytorch3d look_at_view_transform / look_at_rotation restated (init-time host code, float64):
    z = norm(at - eye), x = norm(up x z), y = norm(z x x); R has the axes as columns, T = -R^T eye."""
    eye, at, up = (np.asarray(v, dtype=np.float64).reshape(3) for v in (eye, at, up))

    def norm(v):
        return v / max(np.linalg.norm(v), 1e-5)
    z = norm(at - eye)
    x = norm(np.cross(up, z))
    y = norm(np.cross(z, x))
    if np.allclose(x, 0.0, atol=5e-3):
        x = norm(np.cross(y, z))
    R = np.stack((x, y, z), axis=1)
    return R, -(R.T @ eye)


class PinholeCamera(object):
    """The constants pytorch3d's camera object would hold, plus the C-ABI struct the kernels take."""

    def __init__(self, R, T, focal_length, principal_point, image_size, device):
        self.focal_length, self.principal_point, self.image_size = focal_length, principal_point, image_size
        self._device = device
        self.set_pose(R, T)

    def set_pose(self, R, T):
        self._R, self._T = np.asarray(R, dtype=np.float64), np.asarray(T, dtype=np.float64)
        self.struct = _lib.make_camera(self._R.tolist(), self._T.tolist(), self.focal_length,
                                       self.principal_point[0], self.principal_point[1])

    @property
    def R(self):
        return torch.tensor(self._R, dtype=torch.float32, device=self._device).unsqueeze(0)

    @property
    def T(self):
        return torch.tensor(self._T, dtype=torch.float32, device=self._device).unsqueeze(0)


class P3dPoseProjection(torch.nn.Module):
    def __init__(self, device=None, pedestrian=None, camera_rgb=None, look_at=None, camera_position=None,
                 image_size=(800, 600), fov=90.0, lens_x_size=0.08, distance=3.1, elevation=1.2, **kwargs):
        super().__init__()
        self._device = torch.device(device) if device is not None else torch.device('cuda', 0)
        self._pedestrian = pedestrian
        if camera_rgb is not None:  # a carla.Sensor (or RGBCameraMock-like object) with .attributes
            attrs = camera_rgb.attributes
            image_size = (int(attrs['image_size_x']), int(attrs['image_size_y']))
            fov, lens_x_size = float(attrs['fov']), float(attrs['lens_x_size'])
        self._image_size = tuple(image_size)
        if camera_position is not None and look_at is not None:
            d, shift, elev = camera_position
            self._eye, self._look_at = (d, shift, -elev), tuple(look_at)
        else:
            if pedestrian is None:
                raise AssertionError('Pedestrian must be provided when camera_position or look_at is not.')
            # RGBCameraMock + get_camera_transform defaults: the camera sits `distance` in front of the pedestrian
            # root at `elevation`, looking at the point (0, 0, elevation) (pose_projection.py:67-75)
            self._look_at, self._eye = (0.0, 0.0, -elevation), (distance, 0.0, -elevation)
        R, T = look_at_view_transform(self._eye, self._look_at, (0.0, 0.0, -1.0))
        sensor_width_mm = lens_x_size * 1000
        focal_length_mm = sensor_width_mm / (2 * np.tan(np.deg2rad(fov) / 2))
        self.camera = PinholeCamera(R, T, float(focal_length_mm * 10),           # p3d_pose_projection.py:63
                                    (self._image_size[0] / 2, self._image_size[1] / 2), self._image_size,
                                    self._device)

    @property
    def image_size(self):
        return self._image_size

    def update_camera(self, camera_position):
        """(distance, shift, elevation) -> new eye, same look-at point (p3d_pose_projection.py:71-89)."""
        d, shift, elev = camera_position
        self._eye = (d, shift, -elev)
        self.camera.set_pose(*look_at_view_transform(self._eye, self._look_at, (0.0, 0.0, -1.0)))

    def forward(self, x, loc, rot):
        """x (N,B,3) absolute p3d locations, loc (N,3) / rot (N,3,3) world transform -> (N,B,3) pixels + 1/z
        (p3d_pose_projection.py:115-152)."""
        return ops.project(x, loc, rot, self.camera.struct)

    def current_pose_to_points(self):
        """Projection of the bound pedestrian's current pose (p3d_pose_projection.py:91-113)."""
        ped = self._pedestrian
        root = ped.transform
        absolute = torch.tensor([(t.location.x, t.location.y, -t.location.z)
                                 for t in ped.current_pose.absolute.values()], device=self._device, dtype=torch.float32)
        loc = torch.tensor(((root.location.x, root.location.y, -root.location.z),), device=self._device,
                           dtype=torch.float32)
        rot = torch.tensor(((np.deg2rad(-root.rotation.roll), np.deg2rad(-root.rotation.pitch),
                             np.deg2rad(-root.rotation.yaw)),), device=self._device, dtype=torch.float32)
        pts = self.forward(absolute.unsqueeze(0), loc, ops.euler_angles_to_matrix(rot, 'XYZ'))[0]
        return pts.cpu().numpy()[..., :2]
