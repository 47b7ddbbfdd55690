"""Plain value types standing in for carla.Location / carla.Rotation / carla.Transform on the tensor path
(the reference falls back to carla_utils/mock_carla.py when the CARLA client is not importable; only the field
names matter to walker_control/p3d_pose.py:34-96).  If the real `carla` module is importable it is used instead."""
try:  # pragma: no cover - CARLA is not installed on the build / GPU boxes
    from carla import Location, Rotation, Transform  # noqa: F401
except ImportError:
    class Location(object):
        __slots__ = ('x', 'y', 'z')

        def __init__(self, x=0.0, y=0.0, z=0.0):
            self.x, self.y, self.z = float(x), float(y), float(z)

        def __repr__(self):
            return 'Location(x=%g, y=%g, z=%g)' % (self.x, self.y, self.z)

    class Rotation(object):
        __slots__ = ('pitch', 'yaw', 'roll')

        def __init__(self, pitch=0.0, yaw=0.0, roll=0.0):
            self.pitch, self.yaw, self.roll = float(pitch), float(yaw), float(roll)

        def __repr__(self):
            return 'Rotation(pitch=%g, yaw=%g, roll=%g)' % (self.pitch, self.yaw, self.roll)

    class Transform(object):
        __slots__ = ('location', 'rotation')

        def __init__(self, location=None, rotation=None):
            self.location = location if location is not None else Location()
            self.rotation = rotation if rotation is not None else Rotation()

        def __repr__(self):
            return 'Transform(%r, %r)' % (self.location, self.rotation)
