"""Experiment builds of the library: the same sources with extra -D flags, written beside the product's .so and
selected at run time with P2C_LIB=<path>.  Built HERE (nvcc cross-compiles), they travel to the GPU box.

  python tools/variants.py ieee=P2C_IEEE_MATH dsum=P2C_SUMS_DOUBLE both=P2C_IEEE_MATH,P2C_SUMS_DOUBLE
  P2C_LIB=pedestrians_video_2_carla_b200/libp2c_ieee.so python tools/grad_error_table.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pedestrians_video_2_carla_b200 import _build  # noqa: E402


def main():
    for spec in sys.argv[1:]:
        name, defs = spec.split('=', 1)
        out = os.path.join(_build.PKG_DIR, 'libp2c_%s.so' % name)
        # NAME=DEF1,DEF2 : extra -D flags;  NAME=file.cu:/path/to/other_version.cu : swap one source file
        overrides = {d.split(':', 1)[0]: d.split(':', 1)[1] for d in defs.split(',') if ':' in d}
        print(_build.build(force=True, defines=[d for d in defs.split(',') if d and ':' not in d], out=out,
                           overrides=overrides))


if __name__ == '__main__':
    main()
