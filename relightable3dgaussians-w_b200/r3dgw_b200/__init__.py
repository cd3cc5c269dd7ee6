"""Host-side support for the B200-native splat rasterizer + SH relighting path.

Importable packages in this directory (put `relightable3dgaussians-w_b200/` on PYTHONPATH):
  diff_gaussian_rasterization  drop-in for the reference binding of the same name
  simple_knn                   drop-in (`simple_knn._C.distCUDA2`)
  r3dgw_b200                   native loader, fused shading op / EnvironmentLight, synthetic scenes, render mirror
"""
from . import native  # noqa: F401

__version__ = "0.1"
