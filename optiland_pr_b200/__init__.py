"""optiland_pr_b200 -- fused sequential real-ray tracing on NVIDIA B200 (sm_100a).

One hot path of the Optiland optical-design package, `SurfaceGroup.trace` on `RealRays`
(optiland/surfaces/surface_group.py:232-244), as hand-written CUDA behind the C ABI in
include/olrt.h. See DESIGN.md.
"""
from . import backend  # noqa: F401
from .apertures import (DifferenceAperture, EllipticalAperture, IntersectionAperture,  # noqa: F401
                        OffsetRadialAperture, RadialAperture, RectangularAperture, UnionAperture)
from .coordinate_system import CoordinateSystem  # noqa: F401
from .geometries import (EvenAsphere, OddAsphere, Plane, PolynomialGeometry,  # noqa: F401
                         StandardGeometry, ZernikePolynomialGeometry)
from .materials import IdealMaterial, Material, SellmeierMaterial  # noqa: F401
from .optic import Optic  # noqa: F401
from .prescription import NotEligible, PackedSystem, pack_surface_group  # noqa: F401
from .rays import RealRays  # noqa: F401
from .surfaces import ObjectSurface, Surface, SurfaceGroup  # noqa: F401
from .trace import TraceResult, fused_trace  # noqa: F401

__version__ = "0.1.0"
