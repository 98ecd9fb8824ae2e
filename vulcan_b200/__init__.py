"""vulcan_b200 — B200-native (sm_100a) implementation of Vulcan's TSDF-fusion hot path:
Volume::SetView, Depth/Color/LightIntegrator::Integrate, Tracer::Trace.

Importing this package loads vulcan_b200/libvulcan_b200.so (the C ABI of
include/vulcan_b200.h) and fails loudly if it has not been built.
"""
from ._lib import lib, VulcanError, LIB_PATH  # noqa: F401
from .api import (Volume, Integrator, DepthIntegrator, ColorIntegrator, LightIntegrator, Tracer, Frame,  # noqa: F401
                  Transform, Projection, SetDefaultMathMode, GetDefaultMathMode, DepthTracker, PyramidTracker, Extractor, Exporter, Mesh)
from ._lib import VK_MATH_REFERENCE, VK_MATH_STRICT  # noqa: F401
