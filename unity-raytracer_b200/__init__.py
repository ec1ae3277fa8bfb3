"""unity-raytracer_b200 — B200-native (sm_100a) implementation of the unity-raytracer per-pixel hot path:
ray generation, ray-sphere / ray-triangle intersection, BVH traversal and the shading/light loop, behind the C-ABI of
include/rtb200.h.  Import with importlib (the directory name carries a hyphen):

    rtb = importlib.import_module("unity-raytracer_b200")

Only what the path needs lives here: csrc/ (CUDA kernels + C-ABI), build.py (nvcc recipe), host.py (ctypes binding and
the mirror of the reference's RayTracingSetup interface), scenes.py (the reference scene + synthetic configs).
"""
from . import scenes            # noqa: F401
from . import host              # noqa: F401
from .host import RayTracer, RayTracingSetup, PipelinedRayTracingSetup, CameraData, ImagePlane, Resolution, load_library, LIB_PATH  # noqa: F401
from .build import build        # noqa: F401
