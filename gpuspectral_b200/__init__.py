"""gpuspectral_b200 — B200-native path-tracing core (CUDA sm_100a) behind the render path of GPUSpectral.

Importing this package loads the native libraries; it fails loudly when they are missing.
"""
from ._native import (BSDF_DIFFUSE, BSDF_NAMES, BSDF_ROUGH_CONDUCTOR, BSDF_ROUGH_FLOOR, BSDF_ROUGH_PLASTIC,
                      BSDF_SMOOTH_CONDUCTOR, BSDF_SMOOTH_DIELECTRIC, BSDF_SMOOTH_FLOOR, BSDF_SMOOTH_PLASTIC,
                      BSDF_STRIDES, BSDF_TYPE_COUNT, NO_HIT)
from .api import HIT_DTYPE, RAY_DTYPE, Context, GsbError, PathTracer, Scene, image_diff, make_rays, read_pfm, write_pfm

__all__ = ["Context", "PathTracer", "Scene", "GsbError", "make_rays", "read_pfm", "write_pfm", "image_diff", "RAY_DTYPE", "HIT_DTYPE", "NO_HIT",
           "BSDF_TYPE_COUNT", "BSDF_NAMES", "BSDF_STRIDES"]
