"""raytracingc_b200 -- B200-native render loop of RayTracingC behind a C ABI.

The product is ``librtc_b200.so`` (CUDA for sm_100a, ``include/rtc_b200.h``);
this package only builds it (``build``) and binds it for tests and the
benchmark (``api``).  Nothing here renders on the CPU.
"""
from . import api  # noqa: F401
from .api import (DEFAULT_SPHERES, SPHERE_DTYPE, TRI_DTYPE, Camera, DeviceJob, DeviceScene, RtcError, Scene,  # noqa: F401
                  camera_look_at, device_count, load_library, make_scene, render, shutdown)

__all__ = ["api", "render", "camera_look_at", "make_scene", "DeviceScene", "DeviceJob", "load_library",
           "device_count", "shutdown", "TRI_DTYPE", "SPHERE_DTYPE", "DEFAULT_SPHERES", "Camera", "Scene", "RtcError"]
