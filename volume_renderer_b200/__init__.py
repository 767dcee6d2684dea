"""volume_renderer_b200 - B200-native ray-march engine behind the reference's MATLAB-facing API.

Only the hot path of raphiniert-com/volume_renderer lives here: `csrc/` holds the CUDA kernels
and the C ABI (include/vr_abi.h), `api.py` mirrors the MATLAB classes that call it.
"""
from .api import (HenyeyGreenstein, LightSource, StereoRenderMode, Volume,
                  VolumeRender, timestamp)

__all__ = ["VolumeRender", "Volume", "LightSource", "StereoRenderMode", "HenyeyGreenstein", "timestamp"]
