"""raytracing_gpu_b200 — B200-native path-tracing core behind the scene-description API of
daRoyalCacti/Raytracing_GPU.  The product is the C-ABI shared library ``librt_b200.so`` (include/rt_b200.h);
this package is a thin ctypes binding over it plus helpers for tests and benchmarks.  There is no CPU or
PyTorch fallback: anything that needs the GPU raises when the library or the device is missing."""
from . import capi
from .capi import (Scene, RenderSettings, RenderStats, BuildStats, RtError, lib, lib_path, image_height,
                   quantize_frame, ObjMesh, DrawSettings, average_frames, write_ppm, write_png, write_npy, ProgressiveRender, CheckpointInfo,
                   RENDER_COUNTERS, RENDER_TIMERS, RENDER_ACCUMULATE, SEM_REFERENCE, SEM_SPHERE_FAR_ROOT,
                   SEM_UNIT_TRIANGLE_NORMALS, SEM_FIXED)

__all__ = ["Scene", "RenderSettings", "RenderStats", "BuildStats", "RtError", "lib", "lib_path", "image_height",
           "quantize_frame", "ObjMesh", "DrawSettings", "average_frames", "write_ppm",
           "write_png", "write_npy", "ProgressiveRender", "CheckpointInfo", "RENDER_COUNTERS", "RENDER_TIMERS", "RENDER_ACCUMULATE",
           "SEM_REFERENCE", "SEM_SPHERE_FAR_ROOT", "SEM_UNIT_TRIANGLE_NORMALS", "SEM_FIXED"]
