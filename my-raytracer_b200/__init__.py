"""my-raytracer_b200 — B200-native (sm_100a) implementation of the ray tracer's hot path:
primary rays -> median-split-BVH traversal -> Phong shading with shadow rays and mirror
reflections, behind a C ABI (include/rt_b200.h) that takes the reference's own Data / BVHNodes_SoA
layouts.

The directory name has a hyphen (task layout), so import it through ``rtpkg.load()`` at the repo
root, which registers it as ``my_raytracer_b200``.

Nothing here falls back to the CPU: ``capi.load_library()`` raises if the CUDA extension has not
been built, and every render entry point returns RT_ERR_NO_DEVICE without a GPU.
"""
from . import blobfmt, capi, scene, scene_files, scenes  # noqa: F401
from .capi import load_library  # noqa: F401
from .scene_files import FileScene, read_image, write_image, write_scene  # noqa: F401
from .scene import DeviceScene, HostScene, MeshDesc, SceneDesc, material, quantize_rgb8  # noqa: F401
