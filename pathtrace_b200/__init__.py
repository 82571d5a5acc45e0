"""pathtrace_b200 -- a B200-native wavefront path-tracing backend behind the pathtrace scene/config formats.

Only what the hot path needs: csrc/ (CUDA kernels + C ABI + host scene compiler), the ctypes binding and the
host-side mirror of the reference's Renderer lifecycle.
"""
from ._capi import B200ptError, LIB_PATH, lib  # noqa: F401
from .renderer import Config, Scene, Wavefront, default_config_dict, render_file, renderer_from_config  # noqa: F401

__all__ = ["B200ptError", "Config", "Scene", "Wavefront", "default_config_dict", "render_file", "renderer_from_config",
           "lib", "LIB_PATH"]
