"""mitsuba2_b200 -- B200-native backend for Mitsuba 2's path-tracing hot path.

Python mirror of the reference's user-facing entry points for this path (``set_variant``, ``load_file``,
``Scene.ray_intersect``, ``Integrator.render`` ...) over the C ABI in ``include/mtsb200.h``.
The variant names ``cuda_b200_rgb`` / ``cuda_b200_spectral`` are accepted; only the rgb one is implemented so far.
"""
from . import _abi
from .scenedata import SceneData, Mesh, look_at, parse_fov, perspective_sample_to_camera
from .scene import Scene, pcg32_floats, sampler_seed
from . import xml as _xml

_variant = None
VARIANTS = ("cuda_b200_rgb", "cuda_b200_spectral")


def variants():
    return list(VARIANTS)


def set_variant(name):
    """``mitsuba.set_variant`` (src/python/__init__.py:115-170)."""
    global _variant
    if name not in VARIANTS:
        raise ImportError(f"Requested an unsupported variant \"{name}\". The following variants are available: {', '.join(VARIANTS)}.")
    if name == "cuda_b200_spectral":
        raise NotImplementedError("cuda_b200_spectral: spectral mode is not implemented yet (see DESIGN.md)")
    _variant = name


def variant():
    return _variant


def load_file(path, device=0, **params):
    """``mitsuba.core.xml.load_file``: parse a Mitsuba XML scene and upload it -> :class:`Scene`."""
    return Scene(_xml.load_file(path, **params), device=device)


def load_string(string, device=0, **params):
    return Scene(_xml.load_string(string, **params), device=device)
