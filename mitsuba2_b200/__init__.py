"""mitsuba2_b200 -- B200-native backend for Mitsuba 2's path-tracing hot path.

Python mirror of the reference's user-facing entry points for this path (``set_variant``, ``load_file``,
``Scene.ray_intersect``, ``Integrator.render`` ...) over the C ABI in ``include/mtsb200.h``.
The variants ``cuda_b200_rgb`` (scalar_rgb semantics) and ``cuda_b200_spectral`` (scalar_spectral: 4 wavelengths per path)
select the colour mode of scenes loaded afterwards.
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
    _variant = name


def variant():
    return _variant


def _with_mode(sd):
    sd.mode = "spectral" if _variant == "cuda_b200_spectral" else "rgb"
    return sd


def load_file(path, device=0, **params):
    """``mitsuba.core.xml.load_file``: parse a Mitsuba XML scene and upload it -> :class:`Scene`."""
    return Scene(_with_mode(_xml.load_file(path, **params)), device=device)


def load_string(string, device=0, **params):
    return Scene(_with_mode(_xml.load_string(string, **params)), device=device)
