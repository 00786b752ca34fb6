"""raytracer-challenge_b200 — B200-native ``Camera::render`` hot path of joedane/raytracer-challenge.

Layout (only what the path needs):
  csrc/        hand-written sm_100a CUDA kernels + the C ABI of include/rtc.h (librtc_b200.so)
  transform.py, scene.py, camera.py   host-side mirror of the reference API (World, Camera, ...)
  scenes.py    the workload scenes of BASELINE.json
  ffi.py       ctypes declarations of the C ABI

The directory name has a hyphen (it is the name the task fixes), so import it with
``importlib.import_module("raytracer-challenge_b200")`` or through the root-level alias
``rtc_b200``.
"""
from . import ffi  # noqa: F401
from .camera import Camera, Canvas, Renderer, default_renderer  # noqa: F401
from .scene import (CheckerPattern, Color, Cube, GradientPattern, GridPattern, Light, Material,  # noqa: F401
                    Pattern, Plane, RingPattern, Shape, Sphere, StripePattern, TestPattern, World)
from .transform import EPSILON, Matrix  # noqa: F401
from . import scenes  # noqa: F401
from . import loader  # noqa: F401,E402  (scene front-ends: YAML dialect, Lua-table schema)
