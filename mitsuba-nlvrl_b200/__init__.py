"""mitsuba-nlvrl_b200: B200-native VRL volumetric gather behind the `vrl` integrator surface of
VincentiusIV/mitsuba-NLVRL.  CUDA core + C ABI live in csrc/ (libvrl_b200.so); this package holds the
host-side mirror (ctypes binding, scene PODs, integrator front-end)."""
from . import abi, scenes, api, sharding  # noqa: F401

__all__ = ["abi", "scenes", "api", "sharding"]
