"""Ops: functional surface (``functional``), nn.Module layers (``layers``) and, on a GPU box,
the ctypes bindings of the sm_100a kernels (``cuda``, loaded lazily)."""
from . import functional, layers  # noqa: F401
