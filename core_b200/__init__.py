"""core_b200 — B200-native ray/integrator hot path behind the yafrayInterface_t API."""
from .interface import YafrayInterface  # noqa: F401
