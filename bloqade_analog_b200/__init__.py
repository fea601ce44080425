"""bloqade_analog_b200 - B200-native emulator backend for bloqade-analog's Python-emulator hot path."""
from . import ir  # noqa: F401
from .space import Space, SpaceType  # noqa: F401

__version__ = "0.1.0"
