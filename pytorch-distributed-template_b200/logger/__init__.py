"""Logging plug-ins (reference: logger/__init__.py)."""
from .logger import setup_logging  # noqa: F401
from .visualization import TensorboardWriter  # noqa: F401

__all__ = ["setup_logging", "TensorboardWriter"]
