"""`trainer` package: the hot loop (reference: trainer/__init__.py)."""
from .trainer import Trainer  # noqa: F401

__all__ = ["Trainer"]
