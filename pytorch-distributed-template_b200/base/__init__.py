"""Base classes of the plug-in surface (reference: base/__init__.py)."""
from .base_data_loader import BaseDataLoader  # noqa: F401
from .base_model import BaseModel  # noqa: F401
from .base_trainer import BaseTrainer  # noqa: F401

__all__ = ["BaseDataLoader", "BaseModel", "BaseTrainer"]
