from .base_data_loader import *
from .base_model import *
from .base_trainer import *
