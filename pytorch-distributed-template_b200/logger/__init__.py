from .logger import *
from .visualization import *
