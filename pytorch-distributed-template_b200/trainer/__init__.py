from .trainer import *
