from .util import *
from .dist import *
