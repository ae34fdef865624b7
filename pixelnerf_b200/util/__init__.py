from .util import *  # noqa: F401,F403
from . import util  # noqa: F401
