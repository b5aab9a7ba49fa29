from .shell import *  # noqa: F401,F403
from .grid import *  # noqa: F401,F403
from .snapshot import *  # noqa: F401,F403
