from .io import *  # noqa: F401,F403
