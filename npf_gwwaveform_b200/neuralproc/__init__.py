from .family import *  # noqa: F401,F403
