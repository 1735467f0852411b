from .attenders import *  # noqa: F401,F403
from .blocks import *  # noqa: F401,F403
from .convnets import *  # noqa: F401,F403
from .setconv import *  # noqa: F401,F403
