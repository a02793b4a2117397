from .types import *  # noqa: F401,F403
from .io import *  # noqa: F401,F403
from .grid import *  # noqa: F401,F403
from .functions import *  # noqa: F401,F403
