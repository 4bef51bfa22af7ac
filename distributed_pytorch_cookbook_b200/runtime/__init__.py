from .arena import ParamArena  # noqa: F401
from .graph import GraphedStep  # noqa: F401
