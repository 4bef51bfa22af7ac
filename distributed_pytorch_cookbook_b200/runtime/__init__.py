from .arena import ParamArena  # noqa: F401
from .graph import GraphedStep  # noqa: F401
from .symm import SymmContext, SymmetricBuffer, symmetric_memory_available  # noqa: F401,E402
