from .timing import PhaseTimer  # noqa: F401
from .checkpoint import save_checkpoint, load_checkpoint  # noqa: F401
