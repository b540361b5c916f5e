from .context import DistContext, init_distributed, shutdown  # noqa: F401
from .aggregation import Aggregator  # noqa: F401
