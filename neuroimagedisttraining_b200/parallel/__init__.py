from .context import DistContext, init_distributed, shutdown  # noqa: F401
from .aggregation import Aggregator  # noqa: F401
from .syncbn import SynchronizedBatchNorm, SynchronizedBatchNorm3d, convert_sync_batchnorm  # noqa: F401
