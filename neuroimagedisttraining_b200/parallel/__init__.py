from .context import DistContext, ReplicatedContext, init_distributed, shutdown  # noqa: F401
from .aggregation import Aggregator  # noqa: F401
from .syncbn import SynchronizedBatchNorm, SynchronizedBatchNorm3d, convert_sync_batchnorm  # noqa: F401
from .spatial import (DepthPartition, SlabParallelFeatures, SpatialParallelModel, allreduce_gradients, gather_depth, reslab, scatter_depth,  # noqa: F401
                      slab_avg_pool3d, slab_conv3d, slab_max_pool3d)
