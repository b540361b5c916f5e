from .volumes import VolumeStore, SyntheticABCD, InMemoryVolumeStore, HDF5VolumeStore, ABCD_SHAPE  # noqa: F401
from .partition import (site_partition, rescale_partition, partition_labels, homo_partition, dir_partition,  # noqa: F401
                        n_cls_partition, my_part_partition, hetero_dirichlet_partition,
                        non_iid_partition_with_dirichlet_distribution, record_net_data_stats)
from .loaders import (IndexLoader, TensorBatchLoader, load_partition_data_abcd, load_partition_data_abcd_rescale,  # noqa: F401
                      load_partition_data_images, synthetic_image_dataset)
