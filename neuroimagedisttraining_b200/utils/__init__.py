from .timing import PhaseTimer  # noqa: F401
from .checkpoint import save_checkpoint, load_checkpoint  # noqa: F401
from .flops import count_training_flops, count_inference_flops, count_model_param_flops  # noqa: F401
from .slim import cosine_annealing, model_difference, hamming_distance  # noqa: F401
from .topology import SymmetricTopologyManager, AsymmetricTopologyManager  # noqa: F401
from .robust import RobustAggregator  # noqa: F401
from .logger import logger_config, logging_config, JsonlMetrics  # noqa: F401
from .context import raise_dist_error, raise_MPI_error, Watchdog  # noqa: F401
