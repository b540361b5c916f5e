from .trainer import ModelTrainer, SlabTrainer  # noqa: F401
from .client import Client  # noqa: F401
from .engine import TorchEngine, SpatialTorchEngine, DataParallelTorchEngine, StepHParams, EvalCounters  # noqa: F401
from .fedavg import FedAvgAPI  # noqa: F401
from .salientgrads import SalientGradsAPI, SailentGradsAPI  # noqa: F401
from .local import LocalAPI  # noqa: F401
from .ditto import DittoAPI  # noqa: F401
from .decentralized import DPSGDAPI, DisPFLAPI, FedFomoAPI, choose_neighbors  # noqa: F401
from .subavg import SubAvgAPI, fake_prune, real_prune, dist_masks, print_pruning  # noqa: F401
from .turboaggregate import TurboAggregateAPI  # noqa: F401

dispflAPI = DisPFLAPI   # the reference's class names
dpsgdAPI = DPSGDAPI
FEDFOMOAPI = FedFomoAPI
