from .trainer import ModelTrainer, SlabTrainer  # noqa: F401
from .client import Client  # noqa: F401
from .engine import TorchEngine, StepHParams, EvalCounters  # noqa: F401
from .fedavg import FedAvgAPI  # noqa: F401
from .salientgrads import SalientGradsAPI, SailentGradsAPI  # noqa: F401
