"""Differentiable architecture search (DARTS): search space, searchable network, architect (first / second order),
discrete networks from genotypes and a search loop.  The reference vendors the original DARTS / GDAS scripts as dead code
(``fedml_api/model/cv/darts/*.py``, 1986 lines, ``nn.DataParallel`` + wandb); this is a compact functional equivalent: DARTS and GDAS search spaces, CIFAR / ImageNet genotype networks,
search and training loops, helpers, DOT export."""
from .genotypes import Genotype, PRIMITIVES, DARTS_V2, FedNAS_V1  # noqa: F401
from .operations import OPS, FactorizedReduce, ReLUConvBN  # noqa: F401
from .model_search import Network as SearchNetwork, MixedOp  # noqa: F401
from .model_search_gdas import NetworkGumbelSoftmax, Network_GumbelSoftmax  # noqa: F401
from .model import NetworkCIFAR, NetworkImageNet  # noqa: F401
from .architect import Architect  # noqa: F401
from .train_search import search  # noqa: F401
from .train import train, train_epoch, infer  # noqa: F401
from . import utils, visualize  # noqa: F401
