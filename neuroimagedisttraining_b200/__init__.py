"""B200-native federated-learning simulator for 3-D neuroimaging.

A from-scratch framework with the capabilities of bishalth01/NeuroImageDistTraining
(FedAvg / SalientGrads / DisPFL / Ditto / D-PSGD / FedFomo / Local / Sub-FedAvg on
3-D sMRI classifiers), designed for sm_100a: device-resident flat parameter slabs,
hand-written CUDA kernels (``csrc/``) reached through a C ABI (``include/nidt_b200.h``),
and one process per GPU over NCCL / NVLink symmetric memory.
"""
__version__ = "0.1.0"

from .runtime.slab import ParamSlab, SlabLayout  # noqa: F401
