from .message import Message  # noqa: F401
from .base import (Observer, BaseCommunicationManager, InProcessCommManager, TorchDistCommManager, GrpcCommManager,  # noqa: F401
                   MqttCommManager)
from .managers import (ClientManager, ServerManager, make_comm_manager, FedAvgServerManager, FedAvgClientManager,  # noqa: F401
                       SalientGradsServerManager, SalientGradsClientManager)  # noqa: F401
from .mqtt_mini import MiniMqttBroker, MiniMqttClient  # noqa: F401

# the reference's class names (fedml_core/distributed/communication/{mpi/com_manager.py:13, gRPC/grpc_comm_manager.py:20})
MpiCommunicationManager = TorchDistCommManager      # point-to-point object passing between ranks; torch.distributed instead of mpi4py
GRPCCommManager = GrpcCommManager
