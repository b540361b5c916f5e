from .message import Message  # noqa: F401
from .base import (Observer, BaseCommunicationManager, InProcessCommManager, TorchDistCommManager, GrpcCommManager,  # noqa: F401
                   MqttCommManager)
from .managers import (ClientManager, ServerManager, make_comm_manager, FedAvgServerManager, FedAvgClientManager)  # noqa: F401
from .mqtt_mini import MiniMqttBroker, MiniMqttClient  # noqa: F401
