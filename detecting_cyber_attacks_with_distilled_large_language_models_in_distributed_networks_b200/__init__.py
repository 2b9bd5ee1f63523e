"""B200-native federated DDoS detection: DistilBERT-based ``DDoSClassifier`` fine-tuned by federated clients, FedAvg
over NVSwitch peer memory, hand-written sm_100a kernels for every hot op.  See DESIGN.md / INTEGRATION.md."""
from .config import FedConfig  # noqa: F401
from .models.arena import ModelDims, ParamArena  # noqa: F401

__version__ = "0.1.0"
