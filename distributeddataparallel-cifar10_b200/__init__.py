"""B200-native NetResDeep data-parallel training step (drop-in for the hot path of
BaamPark/DistributedDataParallel-Cifar10).  Importing this package loads libnrd.so; there is no
CPU fallback."""
from . import _lib  # noqa: F401  (fails loudly when the CUDA library is missing)
from .model.resnet import NetResDeep, ResBlock  # noqa: F401
from .nn import CrossEntropyLoss, BCEWithLogitsLoss, binary_cross_entropy_with_logits  # noqa: F401
from .optim import SGD  # noqa: F401
from .parallel import DistributedDataParallel  # noqa: F401
from .data import SyntheticCifar10Loader, DevicePrefetcher, Cifar10DeviceLoader  # noqa: F401
from .graph import TrainStep  # noqa: F401
from . import main, main_no_ddp  # noqa: F401,E402
