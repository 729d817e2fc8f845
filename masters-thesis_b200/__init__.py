"""think-and-tell-b200: the AttemptFour NIC train step and greedy decode on B200."""
from .model import NIC, Adam, image_groups  # noqa: F401
from .dense_encoder import DenseEncoder  # noqa: F401
from .dataset import DeviceDataGenerator, MultiSubjectDeviceDataGenerator, pad_sequences  # noqa: F401
from . import synthetic, dataset  # noqa: F401
