"""B200-native hot path of seang123/Masters-Thesis ("Think and Tell"): the AttemptFour NIC model family's train step and
decoders behind the reference's Python call surface; all arithmetic in libtnt.so (hand-written sm_100a CUDA)."""
from .model import NIC, Adam, image_groups  # noqa: F401
from .dense_encoder import DenseEncoder  # noqa: F401
from .dataset import DeviceDataGenerator, MultiSubjectDeviceDataGenerator, pad_sequences  # noqa: F401
from . import synthetic, dataset, metrics, keras_io  # noqa: F401
