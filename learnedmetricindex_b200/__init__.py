"""B200-native implementation of the LMI query path (``LearnedIndex.search``).

Host code is Python/PyTorch (device memory, streams, ``torch.distributed``); every
arithmetic step runs in hand-written sm_100a CUDA behind the C ABI of
``include/lmi_b200.h`` (``liblmi_b200.so``, loaded with ctypes).  No CPU fallback.
"""
from .LearnedIndex import EMPTY_VALUE, LearnedIndex, plan_scan  # noqa: F401
from .layout import DeviceBucketLayout  # noqa: F401
from .model import NeuralNetwork, data_X_to_torch  # noqa: F401

__all__ = ["LearnedIndex", "NeuralNetwork", "DeviceBucketLayout", "EMPTY_VALUE", "plan_scan", "data_X_to_torch"]
