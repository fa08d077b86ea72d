"""vllm._C.cuda_utils (reference csrc/pybind.cpp:101-111, csrc/cuda_utils_kernels.cu:5-35)."""
from vllm_splitwise_b200 import _lib


def get_device_attribute(attribute: int, device_id: int) -> int:
    return _lib.load().b200_get_device_attribute(int(attribute), int(device_id))


def get_max_shared_memory_per_block_device_attribute(device_id: int) -> int:
    return _lib.load().b200_get_max_shared_memory_per_block_device_attribute(int(device_id))
