import contextlib

import torch

from .. import _lib

_DTYPE = {torch.float32: _lib.F32, torch.float16: _lib.F16, torch.bfloat16: _lib.BF16,
          torch.uint8: _lib.U8_E5M2}


def dtype_code(t: torch.Tensor, what: str) -> int:
    try:
        return _DTYPE[t.dtype]
    except KeyError:
        raise RuntimeError(f"Unsupported data type: {t.dtype} ({what})") from None


def stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def device_ctx(device: torch.device):
    """Equivalent of the reference's at::cuda::OptionalCUDAGuard(device_of(tensor))."""
    if device.index is None or device.index == torch.cuda.current_device():
        return contextlib.nullcontext()
    return torch.cuda.device(device)


def require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (there is no CPU path)")
