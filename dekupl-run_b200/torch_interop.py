"""zero-copy view of library-owned device memory as a torch tensor (for NCCL collectives)"""
import torch

_TYPESTR = {torch.float64: "<f8", torch.int64: "<i8", torch.int32: "<i4", torch.uint8: "|u1"}


class _DevArray:
    def __init__(self, ptr, n, dtype):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": _TYPESTR[dtype], "data": (int(ptr), False), "version": 2}


def tensor_from_ptr(ptr, n, dtype, device):
    """the tensor aliases the library's buffer: use it before the owning object is freed"""
    return torch.as_tensor(_DevArray(ptr, n, dtype), device=device)
