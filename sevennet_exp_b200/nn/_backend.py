"""Library handle + launch plumbing shared by the module wrappers."""
import ctypes

import torch

from .. import _lib
from ..engine import Kernels

_override = None
_kernels = {}


def set_lib(lib):
    """Tests only: inject the CPU emulation build of the kernel sources.  The product path always goes
    through ``_lib.load()``, which raises without the CUDA library or without a GPU."""
    global _override
    _override = lib
    _kernels.clear()


def get_lib():
    return _override if _override is not None else _lib.load()


def kernels(device) -> Kernels:
    device = torch.device(device)
    if device.type == 'cuda' and device.index is None:
        device = torch.device('cuda', torch.cuda.current_device())
    key = (device.type, device.index)
    if key not in _kernels:
        _kernels[key] = Kernels(get_lib(), device)
    return _kernels[key]


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None
