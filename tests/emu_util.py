"""TEST INFRASTRUCTURE: compiles generated kernel text with -DSEVENN_EMU (plain C++, a warp
is a lane loop) so that the generated arithmetic can be checked against the oracle on a box
without a GPU.  Never used by the product."""
import ctypes
import hashlib
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'sevennet_exp_b200', 'csrc')


def build_emu(src_text: str) -> ctypes.CDLL:
    h = hashlib.sha1(src_text.encode()).hexdigest()[:12]
    d = os.path.join(tempfile.gettempdir(), 'sevenn_emu')
    os.makedirs(d, exist_ok=True)
    so = os.path.join(d, f'emu_{h}.so')
    if not os.path.exists(so):
        cu = os.path.join(d, f'emu_{h}.cpp')
        with open(cu, 'w') as f:
            f.write(src_text)
        subprocess.check_call(['g++', '-O1', '-DSEVENN_EMU', '-I', CSRC, '-shared', '-fPIC', '-o', so, cu])
    return ctypes.CDLL(so)


def fptr(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags['C_CONTIGUOUS']
    return a.ctypes.data_as(ctypes.c_void_p)


def iptr(a: np.ndarray):
    assert a.dtype == np.int32 and a.flags['C_CONTIGUOUS']
    return a.ctypes.data_as(ctypes.c_void_p)
