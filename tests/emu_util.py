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


def build_full_emu(conv_specs):
    """All static units + the given conv signatures, emulation flavour, as a sevennet_exp_b200._lib.Lib."""
    from sevennet_exp_b200 import _lib
    from sevennet_exp_b200.codegen import conv_codegen
    units = ['misc.cu', 'nl.cu', 'edge.cu', 'node.cu', 'gemm_simt.cu', 'gemm_tc.cu', 'train.cu']
    text = ''
    for u in units:
        text += f'// ---- {u}\n' + open(os.path.join(CSRC, u)).read() + '\n'
    body = {s.signature(): conv_codegen.generate(s) for s in conv_specs}
    h = hashlib.sha1((text + ''.join(sorted(body.values()))).encode()).hexdigest()[:12]
    d = os.path.join(tempfile.gettempdir(), 'sevenn_emu')
    os.makedirs(d, exist_ok=True)
    so = os.path.join(d, f'full_{h}.so')
    if not os.path.exists(so):
        objs = []
        for k, u in enumerate(units):
            o = os.path.join(d, f'full_{h}_{k}.o')
            subprocess.check_call(['g++', '-O1', '-DSEVENN_EMU', '-x', 'c++', '-I', CSRC, '-fPIC', '-c',
                                   os.path.join(CSRC, u), '-o', o])
            objs.append(o)
        for sig, t in body.items():
            cpp = os.path.join(d, f'full_{h}_{sig}.cpp')
            open(cpp, 'w').write(t)
            o = cpp[:-4] + '.o'
            subprocess.check_call(['g++', '-O1', '-DSEVENN_EMU', '-I', CSRC, '-fPIC', '-c', cpp, '-o', o])
            objs.append(o)
        subprocess.check_call(['g++', '-shared', '-o', so] + objs)
    return _lib.Lib(ctypes.CDLL(so))


def graph_from_edges(types, edge_index, edge_vec, n_local, volume, device='cpu'):
    """Engine Graph from a reference-style (dst-sorted) edge list (host construction, tests only)."""
    import torch
    from sevennet_exp_b200.engine import Graph
    ei = np.asarray(edge_index)
    n = len(types)
    order = np.lexsort((np.arange(ei.shape[1]), ei[0]))
    assert np.all(order == np.arange(ei.shape[1])), 'edge list must be sorted by edge_index[0]'
    rowptr = np.zeros(n_local + 1, np.int64)
    np.add.at(rowptr, ei[0] + 1, 1)
    rowptr = np.cumsum(rowptr)
    corder = np.lexsort((np.arange(ei.shape[1]), ei[1]))
    cptr = np.zeros(n + 1, np.int64)
    np.add.at(cptr, ei[1] + 1, 1)
    cptr = np.cumsum(cptr)
    t = lambda a, dt: torch.tensor(np.ascontiguousarray(a), dtype=dt, device=device)  # noqa: E731
    return Graph(t(types, torch.int32), t(edge_vec, torch.float32), t(rowptr, torch.int32), t(ei[1], torch.int32),
                 t(cptr, torch.int32), t(corder, torch.int32), n_local, volume)
