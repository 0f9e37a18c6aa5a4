"""Builds libsevenn_b200.so in-tree with nvcc for sm_100a (and the generated conv units).

    python -m sevennet_exp_b200.build            # prebuilt signatures (SevenNet-0, l3i5, test models)
"""
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
GEN = os.path.join(CSRC, 'generated')
LIBDIR = os.path.join(HERE, '_lib')
LIB = os.path.join(LIBDIR, 'libsevenn_b200.so')
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '-Xcompiler', '-fPIC']
STATIC_UNITS = ['misc.cu', 'nl.cu', 'edge.cu', 'node.cu', 'gemm_simt.cu', 'gemm_tc.cu', 'model.cu']


def nvcc():
    p = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(p):
        raise RuntimeError('nvcc not found: cannot build the sm_100a kernels')
    return p


def prebuilt_configs():
    """Model configs whose conv signatures ship prebuilt."""
    from .presets import PRESETS
    return PRESETS


def conv_specs(configs):
    from .spec import build_spec
    out = {}
    for cfg in configs.values():
        for L in build_spec(cfg).layers:
            out.setdefault(L.conv.signature(), L.conv)
    return out


def _compile(src, obj):
    if os.path.exists(obj) and os.path.getmtime(obj) >= max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(CSRC, 'sevenn_common.cuh')),
            os.path.getmtime(os.path.join(CSRC, 'conv_common.cuh')),
            os.path.getmtime(os.path.join(CSRC, 'fused_common.cuh')),
            os.path.getmtime(os.path.join(HERE, '..', 'include', 'sevenn_b200.h'))):
        return
    subprocess.check_call([nvcc()] + NVCC_FLAGS + ['-I', CSRC, '-c', src, '-o', obj])


def write_generated(specs):
    from .codegen import conv_codegen, fused_codegen
    os.makedirs(GEN, exist_ok=True)
    paths = []
    for sig, spec in specs.items():
        units = [(f'conv_{sig}.cu', conv_codegen.generate(spec))]
        # experimental fused radial-MLP + convolution forward (correct, but slower than GEMM + conv on B200: see
        # DESIGN.md 6.3); built only on request, nothing in the product path calls it
        if os.environ.get('SEVENN_BUILD_FUSED') == '1' and fused_codegen.supported(spec):
            units.append((f'fused_{sig}.cu', fused_codegen.generate(spec)))
        for name, text in units:
            p = os.path.join(GEN, name)
            if not os.path.exists(p) or open(p).read() != text:
                with open(p, 'w') as f:
                    f.write(text)
            paths.append(p)
    return paths


def build(extra_specs=None, verbose=True):
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, 'obj')
    os.makedirs(objdir, exist_ok=True)
    specs = conv_specs(prebuilt_configs())
    if extra_specs:
        specs.update(extra_specs)
    srcs = [os.path.join(CSRC, u) for u in STATIC_UNITS] + write_generated(specs)
    # previously generated units for other signatures stay part of the library
    for f in sorted(os.listdir(GEN)):
        p = os.path.join(GEN, f)
        if f.endswith('.cu') and p not in srcs:
            srcs.append(p)
    objs = [os.path.join(objdir, os.path.basename(s)[:-3] + '.o') for s in srcs]
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        list(ex.map(lambda a: _compile(*a), zip(srcs, objs)))
    tmp = LIB + '.tmp'
    subprocess.check_call([nvcc(), '-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-o', tmp] + objs + ['-lcudart', '-ldl'])
    os.replace(tmp, LIB)
    if verbose:
        print(f'built {LIB} ({len(srcs)} units, {len(specs)} conv signatures)')
    return LIB


if __name__ == '__main__':
    build()
