"""Builds libsevenn_b200.so in-tree with nvcc for sm_100a (and the generated conv units).

    python -m sevennet_exp_b200.build            # prebuilt signatures (SevenNet-0, l3i5, test models)
"""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
GEN = os.path.join(CSRC, 'generated')
LIBDIR = os.path.join(HERE, '_build')
LIB = os.path.join(LIBDIR, 'libsevenn_b200.so')
TORCH_LIB = os.path.join(LIBDIR, 'libsevenn_b200_torch.so')      # torch.ops.sevenn_b200.* (csrc/torch_op.cpp)
PAIR_HARNESS = os.path.join(LIBDIR, 'sevenn_pair_harness')        # libtorch client like pair_e3gnn.cpp (csrc/pair_harness.cpp)
MD_HOST = os.path.join(LIBDIR, 'sevenn_md_host')                  # torch-free C++ host on the C ABI (csrc/md_host.cpp)
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '-Xcompiler', '-fPIC']
STATIC_UNITS = ['misc.cu', 'nl.cu', 'edge.cu', 'node.cu', 'gemm_simt.cu', 'gemm_tc.cu', 'model.cu', 'train.cu']


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


def build_torch_op(verbose=True):
    """libsevenn_b200_torch.so: the TORCH_LIBRARY registration of torch.ops.sevenn_b200.evaluate (csrc/torch_op.cpp, the one
    unit that includes torch headers) linked against libsevenn_b200.so next to it ($ORIGIN rpath)."""
    import sysconfig
    import torch
    import torch.utils.cpp_extension as ce
    src = os.path.join(CSRC, 'torch_op.cpp')
    hdr = os.path.join(HERE, '..', 'include', 'sevenn_b200.h')
    if os.path.exists(TORCH_LIB) and os.path.getmtime(TORCH_LIB) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return TORCH_LIB
    tlib = os.path.join(os.path.dirname(torch.__file__), 'lib')
    cuda_home = os.environ.get('CUDA_HOME', '/usr/local/cuda')
    cmd = ['g++', '-O2', '-std=c++17', '-fPIC', '-shared', f'-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}']
    for inc in ce.include_paths() + [os.path.join(cuda_home, 'include'), sysconfig.get_paths()['include']]:
        cmd += ['-I', inc]
    cmd += [src, '-o', TORCH_LIB + '.tmp', '-L', LIBDIR, '-lsevenn_b200', '-L', tlib, '-ltorch', '-ltorch_cpu',
            '-lc10', '-ltorch_cuda', '-lc10_cuda', '-Wl,-rpath,$ORIGIN', f'-Wl,-rpath,{tlib}']
    subprocess.check_call(cmd)
    os.replace(TORCH_LIB + '.tmp', TORCH_LIB)
    if verbose:
        print(f'built {TORCH_LIB}')
    return TORCH_LIB


def build_md_host(verbose=True):
    """sevenn_md_host: a C++ program (no torch, no Python) that drives libsevenn_b200.so the way a pair style would."""
    src = os.path.join(CSRC, 'md_host.cpp')
    hdr = os.path.join(HERE, '..', 'include', 'sevenn_b200.h')
    if os.path.exists(MD_HOST) and os.path.getmtime(MD_HOST) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return MD_HOST
    cuda_home = os.environ.get('CUDA_HOME', '/usr/local/cuda')
    subprocess.check_call(['g++', '-O2', '-std=c++17', src, '-o', MD_HOST + '.tmp', '-I', os.path.join(cuda_home, 'include'),
                           '-L', LIBDIR, '-lsevenn_b200', '-L', os.path.join(cuda_home, 'lib64'), '-lcudart',
                           '-Wl,-rpath,$ORIGIN', f'-Wl,-rpath,{os.path.join(cuda_home, "lib64")}'])
    os.replace(MD_HOST + '.tmp', MD_HOST)
    if verbose:
        print(f'built {MD_HOST}')
    return MD_HOST


def build_pair_harness(verbose=True):
    """sevenn_pair_harness: a libtorch program that loads a deployed archive and calls it exactly like the reference's
    LAMMPS pair style does (csrc/pair_harness.cpp); links libtorch only -- the op library is dlopen'ed at run time."""
    import torch
    import torch.utils.cpp_extension as ce
    src = os.path.join(CSRC, 'pair_harness.cpp')
    if os.path.exists(PAIR_HARNESS) and os.path.getmtime(PAIR_HARNESS) >= os.path.getmtime(src):
        return PAIR_HARNESS
    tlib = os.path.join(os.path.dirname(torch.__file__), 'lib')
    cuda_home = os.environ.get('CUDA_HOME', '/usr/local/cuda')
    cmd = ['g++', '-O2', '-std=c++17', f'-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}']
    for inc in ce.include_paths() + [os.path.join(cuda_home, 'include')]:
        cmd += ['-I', inc]
    cmd += [src, '-o', PAIR_HARNESS + '.tmp', '-L', tlib, '-Wl,--no-as-needed', '-ltorch', '-ltorch_cpu', '-lc10',
            '-ltorch_cuda', '-lc10_cuda', '-Wl,--as-needed', '-ldl', f'-Wl,-rpath,{tlib}']
    subprocess.check_call(cmd)
    os.replace(PAIR_HARNESS + '.tmp', PAIR_HARNESS)
    if verbose:
        print(f'built {PAIR_HARNESS}')
    return PAIR_HARNESS


if __name__ == '__main__':
    build()
    build_torch_op()
    build_md_host()
    build_pair_harness()
