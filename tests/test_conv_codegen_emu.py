"""Generated conv kernel text (emulation flavour) vs the oracle's e3nn restatement:
forward message+segment-sum, and the analytic backward (dW, dx, dY) vs autograd."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import e3nn_ref as o3
from oracle import sevenn_ref as S
from sevennet_exp_b200 import irreps as ir
from sevennet_exp_b200 import spec as sp
from sevennet_exp_b200.codegen import conv_codegen as cg
from tests.emu_util import build_emu, fptr, iptr

CASES = {
    'o3_ch4': ('4x0e+4x1o+4x2e', -1, 2, '12x0e+16x1o+8x1e+8x2o+16x2e'),
    'so3_l2': ('64x0e+32x1e+32x2e', 1, 2, '1x0e+1x1e+1x2e'),
    'so3_l3': ('64x0e+32x1e+32x2e+32x3e', 1, 3, '1x0e+1x1e+1x2e+1x3e'),
    'last': ('32x0e+32x1e+32x2e', 1, 2, '1x0e'),
}


def _setup(case):
    xs, p, lmax, outs = CASES[case]
    irx = ir.parse_irreps(xs)
    filt = ir.spherical_harmonics_irreps(lmax, p)
    out = ir.parse_irreps(outs)
    conv = sp.make_conv('c', irx, filt, out, [8, 64, 64])
    oc = S.IrrepsConvolution(o3.Irreps(xs), o3.Irreps([(1, (l, pp)) for _, (l, pp) in filt]),
                             o3.Irreps(outs), [8, 64, 64], torch.nn.functional.silu, denominator=1.7)
    assert oc.weight_numel == conv.wn
    assert str(oc.irreps_mid) == ir.irreps_str(conv.lmid.entries)
    return conv, oc


@pytest.mark.parametrize('case', list(CASES))
def test_conv_fwd_bwd_emu(case):
    conv, oc = _setup(case)
    g = torch.Generator().manual_seed(1)
    n, E = 7, 40
    dst = torch.sort(torch.randint(0, n, (E,), generator=g))[0]
    src = torch.randint(0, n, (E,), generator=g)
    x = torch.randn(n, conv.lx.dim, generator=g, dtype=torch.float64, requires_grad=True)
    Ssh = (conv.lmax_sh + 1) ** 2
    Y = torch.randn(E, Ssh, generator=g, dtype=torch.float64, requires_grad=True)
    w = torch.randn(E, conv.wn, generator=g, dtype=torch.float64, requires_grad=True)
    msg = oc.convolution.double()(x[src], Y, w)
    ref = torch.zeros(n, msg.shape[1], dtype=torch.float64).index_add_(0, dst, msg) / 1.7
    gout = torch.randn(ref.shape, generator=g, dtype=torch.float64)
    gx, gY, gw = torch.autograd.grad(ref, [x, Y, w], gout)

    lib = build_emu(cg.generate(conv))
    sig = cg.signature(conv)
    rowptr = np.zeros(n + 1, np.int32)
    np.add.at(rowptr, dst.numpy() + 1, 1)
    rowptr = np.cumsum(rowptr).astype(np.int32)
    col = src.numpy().astype(np.int32)
    xd = np.ascontiguousarray(conv.lx.from_e3nn(x.detach()).numpy().astype(np.float32))
    Yd = np.ascontiguousarray(Y.detach().numpy().astype(np.float32))
    wd = np.ascontiguousarray(w.detach().numpy().astype(np.float32))
    out = np.zeros((n, conv.lmid.dim), np.float32)
    f = getattr(lib, f'sevenn_conv_fwd_{sig}')
    f.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 4 + [ctypes.c_int, ctypes.c_float,
                                                                            ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    assert f(fptr(xd), conv.lx.dim, fptr(Yd), fptr(wd), iptr(rowptr), iptr(col), n, 1 / 1.7,
             fptr(out), conv.lmid.dim, None) == 0
    got = conv.lmid.to_e3nn(torch.from_numpy(out)).double()
    assert torch.allclose(got, ref, atol=2e-5, rtol=1e-5), (got - ref).abs().max()

    god = np.ascontiguousarray(conv.lmid.from_e3nn(gout).numpy().astype(np.float32))
    dW = np.zeros((E, conv.wn), np.float32)
    dxe = np.zeros((E, conv.lx.dim), np.float32)
    nu = cg.n_units(conv)
    dY = np.zeros((nu, E, Ssh), np.float32)           # one slab per (irrep group, 32-channel chunk)
    b = getattr(lib, f'sevenn_conv_bwd_{sig}')
    b.argtypes = ([ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 4 + [ctypes.c_int, ctypes.c_float,
                  ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                  ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p])
    assert b(fptr(xd), conv.lx.dim, fptr(Yd), fptr(wd), iptr(rowptr), iptr(col), n, 1 / 1.7,
             fptr(god), conv.lmid.dim, fptr(dW), fptr(dxe), conv.lx.dim, fptr(dY), E * Ssh, None) == 0
    assert torch.allclose(torch.from_numpy(dW).double(), gw, atol=5e-5, rtol=1e-4)
    assert torch.allclose(torch.from_numpy(dY).double().sum(0), gY, atol=2e-4, rtol=1e-4)
    dx = torch.zeros(n, conv.lx.dim, dtype=torch.float64).index_add_(0, src, torch.from_numpy(dxe).double())
    assert torch.allclose(conv.lx.to_e3nn(dx), gx, atol=2e-4, rtol=1e-4)


def test_committed_generated_kernels_match_the_generator():
    """csrc/generated/*.cu is generator output kept in the tree (the hot kernels, readable and diffable): regenerating
    every prebuilt signature must give byte-identical files, so the committed text cannot drift from conv_codegen.py."""
    import os
    from sevennet_exp_b200 import build
    from sevennet_exp_b200.codegen import conv_codegen
    specs = build.conv_specs(build.prebuilt_configs())
    assert len(specs) >= 5
    for sig, spec in specs.items():
        path = os.path.join(build.GEN, f'conv_{sig}.cu')
        assert os.path.exists(path), path
        assert open(path).read() == conv_codegen.generate(spec), f'{path} differs from the generator output'
