"""The C-ABI library loads (no GPU needed) and exports every symbol include/sevenn_b200.h declares."""
import os
import re

from sevennet_exp_b200 import _lib, build, presets, spec as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ensure_built():
    if not os.path.exists(_lib.LIB_PATH):
        build.build(verbose=False)


def test_header_symbols_exported():
    _ensure_built()
    hdr = open(os.path.join(ROOT, 'include', 'sevenn_b200.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    names = re.findall(r'^\s*(?:int|long long)\s+(sevenn_\w+)\s*\(', hdr, flags=re.M)
    assert len(names) >= 20
    cdll = _lib.load_for_symbol_check()
    for n in names:
        assert hasattr(cdll, n), f'{n} declared in the header but not exported'
        assert n in _lib.SIGNATURES, f'{n} has no ctypes signature'
    assert cdll.sevenn_abi_version() == 1


def test_prebuilt_conv_signatures_exported():
    _ensure_built()
    cdll = _lib.load_for_symbol_check()
    for name, cfg in presets.PRESETS.items():
        for L in sp.build_spec(cfg).layers:
            sig = L.conv.signature()
            for k in ('fwd', 'bwd', 'info'):
                assert hasattr(cdll, f'sevenn_conv_{k}_{sig}'), (name, L.t, sig)


def test_product_refuses_to_run_without_gpu():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    _lib._cached = None
    with pytest.raises(_lib.SevennError):
        _lib.load()
