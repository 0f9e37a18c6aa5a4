"""Drop-in surface: every class of the reference's module seam (sevenn/nn/*.py), the calculator constructor and the
deploy / model-build entry points exist here with the same parameter names in the same order and the same public
methods (tests/golden/reference_nn_signatures.json, extracted from the reference with `ast` by
tests/golden/make_reference_signatures.py).  Variants that are outside the hot path are listed explicitly."""
import inspect
import json
import os

# nothing of the reference's module seam is left out (the multi-modal pieces: tests/test_nn_modules_emu.py)
OUT_OF_SCOPE_CLASSES = set()
OUT_OF_SCOPE_METHODS = set()


def _params(fn):
    return [p for p in inspect.signature(fn).parameters if p not in ('self', 'kwargs', 'linear_kwargs')]


def test_nn_classes_match_reference_signatures(golden_dir):
    from sevennet_exp_b200.nn import modules as M
    sig = json.load(open(os.path.join(golden_dir, 'reference_nn_signatures.json')))
    checked = 0
    for name, ref in sig['nn'].items():
        if name in OUT_OF_SCOPE_CLASSES:
            assert not hasattr(M, name)
            continue
        cls = getattr(M, name)
        ours = _params(cls.__init__)
        missing = [p for p in ref['params'] if p not in ours]
        assert not missing, (name, missing)
        assert [p for p in ours if p in ref['params']] == ref['params'], (name, ours, ref['params'])
        for m in ref['methods']:
            assert hasattr(cls, m) or (name, m) in OUT_OF_SCOPE_METHODS, (name, m)
        checked += 1
    assert checked >= 20


def test_entry_points_match_reference_signatures(golden_dir):
    from sevennet_exp_b200 import deploy
    from sevennet_exp_b200.nn import build_E3_equivariant_model
    from sevennet_exp_b200.sevennet_calculator import SevenNetCalculator
    sig = json.load(open(os.path.join(golden_dir, 'reference_nn_signatures.json')))
    ours = _params(SevenNetCalculator.__init__)
    assert ours[:len(sig['calculator']['params'])] == sig['calculator']['params']
    for m in sig['calculator']['methods']:
        assert hasattr(SevenNetCalculator, m), m
    assert _params(deploy.deploy) == sig['deploy']['deploy']
    assert _params(deploy.deploy_parallel) == sig['deploy']['deploy_parallel']
    assert _params(build_E3_equivariant_model) == sig['model_build']['build_E3_equivariant_model']
