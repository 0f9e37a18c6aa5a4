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


def test_version_gating_follows_the_reference_string_comparison():
    """scripts/backward_compatibility.py:13-36,39-71,155-176: numeric (major, minor) for the config patches, STRING
    comparison for map_old_model / sort_old_convolution, ValueError without a version, old module names mapped."""
    import pytest
    import torch
    from sevennet_exp_b200 import checkpoint as C
    assert C.needs_sort('0.10.0') and C.needs_sort('0.10.4') and C.needs_sort('0.11.0.dev0')
    assert not C.needs_sort('0.11.0') and not C.needs_sort('0.11.1')
    assert not C.needs_sort('0.9.5') and not C.needs_map('0.9.5')          # '0.9.5' > '0.10.0' as strings, like the reference
    assert C.needs_map('0.1.0') and not C.needs_map('0.10.0')
    with pytest.raises(ValueError):
        C.patch_old_config({'cutoff': 5.0})
    cfg = C.patch_old_config({'version': '0.9.2', 'cutoff_function': {'cutoff_function_name': 'XPLOR', 'poly_cut_p_value': 6},
                              'train_avg_num_neigh': True})
    assert cfg['_normalize_sph'] is False and cfg['conv_denominator'] == 0.0 and cfg['train_denominator'] is True
    assert 'poly_cut_p_value' not in cfg['cutoff_function']
    with pytest.raises(ValueError):
        C.patch_old_config({'version': '0.9.2', 'optimize_by_reduce': False, 'cutoff_function': {'cutoff_function_name': 'poly_cut'}})
    sd = C.map_old_model({'EdgeEmbedding.basis_function.coeffs': torch.zeros(8), '1 convolution.denumerator': torch.ones(1),
                          'reducing nn hidden to energy.linear.weight': torch.zeros(3), 'onehot_to_feature_x.linear.weight': torch.zeros(2)})
    assert set(sd) == {'edge_embedding.basis_function.coeffs', '1_convolution.denominator',
                       'reduce_hidden_to_energy.linear.weight', 'onehot_to_feature_x.linear.weight'}
