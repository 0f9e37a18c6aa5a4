"""Trainable-parameter counts of the ten config variants pinned by the reference's
tests/unit_tests/test_model.py:164-182 (and the modal variants, :185-213) -- for the oracle's builder and for
sevennet_exp_b200.nn.build_E3_equivariant_model.  The only reference-held pin that touches lmax = 3 (26866)."""
import pytest

from oracle import sevenn_ref as S


def base_config():
    # reference tests/unit_tests/test_model.py:54-86 (species of NaCl + H2O + H, sorted: Cl H Na O)
    return {
        'cutoff': 4.0, 'channel': 4, 'radial_basis': {'radial_basis_name': 'bessel'},
        'cutoff_function': {'cutoff_function_name': 'poly_cut'}, 'interaction_type': 'nequip', 'lmax': 2,
        'is_parity': True, 'num_convolution_layer': 3, 'weight_nn_hidden_neurons': [64, 64], 'act_radial': 'silu',
        'act_scalar': {'e': 'silu', 'o': 'tanh'}, 'act_gate': {'e': 'silu', 'o': 'tanh'}, 'conv_denominator': 30.0,
        'train_denominator': False, 'self_connection_type': 'nequip', 'shift': -10.0, 'scale': 10.0,
        'train_shift_scale': False, 'irreps_manual': False, 'lmax_edge': -1, 'lmax_node': -1, 'readout_as_fcn': False,
        'use_bias_in_linear': False, '_normalize_sph': True,
        'chemical_species': ['Cl', 'H', 'Na', 'O'], '_number_of_species': 4, '_type_map': {17: 0, 1: 1, 11: 2, 8: 3},
    }


N_PARAM = [
    ({}, 20642),
    ({'train_denominator': True}, 20642 + 3),
    ({'train_shift_scale': True}, 20642 + 2),
    ({'shift': [1.0] * 4}, 20642),
    ({'scale': [1.0] * 4, 'train_shift_scale': True}, 20642 + 8),
    ({'num_convolution_layer': 4}, 33458),
    ({'lmax': 3}, 26866),
    ({'channel': 2}, 16883),
    ({'is_parity': False}, 20386),
    ({'self_connection_type': 'linear'}, 20114),
]
N_PARAM_MODAL = [
    ({}, 20642),
    ({'use_modal_node_embedding': True}, 20642 + 8),
    ({'use_modal_self_inter_intro': True}, 20642 + 2 * 4 * 3),
    ({'use_modal_self_inter_outro': True}, 20642 + 2 * (12 + 20 + 4)),
    ({'use_modal_output_block': True}, 20642 + 2 * 4 // 2),
]
MODAL = {'use_modality': True, '_number_of_modalities': 2, '_modal_map': {'x1': 0, 'x2': 1},
         'use_modal_node_embedding': False, 'use_modal_self_inter_intro': False, 'use_modal_self_inter_outro': False,
         'use_modal_output_block': False, 'use_modal_wise_shift': False, 'use_modal_wise_scale': False}


def _count(model):
    return sum(p.numel() for p in model.parameters() if p.requires_grad)


@pytest.mark.parametrize('cf,ref', N_PARAM + [({**MODAL, **c}, r) for c, r in N_PARAM_MODAL])
def test_oracle_num_params(cf, ref):
    cfg = base_config()
    cfg.update(cf)
    assert _count(S.build_model(cfg)) == ref


@pytest.mark.parametrize('cf,ref', N_PARAM + [({**MODAL, **c}, r) for c, r in N_PARAM_MODAL])
def test_nn_seam_num_params(cf, ref):
    from sevennet_exp_b200 import nn as bnn
    cfg = base_config()
    cfg.update(cf)
    assert _count(bnn.build_E3_equivariant_model(cfg)) == ref
