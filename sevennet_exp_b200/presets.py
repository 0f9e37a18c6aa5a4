"""Model hyper-parameters of the reference's published potentials (sevenn/presets/sevennet-0.yaml:1-31,
sevennet-l3i5.yaml:1-39, pretrained_potentials/SevenNet_0__11Jul2024/pre_train.yaml:28) and of the two
small models the reference ships in-tree (tests/data/checkpoints/cp_0.pth, deployed_serial.pt)."""

MPTRJ_Z = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27,
           28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 44, 45, 46, 47, 48, 49, 50, 51, 52,
           53, 54, 55, 56, 57, 58, 59, 60, 61, 62, 63, 64, 65, 66, 67, 68, 69, 70, 71, 72, 73, 74, 75, 76, 77,
           78, 79, 80, 81, 82, 83, 89, 90, 91, 92, 93, 94]   # 89 species


def _common():
    return {
        'cutoff': 5.0, 'is_parity': False, 'num_convolution_layer': 5,
        'weight_nn_hidden_neurons': [64, 64],
        'radial_basis': {'radial_basis_name': 'bessel', 'bessel_basis_num': 8},
        'act_gate': {'e': 'silu', 'o': 'tanh'}, 'act_scalar': {'e': 'silu', 'o': 'tanh'},
        'act_radial': 'silu', 'train_shift_scale': False, 'train_denominator': False,
        'self_connection_type': 'linear', 'use_bias_in_linear': False, 'readout_as_fcn': False,
        '_number_of_species': len(MPTRJ_Z), '_type_map': {z: i for i, z in enumerate(MPTRJ_Z)},
        'shift': [0.0] * len(MPTRJ_Z), 'scale': [1.0] * len(MPTRJ_Z),
    }


SEVENNET_0 = {
    **_common(), 'channel': 128, 'lmax': 2,
    'irreps_manual': ['128x0e'] + ['128x0e+64x1e+32x2e'] * 4 + ['128x0e'],
    'cutoff_function': {'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5},
    'conv_denominator': 35.989574, '_normalize_sph': False, 'version': '0.9.0',
}

SEVENNET_L3I5 = {
    **_common(), 'channel': 128, 'lmax': 3,
    'irreps_manual': ['128x0e'] + ['128x0e+64x1e+32x2e+32x3e'] * 4 + ['128x0e'],
    'cutoff_function': {'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6},
    'conv_denominator': 35.989574, '_normalize_sph': True, 'version': '0.10.0',
}

# tests/data/checkpoints/cp_0.pth (Hf/O, O3 parity, nequip self connection)
CP0_LIKE = {
    'cutoff': 4.0, 'channel': 4, 'lmax': 2, 'is_parity': True, 'num_convolution_layer': 3,
    'weight_nn_hidden_neurons': [64, 64], 'radial_basis': {'radial_basis_name': 'bessel'},
    'cutoff_function': {'cutoff_function_name': 'poly_cut'}, 'conv_denominator': 5.0,
    'self_connection_type': 'nequip', '_normalize_sph': True, '_number_of_species': 2,
    '_type_map': {72: 0, 8: 1}, 'shift': 0.0, 'scale': 1.0, 'irreps_manual': False,
}

# example_inputs/md_serial_example/deployed_serial.pt (sevenn 0.8.6 artefact: Hf/O, lmax 1, O3 parity, 4 layers,
# unnormalised SH); weights transplanted by tests/golden/make_deployed_transplant.py
DEPLOYED_SERIAL_LIKE = {
    'cutoff': 4.0, 'channel': 4, 'lmax': 1, 'is_parity': True, 'num_convolution_layer': 4,
    'weight_nn_hidden_neurons': [64, 64], 'radial_basis': {'radial_basis_name': 'bessel', 'bessel_basis_num': 8},
    'cutoff_function': {'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6},
    'conv_denominator': 4.793977737426758, 'self_connection_type': 'nequip', '_normalize_sph': False,
    '_number_of_species': 2, '_type_map': {72: 0, 8: 1}, 'shift': 0.0, 'scale': 1.0, 'irreps_manual': False,
}

# small SO3 models used by the GPU parity tests (same structure as SevenNet-0 / l3i5, 32 channels)
TEST_SO3_L2 = {
    **_common(), 'channel': 32, 'lmax': 2, 'num_convolution_layer': 3,
    'irreps_manual': ['32x0e', '32x0e+32x1e+32x2e', '32x0e+32x1e+32x2e', '32x0e'],
    'cutoff_function': {'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5},
    'conv_denominator': 20.0, '_normalize_sph': False, '_number_of_species': 3,
    '_type_map': {3: 0, 15: 1, 16: 2}, 'shift': [0.1, -0.2, 0.3], 'scale': [1.0, 1.5, 0.7],
}
TEST_SO3_L3 = {
    **_common(), 'channel': 32, 'lmax': 3, 'num_convolution_layer': 3,
    'irreps_manual': ['32x0e', '32x0e+32x1e+32x2e+32x3e', '32x0e+32x1e+32x2e+32x3e', '32x0e'],
    'cutoff_function': {'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6},
    'conv_denominator': 20.0, '_normalize_sph': True, '_number_of_species': 3,
    '_type_map': {3: 0, 15: 1, 16: 2}, 'shift': [0.1, -0.2, 0.3], 'scale': [1.0, 1.5, 0.7],
}

PRESETS = {'7net-0': SEVENNET_0, '7net-l3i5': SEVENNET_L3I5, 'cp0': CP0_LIKE,
           'deployed-serial': DEPLOYED_SERIAL_LIKE,
           'test-so3-l2': TEST_SO3_L2, 'test-so3-l3': TEST_SO3_L3}
