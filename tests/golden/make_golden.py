"""Generates the committed fixtures under tests/golden/ (run in the authoring
container only; needs /root/reference, which does not exist on the GPU box).

  cp0_model.pt   : config + model_state_dict of the reference's own test
                   checkpoint tests/data/checkpoints/cp_0.pth (data, not source)
  hfo2_frames.npz: the two labelled frames of tests/data/systems/hfo2.extxyz
  w3j_cp0.pt     : the 8 Wigner-3j golden buffers stored inside cp_0.pth
  deployed_serial_hfo2.npz : E/F/S of the reference's frozen TorchScript artefact
                   example_inputs/md_serial_example/deployed_serial.pt on frame 0
                   (real reference code path, e3nn inlined) + its inputs
"""
import os, sys
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, '..', '..'))
from oracle import sevenn_ref as S  # noqa: E402

REF = '/root/reference'


def main():
    cp = torch.load(f'{REF}/tests/data/checkpoints/cp_0.pth', weights_only=False, map_location='cpu')
    keep = ['cutoff', 'channel', 'radial_basis', 'cutoff_function', 'interaction_type', 'lmax',
            'is_parity', 'num_convolution_layer', 'weight_nn_hidden_neurons', 'act_radial',
            'act_scalar', 'act_gate', 'conv_denominator', 'train_denominator',
            'self_connection_type', 'train_shift_scale', 'irreps_manual', 'lmax_edge', 'lmax_node',
            'readout_as_fcn', 'use_bias_in_linear', '_normalize_sph', 'chemical_species',
            '_chemical_species_by_atomic_number', '_number_of_species', '_type_map', 'shift',
            'scale', 'version']
    cfg = {k: cp['config'][k] for k in keep}
    sd = {k: v.clone() for k, v in cp['model_state_dict'].items()}
    torch.save({'config': cfg, 'model_state_dict': sd}, f'{HERE}/cp0_model.pt')
    w3j = {k.split('.')[-1]: v.clone() for k, v in sd.items()
           if k.startswith('1_convolution') and '_w3j_' in k}
    torch.save(w3j, f'{HERE}/w3j_cp0.pt')

    frames = S.read_extxyz(f'{REF}/tests/data/systems/hfo2.extxyz')
    out = {}
    for k, fr in enumerate(frames):
        for name, v in fr.items():
            out[f'f{k}_{name}'] = np.asarray(v)
    np.savez(f'{HERE}/hfo2_frames.npz', **out)

    # reference TorchScript artefact on frame 0 (brute-force neighbor list from the oracle)
    extra = {'cutoff': b'', 'chemical_symbols_to_index': b''}
    m = torch.jit.load(f'{REF}/example_inputs/md_serial_example/deployed_serial.pt',
                       _extra_files=extra, map_location='cpu')
    fr = frames[0]
    rc = float(extra['cutoff'])
    i, j, D, Sh = S.neighbor_list_bruteforce(fr['pos'], fr['cell'], fr['pbc'], rc)
    tmap = {72: 0, 8: 1}
    pos = torch.tensor(fr['pos'], dtype=torch.float32, requires_grad=True)
    data = {
        'x': torch.tensor([tmap[int(z)] for z in fr['numbers']], dtype=torch.int64),
        'pos': pos,
        'edge_index': torch.tensor(np.stack([i, j]), dtype=torch.int64),
        'num_atoms': torch.tensor([len(fr['numbers'])], dtype=torch.int64),
        'cell_lattice_vectors': torch.tensor(fr['cell'], dtype=torch.float32),
        'cell_volume': torch.tensor(abs(np.linalg.det(fr['cell'])), dtype=torch.float32),
        'pbc_shift': torch.tensor(Sh, dtype=torch.float32),
    }
    res = m(data)
    np.savez(f'{HERE}/deployed_serial_hfo2.npz',
             energy=res['inferred_total_energy'].detach().numpy(),
             force=res['inferred_force'].detach().numpy(),
             stress=res['inferred_stress'].detach().numpy(),
             atomic_energy=res['atomic_energy'].detach().numpy(),
             edge_index=np.stack([i, j]), pbc_shift=Sh, cutoff=rc)
    print('E', res['inferred_total_energy'].item(), 'edges', len(i))


if __name__ == '__main__':
    main()
