"""Transplants the weights frozen inside the reference's deployed TorchScript artefact
example_inputs/md_serial_example/deployed_serial.pt (sevenn 0.8.6, Hf/O, 4 channels, lmax 1, O3 parity, 4
convolution layers, rc = 4, poly_cut p = 6, unnormalised SH, nequip self-connection) into a state_dict keyed like
a sevenn checkpoint, so that the oracle (and the CUDA engine) can be checked per atom against outputs of the
reference's own code path (tests/golden/deployed_serial_hfo2.npz, written by make_golden.py).

Run in the authoring container only (needs /root/reference).  Output: tests/golden/deployed_serial_transplant.pt
(numbers only: config + weights; no TorchScript code is stored).

How: `torch.jit.freeze` turned every parameter into a constant `CONSTANTS.cN`; their first uses appear in
execution order in `module.code`, which is the module order of `AtomGraphSequential`.  Weights are e3nn's flat
parameter vectors reshaped per instruction ([mul_in, mul_out] / [mul_1, mul_2, mul_out]), so concatenating the
flattened constants of one module restores its flat vector.  The 0.8.6 architecture keeps the full irreps in the
last interaction layer although the readout only reads its 0e block; the current architecture (model_build.py:
498-540, last layer lmax 0 / even) is exactly the live sub-network, so the last layer takes the matching slices.
The last radial-MLP layer is stored in instruction-creation order (what a < 0.11.0 checkpoint holds); the
loader's `sort_old_convolution` brings it to the sorted order.
"""
import math
import os
import re

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference/example_inputs/md_serial_example/deployed_serial.pt'


def main():
    m = torch.jit.load(REF, map_location='cpu')
    code, consts = m.code_with_constants
    consts = consts.const_mapping if hasattr(consts, 'const_mapping') else consts
    order, seen = [], set()
    lines = code.split('\n')
    defs = {}
    for line in lines:
        mm = re.match(r'\s*(\w+) = (.*)$', line)
        if mm:
            defs[mm.group(1)] = mm.group(2)
    folded = {}
    for line in lines:
        for name in re.findall(r'CONSTANTS\.(c\d+)', line):
            if name in seen:
                continue
            seen.add(name)
            kind = 'weight' if ('tensordot' in line or 'matmul' in line) else 'other'
            order.append((name, kind, line.strip()))
            mm = re.search(r'torch\.tensordot\((\w+), CONSTANTS', line)
            if mm:
                # e3nn writes the 1/sqrt(fan_in) of a Linear / FullyConnectedTensorProduct path either as an explicit
                # `torch.mul(operand, literal)` or (opt_einsum_fx, whichever operand is smaller) into the weight, where
                # freezing folded it into the constant: no explicit scaling of the operand => undo it on the constant
                d = defs.get(mm.group(1), '')
                if d.startswith('torch.einsum'):        # FCTP: x (x) (one-hot * literal), one level down
                    d = ' '.join(defs.get(a, '') for a in re.findall(r'\[(\w+), (\w+)\]', d)[0])
                folded[name] = re.search(r'torch\.mul\(\w+, [0-9.e+-]+\)', d) is None
    def raw(n):
        w = consts[n]
        if folded.get(n, False):
            fan_in = w.shape[0] * (w.shape[1] if w.dim() == 3 else 1)
            w = w * math.sqrt(fan_in)
        return w

    weights = [raw(n) for n, kind, _ in order if kind == 'weight' and tuple(consts[n].shape) != (3, 3, 3)]
    print('normalisation folded into the constant:', [n for n, f in folded.items() if f and tuple(consts[n].shape) != (3, 3, 3)])
    w3j_111 = [consts[n] for n, kind, _ in order if tuple(consts[n].shape) == (3, 3, 3)][0]
    scal = {n: consts[n] for n, kind, _ in order if kind == 'other'}
    coeffs = [v for v in scal.values() if tuple(v.shape) == (8,)][0]
    one = [(n, v) for n, v in scal.items() if tuple(v.shape) == (1,)]
    # order of first use: conv denominator (div), then scale (mul), then shift (add)
    denom, scale, shift = [float(v) for _, v in one]
    acts = sorted(float(v) for v in scal.values() if v.dim() == 0 and v.dtype == torch.float64)
    assert abs(acts[1] - 1.6791767923989418) < 1e-12, acts      # normalize2mom(silu), tanh is acts[0]

    cfg = dict(cutoff=4.0, channel=4, radial_basis={'radial_basis_name': 'bessel', 'bessel_basis_num': 8},
               cutoff_function={'cutoff_function_name': 'poly_cut', 'poly_cut_p_value': 6}, lmax=1,
               is_parity=True, num_convolution_layer=4, weight_nn_hidden_neurons=[64, 64], act_radial='silu',
               act_scalar={'e': 'silu', 'o': 'tanh'}, act_gate={'e': 'silu', 'o': 'tanh'},
               conv_denominator=denom, train_denominator=False, self_connection_type='nequip',
               train_shift_scale=False, irreps_manual=False, lmax_edge=-1, lmax_node=-1, readout_as_fcn=False,
               use_bias_in_linear=False, _normalize_sph=False, chemical_species=['Hf', 'O'],
               _chemical_species_by_atomic_number=[72, 8], _number_of_species=2, _type_map={72: 0, 8: 1},
               shift=shift, scale=scale, version='0.8.6')

    it = iter(weights)

    def take(shapes):
        out = []
        for s in shapes:
            w = next(it)
            assert tuple(w.shape) == tuple(s), (tuple(w.shape), s)
            out.append(w)
        return out

    def flat(ws):
        return torch.cat([w.reshape(-1) for w in ws]).clone()

    sd = {'edge_embedding.basis_function.coeffs': coeffs.clone()}
    sd['onehot_to_feature_x.linear.weight'] = flat(take([(2, 4)]))
    # per layer: sc intro (FCTP blocks), si1 blocks, radial MLP, si2 blocks -- shapes as printed from the artefact
    layers = [
        dict(sc=[(4, 2, 8)], si1=[(4, 4)], wn=8, si2=[(4, 8), (4, 4)]),
        dict(sc=[(4, 2, 12), (4, 2, 4)], si1=[(4, 4)] * 2, wn=20, si2=[(8, 12), (8, 4), (4, 4)]),
        dict(sc=[(4, 2, 12), (4, 2, 4), (4, 2, 4)], si1=[(4, 4)] * 3, wn=32,
             si2=[(4, 4), (8, 12), (12, 4), (8, 4)]),
        dict(sc=[(4, 2, 4), (4, 2, 12), (4, 2, 4), (4, 2, 4)], si1=[(4, 4)] * 4, wn=40,
             si2=[(8, 4), (8, 12), (12, 4), (12, 4)]),
    ]
    for t, L in enumerate(layers):
        sc, si1 = take(L['sc']), take(L['si1'])
        # e3nn's FullyConnectedNet divides by sqrt(h_in) inside forward; freezing folded that into the constants
        mlp = [w * math.sqrt(w.shape[0]) for w in take([(8, 64), (64, 64), (64, L['wn'])])]
        si2 = take(L['si2'])
        if t == 3:
            # live sub-network of the last layer (see module docstring).  x = 4x0o+4x0e+4x1o+4x1e; creation order
            # of the 10 paths: 0o0e>0o, 0o1o>1e, 0e0e>0e, 0e1o>1o, 1o0e>1o, 1o1o>0e, 1o1o>1e, 1e0e>1e, 1e1o>0o, 1e1o>1o
            sc = [sc[1][:, :, :4]]                                    # 0e (x) onehot -> 4x0e scalars
            mlp[2] = torch.cat([mlp[2][:, 8:12], mlp[2][:, 20:24]], dim=1)
            si2 = [si2[1][:, :4]]                                     # 8x0e -> first 4 of (4 scalars + 8 gates)
        sd[f'{t}_self_connection_intro.fc_tensor_product.weight'] = flat(sc)
        sd[f'{t}_self_interaction_1.linear.weight'] = flat(si1)
        sd[f'{t}_convolution.denominator'] = torch.tensor([denom])
        for k in range(3):
            sd[f'{t}_convolution.weight_nn.layer{k}.weight'] = mlp[k].clone()
        sd[f'{t}_self_interaction_2.linear.weight'] = flat(si2)
    sd['reduce_input_to_hidden.linear.weight'] = flat(take([(4, 2)]))
    sd['reduce_hidden_to_energy.linear.weight'] = flat(take([(2, 1)]))
    assert next(it, None) is None
    sd['rescale_atomic_energy.shift'] = torch.tensor([shift])
    sd['rescale_atomic_energy.scale'] = torch.tensor([scale])
    # the only l1,l2,l3 > 0 path is 1o x 1o -> 1e: keep the artefact's Wigner-3j so the loader can sign-check it
    for t in (1, 2):
        sd[f'{t}_convolution.convolution._compiled_main_left_right._w3j_1_1_1'] = w3j_111.clone()
    torch.save({'config': cfg, 'model_state_dict': sd}, f'{HERE}/deployed_serial_transplant.pt')
    print('denominator', denom, 'scale', scale, 'shift', shift, 'tensors', len(sd))


if __name__ == '__main__':
    main()
