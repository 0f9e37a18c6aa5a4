"""CPU restatement of SevenNet's E/F/S evaluation path (TEST INFRASTRUCTURE).

Follows, module by module, the reference files cited in each docstring
(/root/reference/sevenn/...).  Same module names, same state_dict keys, same
data-dict keys, autograd forces exactly like the reference.  Built on
oracle/e3nn_ref.py instead of e3nn.
"""
import copy
import math
from collections import OrderedDict
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn as nn

from . import e3nn_ref as o3
from .e3nn_ref import Irreps

# data keys, sevenn/_keys.py:26-86
EDGE_VEC, EDGE_IDX, EDGE_LEN = 'edge_vec', 'edge_index', 'edge_length'
X, X_GHOST, NODE_ATTR, ATOM_TYPE = 'x', 'x_ghost', 'node_attr', 'atom_type'
EDGE_ATTR, EDGE_EMB = 'edge_attr', 'edge_embedding'
SC_TMP = 'self_cont_tmp'
SCALED_AE, AE, E_TOT = 'scaled_atomic_energy', 'atomic_energy', 'inferred_total_energy'
FORCE, STRESS = 'inferred_force', 'inferred_stress'
POS, CELL, SHIFT, VOL, NATOMS, BATCH, NLOCAL = (
    'pos', 'cell_lattice_vectors', 'pbc_shift', 'cell_volume', 'num_atoms', 'batch', 'nlocal')
Z = 'atomic_numbers'
MODAL_TYPE, MODAL_ATTR = 'modal_type', 'modal_attr'


def shifted_softplus(x):
    return torch.nn.functional.softplus(x) - math.log(2.0)


ACT = {  # sevenn/_const.py:34-48
    'relu': torch.nn.functional.relu, 'silu': torch.nn.functional.silu, 'tanh': torch.tanh,
    'abs': torch.abs, 'ssp': shifted_softplus, 'sigmoid': torch.sigmoid,
    'elu': torch.nn.functional.elu,
}


# ------------------------------------------------------------------ edge --
class BesselBasis(nn.Module):
    """sevenn/nn/edge_embedding.py:81-104"""

    def __init__(self, cutoff_length, bessel_basis_num=8, trainable_coeff=True):
        super().__init__()
        self.num_basis, self.cutoff_length = bessel_basis_num, cutoff_length
        self.prefactor = 2.0 / cutoff_length
        c = torch.FloatTensor([n * math.pi / cutoff_length for n in range(1, bessel_basis_num + 1)])
        self.coeffs = nn.Parameter(c) if trainable_coeff else c

    def forward(self, r):
        ur = r.unsqueeze(-1)
        return self.prefactor * torch.sin(self.coeffs * ur) / ur


class PolynomialCutoff(nn.Module):
    """sevenn/nn/edge_embedding.py:107-133"""

    def __init__(self, cutoff_length, poly_cut_p_value=6):
        super().__init__()
        p = poly_cut_p_value
        self.r_cut, self.p = cutoff_length, p
        self.c0, self.c1, self.c2 = (p + 1.0) * (p + 2.0) / 2.0, p * (p + 2.0), p * (p + 1.0) / 2.0

    def forward(self, r):
        r = r / self.r_cut
        return (1 - self.c0 * torch.pow(r, self.p) + self.c1 * torch.pow(r, self.p + 1.0)
                - self.c2 * torch.pow(r, self.p + 2.0))


class XPLORCutoff(nn.Module):
    """sevenn/nn/edge_embedding.py:136-161"""

    def __init__(self, cutoff_length, cutoff_on):
        super().__init__()
        self.r_on, self.r_cut = cutoff_on, cutoff_length

    def forward(self, r):
        r_sq, on2, c2 = r * r, self.r_on ** 2, self.r_cut ** 2
        return torch.where(r < self.r_on, 1.0,
                           (c2 - r_sq) ** 2 * (c2 + 2 * r_sq - 3 * on2) / (c2 - on2) ** 3)


class SphericalEncoding(nn.Module):
    """sevenn/nn/edge_embedding.py:165-186"""

    def __init__(self, lmax, parity=-1, normalization='component', normalize=True):
        super().__init__()
        self.lmax, self.normalize, self.normalization = lmax, normalize, normalization
        self.irreps_out = Irreps.spherical_harmonics(lmax, parity)

    def forward(self, r):
        return o3.spherical_harmonics(self.lmax, r, self.normalize, self.normalization)


class EdgeEmbedding(nn.Module):
    """sevenn/nn/edge_embedding.py:190-217"""

    def __init__(self, basis_module, cutoff_module, spherical_module):
        super().__init__()
        self.basis_function, self.cutoff_function, self.spherical = (
            basis_module, cutoff_module, spherical_module)

    def forward(self, data):
        rvec = data[EDGE_VEC]
        r = torch.linalg.norm(rvec, dim=-1)
        data[EDGE_LEN] = r
        data[EDGE_EMB] = self.basis_function(r) * self.cutoff_function(r).unsqueeze(-1)
        data[EDGE_ATTR] = self.spherical(rvec)
        return data


class MultiCutoffEdgeEmbedding(nn.Module):
    """sevenn/nn/edge_embedding.py:220-257 (the fork's per-layer cutoffs): for every distinct cutoff c the edges
    with r <= c, their Bessel(c) x envelope(c) embedding and harmonics are stored under keys suffixed str(c)."""

    def __init__(self, basis_module, cutoff_module, spherical_module):
        super().__init__()
        self.basis_functions = nn.ModuleList(basis_module)
        self.cutoff_function, self.spherical = cutoff_module, spherical_module

    def forward(self, data):
        rvec = data[EDGE_VEC]
        r = torch.linalg.norm(rvec, dim=-1)
        data[EDGE_LEN] = r
        for basis in self.basis_functions:
            c = basis.cutoff_length
            self.cutoff_function.r_cut = c
            if isinstance(self.cutoff_function, XPLORCutoff):
                self.cutoff_function.r_on = c - 1
            keep = r <= c
            rr = r[keep]
            data[EDGE_EMB + str(c)] = basis(rr) * self.cutoff_function(rr).unsqueeze(-1)
            data[EDGE_ATTR + str(c)] = self.spherical(rvec[keep])
            data[EDGE_IDX + str(c)] = data[EDGE_IDX][:, keep]
        return data


class EdgePreprocess(nn.Module):
    """sevenn/nn/edge_embedding.py:13-78 (non-batched branch)"""

    def __init__(self, is_stress=True):
        super().__init__()
        self.is_stress = is_stress

    def forward(self, data):
        cell, shift, pos = data[CELL].view(3, 3), data[SHIFT], data[POS]
        if self.is_stress:
            strain = torch.zeros(3, 3, dtype=pos.dtype, requires_grad=True)
            data['_strain'] = strain
            sym = 0.5 * (strain + strain.T)
            pos = pos + pos @ sym
            cell = cell + cell @ sym
        i0, i1 = data[EDGE_IDX][0], data[EDGE_IDX][1]
        ev = pos[i1] - pos[i0] + shift @ cell
        data[EDGE_VEC] = ev
        data[EDGE_LEN] = torch.linalg.norm(ev, dim=-1)
        return data


# ------------------------------------------------------------------ node --
class OnehotEmbedding(nn.Module):
    """sevenn/nn/node_embedding.py:15-53"""

    def __init__(self, num_classes, data_key_x=X, data_key_out=None, data_key_save=None,
                 data_key_additional=None):
        super().__init__()
        self.num_classes, self.key_x = num_classes, data_key_x
        self.key_output = data_key_out or data_key_x
        self.key_save, self.key_add = data_key_save, data_key_additional

    def forward(self, data):
        inp = data[self.key_x]
        embd = torch.nn.functional.one_hot(inp, self.num_classes).to(torch.get_default_dtype())
        data[self.key_output] = embd
        if self.key_add is not None:
            data[self.key_add] = embd
        if self.key_save is not None:
            data[self.key_save] = inp
        return data


class IrrepsLinear(nn.Module):
    """sevenn/nn/linear.py:14-100.  num_modalities > 1 (:65-92): the modality one-hot (data['modal_attr'], one row
    for the whole structure) is appended to the input as extra 0e channels."""

    def __init__(self, irreps_in, irreps_out, data_key_in, data_key_out=None, biases=False, num_modalities=0):
        super().__init__()
        self.key_input, self.key_output = data_key_in, data_key_out or data_key_in
        self.num_modalities = num_modalities
        irreps_in = Irreps(irreps_in)
        if num_modalities > 1:
            irreps_in = irreps_in + Irreps(f'{num_modalities}x0e')
        self.irreps_in, self.irreps_out = irreps_in, Irreps(irreps_out)
        self.linear = o3.Linear(self.irreps_in, irreps_out, biases=biases)

    def forward(self, data):
        x = data[self.key_input]
        if self.num_modalities > 1:
            onehot = data[MODAL_ATTR].reshape(1, self.num_modalities).to(x.dtype).expand(len(x), -1)
            x = torch.cat([x, onehot], dim=1)
        data[self.key_output] = self.linear(x)
        return data


class FCN_e3nn(nn.Module):
    """sevenn/nn/linear.py:145-180: e3nn FullyConnectedNet on scalar node features (readout_as_fcn)."""

    def __init__(self, irreps_in, dim_out, hidden_neurons, activation, data_key_in, data_key_out=None):
        super().__init__()
        self.key_input, self.key_output = data_key_in, data_key_out or data_key_in
        irreps_in = Irreps(irreps_in)
        assert all(ir.is_scalar() for _, ir in irreps_in)
        self.fcn = o3.FullyConnectedNet([irreps_in.dim] + list(hidden_neurons) + [dim_out], activation)

    def forward(self, data):
        data[self.key_output] = self.fcn(data[self.key_input])
        return data


class SelfConnectionIntro(nn.Module):
    """sevenn/nn/self_connection.py:10-64 ('nequip': FCTP of x with one-hot)"""

    def __init__(self, irreps_in, irreps_operand, irreps_out):
        super().__init__()
        self.fc_tensor_product = o3.FullyConnectedTensorProduct(irreps_in, irreps_operand, irreps_out)

    def forward(self, data):
        data[SC_TMP] = self.fc_tensor_product(data[X], data[NODE_ATTR])
        return data


class SelfConnectionLinearIntro(nn.Module):
    """sevenn/nn/self_connection.py:68-108"""

    def __init__(self, irreps_in, irreps_operand, irreps_out):
        super().__init__()
        self.linear = o3.Linear(irreps_in, irreps_out)

    def forward(self, data):
        data[SC_TMP] = self.linear(data[X])
        return data


class SelfConnectionOutro(nn.Module):
    """sevenn/nn/self_connection.py:112-128"""

    def forward(self, data):
        data[X] = data[X] + data[SC_TMP]
        del data[SC_TMP]
        return data


class EquivariantGate(nn.Module):
    """sevenn/nn/equivariant_gate.py:13-61"""

    def __init__(self, irreps_x, act_scalar_dict, act_gate_dict):
        super().__init__()
        pm = {'e': 1, 'o': -1}
        act_s = {pm[k]: v for k, v in act_scalar_dict.items()}
        act_g = {pm[k]: v for k, v in act_gate_dict.items()}
        gated = Irreps([(m, ir) for m, ir in irreps_x if ir.l > 0])
        scalars = Irreps([(m, ir) for m, ir in irreps_x if ir.l == 0])
        gp = 1 if '0e' in scalars else -1
        gates = Irreps([(m, (0, gp)) for m, _ in gated])
        self.gate = o3.Gate(scalars, [act_s[ir.p] for _, ir in scalars],
                            gates, [act_g[ir.p] for _, ir in gates], gated)

    def get_gate_irreps_in(self):
        return self.gate.irreps_in

    def forward(self, data):
        data[X] = self.gate(data[X])
        return data


# ---------------------------------------------------------- convolution --
class IrrepsConvolution(nn.Module):
    """sevenn/nn/convolution.py:29-137"""

    def __init__(self, irreps_x, irreps_filter, irreps_out, weight_layer_input_to_hidden,
                 weight_layer_act, denominator=1.0, train_denominator=False, is_parallel=False,
                 data_key_filter=EDGE_ATTR, data_key_weight_input=EDGE_EMB, data_key_edge_idx=EDGE_IDX):
        super().__init__()
        self.key_filter, self.key_weight_input, self.key_edge_idx = (
            data_key_filter, data_key_weight_input, data_key_edge_idx)
        irreps_x, irreps_filter, irreps_out = Irreps(irreps_x), Irreps(irreps_filter), Irreps(irreps_out)
        self.denominator = nn.Parameter(torch.FloatTensor([denominator]),
                                        requires_grad=train_denominator)
        self.is_parallel = is_parallel
        instructions, mid, wn = [], [], 0
        for i, (mul_x, ir_x) in enumerate(irreps_x):
            for j, (_, ir_f) in enumerate(irreps_filter):
                for ir_o in ir_x * ir_f:
                    if ir_o in irreps_out:
                        k = len(mid)
                        wn += mul_x
                        mid.append((mul_x, ir_o))
                        instructions.append((i, j, k, 'uvu', True))
        mid, p, _ = Irreps(mid).sort()
        instructions = [(a, b, p[c], m, t) for a, b, c, m, t in instructions]
        self._instructions_before_sort = instructions
        instructions = sorted(instructions, key=lambda t: t[2])
        self.irreps_x, self.irreps_filter, self.irreps_mid = irreps_x, irreps_filter, mid
        self.weight_numel = wn
        self.convolution = o3.TensorProduct(irreps_x, irreps_filter, mid, instructions,
                                            shared_weights=False, internal_weights=False)
        self.weight_nn = o3.FullyConnectedNet(list(weight_layer_input_to_hidden) + [wn],
                                              weight_layer_act)

    def forward(self, data):
        weight = self.weight_nn(data[self.key_weight_input])
        x = data[X]
        if self.is_parallel:
            x = torch.cat([x, data[X_GHOST]])
        edge_src, edge_dst = data[self.key_edge_idx][1], data[self.key_edge_idx][0]
        message = self.convolution(x[edge_src], data[self.key_filter], weight)
        out = torch.zeros(len(x), message.shape[1], dtype=x.dtype)
        out.index_add_(0, edge_dst, message)
        out = out.div(self.denominator)
        if self.is_parallel:
            out = out[: int(data[NLOCAL])]
        data[X] = out
        return data


# --------------------------------------------------------------- output --
class Rescale(nn.Module):
    """sevenn/nn/scale.py:22-56"""

    def __init__(self, shift, scale, train_shift_scale=False):
        super().__init__()
        self.shift = nn.Parameter(torch.FloatTensor([shift]), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor([scale]), requires_grad=train_shift_scale)

    def forward(self, data):
        data[AE] = data[SCALED_AE] * self.scale + self.shift
        return data


class SpeciesWiseRescale(nn.Module):
    """sevenn/nn/scale.py:60-161"""

    def __init__(self, shift, scale, train_shift_scale=False):
        super().__init__()
        n = len(shift) if isinstance(shift, list) else len(scale)
        shift = [shift] * n if isinstance(shift, float) else shift
        scale = [scale] * n if isinstance(scale, float) else scale
        self.shift = nn.Parameter(torch.FloatTensor(shift), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor(scale), requires_grad=train_shift_scale)

    def forward(self, data):
        idx = data[ATOM_TYPE]
        data[AE] = data[SCALED_AE] * self.scale[idx].view(-1, 1) + self.shift[idx].view(-1, 1)
        return data


class ModalWiseRescale(nn.Module):
    """sevenn/nn/scale.py:165-330: shift / scale are [n_modal, n_types] where modal wise, else [n_types]."""

    def __init__(self, shift, scale, use_modal_wise_shift, use_modal_wise_scale, train_shift_scale=False):
        super().__init__()
        self.shift = nn.Parameter(torch.FloatTensor(shift), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor(scale), requires_grad=train_shift_scale)
        self.mw_shift, self.mw_scale = use_modal_wise_shift, use_modal_wise_scale

    def forward(self, data):
        idx, m = data[ATOM_TYPE], int(data[MODAL_TYPE])
        shift = (self.shift[m] if self.mw_shift else self.shift)[idx]
        scale = (self.scale[m] if self.mw_scale else self.scale)[idx]
        data[AE] = data[SCALED_AE] * scale.view(-1, 1) + shift.view(-1, 1)
        return data


class AtomReduce(nn.Module):
    """sevenn/nn/linear.py:104-141"""

    def __init__(self):
        super().__init__()
        self._is_batch_data = False

    def forward(self, data):
        if self._is_batch_data:
            src = data[AE].squeeze(1)
            size = int(data[BATCH].max()) + 1
            out = torch.zeros(size, dtype=src.dtype)
            out.index_add_(0, data[BATCH], src)
            data[E_TOT] = out
        else:
            data[E_TOT] = torch.sum(data[AE])
        return data


class ForceStressOutputFromEdge(nn.Module):
    """sevenn/nn/force_output.py:140-224"""

    def __init__(self):
        super().__init__()
        self._is_batch_data = False

    def forward(self, data):
        tot = int(torch.sum(data[NATOMS]))
        rij, ei = data[EDGE_VEC], data[EDGE_IDX]
        fij = torch.autograd.grad([data[E_TOT].sum()], [rij], create_graph=self.training,
                                  allow_unused=True)[0]
        if fij is None:
            return data
        pf = torch.zeros(tot, 3, dtype=fij.dtype).index_add_(0, ei[0], fij)
        nf = torch.zeros(tot, 3, dtype=fij.dtype).index_add_(0, ei[1], fij)
        data[FORCE] = pf - nf
        vir = torch.cat([rij * fij, (rij[:, 0] * fij[:, 1]).unsqueeze(-1),
                         (rij[:, 1] * fij[:, 2]).unsqueeze(-1),
                         (rij[:, 2] * fij[:, 0]).unsqueeze(-1)], dim=-1)
        s = torch.zeros(tot, 6, dtype=fij.dtype).index_add_(0, ei[1], vir)
        if self._is_batch_data:
            b = data[BATCH]
            sout = torch.zeros(int(b.max()) + 1, 6, dtype=fij.dtype).index_add_(0, b, s)
        else:
            sout = s.sum(dim=0)
        data[STRESS] = torch.neg(sout) / data[VOL].unsqueeze(-1)
        return data


class ForceStressOutput(nn.Module):
    """sevenn/nn/force_output.py:55-136 (non-batched): pos + strain autograd."""

    def forward(self, data):
        g = torch.autograd.grad([data[E_TOT].sum()], [data[POS], data['_strain']],
                                allow_unused=True)
        if g[0] is not None:
            data[FORCE] = -g[0]
        if g[1] is not None:
            vol = data[VOL]
            vol = vol if vol >= 1e-3 else torch.tensor(1e-3)
            st = -(g[1] / vol)
            data[STRESS] = torch.stack((st[0, 0], st[1, 1], st[2, 2], st[0, 1], st[1, 2], st[0, 2]))
        return data


# ----------------------------------------------------------- sequential --
class AtomGraphSequential(nn.Sequential):
    """sevenn/nn/sequential.py:41-182"""

    def __init__(self, modules, cutoff=0.0, type_map=None, eval_type_map=True, data_key_grad=None):
        super().__init__(OrderedDict(modules))
        self.cutoff, self.type_map, self.key_grad = cutoff, type_map, data_key_grad
        self.eval_type_map = eval_type_map and type_map is not None
        if type_map is not None:
            lut = -torch.ones(120, dtype=torch.long)
            for z, i in type_map.items():
                lut[z] = i
            self.z_to_onehot_tensor = lut

    def set_is_batch_data(self, flag):
        for m in self:
            if hasattr(m, '_is_batch_data'):
                m._is_batch_data = flag

    def forward(self, data):
        if self.eval_type_map:
            data[X] = self.z_to_onehot_tensor[data[Z]]
        if self.key_grad is not None:
            data[self.key_grad].requires_grad_(True)
        for m in self:
            data = m(data)
        return data


def infer_irreps_out(irreps_x, irreps_operand, drop_l=False, parity_mode='full',
                     fix_multiplicity=False):
    """sevenn/util.py:163-185"""
    out = o3.full_tensor_product_irreps_out(irreps_x, irreps_operand).simplify()
    new = []
    for mul, (l, p) in out:
        if drop_l is not False and l > drop_l:
            continue
        if parity_mode == 'even' and p == -1:
            continue
        if parity_mode == 'sph' and p != (-1) ** l:
            continue
        new.append((fix_multiplicity if fix_multiplicity else mul, (l, p)))
    return Irreps(new)


DEFAULTS = {  # sevenn/_const.py:87-128
    'cutoff': 4.5, 'channel': 32, 'irreps_manual': False, 'lmax': 1, 'lmax_edge': -1,
    'lmax_node': -1, 'is_parity': True, 'num_convolution_layer': 3,
    'radial_basis': {'radial_basis_name': 'bessel'},
    'cutoff_function': {'cutoff_function_name': 'poly_cut'},
    'act_radial': 'silu', 'weight_nn_hidden_neurons': [64, 64],
    'act_scalar': {'e': 'silu', 'o': 'tanh'}, 'act_gate': {'e': 'silu', 'o': 'tanh'},
    'conv_denominator': 1.0, 'train_denominator': False, 'train_shift_scale': False,
    'use_bias_in_linear': False, 'readout_as_fcn': False, 'self_connection_type': 'nequip',
    'interaction_type': 'nequip', '_normalize_sph': True, 'shift': 0.0, 'scale': 1.0,
}


def init_edge_embedding(cfg):
    """sevenn/model_build.py:69-116 (single cutoff)"""
    env = dict(cfg['cutoff_function'])
    name = env.pop('cutoff_function_name')
    if name == 'poly_cut':
        cut = PolynomialCutoff(cfg['cutoff'], **env)
    elif name == 'XPLOR':
        cut = XPLORCutoff(cfg['cutoff'], **env)
    else:
        raise ValueError(name)
    lmax_edge = cfg['lmax_edge'] if cfg['lmax_edge'] > 0 else cfg['lmax']
    sph = SphericalEncoding(lmax_edge, -1 if cfg['is_parity'] else 1, normalize=cfg['_normalize_sph'])
    rb = dict(cfg['radial_basis'])
    assert rb.pop('radial_basis_name') == 'bessel'
    if cfg.get('multi_cutoff'):
        # model_build.py:87-103; the reference iterates list(set(...)), i.e. ascending for small numbers
        return MultiCutoffEdgeEmbedding([BesselBasis(c, **rb) for c in sorted(set(cfg['multi_cutoff']))], cut, sph)
    return EdgeEmbedding(BesselBasis(cfg['cutoff'], **rb), cut, sph)


def build_model(config, parallel=False):
    """sevenn/model_build.py:384-576 + nn/interaction_blocks.py:12-88.
    parallel=True returns the list of per-layer segments of
    model_build.py:299-367."""
    cfg = {**DEFAULTS, **copy.deepcopy(config)}
    layers = OrderedDict()
    S, C, L = cfg['_number_of_species'], cfg['channel'], cfg['num_convolution_layer']
    lmax_node = cfg['lmax_node'] if cfg['lmax_node'] > 0 else cfg['lmax']
    act_radial = ACT[cfg['act_radial']]
    sc_type = cfg['self_connection_type']
    sc_cls = {'nequip': SelfConnectionIntro, 'linear': SelfConnectionLinearIntro}[sc_type]
    manual = None
    if cfg['irreps_manual'] is not False:
        manual = [Irreps(s) for s in cfg['irreps_manual']]
        assert len(manual) == L + 1
    denom = cfg['conv_denominator']
    denom = denom if isinstance(denom, list) else [denom] * L
    ee = init_edge_embedding(cfg)
    irreps_filter = ee.spherical.irreps_out
    layers['edge_embedding'] = ee
    irreps_x = Irreps(f'{C}x0e') if manual is None else manual[0]
    layers['onehot_idx_to_onehot'] = OnehotEmbedding(S, X, X, ATOM_TYPE, NODE_ATTR)
    nm = cfg.get('_number_of_modalities', 0) if cfg.get('use_modality', False) else 0     # model_build.py:198-244
    m_emb, m_si1 = (nm if cfg.get(k, False) else 0 for k in ('use_modal_node_embedding', 'use_modal_self_inter_intro'))
    m_si2, m_out = (nm if cfg.get(k, False) else 0 for k in ('use_modal_self_inter_outro', 'use_modal_output_block'))
    if nm:
        layers['one_hot_modality'] = OnehotEmbedding(nm, MODAL_TYPE, MODAL_ATTR)
    layers['onehot_to_feature_x'] = IrrepsLinear(Irreps(f'{S}x0e'), irreps_x, X,
                                                 biases=cfg['use_bias_in_linear'], num_modalities=m_emb)
    multi = cfg.get('multi_cutoff') or None
    basis0 = ee.basis_functions[0] if multi else ee.basis_function
    wnn = [basis0.num_basis] + list(cfg['weight_nn_hidden_neurons'])
    act_s = {k: ACT[v] for k, v in cfg['act_scalar'].items()}
    act_g = {k: ACT[v] for k, v in cfg['act_gate'].items()}
    for t in range(L):
        parity_mode, lm = 'full', lmax_node
        if t == L - 1:
            lm, parity_mode = 0, 'even'
        irreps_out = (infer_irreps_out(irreps_x, irreps_filter, lm, parity_mode, C)
                      if manual is None else manual[t + 1])
        irreps_out_tp = infer_irreps_out(irreps_x, irreps_filter, irreps_out.lmax, parity_mode, False)
        gate = EquivariantGate(irreps_out, act_s, act_g)
        gate_in = gate.get_gate_irreps_in()
        layers[f'{t}_self_connection_intro'] = sc_cls(irreps_x, Irreps(f'{S}x0e'), gate_in)
        layers[f'{t}_self_interaction_1'] = IrrepsLinear(irreps_x, irreps_x, X,
                                                         biases=cfg['use_bias_in_linear'], num_modalities=m_si1)
        ckeys = {}
        if multi:                                        # interaction_blocks.py:54-60
            ckeys = dict(data_key_filter=f'{EDGE_ATTR}{multi[t]}', data_key_weight_input=f'{EDGE_EMB}{multi[t]}',
                         data_key_edge_idx=f'{EDGE_IDX}{multi[t]}')
        layers[f'{t}_convolution'] = IrrepsConvolution(
            irreps_x, irreps_filter, irreps_out_tp, wnn, act_radial, denom[t],
            cfg['train_denominator'], is_parallel=parallel, **ckeys)
        layers[f'{t}_self_interaction_2'] = IrrepsLinear(
            layers[f'{t}_convolution'].irreps_mid, gate_in, X, biases=cfg['use_bias_in_linear'], num_modalities=m_si2)
        layers[f'{t}_self_connection_outro'] = SelfConnectionOutro()
        layers[f'{t}_equivariant_gate'] = gate
        irreps_x = irreps_out
    if cfg['readout_as_fcn']:                            # model_build.py:141-155
        layers['readout_FCN'] = FCN_e3nn(irreps_x, 1, cfg.get('readout_fcn_hidden_neurons', [30, 30]),
                                         ACT[cfg.get('readout_fcn_activation', 'relu')], X, SCALED_AE)
    else:
        hidden = Irreps([(irreps_x.dim // 2, (0, 1))])
        layers['reduce_input_to_hidden'] = IrrepsLinear(irreps_x, hidden, X, biases=cfg['use_bias_in_linear'],
                                                        num_modalities=m_out)
        layers['reduce_hidden_to_energy'] = IrrepsLinear(hidden, Irreps('1x0e'), X, SCALED_AE,
                                                         biases=cfg['use_bias_in_linear'])
    shift, scale = cfg['shift'], cfg['scale']
    shift = shift.tolist() if hasattr(shift, 'tolist') else shift
    scale = scale.tolist() if hasattr(scale, 'tolist') else scale
    if nm:                                               # model_build.py:173-184
        # ModalWiseRescale.from_mappers (nn/scale.py:238-330): a float broadcasts to [n_modal, n_types] (modal wise) or
        # [n_types]; a per-modality list broadcasts over types; everything else is taken as it is
        def _bcast(v, mw):
            if isinstance(v, float):
                return [[v] * S] * nm if mw else [v] * S
            if mw and isinstance(v, list) and isinstance(v[0], float) and len(v) == nm:
                return [[t] * S for t in v]
            return v
        shift = _bcast(shift, cfg.get('use_modal_wise_shift', False))
        scale = _bcast(scale, cfg.get('use_modal_wise_scale', False))
        layers['rescale_atomic_energy'] = ModalWiseRescale(
            shift, scale, cfg.get('use_modal_wise_shift', False), cfg.get('use_modal_wise_scale', False),
            cfg['train_shift_scale'])
    elif isinstance(shift, float) and isinstance(scale, float):
        layers['rescale_atomic_energy'] = Rescale(shift, scale, cfg['train_shift_scale'])
    else:
        shift = shift * S if isinstance(shift, list) and len(shift) == 1 else shift
        scale = scale * S if isinstance(scale, list) and len(scale) == 1 else scale
        layers['rescale_atomic_energy'] = SpeciesWiseRescale(shift, scale, cfg['train_shift_scale'])
    layers['reduce_total_enegy'] = AtomReduce()
    layers['force_output'] = ForceStressOutputFromEdge()
    common = dict(cutoff=cfg['cutoff'], type_map=cfg['_type_map'], data_key_grad=EDGE_VEC)
    if not parallel:
        return AtomGraphSequential(layers, **common)
    # ---- model_build.py:299-367: split at every {t}_self_interaction_1, t>=1
    items = list(layers.items())

    def insert_after(name, pair):
        idx = [k for k, _ in items].index(name)
        items.insert(idx + 1, pair)

    irreps0 = layers['onehot_to_feature_x'].irreps_out
    insert_after('onehot_to_feature_x', ('one_hot_ghost', OnehotEmbedding(S, X_GHOST)))
    insert_after('one_hot_ghost', ('ghost_onehot_to_feature_x',
                                   IrrepsLinear(Irreps(f'{S}x0e'), irreps0, X_GHOST)))
    insert_after('0_self_interaction_1', ('ghost_0_self_interaction_1',
                                          IrrepsLinear(irreps0, irreps0, X_GHOST)))
    segs = []
    for i in range(1, L):
        idx = [k for k, _ in items].index(f'{i}_self_interaction_1')
        segs.append(OrderedDict(items[: idx + 1]))
        items = [('edge_embedding', init_edge_embedding(cfg))] + items[idx + 1:]
    last = OrderedDict(items)
    del last['force_output']
    segs.append(last)
    return [AtomGraphSequential(s, eval_type_map=False, **common) for s in segs]


# ---------------------------------------------------------- checkpoints --
def sort_old_convolution(model, state_dict):
    """sevenn/scripts/backward_compatibility.py:74-154: permute (and sign-fix)
    the last radial-MLP layer columns from creation order to sorted-instruction
    order for checkpoints written before 0.11.0."""
    sd = dict(state_dict)
    for name, conv in model.named_children():
        if not isinstance(conv, IrrepsConvolution):
            continue
        old = [(a, b, c) for a, b, c, _, _ in conv._instructions_before_sort]
        new = sorted(old, key=lambda t: t[2])
        n = len(conv.weight_nn.hs) - 2
        key = f'{name}.weight_nn.layer{n}.weight'
        ww = sd[key]
        cols, off = [None] * len(old), 0
        for ins in old:
            mul = conv.irreps_x[ins[0]][0]
            blk = ww[:, off:off + mul]
            off += mul
            l1, l2, l3 = conv.irreps_x[ins[0]][1].l, conv.irreps_filter[ins[1]][1].l, conv.irreps_mid[ins[2]][1].l
            wkey = f'{name}.convolution._compiled_main_left_right._w3j_{l1}_{l2}_{l3}'
            if l1 > 0 and l2 > 0 and l3 > 0 and wkey in sd:
                now = o3.wigner_3j(l1, l2, l3)
                if not torch.allclose(sd[wkey], now):
                    assert torch.allclose(sd[wkey], -now)
                    blk = -blk
            cols[new.index(ins)] = blk
        sd[key] = torch.cat(cols, dim=1).clone()
    return sd


def load_checkpoint(path, parallel=False):
    """sevenn/checkpoint.py:258-311 (+ backward_compatibility.patch_state_dict_if_old)"""
    cp = torch.load(path, weights_only=False, map_location='cpu')
    cfg = dict(cp['config'])
    model = build_model(cfg)
    sd = cp['model_state_dict']
    ver = '.'.join(str(cfg.get('version', '0.0.0')).split('.')[:3])
    if ver < '0.11.0':
        sd = sort_old_convolution(model, sd)
    own = model.state_dict()
    missing = [k for k in own if k not in sd]
    assert not missing, missing
    model.load_state_dict({k: v for k, v in sd.items() if k in own}, strict=False)
    return model, cfg


# -------------------------------------------------------- neighbor list --
def neighbor_list_bruteforce(pos, cell, pbc, cutoff):
    """Edge-set oracle for sevenn/train/dataload.py:32-88: all (i, j, S) with
    |pos_j - pos_i + S.cell| < cutoff, minus the zero-vector self pair; float64."""
    pos, cell = np.asarray(pos, dtype=np.float64), np.asarray(cell, dtype=np.float64)
    pbc = np.asarray(pbc, dtype=bool)
    n = len(pos)
    reps = []
    for d in range(3):
        if not pbc[d]:
            reps.append(0)
            continue
        a, b = cell[(d + 1) % 3], cell[(d + 2) % 3]
        nrm = np.cross(a, b)
        h = abs(np.dot(cell[d], nrm)) / np.linalg.norm(nrm)
        # positions need not be wrapped: allow for |frac| spread as well
        frac = pos @ np.linalg.inv(cell)
        spread = int(np.ceil(frac[:, d].max() - frac[:, d].min())) if n else 0
        reps.append(int(np.ceil(cutoff / h)) + spread)
    src, dst, vec, shf = [], [], [], []
    rng = [range(-r, r + 1) for r in reps]
    for sx in rng[0]:
        for sy in rng[1]:
            for sz in rng[2]:
                S = np.array([sx, sy, sz], dtype=np.float64)
                off = S @ cell
                D = pos[None, :, :] + off[None, None, :] - pos[:, None, :]   # [i, j, 3]
                d2 = (D ** 2).sum(-1)
                mask = d2 < cutoff * cutoff
                if sx == 0 and sy == 0 and sz == 0:
                    mask &= ~np.eye(n, dtype=bool)
                i, j = np.nonzero(mask)
                src.append(i); dst.append(j); vec.append(D[i, j]); shf.append(np.tile(S, (len(i), 1)))
    src, dst = np.concatenate(src), np.concatenate(dst)
    vec, shf = np.concatenate(vec), np.concatenate(shf)
    order = np.lexsort((shf[:, 2], shf[:, 1], shf[:, 0], dst, src))
    return src[order].astype(np.int64), dst[order].astype(np.int64), vec[order], shf[order]


def atoms_to_graph(numbers, pos, cell, pbc, cutoff, dtype=torch.float32):
    """sevenn/train/dataload.py:102-130 + atom_graph_data.py:69-75."""
    i, j, D, S = neighbor_list_bruteforce(pos, cell, pbc, cutoff)
    vol = abs(float(np.linalg.det(np.asarray(cell, dtype=np.float64))))
    if vol == 0:
        vol = float(np.finfo(float).eps)
    return {
        X: torch.as_tensor(np.asarray(numbers), dtype=torch.int64),
        Z: torch.as_tensor(np.asarray(numbers), dtype=torch.int64),
        POS: torch.as_tensor(np.asarray(pos), dtype=dtype),
        EDGE_IDX: torch.as_tensor(np.stack([i, j]), dtype=torch.int64),
        EDGE_VEC: torch.as_tensor(D, dtype=dtype),
        CELL: torch.as_tensor(np.asarray(cell), dtype=dtype),
        SHIFT: torch.as_tensor(S, dtype=dtype),
        VOL: torch.tensor(vol, dtype=dtype),
        NATOMS: torch.tensor(len(numbers), dtype=torch.int64),
    }


def read_extxyz(path):
    """Minimal extxyz reader for tests/data/systems/hfo2.extxyz (species, pos,
    forces; Lattice, stress, energy, pbc in the comment line)."""
    import re
    from . import periodic_table
    frames, lines, k = [], open(path).read().splitlines(), 0
    while k < len(lines) and lines[k].strip():
        n = int(lines[k]); com = lines[k + 1]
        kv = dict(re.findall(r'(\w+)=("[^"]*"|\S+)', com))
        kv = {a: b.strip('"') for a, b in kv.items()}
        lat = np.array(kv['Lattice'].split(), dtype=float).reshape(3, 3)
        props = kv['Properties'].split(':')
        cols, c = {}, 0
        for a in range(0, len(props), 3):
            cols[props[a]] = (c, c + int(props[a + 2])); c += int(props[a + 2])
        rows = [l.split() for l in lines[k + 2:k + 2 + n]]
        fr = {
            'numbers': np.array([periodic_table.Z_OF[r[0]] for r in rows]),
            'pos': np.array([[float(v) for v in r[cols['pos'][0]:cols['pos'][1]]] for r in rows]),
            'cell': lat, 'pbc': np.array([t == 'T' for t in kv.get('pbc', 'T T T').split()]),
        }
        if 'forces' in cols:
            fr['forces'] = np.array([[float(v) for v in r[cols['forces'][0]:cols['forces'][1]]] for r in rows])
        if 'stress' in kv:
            fr['stress3x3'] = np.array(kv['stress'].split(), dtype=float).reshape(3, 3)
        e = kv.get('free_energy', kv.get('energy'))
        if e is not None:
            fr['energy'] = float(e)
        frames.append(fr)
        k += 2 + n
    return frames
