"""Structural description of a SevenNet model in device form.

Mirrors the shape/ordering bookkeeping of ``sevenn/model_build.py:384-576``,
``sevenn/nn/interaction_blocks.py:12-88``, ``sevenn/nn/convolution.py:60-81``
(uvu instruction list sorted by output index), ``sevenn/nn/equivariant_gate.py``
and e3nn's ``o3.Linear`` / ``nn.Gate`` weight layouts, and turns them into the
flat tables the CUDA kernels consume (see irreps.Layout for the ir_mul layout).
"""
import copy
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import irreps as ir
from .irreps import Layout

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

ACT_ID = {'silu': 0, 'tanh': 1, 'abs': 2, 'relu': 3, 'sigmoid': 4, 'ssp': 5}
ACT_PARITY = {'silu': 0, 'tanh': -1, 'abs': 1, 'relu': 0, 'sigmoid': 0, 'ssp': 0}  # 0 = neither


@dataclass
class LinearBlock:
    l: int
    g_in: int
    g_out: int
    K: int                    # sum of contributing input multiplicities (= fan_in)
    N: int                    # output multiplicity of this output entry
    u_off_out: int            # column offset inside the output group
    segs: List[Tuple[int, int]]   # (flat weight offset, mul_in) stacked along K
    fan_extra: int = 0        # modality one-hot channels that also feed this (0e) output: fan_in = K + fan_extra
    modal_off: int = -1       # flat offset of their [n_modal, N] weights
    bias_off: int = -1        # offset of this (0e) output entry inside o3.Linear's bias vector
    o_entry: int = -1         # index of the output entry (e3nn order) this block writes


@dataclass
class LinearSpec:
    key: str                  # state_dict key of the flat e3nn weight
    lin: Layout
    lout: Layout
    blocks: List[LinearBlock]
    numel: int
    n_modal: int = 0          # sevenn/nn/linear.py:65-92: the input carries n_modal extra 0e one-hot channels
    n_bias: int = 0           # o3.Linear(biases=True): one bias per 0e output channel (key + '.bias' next to '.weight')

    def dense(self, flat: np.ndarray, block: LinearBlock) -> np.ndarray:
        """[K, N] weight of one block, 1/sqrt(fan_in) folded in."""
        rows = [flat[o:o + m * block.N].reshape(m, block.N) for o, m in block.segs]
        return np.concatenate(rows, axis=0) / math.sqrt(block.K + block.fan_extra)

    @property
    def bias_key(self) -> str:
        return self.key[:-len('weight')] + 'bias'

    def constant_row(self, flat: np.ndarray, block: LinearBlock, modal: int = 0, bias=None):
        """What a fixed modality and the bias add to every row of this block's output ([N], or None): the one-hot
        picks row ``modal`` of the modality weights (normalised like all weights); the bias is added as is."""
        row = None
        if block.modal_off >= 0:
            o = block.modal_off + modal * block.N
            row = flat[o:o + block.N] / math.sqrt(block.K + block.fan_extra)
        if block.bias_off >= 0 and bias is not None:
            b = np.asarray(bias[block.bias_off:block.bias_off + block.N], dtype=np.float64)
            row = b if row is None else row + b
        return row


def make_linear(key: str, irreps_in, irreps_out, n_modal: int = 0, biases: bool = False) -> LinearSpec:
    lin, lout = Layout(irreps_in), Layout(irreps_out)
    # e3nn instruction order: for i_in, for i_out, if same irrep
    off, table = 0, {}
    for i, (mi, iri) in enumerate(lin.entries):
        for o, (mo, iro) in enumerate(lout.entries):
            if iri == iro:
                table[(i, o)] = off
                off += mi * mo
    # the modality one-hot is the LAST input entry (irreps_in + 'Mx0e'): its instructions close the weight vector
    modal_tab, bias_tab, b_off = {}, {}, 0
    for o, (mo, iro) in enumerate(lout.entries):
        if tuple(iro) == (0, 1):
            if n_modal > 1:
                modal_tab[o] = off
                off += n_modal * mo
            if biases:
                bias_tab[o] = b_off
                b_off += mo
    blocks = []
    for o, (mo, iro) in enumerate(lout.entries):
        segs = [(table[(i, o)], mi) for i, (mi, iri) in enumerate(lin.entries) if iri == iro]
        if not segs or mo == 0:
            continue
        K = sum(m for _, m in segs)
        g_in = lin.group_of(iro)
        assert lin.groups[g_in][0] == K, 'inputs of one irrep must be contiguous'
        blocks.append(LinearBlock(iro[0], g_in, lout.entry_group[o], K, mo, lout.entry_uoff[o], segs,
                                  fan_extra=n_modal if o in modal_tab else 0, modal_off=modal_tab.get(o, -1),
                                  bias_off=bias_tab.get(o, -1), o_entry=o))
    assert all(any(b.modal_off == v for b in blocks) for v in modal_tab.values()), 'modal-only outputs unsupported'
    return LinearSpec(key, lin, lout, blocks, off, n_modal=n_modal if n_modal > 1 else 0, n_bias=b_off)


@dataclass
class ConvPath:
    gx: int        # input group (in lx)
    l1: int
    l2: int
    l3: int
    gm: int        # output group in lmid
    u_off: int     # channel offset inside the output group
    mul: int
    w_off: int     # offset inside the per-edge weight row


@dataclass
class ConvSpec:
    name: str
    lx: Layout
    lmid: Layout
    lmax_sh: int
    paths: List[ConvPath]
    wn: int
    hidden: List[int]          # radial MLP sizes [nb, h1, h2, wn]
    paths_creation_order: List[Tuple[int, int, int, int]] = field(default_factory=list)
    # (w_off_old, mul, new path index, (l1,l2,l3)) for checkpoints < 0.11.0

    def signature(self) -> str:
        gx = '_'.join(f'{m}x{l}' for m, (l, _) in self.lx.groups)
        gm = '_'.join(f'{m}x{l}' for m, (l, _) in self.lmid.groups)
        ps = '_'.join(f'{p.gx}{p.l1}{p.l2}{p.l3}{p.gm}' for p in self.paths)
        import hashlib
        h = hashlib.sha1(f'{gx}|{gm}|{ps}|{self.lmax_sh}'.encode()).hexdigest()[:10]
        return f'c{h}'


def make_conv(name, irreps_x, irreps_filter, irreps_out, hidden) -> ConvSpec:
    """sevenn/nn/convolution.py:60-81."""
    lx = Layout(irreps_x)
    assert lx.groups == list(irreps_x), 'conv input irreps must be simplified'
    instr, mid = [], []
    out_set = {i for _, i in irreps_out}
    for i, (mul_x, ir_x) in enumerate(irreps_x):
        for j, (_, ir_f) in enumerate(irreps_filter):
            for ir_o in ir.tp_irreps(ir_x, ir_f):
                if ir_o in out_set:
                    instr.append((i, j, len(mid)))
                    mid.append((mul_x, ir_o))
    mid_sorted, perm = ir.sort_irreps(mid)
    instr = [(a, b, perm[c]) for a, b, c in instr]
    lmid = Layout(mid_sorted)
    # creation order bookkeeping (old checkpoints), then sort by output index
    creation, off = [], 0
    for a, b, c in instr:
        creation.append((off, irreps_x[a][0], c))
        off += irreps_x[a][0]
    order = sorted(range(len(instr)), key=lambda t: instr[t][2])
    paths, w_off = [], 0
    for t in order:
        a, b, c = instr[t]
        mul, (l1, _) = irreps_x[a]
        l2 = irreps_filter[b][1][0]
        l3 = mid_sorted[c][1][0]
        paths.append(ConvPath(a, l1, l2, l3, lmid.entry_group[c], lmid.entry_uoff[c], mul, w_off))
        w_off += mul
    pco = []
    for (o_old, mul, c) in creation:
        a, b, _ = instr[[i for i, t in enumerate(instr) if t[2] == c][0]]
        pco.append((o_old, mul, c, (irreps_x[a][1][0], irreps_filter[b][1][0], mid_sorted[c][1][0])))
    lmax_sh = max(l for _, (l, _) in irreps_filter)
    return ConvSpec(name, lx, lmid, lmax_sh, paths, w_off, list(hidden) + [w_off], pco)


@dataclass
class GatePiece:
    kind: int        # 0 = scalar, 1 = gated
    l: int
    mul: int
    in_off: int      # float offset of the (first row of the) source block inside the input row
    in_stride: int   # row stride (= group multiplicity) of the source block
    gate_off: int    # float offset of the gate scalars (kind 1)
    out_off: int
    out_stride: int
    act: int
    cst: float


@dataclass
class GateSpec:
    lin: Layout
    lout: Layout
    pieces: List[GatePiece]


def make_gate(irreps_x, act_scalar: Dict[str, str], act_gate: Dict[str, str]):
    """sevenn/nn/equivariant_gate.py:13-61 + e3nn.nn.Gate (sort + simplify of its input)."""
    pm = {1: 'e', -1: 'o'}
    scalars = [(m, i) for m, i in irreps_x if i[0] == 0]
    gated = [(m, i) for m, i in irreps_x if i[0] > 0]
    gp = 1 if (0, 1) in [i for _, i in scalars] else -1
    gates = [(m, (0, gp)) for m, _ in gated]
    cat = scalars + gates + gated
    srt, perm = ir.sort_irreps(cat)
    lin = Layout(srt)

    def out_parity(act, p):
        if p == 1:
            return 1
        q = ACT_PARITY[act]
        assert q != 0, 'activation on odd scalars must be even or odd'
        return q

    out_sc = [(m, (0, out_parity(act_scalar[pm[i[1]]], i[1]))) for m, i in scalars]
    gate_par = [out_parity(act_gate[pm[i[1]]], i[1]) for _, i in gates]
    out_gd = [(m, (i[0], i[1] * gpar)) for (m, i), gpar in zip(gated, gate_par)]
    irreps_out = out_sc + out_gd
    lout = Layout(irreps_out)
    pieces = []

    def src(k_cat):
        k = perm[k_cat]
        g = lin.entry_group[k]
        return lin.group_off[g] + lin.entry_uoff[k], lin.groups[g][0]

    for k, (m, i) in enumerate(scalars):
        a = act_scalar[pm[i[1]]]
        off, st = src(k)
        og = lout.entry_group[k]
        pieces.append(GatePiece(0, 0, m, off, st, 0, lout.group_off[og] + lout.entry_uoff[k],
                                lout.groups[og][0], ACT_ID[a], ir.normalize2mom_const(a)))
    ns = len(scalars)
    for k, (m, i) in enumerate(gated):
        a = act_gate[pm[gates[k][1][1]]]
        goff, _ = src(ns + k)
        off, st = src(ns + len(gates) + k)
        og = lout.entry_group[ns + k]
        pieces.append(GatePiece(1, i[0], m, off, st, goff, lout.group_off[og] + lout.entry_uoff[ns + k],
                                lout.groups[og][0], ACT_ID[a], ir.normalize2mom_const(a)))
    return GateSpec(lin, lout, pieces), srt, irreps_out


@dataclass
class LayerSpec:
    t: int
    sc_type: str                     # 'linear' | 'nequip'
    sc: LinearSpec                   # for 'nequip' the flat weight is [mul_in, S, mul_out] per block
    si1: LinearSpec
    conv: ConvSpec
    si2: LinearSpec
    gate: GateSpec
    denominator: float


@dataclass
class ModelSpec:
    cfg: dict
    n_species: int
    type_map: Dict[int, int]
    cutoff: float
    lmax_sh: int
    sh_parity: int
    normalize_sph: bool
    n_basis: int
    cutoff_fn: str                   # 'poly_cut' | 'XPLOR'
    cutoff_p: float
    cutoff_on: float
    act_radial: str
    embed: LinearSpec
    layers: List[LayerSpec]
    read1: LinearSpec
    read2: LinearSpec
    rescale_species_wise: bool
    multi_cutoff: Optional[List[float]] = None     # per-layer cutoffs (MultiCutoffEdgeEmbedding), else None
    n_modal: int = 0                                # number of modalities (model_build.py:198-244), 0 = none
    modal_map: Optional[Dict[str, int]] = None      # modality name -> index
    modal_wise_shift: bool = False                  # rescale_atomic_energy.shift is [n_modal, n_species]
    modal_wise_scale: bool = False
    readout_fcn: Optional[List[int]] = None         # readout_as_fcn (model_build.py:141-155): [in, hidden..., 1]
    readout_act: str = 'relu'

    def linears(self) -> List['LinearSpec']:
        out = [self.embed]
        for L in self.layers:
            out += [L.sc, L.si1, L.si2]
        return out + ([] if self.readout_fcn else [self.read1, self.read2])

    def coeff_key(self, t: int) -> str:
        """state_dict key of the Bessel coefficients convolution layer ``t`` uses (model_build.py:87-103: one
        BesselBasis per distinct cutoff, ascending)."""
        if not self.multi_cutoff:
            return 'edge_embedding.basis_function.coeffs'
        return f'edge_embedding.basis_functions.{sorted(set(self.multi_cutoff)).index(self.multi_cutoff[t])}.coeffs'

    def expected_keys(self) -> Dict[str, Tuple[int, ...]]:
        out = {self.coeff_key(t): (self.n_basis,) for t in range(len(self.layers))}
        out[self.embed.key] = (self.embed.numel,)
        for L in self.layers:
            out[L.sc.key] = (L.sc.numel,)
            out[L.si1.key] = (L.si1.numel,)
            h = L.conv.hidden
            for k in range(len(h) - 1):
                out[f'{L.conv.name}.weight_nn.layer{k}.weight'] = (h[k], h[k + 1])
            out[f'{L.conv.name}.denominator'] = (1,)
            out[L.si2.key] = (L.si2.numel,)
        if self.readout_fcn:
            for k, (a, b) in enumerate(zip(self.readout_fcn, self.readout_fcn[1:])):
                out[f'readout_FCN.fcn.layer{k}.weight'] = (a, b)
        else:
            out[self.read1.key] = (self.read1.numel,)
            out[self.read2.key] = (self.read2.numel,)
        for lin in self.linears():
            if lin.n_bias:
                out[lin.bias_key] = (lin.n_bias,)
        n = self.n_species if self.rescale_species_wise else 1
        out['rescale_atomic_energy.shift'] = (self.n_modal, n) if self.modal_wise_shift else (n,)
        out['rescale_atomic_energy.scale'] = (self.n_modal, n) if self.modal_wise_scale else (n,)
        return out


def build_spec(config: dict) -> ModelSpec:
    cfg = {**DEFAULTS, **copy.deepcopy(config)}
    bias = bool(cfg['use_bias_in_linear'])
    nm = int(cfg.get('_number_of_modalities', 0)) if cfg.get('use_modality', False) else 0
    m_emb, m_si1, m_si2, m_out = (nm if cfg.get(k, False) else 0 for k in (
        'use_modal_node_embedding', 'use_modal_self_inter_intro', 'use_modal_self_inter_outro',
        'use_modal_output_block'))
    assert cfg['radial_basis']['radial_basis_name'] == 'bessel'
    S, C, L = cfg['_number_of_species'], cfg['channel'], cfg['num_convolution_layer']
    lmax_node = cfg['lmax_node'] if cfg['lmax_node'] > 0 else cfg['lmax']
    lmax_edge = cfg['lmax_edge'] if cfg['lmax_edge'] > 0 else cfg['lmax']
    sh_p = -1 if cfg['is_parity'] else 1
    irreps_filter = ir.spherical_harmonics_irreps(lmax_edge, sh_p)
    manual = None
    if cfg['irreps_manual'] is not False:
        manual = [ir.parse_irreps(s) for s in cfg['irreps_manual']]
        assert len(manual) == L + 1
    denom = cfg['conv_denominator']
    denom = list(denom) if isinstance(denom, (list, tuple)) else [denom] * L
    nb = cfg['radial_basis'].get('bessel_basis_num', 8)
    hidden = [nb] + list(cfg['weight_nn_hidden_neurons'])
    one_hot = [(S, (0, 1))]
    irreps_x = [(C, (0, 1))] if manual is None else manual[0]
    embed = make_linear('onehot_to_feature_x.linear.weight', one_hot, irreps_x, m_emb, bias)
    layers = []
    for t in range(L):
        parity_mode, lm = 'full', lmax_node
        if t == L - 1:
            lm, parity_mode = 0, 'even'
        irreps_out = (ir.infer_irreps_out(irreps_x, irreps_filter, lm, parity_mode, C)
                      if manual is None else manual[t + 1])
        out_lmax = max(l for _, (l, _) in irreps_out)
        irreps_out_tp = ir.infer_irreps_out(irreps_x, irreps_filter, out_lmax, parity_mode, False)
        gate, gate_in, gate_out = make_gate(irreps_out, cfg['act_scalar'], cfg['act_gate'])
        gate_in_s = ir.simplify(gate_in)
        sc_type = cfg['self_connection_type']
        if sc_type == 'linear':
            sc = make_linear(f'{t}_self_connection_intro.linear.weight', irreps_x, gate_in_s)
        elif sc_type == 'nequip':
            sc = make_linear(f'{t}_self_connection_intro.fc_tensor_product.weight', irreps_x, gate_in_s)
            sc.numel *= S
        else:
            raise ValueError(sc_type)
        si1 = make_linear(f'{t}_self_interaction_1.linear.weight', irreps_x, irreps_x, m_si1, bias)
        conv = make_conv(f'{t}_convolution', irreps_x, irreps_filter, irreps_out_tp, hidden)
        si2 = make_linear(f'{t}_self_interaction_2.linear.weight', conv.lmid.entries, gate_in_s, m_si2, bias)
        layers.append(LayerSpec(t, sc_type, sc, si1, conv, si2, gate, float(denom[t])))
        irreps_x = ir.simplify(gate_out)
        assert irreps_x == list(gate_out), 'gate output irreps must be simplified'
    hid = [(ir.irreps_dim(irreps_x) // 2, (0, 1))]
    read1 = make_linear('reduce_input_to_hidden.linear.weight', irreps_x, hid, m_out, bias)
    read2 = make_linear('reduce_hidden_to_energy.linear.weight', hid, [(1, (0, 1))], 0, bias)
    readout_fcn = None
    if cfg['readout_as_fcn']:
        assert all(tuple(i) == (0, 1) for _, i in irreps_x), 'readout_as_fcn needs scalar node features'
        readout_fcn = [ir.irreps_dim(irreps_x)] + list(cfg.get('readout_fcn_hidden_neurons', [30, 30])) + [1]
    shift, scale = cfg['shift'], cfg['scale']
    shift = shift.tolist() if hasattr(shift, 'tolist') else shift
    scale = scale.tolist() if hasattr(scale, 'tolist') else scale
    species_wise = bool(nm) or not (isinstance(shift, float) and isinstance(scale, float))
    cf = dict(cfg['cutoff_function'])
    name = cf['cutoff_function_name']
    return ModelSpec(
        cfg=cfg, n_species=S, type_map=dict(cfg.get('_type_map') or {}), cutoff=float(cfg['cutoff']),
        lmax_sh=lmax_edge, sh_parity=sh_p, normalize_sph=bool(cfg['_normalize_sph']), n_basis=nb,
        cutoff_fn=name, cutoff_p=float(cf.get('poly_cut_p_value', 6)),
        cutoff_on=float(cf.get('cutoff_on', 0.0)), act_radial=cfg['act_radial'], embed=embed,
        layers=layers, read1=read1, read2=read2, rescale_species_wise=species_wise,
        multi_cutoff=[float(c) for c in cfg['multi_cutoff']] if cfg.get('multi_cutoff') else None,
        n_modal=nm, modal_map=dict(cfg.get('_modal_map') or {}) or None,
        modal_wise_shift=bool(nm and cfg.get('use_modal_wise_shift', False)),
        modal_wise_scale=bool(nm and cfg.get('use_modal_wise_scale', False)),
        readout_fcn=readout_fcn, readout_act=cfg.get('readout_fcn_activation', 'relu'))


def nequip_sc_dense(spec: LinearSpec, flat: np.ndarray, block: LinearBlock, S: int) -> np.ndarray:
    """FullyConnectedTensorProduct(x, one_hot) as a species-indexed linear:
    returns [S, K, N] with the e3nn path factor folded in
    (sevenn/nn/self_connection.py:10-64; SURVEY Appendix A rule 6).
    e3nn 'uvw' weight of one path is [mul_in, S, mul_out]; paths are ordered
    (i_in, i_operand=0, i_out) like o3.Linear's (i_in, i_out)."""
    rows = []
    for o, m in block.segs:
        o3 = o * S
        rows.append(flat[o3:o3 + m * S * block.N].reshape(m, S, block.N))
    W = np.concatenate(rows, axis=0)            # [K, S, N]
    l = block.l
    c = abs(float(ir.wigner_3j(l, 0, l)[0, 0, 0])) * math.sqrt(2 * l + 1)   # = 1
    sign = 1.0 if ir.wigner_3j(l, 0, l)[0, 0, 0] > 0 else -1.0
    return np.ascontiguousarray(W.transpose(1, 0, 2)) * (sign * c / math.sqrt(block.K * S))
