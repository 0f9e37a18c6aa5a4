"""Drop-in mirrors of the reference's ``sevenn.nn`` modules on the B200 kernels.

Same class names, constructor arguments, ``forward(data) -> data`` contract, data-dict keys
(sevenn/_keys.py) and parameter names / shapes as the reference, so a reference ``state_dict`` loads with
``load_state_dict`` unchanged.  Node features cross every module boundary in e3nn's ``[mul][2l+1]``
layout and are permuted to the library's ``[2l+1][mul]`` layout inside.  Gradients flow to activations
only (see functional.py)."""
import math
from collections import OrderedDict
from typing import Dict

import numpy as np
import torch
import torch.nn as nn

from .. import _lib
from .. import irreps as ir
from .. import spec as sp
from ..codegen import conv_codegen
from ..engine import _Lin
from . import functional as F
from . import train_functional as T
from ._backend import get_lib, kernels, ptr

# data keys (sevenn/_keys.py:26-86)
EDGE_VEC, EDGE_IDX, EDGE_LEN = 'edge_vec', 'edge_index', 'edge_length'
X, X_GHOST, NODE_ATTR, ATOM_TYPE = 'x', 'x_ghost', 'node_attr', 'atom_type'
EDGE_ATTR, EDGE_EMB = 'edge_attr', 'edge_embedding'
SC_TMP = 'self_cont_tmp'
SCALED_AE, AE, E_TOT = 'scaled_atomic_energy', 'atomic_energy', 'inferred_total_energy'
FORCE, STRESS = 'inferred_force', 'inferred_stress'
POS, CELL, SHIFT, VOL, NATOMS, BATCH, NLOCAL = (
    'pos', 'cell_lattice_vectors', 'pbc_shift', 'cell_volume', 'num_atoms', 'batch', 'nlocal')
Z = 'atomic_numbers'
GRAPH = '_b200_graph'          # cached CSR/CSC of data[EDGE_IDX] (private key, removed by force_output)


def _irreps(x):
    return ir.parse_irreps(x if isinstance(x, (str, list, tuple)) else str(x))


_PERM_CACHE = {}


def _perm(layout: ir.Layout, device):
    key = (id(layout), str(device))
    hit = _PERM_CACHE.get(key)
    if hit is None or hit[0] is not layout:
        p = torch.as_tensor(layout.perm_to_e3nn(), device=device)
        hit = _PERM_CACHE[key] = (layout, p, torch.argsort(p))
    return hit[1], hit[2]


class _PermCols(torch.autograd.Function):
    """x[:, p] for a PERMUTATION p of the columns (e3nn <-> device irreps layout).  torch's own backward of an index is a
    sort-based index_put (0.7 ms per call, half the GPU time of a training step); for a permutation it is simply the
    inverse gather, and the Function is its own derivative family (second order for the training path)."""

    @staticmethod
    def forward(ctx, x, p, inv):
        ctx.p, ctx.inv = p, inv
        return x.index_select(1, p)

    @staticmethod
    def backward(ctx, g):
        return _PermCols.apply(g, ctx.inv, ctx.p), None, None


def _pcols(x, layout, device, to_e3nn: bool):
    p, inv = _perm(layout, device)
    return _PermCols.apply(x, p, inv) if to_e3nn else _PermCols.apply(x, inv, p)


SIZES = '_b200_sizes'      # optional (total atoms, number of graphs) python ints in the data dict: no host sync needed


def _n_graphs(data):
    sz = data.get(SIZES)
    return int(sz[1]) if sz is not None else int(data[BATCH].max()) + 1


def _n_atoms_total(data):
    sz = data.get(SIZES)
    return int(sz[0]) if sz is not None else int(torch.sum(data[NATOMS]))


def _train_mode(module) -> bool:
    """Parameter gradients / double backward wanted: the module is in training mode and autograd is recording."""
    return module.training and torch.is_grad_enabled()


class _Weight(nn.Module):
    """Stand-in for the e3nn submodule that owns a flat ``weight`` (o3.Linear / FullyConnectedTensorProduct)."""

    def __init__(self, numel, n_bias=0):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(numel))
        if n_bias:
            self.bias = nn.Parameter(torch.zeros(n_bias))


class _LinearCore(nn.Module):
    """Flat e3nn weight -> device blocks (rebuilt when the parameter changes)."""

    def _setup(self, spec: sp.LinearSpec, owner: str, by_type_S: int = 0):
        self._spec, self._owner, self._S = spec, owner, by_type_S
        self._cache = None

    def _lin(self, device) -> _Lin:
        w = getattr(self, self._owner).weight
        key = (w._version, w.data_ptr(), str(device))
        if self._cache is None or self._cache[0] != key:
            flat = w.detach().cpu().numpy().astype(np.float64)
            lin = _Lin(self._spec, flat, device, self._S)
            lin.has_const = False          # modality / bias terms are added by the module (they depend on the data)
            self._cache = (key, lin)
        return self._cache[1]

    def _dense_blocks(self):
        """Dense per-block weights as differentiable torch ops on the flat e3nn parameter (training path): o3.Linear
        blocks [K, N] / sqrt(fan_in); FullyConnectedTensorProduct(x, one_hot) blocks [S, K, N] (spec.nequip_sc_dense)."""
        w, spec, S = getattr(self, self._owner).weight, self._spec, self._S
        out = []
        for b in spec.blocks:
            if S:
                rows = [w[o * S:o * S + m * S * b.N].reshape(m, S, b.N) for o, m in b.segs]
                sign = 1.0 if ir.wigner_3j(b.l, 0, b.l)[0, 0, 0] > 0 else -1.0
                c = abs(float(ir.wigner_3j(b.l, 0, b.l)[0, 0, 0])) * math.sqrt(2 * b.l + 1)
                out.append((torch.cat(rows, 0).permute(1, 0, 2) * (sign * c / math.sqrt(b.K * S))).contiguous())
            else:
                rows = [w[o:o + m * b.N].reshape(m, b.N) for o, m in b.segs]
                out.append((torch.cat(rows, 0) / math.sqrt(b.K + b.fan_extra)).contiguous())
        return out

    def _apply_linear(self, x, types=None):
        dev = x.device
        xin = _pcols(x.float(), self._spec.lin, dev, False)
        if _train_mode(self):
            y = T.BlockMM.apply(T._c(xin), self._spec, False, types, self._S, *self._dense_blocks())
        else:
            y = F.LinearFn.apply(xin, self._lin(dev), types)
        return _pcols(y, self._spec.lout, dev, True)


# ------------------------------------------------------------------ edge --
class BesselBasis(nn.Module):
    """sevenn/nn/edge_embedding.py:81-104 (parameter holder; evaluated inside EdgeEmbedding)."""

    def __init__(self, cutoff_length, bessel_basis_num=8, trainable_coeff=True):
        super().__init__()
        self.num_basis, self.cutoff_length = bessel_basis_num, cutoff_length
        c = torch.FloatTensor([n * math.pi / cutoff_length for n in range(1, bessel_basis_num + 1)])
        self.coeffs = nn.Parameter(c) if trainable_coeff else c      # trainable, edge_embedding.py:96-100


class PolynomialCutoff(nn.Module):
    """sevenn/nn/edge_embedding.py:107-133"""

    def __init__(self, cutoff_length, poly_cut_p_value=6):
        super().__init__()
        self.cutoff_length, self.p = float(cutoff_length), float(poly_cut_p_value)
        self.kind, self.r_on = 0, 0.0


class XPLORCutoff(nn.Module):
    """sevenn/nn/edge_embedding.py:136-161"""

    def __init__(self, cutoff_length, cutoff_on):
        super().__init__()
        self.cutoff_length, self.r_on = float(cutoff_length), float(cutoff_on)
        self.kind, self.p = 1, 6.0


class SphericalEncoding(nn.Module):
    """sevenn/nn/edge_embedding.py:165-186"""

    def __init__(self, lmax, parity=-1, normalization='component', normalize=True):
        super().__init__()
        assert normalization == 'component' and 0 <= lmax <= 3
        self.lmax, self.parity, self.normalize = lmax, parity, normalize
        self.irreps_out = ir.spherical_harmonics_irreps(lmax, parity)


class EdgeEmbedding(nn.Module):
    """sevenn/nn/edge_embedding.py:190-217: one fused kernel for r, Bessel x cutoff and the harmonics."""

    def __init__(self, basis_module, cutoff_module, spherical_module):
        super().__init__()
        self.basis_function, self.cutoff_function, self.spherical = basis_module, cutoff_module, spherical_module

    def forward(self, data):
        rvec = data[EDGE_VEC]
        b, c, s = self.basis_function, self.cutoff_function, self.spherical
        args = (s.lmax, bool(s.normalize), b.num_basis, c.kind, c.cutoff_length, c.p, c.r_on)
        if _train_mode(self):
            Y, emb = T.EdgeFn.apply(T._c(rvec), T._c(b.coeffs.to(rvec.device)), args)
        else:
            Y, emb = F.EdgeEmbedFn.apply(rvec, b.coeffs.to(rvec.device), args)
        data[EDGE_LEN] = torch.linalg.norm(rvec.detach(), dim=-1)
        data[EDGE_EMB], data[EDGE_ATTR] = emb, Y
        return data


class MultiCutoffEdgeEmbedding(nn.Module):
    """sevenn/nn/edge_embedding.py:220-257, the fork's per-layer cutoffs: one fused edge kernel launch per distinct
    cutoff c over the edges with r <= c (Bessel and envelope at cutoff c; XPLOR switches on at c - 1), stored
    under `edge_embedding{c}` / `edge_attr{c}` / `edge_index{c}` for the convolutions that use that cutoff."""

    def __init__(self, basis_module, cutoff_module, spherical_module):
        super().__init__()
        self.basis_functions = nn.ModuleList(basis_module)
        self.cutoff_function, self.spherical = cutoff_module, spherical_module

    def forward(self, data):
        rvec = data[EDGE_VEC]
        r = torch.linalg.norm(rvec.detach(), dim=-1)
        data[EDGE_LEN] = r
        c, s = self.cutoff_function, self.spherical
        for b in self.basis_functions:
            rc = b.cutoff_length
            keep = r <= rc
            r_on = rc - 1 if isinstance(c, XPLORCutoff) else c.r_on
            args = (s.lmax, bool(s.normalize), b.num_basis, c.kind, float(rc), c.p, float(r_on))
            Y, emb = F.EdgeEmbedFn.apply(rvec[keep], b.coeffs.to(rvec.device), args)
            data[EDGE_EMB + str(rc)], data[EDGE_ATTR + str(rc)] = emb, Y
            data[EDGE_IDX + str(rc)] = data[EDGE_IDX][:, keep]
        return data


class EdgePreprocess(nn.Module):
    """sevenn/nn/edge_embedding.py:13-78, the first module of a serial deployed model: edge vectors from
    ``pos``, ``pbc_shift`` and the cell.  The reference also threads a zero strain tensor through ``pos`` and the
    cell so that autograd can return the stress; here the stress is the analytic edge virial (ForceStressOutput
    below), so no strain tensor is needed -- ``is_stress`` is kept for signature compatibility."""

    def __init__(self, is_stress: bool = True):
        super().__init__()
        self.is_stress = is_stress
        self._is_batch_data = False

    def forward(self, data):
        if self._is_batch_data:
            raise NotImplementedError('batched graphs are SURVEY 8f-1 (training path)')
        pos, ei = data[POS], data[EDGE_IDX]
        cell = data[CELL].reshape(3, 3).to(pos.dtype)
        ev = pos.detach()[ei[1]] - pos.detach()[ei[0]] + data[SHIFT].to(pos.dtype) @ cell
        data[EDGE_VEC] = ev.float().requires_grad_(True)
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
        embd = torch.nn.functional.one_hot(inp, self.num_classes).float()
        data[self.key_output] = embd
        if self.key_add is not None:
            data[self.key_add] = embd
        if self.key_save is not None:
            data[self.key_save] = inp
        return data


class IrrepsLinear(_LinearCore):
    """sevenn/nn/linear.py:14-100.  Same constructor and lazy-layer seam (`linear_cls`, `linear_kwargs`,
    `instantiate()`): `linear_cls(irreps_in, irreps_out, **linear_kwargs)` must return a module owning a flat e3nn
    `weight` (and `bias` with biases=True); the default is the B200 parameter holder.  Modality (:65-92): the one-hot
    the reference concatenates to the input only reaches the scalar outputs, so its contribution (and the bias) is
    added to the 0e output entries after the kernel: onehot @ W_modal / sqrt(fan_in + n_modal)."""

    def __init__(self, irreps_in, irreps_out, data_key_in, data_key_out=None, data_key_modal_attr='modal_attr',
                 num_modalities=0, lazy_layer_instantiate=False, **linear_kwargs):
        super().__init__()
        self.key_input, self.key_output = data_key_in, data_key_out or data_key_in
        self.key_modal_attr, self.num_modalities = data_key_modal_attr, 0
        self._irreps_in_wo_modal = _irreps(irreps_in)
        self.irreps_in, self.irreps_out = _irreps(irreps_in), _irreps(irreps_out)
        self.linear_kwargs, self.linear_cls = linear_kwargs, None
        self.linear, self.layer_instantiated = None, False
        self._is_batch_data = True
        self._lspec = sp.make_linear('linear.weight', self._irreps_in_wo_modal, self.irreps_out, 0,
                                     bool(linear_kwargs.get('biases', False)))
        if num_modalities > 1:
            self.set_num_modalities(num_modalities)
        if not lazy_layer_instantiate:
            self.instantiate()

    def set_num_modalities(self, num_modalities):
        """sevenn/nn/linear.py:65-70."""
        if self.layer_instantiated:
            raise ValueError('Layer already instantiated, can not change modalities')
        self.num_modalities = num_modalities
        self.irreps_in = self._irreps_in_wo_modal + [(num_modalities, (0, 1))]
        self._lspec = sp.make_linear('linear.weight', self._irreps_in_wo_modal, self.irreps_out, num_modalities,
                                     bool(self.linear_kwargs.get('biases', False)))

    def instantiate(self):
        if self.linear is not None:
            raise ValueError('Linear layer already exists')
        if self.linear_cls is not None:
            self.linear = self.linear_cls(self.irreps_in, self.irreps_out, **self.linear_kwargs)
        else:
            self.linear = _Weight(self._lspec.numel, self._lspec.n_bias)
        self._setup(self._lspec, 'linear')
        self.layer_instantiated = True

    def _constant_rows(self, data, y):
        """Modality and bias terms of the scalar output entries (e3nn column order), small torch ops."""
        spec, w = self._lspec, self.linear.weight
        offs = np.concatenate([[0], np.cumsum([m * (2 * l + 1) for m, (l, _) in self.irreps_out])])
        onehot = None
        if spec.n_modal:
            onehot = data[self.key_modal_attr].reshape(-1, spec.n_modal).to(y.dtype)
            onehot = onehot[data['batch']] if (self._is_batch_data and 'batch' in data) else onehot.expand(len(y), -1)
        for b in spec.blocks:
            add = None
            if b.modal_off >= 0:
                Wm = w[b.modal_off:b.modal_off + spec.n_modal * b.N].reshape(spec.n_modal, b.N)
                add = onehot @ (Wm.to(y.dtype) / math.sqrt(b.K + b.fan_extra))
            if b.bias_off >= 0:
                bias = self.linear.bias[b.bias_off:b.bias_off + b.N].to(y.dtype)
                add = bias if add is None else add + bias
            if add is not None:
                a = int(offs[b.o_entry])
                y = torch.cat([y[:, :a], y[:, a:a + b.N] + add, y[:, a + b.N:]], dim=1)
        return y

    def forward(self, data):
        assert self.linear is not None, 'Layer is not instantiated'
        y = self._apply_linear(data[self.key_input])
        if self._lspec.n_modal or self._lspec.n_bias:
            y = self._constant_rows(data, y)
        data[self.key_output] = y
        return data


class FCN_e3nn(nn.Module):
    """sevenn/nn/linear.py:145-180: e3nn FullyConnectedNet on scalar node features (readout_as_fcn); parameters
    `fcn.layer{k}.weight`; evaluated by the MLP kernels (GEMM + fused activation, activation-gradient epilogue)."""

    def __init__(self, irreps_in, dim_out, hidden_neurons, activation, data_key_in, data_key_out=None, **e3nn_kwargs):
        super().__init__()
        self.key_input, self.key_output = data_key_in, data_key_out or data_key_in
        self.irreps_in = _irreps(irreps_in)
        assert all(tuple(i) == (0, 1) for _, i in self.irreps_in), 'FCN_e3nn takes scalar features'
        act = activation if isinstance(activation, str) else getattr(activation, '__name__', '')
        self.fcn = _FCN([ir.irreps_dim(self.irreps_in)] + list(hidden_neurons) + [dim_out])
        self._act, self._cst = sp.ACT_ID[act], ir.normalize2mom_const(act)
        self._cache = None

    def _layer(self, device):
        ws = [getattr(self.fcn, f'layer{k}').weight for k in range(len(self.fcn.hs) - 1)]
        key = tuple((w._version, w.data_ptr()) for w in ws) + (str(device),)
        if self._cache is None or self._cache[0] != key:
            Wm = [w.detach().cpu().double().numpy() / math.sqrt(self.fcn.hs[k]) for k, w in enumerate(ws)]
            f32 = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=device)  # noqa: E731
            self._cache = (key, {'mlp': [f32(w) for w in Wm], 'mlpT': [f32(w.T) for w in Wm],
                                 'act': self._act, 'cst': self._cst})
        return self._cache[1]

    def forward(self, data):
        x = data[self.key_input]
        if _train_mode(self):
            data[self.key_output] = _fcn_train(self.fcn, x, self._act, self._cst)
        else:
            data[self.key_output] = F.RadialMLPFn.apply(x.float(), self._layer(x.device))
        return data


class SelfConnectionIntro(_LinearCore):
    """sevenn/nn/self_connection.py:10-64: FullyConnectedTensorProduct(x, one_hot) = species-indexed linear."""

    def __init__(self, irreps_in, irreps_operand, irreps_out, data_key_x=X, data_key_operand=NODE_ATTR,
                 lazy_layer_instantiate=False, **kwargs):
        super().__init__()
        self.key_x, self.key_operand = data_key_x, data_key_operand
        self.irreps_in1, self.irreps_in2, self.irreps_out = _irreps(irreps_in), _irreps(irreps_operand), _irreps(irreps_out)
        self._S = ir.irreps_dim(self.irreps_in2)
        self._lspec = sp.make_linear('fc_tensor_product.weight', self.irreps_in1, ir.simplify(self.irreps_out))
        self._lspec.numel *= self._S
        self.fc_tensor_product, self.layer_instantiated = None, False
        self.fc_tensor_product_cls, self.fc_tensor_product_kwargs = None, kwargs
        if not lazy_layer_instantiate:
            self.instantiate()

    def instantiate(self):
        if self.fc_tensor_product is not None:
            raise ValueError('fc_tensor_product layer already exists')
        self.fc_tensor_product = _Weight(self._lspec.numel)
        self._setup(self._lspec, 'fc_tensor_product', by_type_S=self._S)
        self.layer_instantiated = True

    def forward(self, data):
        assert self.fc_tensor_product is not None, 'Layer is not instantiated'
        types = data[self.key_operand].argmax(dim=1).to(torch.int32).contiguous()
        data[SC_TMP] = self._apply_linear(data[self.key_x], types)
        return data


class SelfConnectionLinearIntro(_LinearCore):
    """sevenn/nn/self_connection.py:68-108 (`irreps_operand` arrives as a keyword from the interaction block and is
    dropped, as in the reference)."""

    def __init__(self, irreps_in, irreps_out=None, data_key_x=X, lazy_layer_instantiate=False, **kwargs):
        super().__init__()
        kwargs.pop('irreps_operand', None)
        self.key_x = data_key_x
        self.irreps_in, self.irreps_out = _irreps(irreps_in), _irreps(irreps_out)
        self._lspec = sp.make_linear('linear.weight', self.irreps_in, ir.simplify(self.irreps_out))
        self.linear, self.layer_instantiated = None, False
        self.linear_cls, self.linear_kwargs = None, kwargs
        if not lazy_layer_instantiate:
            self.instantiate()

    def instantiate(self):
        if self.linear is not None:
            raise ValueError('Linear layer already exists')
        self.linear = _Weight(self._lspec.numel)
        self._setup(self._lspec, 'linear')
        self.layer_instantiated = True

    def forward(self, data):
        assert self.linear is not None, 'Layer is not instantiated'
        data[SC_TMP] = self._apply_linear(data[self.key_x])
        return data


class SelfConnectionOutro(nn.Module):
    """sevenn/nn/self_connection.py:112-128"""

    def __init__(self, data_key_x=X):
        super().__init__()
        self.key_x = data_key_x

    def forward(self, data):
        data[self.key_x] = data[self.key_x] + data[SC_TMP]
        del data[SC_TMP]
        return data


class EquivariantGate(nn.Module):
    """sevenn/nn/equivariant_gate.py:13-61 (activations by name: 'silu', 'tanh', ...)."""

    def __init__(self, irreps_x, act_scalar_dict, act_gate_dict, data_key_x=X):
        super().__init__()
        self.key_x = data_key_x
        name = lambda a: a if isinstance(a, str) else getattr(a, '__name__', str(a))  # noqa: E731
        self._gspec, self._irreps_in, self._irreps_out = sp.make_gate(
            _irreps(irreps_x), {k: name(v) for k, v in act_scalar_dict.items()},
            {k: name(v) for k, v in act_gate_dict.items()})
        g = self._gspec
        self._gate = dict(din=g.lin.dim, dout=g.lout.dim, n_pieces=len(g.pieces),
                          pieces=(_lib.GatePiece * len(g.pieces))(*[
                              _lib.GatePiece(p.kind, p.l, p.mul, p.in_off, p.in_stride, p.gate_off, p.out_off,
                                             p.out_stride, p.act, p.cst) for p in g.pieces]))

    def get_gate_irreps_in(self):
        return self._irreps_in

    def forward(self, data):
        x = data[self.key_x]
        xin = _pcols(x, self._gspec.lin, x.device, False)
        if _train_mode(self):
            out = T.GateFn.apply(T._c(xin), self._gate)
        else:
            out = F.GateFn.apply(xin, self._gate)
        data[self.key_x] = _pcols(out, self._gspec.lout, x.device, True)
        return data


# ---------------------------------------------------------- convolution --
def _fcn_train(fcn, x, act, cst):
    """e3nn FullyConnectedNet with parameter gradients and double backward: BlockMM + ActMul per layer."""
    hs = fcn.hs
    if not hasattr(fcn, '_specs'):
        fcn._specs = [sp.make_linear('w', [(hs[k], (0, 1))], [(hs[k + 1], (0, 1))]) for k in range(len(hs) - 1)]
    h = T._c(x)
    for k in range(len(hs) - 1):
        W = (getattr(fcn, f'layer{k}').weight / math.sqrt(hs[k])).contiguous()
        h = T.BlockMM.apply(h, fcn._specs[k], False, None, 0, W)
        if k < len(hs) - 2:
            h = T.ActMul.apply(h, None, act, cst, 0)
    return h


class _FCN(nn.Module):
    """Parameter holder mirroring e3nn.nn.FullyConnectedNet: layer{k}.weight of shape [h_in, h_out]."""

    def __init__(self, hs):
        super().__init__()
        for k in range(len(hs) - 1):
            lay = nn.Module()
            lay.weight = nn.Parameter(torch.randn(hs[k], hs[k + 1]))
            self.add_module(f'layer{k}', lay)
        self.hs = list(hs)


class IrrepsConvolution(nn.Module):
    """sevenn/nn/convolution.py:28-139: radial MLP (tcgen05 GEMMs) + fused gather/TP/segment-sum kernel."""

    def __init__(self, irreps_x, irreps_filter, irreps_out, weight_layer_input_to_hidden,
                 weight_layer_act='silu', denominator=1.0, train_denominator=False, data_key_x=X,
                 data_key_filter=EDGE_ATTR, data_key_weight_input=EDGE_EMB, data_key_edge_idx=EDGE_IDX,
                 lazy_layer_instantiate=False, is_parallel=False):
        super().__init__()
        self.key_x, self.key_filter = data_key_x, data_key_filter
        self.key_weight_input, self.key_edge_idx = data_key_weight_input, data_key_edge_idx
        self.layer_instantiated = True       # the fused kernel has no separate e3nn TensorProduct to build lazily
        act = weight_layer_act if isinstance(weight_layer_act, str) else getattr(weight_layer_act, '__name__', '')
        self.denominator = nn.Parameter(torch.FloatTensor([denominator]), requires_grad=train_denominator)
        self.is_parallel = is_parallel
        hidden = list(weight_layer_input_to_hidden)
        self._conv = sp.make_conv('convolution', ir.simplify(_irreps(irreps_x)), _irreps(irreps_filter),
                                  _irreps(irreps_out), hidden)
        self.weight_numel = self._conv.wn
        self._comm_size = self._conv.lx.dim        # ghost row width, convolution.py:55 / deploy.py:121-126
        self.irreps_x, self.irreps_filter = _irreps(irreps_x), _irreps(irreps_filter)
        self.irreps_mid = self._conv.lmid.entries
        self.weight_nn = _FCN(hidden + [self._conv.wn])
        self._act, self._cst = sp.ACT_ID[act], ir.normalize2mom_const(act)
        self._sig = self._conv.signature()
        self._cache = None

    def _layer(self, device):
        ws = [getattr(self.weight_nn, f'layer{k}').weight for k in range(len(self.weight_nn.hs) - 1)]
        key = tuple((w._version, w.data_ptr()) for w in ws) + (str(device),)
        if self._cache is None or self._cache[0] != key:
            hs = self.weight_nn.hs
            Wm = [w.detach().cpu().double().numpy() / math.sqrt(hs[k]) for k, w in enumerate(ws)]
            f32 = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=device)  # noqa: E731
            self._cache = (key, {'mlp': [f32(w) for w in Wm], 'mlpT': [f32(w.T) for w in Wm],
                                 'act': self._act, 'cst': self._cst})
        return self._cache[1]

    def _denominator_value(self) -> float:
        """host copy of the denominator, refreshed when the parameter changes (no device->host read per call)"""
        d = self.denominator
        key = (d._version, d.data_ptr())
        if getattr(self, '_denom_cache', (None, None))[0] != key:
            self._denom_cache = (key, float(d.item()))
        return self._denom_cache[1]

    def instantiate(self):
        """convolution.py:103-116: nothing to build lazily here (kept for the reference's builder loop)."""
        self.layer_instantiated = True

    def forward(self, data):
        x = data[self.key_x]
        dev = x.device
        lib = get_lib()
        if not lib.has_conv(self._sig):
            raise _lib.SevennError(f'no prebuilt conv kernel for signature {self._sig}; build it with '
                                   'sevennet_exp_b200.build.build(extra_specs=...)')
        if self.is_parallel:
            x = torch.cat([x, data[X_GHOST]])
        n_nodes = x.shape[0]
        n_dst = int(data[NLOCAL]) if self.is_parallel else n_nodes
        gkey = GRAPH + ('' if self.key_edge_idx == EDGE_IDX else self.key_edge_idx)
        cached = data.get(gkey)
        if cached is None or cached[0]['n_nodes'] != n_nodes or cached[0]['n_dst'] != n_dst:
            cached = F.build_graph(data[self.key_edge_idx], n_nodes, n_dst, dev)
            data[gkey] = cached
        graph, perm = cached
        emb, Y = data[self.key_weight_input], data[self.key_filter]
        if perm is not None:
            emb, Y = emb[perm], Y[perm]
        xin = _pcols(x.float(), self._conv.lx, dev, False)
        meta = dict(sig=self._sig, dx=self._conv.lx.dim, dmid=self._conv.lmid.dim, wn=self._conv.wn,
                    units=conv_codegen.n_units(self._conv), inv_denom=1.0 / self._denominator_value())
        if _train_mode(self):
            # training: radial MLP and convolution through the second-order-capable Functions (train_functional.py);
            # a trainable denominator divides outside the kernel so that it receives its gradient
            w = _fcn_train(self.weight_nn, emb, self._act, self._cst)
            if self.denominator.requires_grad:
                meta['inv_denom'] = 1.0
            out = T.ConvT.apply(T._c(xin), T._c(Y), T._c(w), graph, meta)
            if self.denominator.requires_grad:
                out = out / self.denominator
        else:
            w = F.RadialMLPFn.apply(emb, self._layer(dev))
            out = F.ConvFn.apply(xin, Y, w, graph, meta)
        data[self.key_x] = _pcols(out, self._conv.lmid, dev, True)
        return data


# --------------------------------------------------------------- output --
class Rescale(nn.Module):
    """sevenn/nn/scale.py:22-56"""

    def __init__(self, shift, scale, data_key_in=SCALED_AE, data_key_out=AE, train_shift_scale=False):
        super().__init__()
        self.shift = nn.Parameter(torch.FloatTensor([shift]), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor([scale]), requires_grad=train_shift_scale)
        self.key_input, self.key_output = data_key_in, data_key_out

    def get_shift(self):
        return self.shift.detach().cpu().tolist()[0]

    def get_scale(self):
        return self.scale.detach().cpu().tolist()[0]

    def forward(self, data):
        data[self.key_output] = data[self.key_input] * self.scale + self.shift
        return data


class SpeciesWiseRescale(nn.Module):
    """sevenn/nn/scale.py:60-161"""

    def __init__(self, shift, scale, data_key_in=SCALED_AE, data_key_out=AE, data_key_indices=ATOM_TYPE,
                 train_shift_scale=False):
        super().__init__()
        self.key_input, self.key_output, self.key_indices = data_key_in, data_key_out, data_key_indices
        n = len(shift) if isinstance(shift, list) else len(scale)
        shift = [shift] * n if isinstance(shift, float) else shift
        scale = [scale] * n if isinstance(scale, float) else scale
        self.shift = nn.Parameter(torch.FloatTensor(shift), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor(scale), requires_grad=train_shift_scale)

    def get_shift(self, type_map=None):
        v = self.shift.detach().cpu().tolist()
        return v if type_map is None else {z: v[i] for z, i in type_map.items()}

    def get_scale(self, type_map=None):
        v = self.scale.detach().cpu().tolist()
        return v if type_map is None else {z: v[i] for z, i in type_map.items()}

    @staticmethod
    def from_mappers(shift, scale, type_map, **kwargs):
        """scale.py:112-147: per-species lists from floats, lists in type order, or {atomic number: value} dicts."""
        def as_list(v):
            if isinstance(v, dict):
                return [float(v[z]) for z, _ in sorted(type_map.items(), key=lambda kv: kv[1])]
            if isinstance(v, (list, tuple)):
                return [float(t) for t in v] if len(v) != 1 else [float(v[0])] * len(type_map)
            return [float(v)] * len(type_map)
        return SpeciesWiseRescale(as_list(shift), as_list(scale), **kwargs)

    def forward(self, data):
        idx = data[self.key_indices]
        data[self.key_output] = (data[self.key_input] * self.scale[idx].view(-1, 1)
                                 + self.shift[idx].view(-1, 1))
        return data


class ModalWiseRescale(nn.Module):
    """sevenn/nn/scale.py:165-360: shift / scale per (modality, species) where modal wise, else per species."""

    def __init__(self, shift, scale, data_key_in=SCALED_AE, data_key_out=AE, data_key_modal_indices='modal_type',
                 data_key_atom_indices=ATOM_TYPE, use_modal_wise_shift=False, use_modal_wise_scale=False,
                 train_shift_scale=False):
        super().__init__()
        self.shift = nn.Parameter(torch.FloatTensor(shift), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor(scale), requires_grad=train_shift_scale)
        self.key_input, self.key_output = data_key_in, data_key_out
        self.key_atom_indices, self.key_modal_indices = data_key_atom_indices, data_key_modal_indices
        self.use_modal_wise_shift, self.use_modal_wise_scale = use_modal_wise_shift, use_modal_wise_scale
        self._is_batch_data = True

    def _resolved(self, which, modal_wise, type_map, modal_map):
        v = getattr(self, which).detach().cpu().tolist()
        per_type = lambda row: row if type_map is None else {z: row[i] for z, i in type_map.items()}  # noqa: E731
        if not modal_wise:
            return per_type(v)
        if modal_map is None:
            return [per_type(row) for row in v]
        return {m: per_type(v[i]) for m, i in modal_map.items()}

    def get_shift(self, type_map=None, modal_map=None):
        return self._resolved('shift', self.use_modal_wise_shift, type_map, modal_map)

    def get_scale(self, type_map=None, modal_map=None):
        return self._resolved('scale', self.use_modal_wise_scale, type_map, modal_map)

    @staticmethod
    def from_mappers(shift, scale, use_modal_wise_shift, use_modal_wise_scale, type_map, modal_map, **kwargs):
        """scale.py:238-338: floats broadcast; {modality: value | per-species list} dicts and lists in modality /
        type order resolve to [n_modal][n_types] (modal wise) or [n_types]."""
        nt, nmod = len(type_map), len(modal_map)

        def species(v):
            if isinstance(v, dict):
                return [float(v[z]) for z, _ in sorted(type_map.items(), key=lambda kv: kv[1])]
            if isinstance(v, (list, tuple)):
                assert len(v) == nt, f'Invalid shift or scale {v}'
                return [float(t) for t in v]
            return [float(v)] * nt

        def solve(v, mw):
            if not mw:
                return species(v)
            if isinstance(v, dict):
                return [species(v[m]) for m, _ in sorted(modal_map.items(), key=lambda kv: kv[1])]
            if isinstance(v, (list, tuple)) and len(v) == nmod and isinstance(v[0], (list, tuple)):
                return [species(row) for row in v]
            if isinstance(v, (list, tuple)) and len(v) == nmod and nmod != nt:
                return [[float(t)] * nt for t in v]
            return [species(v)] * nmod
        return ModalWiseRescale(solve(shift, use_modal_wise_shift), solve(scale, use_modal_wise_scale),
                                use_modal_wise_shift=use_modal_wise_shift, use_modal_wise_scale=use_modal_wise_scale,
                                **kwargs)

    def forward(self, data):
        modal = data[self.key_modal_indices]
        if self._is_batch_data and 'batch' in data:
            modal = modal.reshape(-1)[data['batch']]
        atom = data[self.key_atom_indices]
        shift = self.shift[modal, atom] if self.use_modal_wise_shift else self.shift[atom]
        scale = self.scale[modal, atom] if self.use_modal_wise_scale else self.scale[atom]
        data[self.key_output] = data[self.key_input] * scale.view(-1, 1) + shift.view(-1, 1)
        return data


class AtomReduce(nn.Module):
    """sevenn/nn/linear.py:104-141 (single graph; fp64 accumulation)."""

    def __init__(self, data_key_in=AE, data_key_out=E_TOT, reduce='sum', constant=1.0):
        super().__init__()
        self.key_input, self.key_output, self.reduce, self.constant = data_key_in, data_key_out, reduce, constant
        self._is_batch_data = False

    def forward(self, data):
        src = data[self.key_input]
        if self._is_batch_data:
            b = data[BATCH]
            out = torch.zeros(_n_graphs(data), dtype=torch.float64, device=src.device)
            out.index_add_(0, b, src.reshape(-1).double())
            data[self.key_output] = out.to(src.dtype) * self.constant
        else:
            data[self.key_output] = torch.sum(src.double()).to(src.dtype) * self.constant
        return data


class ForceStressOutputFromEdge(nn.Module):
    """sevenn/nn/force_output.py:140-224: f_ij = dE/d edge_vec (reverse kernels via autograd), then one
    kernel for F_i = sum_{centre i} f - sum_{neighbour i} f and the per-atom virial."""

    def __init__(self, data_key_edge=EDGE_VEC, data_key_edge_idx=EDGE_IDX, data_key_energy=E_TOT,
                 data_key_force=FORCE, data_key_stress=STRESS, data_key_cell_volume=VOL):
        super().__init__()
        self._is_batch_data = False

    def get_grad_key(self):
        return EDGE_VEC

    def _forward_train(self, data):
        """force_output.py:173-222 verbatim in torch ops: `create_graph=True` keeps dE/d(edge_vec) in the graph so that a
        loss on forces / stress back-propagates through the reverse kernels (nn/train_functional.py)."""
        tot = _n_atoms_total(data)
        rij, ei = data[EDGE_VEC], data[EDGE_IDX]
        fij = torch.autograd.grad([data[E_TOT].sum()], [rij], create_graph=True, allow_unused=True)[0]
        for k in [k for k in data if isinstance(k, str) and k.startswith(GRAPH)]:
            del data[k]
        if fij is None:
            return data
        pf = torch.zeros(tot, 3, dtype=fij.dtype, device=fij.device).index_add(0, ei[0], fij)
        nf = torch.zeros(tot, 3, dtype=fij.dtype, device=fij.device).index_add(0, ei[1], fij)
        data[FORCE] = pf - nf
        vir = torch.cat([rij * fij, (rij[:, 0] * fij[:, 1]).unsqueeze(-1), (rij[:, 1] * fij[:, 2]).unsqueeze(-1),
                         (rij[:, 2] * fij[:, 0]).unsqueeze(-1)], dim=-1)
        s = torch.zeros(tot, 6, dtype=fij.dtype, device=fij.device).index_add(0, ei[1], vir)
        if self._is_batch_data:
            b = data[BATCH]
            sout = torch.zeros(_n_graphs(data), 6, dtype=fij.dtype, device=fij.device).index_add(0, b, s)
        else:
            sout = s.sum(dim=0)
        data[STRESS] = torch.neg(sout) / data[VOL].to(fij.dtype).unsqueeze(-1)
        return data

    def forward(self, data):
        if _train_mode(self):
            return self._forward_train(data)
        tot = _n_atoms_total(data)
        rij = data[EDGE_VEC]
        fij = torch.autograd.grad([data[E_TOT].sum()], [rij], allow_unused=True)[0]
        cached = data.pop(GRAPH, None)
        for k in [k for k in data if isinstance(k, str) and k.startswith(GRAPH)]:
            del data[k]                     # per-cutoff graphs of MultiCutoffEdgeEmbedding models
        if fij is None:
            return data
        dev = rij.device
        k = kernels(dev)
        if cached is None or cached[0]['n_nodes'] != tot or cached[0]['n_dst'] != tot:
            cached = F.build_graph(data[EDGE_IDX], tot, tot, dev)
        graph, perm = cached
        f, r = fij.detach().float(), rij.detach().float()
        if perm is not None:
            f, r = f[perm], r[perm]
        f, r = f.contiguous(), r.contiguous()
        Fo, vir = k._empty(tot, 3), k._empty(tot, 6)
        k.lib.call('sevenn_force_virial', ptr(f), ptr(r), ptr(graph['rowptr']), tot, ptr(graph['cptr']),
                   ptr(graph['cedge']), tot, ptr(Fo), ptr(vir), k._stream())
        k.n_launch += 1
        data[FORCE] = Fo
        vol = data[VOL]
        if self._is_batch_data:               # per-graph stress, force_output.py:210-219
            b = data[BATCH]
            sout = torch.zeros(_n_graphs(data), 6, dtype=torch.float64, device=dev).index_add_(0, b, vir.double())
            data[STRESS] = (torch.neg(sout) / vol.double().reshape(-1, 1)).to(f.dtype)
        else:
            data[STRESS] = (torch.neg(vir.double().sum(dim=0)) / vol.double().reshape(-1)[0]).to(f.dtype)
        return data


class ForceStressOutput(ForceStressOutputFromEdge):
    """sevenn/nn/force_output.py:55-136 (serial deployed models).  The reference differentiates E w.r.t. ``pos`` and
    the strain of EdgePreprocess; since edge_vec = pos_j - pos_i + shift.cell, dE/dpos_k is the same signed sum of
    dE/d edge_vec that ForceStressOutputFromEdge forms and -dE/d(strain)/V is the edge virial, so both classes run
    the same force/virial kernel (checked against the reference's own TorchScript outputs, tests/golden)."""

    def __init__(self, data_key_pos=POS, data_key_energy=E_TOT, data_key_force=FORCE, data_key_stress=STRESS,
                 data_key_cell_volume=VOL):
        super().__init__()


class ForceOutput(ForceStressOutputFromEdge):
    """sevenn/nn/force_output.py:12-51: forces only."""

    def __init__(self, data_key_pos=POS, data_key_energy=E_TOT, data_key_force=FORCE):
        super().__init__()

    def forward(self, data):
        had_vol = VOL in data
        if not had_vol:
            data[VOL] = torch.ones((), device=data[EDGE_VEC].device)
        data = super().forward(data)
        data.pop(STRESS, None)
        if not had_vol:
            data.pop(VOL)
        return data


# ----------------------------------------------------------- sequential --
class _ModalInputPrepare(nn.Module):
    """sevenn/nn/sequential.py:22-38: writes the frozen modality index of a deployed single-modality model."""

    def __init__(self, modal_idx: int):
        super().__init__()
        self.modal_idx = modal_idx

    def forward(self, data):
        data['modal_type'] = torch.tensor(self.modal_idx, dtype=torch.int64, device=data['x'].device)
        return data


class AtomGraphSequential(nn.Sequential):
    """sevenn/nn/sequential.py:41-205: the model container (same constructor, same helper methods)."""

    def __init__(self, modules, cutoff=0.0, type_map=None, modal_map=None, eval_type_map=True,
                 eval_modal_map=False, data_key_atomic_numbers=Z, data_key_node_feature=X, data_key_grad=None):
        modules = modules if isinstance(modules, OrderedDict) else OrderedDict(modules)
        if eval_modal_map and modal_map is None:
            raise ValueError('eval_modal_map is True but modal_map is None')          # sequential.py:86-87
        for m in modules.values():                        # sequential.py:_instantiate_modules (lazy-layer seam)
            if not getattr(m, 'layer_instantiated', True):
                m.instantiate()
        super().__init__(modules)
        self.cutoff, self.type_map, self.modal_map, self.key_grad = cutoff, type_map, modal_map, data_key_grad
        self.key_atomic_numbers, self.key_node_feature = data_key_atomic_numbers, data_key_node_feature
        self.eval_type_map = eval_type_map and type_map is not None
        self.eval_modal_map, self.is_batch_data = eval_modal_map, False
        if type_map is not None:
            lut = -torch.ones(120, dtype=torch.long)
            for z, i in type_map.items():
                lut[z] = i
            self.register_buffer('z_to_onehot_tensor', lut, persistent=False)

    def set_is_batch_data(self, flag):
        for m in self:
            if hasattr(m, '_is_batch_data'):
                m._is_batch_data = flag
        self.is_batch_data = flag

    def get_irreps_in(self, modlue_name, attr_key='irreps_in'):
        for m in self._modules[modlue_name].modules():
            if hasattr(m, attr_key):
                return ir.irreps_str(getattr(m, attr_key))
        return None

    def prepand_module(self, key, module):
        rest = [(k, v) for k, v in self._modules.items() if k != key]
        self._modules.clear()                 # torch keeps `_modules` as a plain (insertion-ordered) dict
        self._modules[key] = module
        for k, v in rest:
            self._modules[k] = v

    def replace_module(self, key, module):
        self._modules[key] = module

    def delete_module_by_key(self, key):
        self._modules.pop(key, None)

    def _atomic_numbers_to_onehot(self, atomic_numbers):
        return self.z_to_onehot_tensor.to(atomic_numbers.device)[atomic_numbers]

    def _preprocess(self, data):
        if self.eval_type_map:
            data[self.key_node_feature] = self._atomic_numbers_to_onehot(data[self.key_atomic_numbers])
        if self.eval_modal_map:                              # sequential.py:139-154: data_modality name(s) -> index
            name = data['data_modality']
            idx = self.modal_map[name] if isinstance(name, str) else [self.modal_map[n] for n in name]
            data['modal_type'] = torch.tensor(idx, dtype=torch.int64, device=data[self.key_node_feature].device)
        if self.key_grad is not None:
            data[self.key_grad].requires_grad_(True)
        return data

    def prepare_modal_deploy(self, modal: str):
        """sequential.py:170-176: freeze the model to one modality (single-structure data from then on)."""
        if self.modal_map is None:
            return
        self.eval_modal_map = False
        self.set_is_batch_data(False)
        self.prepand_module('modal_input_prepare', _ModalInputPrepare(self.modal_map[modal]))

    def forward(self, data: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        data = self._preprocess(data)
        for m in self:
            data = m(data)
        return data


def init_edge_embedding(cfg):
    """sevenn/model_build.py:69-116 (single cutoff)."""
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
    if rb.pop('radial_basis_name') != 'bessel':
        raise ValueError('only the bessel radial basis is on the hot path')
    if cfg.get('multi_cutoff'):
        return MultiCutoffEdgeEmbedding([BesselBasis(c, **rb) for c in sorted(set(cfg['multi_cutoff']))], cut, sph)
    return EdgeEmbedding(BesselBasis(cfg['cutoff'], **rb), cut, sph)


def build_E3_equivariant_model(config: dict, parallel: bool = False) -> AtomGraphSequential:
    """sevenn/model_build.py:384-576 + nn/interaction_blocks.py:12-88: same layer names, same order."""
    import copy
    cfg = {**sp.DEFAULTS, **copy.deepcopy(config)}
    layers = OrderedDict()
    S, C, L = cfg['_number_of_species'], cfg['channel'], cfg['num_convolution_layer']
    lmax_node = cfg['lmax_node'] if cfg['lmax_node'] > 0 else cfg['lmax']
    sc_cls = {'nequip': SelfConnectionIntro, 'linear': SelfConnectionLinearIntro}[cfg['self_connection_type']]
    manual = None
    if cfg['irreps_manual'] is not False:
        manual = [ir.parse_irreps(s) for s in cfg['irreps_manual']]
    denom = cfg['conv_denominator']
    denom = list(denom) if isinstance(denom, (list, tuple)) else [denom] * L
    ee = init_edge_embedding(cfg)
    irreps_filter = ee.spherical.irreps_out
    layers['edge_embedding'] = ee
    one_hot = [(S, (0, 1))]
    irreps_x = [(C, (0, 1))] if manual is None else manual[0]
    nm = int(cfg.get('_number_of_modalities', 0)) if cfg.get('use_modality', False) else 0   # model_build.py:198-244
    if nm and parallel:
        raise NotImplementedError('multi-modal checkpoints: serial models only')
    m_emb, m_si1, m_si2, m_out = (nm if cfg.get(k, False) else 0 for k in (
        'use_modal_node_embedding', 'use_modal_self_inter_intro', 'use_modal_self_inter_outro',
        'use_modal_output_block'))
    bias = bool(cfg['use_bias_in_linear'])               # interaction_blocks.py:50,82; model_build.py:130,137,457
    layers['onehot_idx_to_onehot'] = OnehotEmbedding(S, X, X, ATOM_TYPE, NODE_ATTR)
    if nm:
        layers['one_hot_modality'] = OnehotEmbedding(nm, 'modal_type', 'modal_attr', None, None)
    layers['onehot_to_feature_x'] = IrrepsLinear(one_hot, irreps_x, X, biases=bias, num_modalities=m_emb)
    multi = cfg.get('multi_cutoff') or None
    wnn = [(ee.basis_functions[0] if multi else ee.basis_function).num_basis] + list(cfg['weight_nn_hidden_neurons'])
    for t in range(L):
        parity_mode, lm = 'full', lmax_node
        if t == L - 1:
            lm, parity_mode = 0, 'even'
        irreps_out = (ir.infer_irreps_out(irreps_x, irreps_filter, lm, parity_mode, C)
                      if manual is None else manual[t + 1])
        out_lmax = max(l for _, (l, _) in irreps_out)
        irreps_out_tp = ir.infer_irreps_out(irreps_x, irreps_filter, out_lmax, parity_mode, False)
        gate = EquivariantGate(irreps_out, cfg['act_scalar'], cfg['act_gate'])
        gate_in = gate.get_gate_irreps_in()
        layers[f'{t}_self_connection_intro'] = sc_cls(irreps_x, irreps_operand=one_hot, irreps_out=gate_in)
        layers[f'{t}_self_interaction_1'] = IrrepsLinear(irreps_x, irreps_x, X, biases=bias, num_modalities=m_si1)
        layers[f'{t}_convolution'] = IrrepsConvolution(irreps_x, irreps_filter, irreps_out_tp, wnn,
                                                       cfg['act_radial'], denom[t], cfg['train_denominator'],
                                                       is_parallel=parallel,
                                                       **({} if not multi else dict(
                                                           data_key_filter=f'{EDGE_ATTR}{multi[t]}',
                                                           data_key_weight_input=f'{EDGE_EMB}{multi[t]}',
                                                           data_key_edge_idx=f'{EDGE_IDX}{multi[t]}')))
        layers[f'{t}_self_interaction_2'] = IrrepsLinear(layers[f'{t}_convolution'].irreps_mid,
                                                         ir.simplify(gate_in), X, biases=bias,
                                                         num_modalities=m_si2)
        layers[f'{t}_self_connection_outro'] = SelfConnectionOutro()
        layers[f'{t}_equivariant_gate'] = gate
        irreps_x = ir.simplify(gate._irreps_out)
    if cfg['readout_as_fcn']:                            # model_build.py:141-155
        layers['readout_FCN'] = FCN_e3nn(irreps_x, 1, cfg.get('readout_fcn_hidden_neurons', [30, 30]),
                                         cfg.get('readout_fcn_activation', 'relu'), X, SCALED_AE)
    else:
        hidden = [(ir.irreps_dim(irreps_x) // 2, (0, 1))]
        layers['reduce_input_to_hidden'] = IrrepsLinear(irreps_x, hidden, X, biases=bias, num_modalities=m_out)
        layers['reduce_hidden_to_energy'] = IrrepsLinear(hidden, [(1, (0, 1))], X, SCALED_AE, biases=bias)
    shift, scale = cfg['shift'], cfg['scale']
    shift = shift.tolist() if hasattr(shift, 'tolist') else shift
    scale = scale.tolist() if hasattr(scale, 'tolist') else scale
    if nm:                                               # model_build.py:173-184
        layers['rescale_atomic_energy'] = ModalWiseRescale.from_mappers(
            shift, scale, cfg.get('use_modal_wise_shift', False), cfg.get('use_modal_wise_scale', False),
            type_map=cfg['_type_map'], modal_map=cfg['_modal_map'], train_shift_scale=cfg['train_shift_scale'])
    elif isinstance(shift, float) and isinstance(scale, float):
        layers['rescale_atomic_energy'] = Rescale(shift, scale, train_shift_scale=cfg['train_shift_scale'])
    else:
        shift = shift * S if isinstance(shift, list) and len(shift) == 1 else shift
        scale = scale * S if isinstance(scale, list) and len(scale) == 1 else scale
        layers['rescale_atomic_energy'] = SpeciesWiseRescale(shift, scale, train_shift_scale=cfg['train_shift_scale'])
    layers['reduce_total_enegy'] = AtomReduce()
    layers['force_output'] = ForceStressOutputFromEdge()
    common = dict(cutoff=cfg['cutoff'], type_map=cfg['_type_map'], data_key_grad=EDGE_VEC)
    if not parallel:
        model = AtomGraphSequential(layers, modal_map=cfg.get('_modal_map') if nm else None, **common)
        model._sevenn_config = copy.deepcopy(config)      # lets SevenNetCalculator(file_type='model_instance') take the module
        return model
    return [AtomGraphSequential(seg, eval_type_map=False, **common) for seg in _parallel_segments(layers, cfg, S, L)]


def _parallel_segments(layers, cfg, n_species, n_layers):
    """sevenn/model_build.py:299-367 (`_to_parallel_model`): what LAMMPS `pair_style e3gnn/parallel` loads, one
    module list per interaction layer.  Ghost atoms get their own embedding + first self-interaction (their types
    are known locally, so layer 0 needs no communication); the chain is cut after every `{t}_self_interaction_1`
    with t >= 1, because that is the tensor the next convolution gathers from ghosts (`x_ghost`, refreshed by the
    caller between segments); each later segment starts with its own edge embedding (edge_vec is a fresh autograd
    leaf per segment on the LAMMPS side); forces are assembled by the caller, so `force_output` is dropped.
    Ghost modules share the parameters of their local twins through the state_dict keys the reference uses
    (`ghost_onehot_to_feature_x`, `ghost_0_self_interaction_1`)."""
    emb_irreps = layers['onehot_to_feature_x'].irreps_out
    seq = []
    for name, mod in layers.items():
        if name == 'force_output':
            continue
        seq.append((name, mod))
        if name == 'onehot_to_feature_x':
            seq.append(('one_hot_ghost', OnehotEmbedding(n_species, X_GHOST)))
            seq.append(('ghost_onehot_to_feature_x', IrrepsLinear([(n_species, (0, 1))], emb_irreps, X_GHOST)))
        elif name == '0_self_interaction_1':
            seq.append(('ghost_0_self_interaction_1', IrrepsLinear(emb_irreps, emb_irreps, X_GHOST)))
    cuts = [k for k, (name, _) in enumerate(seq)
            if name.endswith('_self_interaction_1') and name[0].isdigit() and int(name.split('_')[0]) >= 1]
    segments, start = [], 0
    for c in cuts + [len(seq) - 1]:
        part = seq[start:c + 1]
        if start > 0:
            part = [('edge_embedding', init_edge_embedding(cfg))] + part
        segments.append(OrderedDict(part))
        start = c + 1
    return segments
