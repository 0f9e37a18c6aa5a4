"""Restatement (pure PyTorch, CPU, fp32/fp64) of the e3nn pieces SevenNet uses.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  The arithmetic of the reference's
hot path is not in /root/reference: it lives in the third-party dependency
``e3nn`` (``pyproject.toml:24``: ``e3nn>=0.5.0``, unpinned, un-vendored).  This
file restates e3nn's published semantics for exactly the call sites of the
reference:

  o3.SphericalHarmonics          sevenn/nn/edge_embedding.py:178-183
  o3.TensorProduct ('uvu')       sevenn/nn/convolution.py:100,114
  e3nn.nn.FullyConnectedNet      sevenn/nn/convolution.py:101,115
  o3.Linear                      sevenn/nn/linear.py:49,60-62
  o3.FullyConnectedTensorProduct sevenn/nn/self_connection.py:40,49-56
  e3nn.nn.Gate                   sevenn/nn/equivariant_gate.py:49-51
  o3.FullTensorProduct/Irreps    sevenn/util.py:172
  o3.wigner_3j                   sevenn/scripts/backward_compatibility.py:75,126

Layout convention everywhere: ``mul_ir`` -- an irreps tensor stores each
(mul, ir) entry as a row-major [mul, 2l+1] block.
"""
import math
from fractions import Fraction
from functools import lru_cache
from typing import List, Tuple

import torch

# --------------------------------------------------------------------------
# Irreps bookkeeping
# --------------------------------------------------------------------------


class Irrep(tuple):
    """(l, p) with p in {+1, -1}.  Ordering is plain tuple ordering (l, then p),
    which is what e3nn's ``Irreps.sort`` uses (odd before even within one l)."""

    def __new__(cls, l, p=None):
        if p is None:
            if isinstance(l, Irrep):
                return l
            if isinstance(l, str):
                s = l.strip()
                return tuple.__new__(cls, (int(s[:-1]), {'e': 1, 'o': -1}[s[-1]]))
            l, p = l
        assert p in (1, -1) and l >= 0
        return tuple.__new__(cls, (int(l), int(p)))

    @property
    def l(self):
        return self[0]

    @property
    def p(self):
        return self[1]

    @property
    def dim(self):
        return 2 * self[0] + 1

    def is_scalar(self):
        return self[0] == 0 and self[1] == 1

    def __mul__(self, other):
        other = Irrep(other)
        p = self.p * other.p
        return [Irrep(l, p) for l in range(abs(self.l - other.l), self.l + other.l + 1)]

    def __repr__(self):
        return f"{self.l}{'e' if self.p == 1 else 'o'}"


class Irreps(tuple):
    """Tuple of (mul, Irrep)."""

    def __new__(cls, spec=None):
        if spec is None:
            spec = []
        if isinstance(spec, Irreps):
            return spec
        items = []
        if isinstance(spec, str):
            for term in spec.split('+'):
                term = term.strip()
                if not term:
                    continue
                if 'x' in term:
                    mul, ir = term.split('x')
                    items.append((int(mul), Irrep(ir)))
                else:
                    items.append((1, Irrep(term)))
        else:
            for it in spec:
                if isinstance(it, Irrep):
                    items.append((1, it))
                else:
                    mul, ir = it
                    items.append((int(mul), Irrep(ir)))
        return tuple.__new__(cls, tuple(items))

    @staticmethod
    def spherical_harmonics(lmax, p=-1):
        return Irreps([(1, (l, p ** l)) for l in range(lmax + 1)])

    @property
    def dim(self):
        return sum(mul * ir.dim for mul, ir in self)

    @property
    def num_irreps(self):
        return sum(mul for mul, _ in self)

    @property
    def lmax(self):
        return max(ir.l for _, ir in self)

    def offsets(self):
        out, o = [], 0
        for mul, ir in self:
            out.append(o)
            o += mul * ir.dim
        return out

    def slices(self):
        offs = self.offsets()
        return [slice(o, o + mul * ir.dim) for o, (mul, ir) in zip(offs, self)]

    def sort(self):
        """Stable sort by (l, p); returns (irreps, p, inv) like e3nn."""
        out = sorted([(ir, i, mul) for i, (mul, ir) in enumerate(self)])
        inv = tuple(i for _, i, _ in out)
        p = [0] * len(inv)
        for new, old in enumerate(inv):
            p[old] = new
        return Irreps([(mul, ir) for ir, _, mul in out]), tuple(p), inv

    def simplify(self):
        out = []
        for mul, ir in self:
            if out and out[-1][1] == ir:
                out[-1] = (out[-1][0] + mul, ir)
            elif mul > 0:
                out.append((mul, ir))
        return Irreps(out)

    def __contains__(self, ir):
        ir = Irrep(ir)
        return any(ir == i for _, i in self)

    def __add__(self, other):
        return Irreps(list(self) + list(Irreps(other)))

    def __repr__(self):
        return '+'.join(f'{mul}x{ir}' for mul, ir in self)


def full_tensor_product_irreps_out(irreps_in1: Irreps, irreps_in2: Irreps) -> Irreps:
    """``FullTensorProduct(a, b).irreps_out`` (sorted, not simplified)."""
    a, b = Irreps(irreps_in1).simplify(), Irreps(irreps_in2).simplify()
    out = []
    for mul1, ir1 in a:
        for mul2, ir2 in b:
            for ir in ir1 * ir2:
                out.append((mul1 * mul2, ir))
    return Irreps(out).sort()[0]


# --------------------------------------------------------------------------
# Wigner 3j in e3nn's real basis (SURVEY Appendix A rule 3)
# --------------------------------------------------------------------------


def _fact(n):
    return math.factorial(n)


def _su2_cg(j1, m1, j2, m2, j3, m3) -> Fraction:
    """<j1 m1 j2 m2 | j3 m3>^2 carrier: returns the signed coefficient as
    (sign, Fraction of the square) folded into a float later; here we return a
    float computed from exact rationals."""
    if m3 != m1 + m2:
        return 0.0
    vmin = max(-j1 + j2 + m3, -j1 + m1, 0)
    vmax = min(j2 + j3 + m1, j3 - j1 + j2, j3 + m3)
    C2 = Fraction(
        (2 * j3 + 1) * _fact(j3 + j1 - j2) * _fact(j3 - j1 + j2) * _fact(j1 + j2 - j3)
        * _fact(j3 + m3) * _fact(j3 - m3),
        _fact(j1 + j2 + j3 + 1) * _fact(j1 - m1) * _fact(j1 + m1) * _fact(j2 - m2) * _fact(j2 + m2),
    )
    S = Fraction(0)
    for v in range(vmin, vmax + 1):
        S += Fraction(
            (-1) ** (v + j2 + m2) * _fact(j2 + j3 + m1 - v) * _fact(j1 - m1 + v),
            _fact(v) * _fact(j3 - j1 + j2 - v) * _fact(j3 + m3 - v) * _fact(v + j1 - j2 - m3),
        )
    return math.sqrt(float(C2)) * float(S)


def _real_to_complex(l) -> torch.Tensor:
    """Change of basis real->complex SH, e3nn convention, times (-i)^l."""
    q = torch.zeros(2 * l + 1, 2 * l + 1, dtype=torch.complex128)
    s2 = 1 / math.sqrt(2)
    for m in range(-l, 0):
        q[l + m, l + abs(m)] = s2
        q[l + m, l - abs(m)] = -1j * s2
    q[l, l] = 1
    for m in range(1, l + 1):
        q[l + m, l + abs(m)] = (-1) ** m * s2
        q[l + m, l - abs(m)] = 1j * (-1) ** m * s2
    return (-1j) ** l * q


@lru_cache(maxsize=None)
def _wigner_3j_f64(l1, l2, l3) -> torch.Tensor:
    assert abs(l2 - l3) <= l1 <= l2 + l3
    C = torch.zeros(2 * l1 + 1, 2 * l2 + 1, 2 * l3 + 1, dtype=torch.complex128)
    for m1 in range(-l1, l1 + 1):
        for m2 in range(-l2, l2 + 1):
            m3 = m1 + m2
            if abs(m3) <= l3:
                C[l1 + m1, l2 + m2, l3 + m3] = _su2_cg(l1, m1, l2, m2, l3, m3)
    Q1, Q2, Q3 = _real_to_complex(l1), _real_to_complex(l2), _real_to_complex(l3)
    C = torch.einsum('ij,kl,mn,ikn->jlm', Q1, Q2, torch.conj(Q3.T), C)
    assert C.imag.abs().max() < 1e-10
    C = C.real
    return C / C.norm()


def wigner_3j(l1, l2, l3, dtype=torch.float32) -> torch.Tensor:
    return _wigner_3j_f64(l1, l2, l3).to(dtype).clone()


# --------------------------------------------------------------------------
# Real spherical harmonics, l <= 3 closed forms (SURVEY Appendix A rule 2)
# plus a generic recursion used to cross-check them in tests.
# --------------------------------------------------------------------------


def spherical_harmonics(lmax: int, vec: torch.Tensor, normalize: bool,
                        normalization: str = 'component') -> torch.Tensor:
    """e3nn ``o3.SphericalHarmonics(0..lmax, normalize, normalization)(vec)``."""
    assert lmax <= 3 and normalization == 'component'
    if normalize:
        vec = torch.nn.functional.normalize(vec, dim=-1)
    x, y, z = vec[..., 0], vec[..., 1], vec[..., 2]
    out = [torch.ones_like(x)]
    if lmax >= 1:
        s = math.sqrt(3.0)
        out += [s * x, s * y, s * z]
    if lmax >= 2:
        s = math.sqrt(5.0)
        s3 = math.sqrt(3.0)
        out += [
            s * s3 * x * z,
            s * s3 * x * y,
            s * (y * y - 0.5 * (x * x + z * z)),
            s * s3 * y * z,
            s * (s3 / 2) * (z * z - x * x),
        ]
    if lmax >= 3:
        s = math.sqrt(7.0)
        x2, y2, z2 = x * x, y * y, z * z
        out += [
            s * (math.sqrt(10.0) / 4) * x * (3 * z2 - x2),
            s * math.sqrt(15.0) * x * y * z,
            s * (math.sqrt(6.0) / 4) * x * (4 * y2 - x2 - z2),
            s * 0.5 * y * (2 * y2 - 3 * x2 - 3 * z2),
            s * (math.sqrt(6.0) / 4) * z * (4 * y2 - x2 - z2),
            s * (math.sqrt(15.0) / 2) * y * (z2 - x2),
            s * (math.sqrt(10.0) / 4) * z * (z2 - 3 * x2),
        ]
    return torch.stack(out, dim=-1)


def spherical_harmonics_recursive(lmax: int, vec: torch.Tensor) -> torch.Tensor:
    """Generic construction Y_{l+1} ~ w3j(l,1,l+1) . Y_l . Y_1 on raw vectors,
    scaled so that block l has squared norm (2l+1)|r|^{2l} ('component')."""
    vec = vec.to(torch.float64)
    blocks = [torch.ones_like(vec[..., :1])]
    y1 = vec  # (x, y, z) order == e3nn l=1 order
    cur = vec
    if lmax >= 1:
        blocks.append(math.sqrt(3.0) * vec)
    for l in range(1, lmax):
        C = _wigner_3j_f64(l, 1, l + 1)
        nxt = torch.einsum('ijk,...i,...j->...k', C, cur, y1)
        # fix the scale on the polar axis (y): Y_l(y_hat)[m=l] = 1 for the raw poly
        e = torch.zeros(3, dtype=torch.float64)
        e[1] = 1.0
        ref = spherical_harmonics_recursive_raw_on_axis(l + 1)
        scale = ref
        nxt = nxt * scale
        cur = nxt
        blocks.append(math.sqrt(2 * (l + 1) + 1) * nxt)
    return torch.cat(blocks, dim=-1)


@lru_cache(maxsize=None)
def spherical_harmonics_recursive_raw_on_axis(l: int) -> float:
    """Scale c_{l-1} such that the raw recursion has unit norm on the sphere."""
    e = torch.zeros(3, dtype=torch.float64)
    e[1] = 1.0
    cur = e.clone()
    for k in range(1, l):
        C = _wigner_3j_f64(k, 1, k + 1)
        nxt = torch.einsum('ijk,i,j->k', C, cur, e)
        c = 1.0 / nxt.norm().item()
        if k + 1 == l:
            return c
        cur = nxt * c
    return 1.0


# --------------------------------------------------------------------------
# Activations: normalize2mom (SURVEY Appendix A rule 7)
# --------------------------------------------------------------------------


@lru_cache(maxsize=None)
def _second_moment_cst(f) -> float:
    gen = torch.Generator(device='cpu').manual_seed(0)
    z = torch.randn(1_000_000, generator=gen, dtype=torch.float64)
    return f(z).pow(2).mean().pow(-0.5).item()


class Normalize2Mom:
    def __init__(self, f):
        self.f = f
        self.cst = _second_moment_cst(f)
        self.is_id = abs(self.cst - 1) < 1e-4

    def __call__(self, x):
        return self.f(x) if self.is_id else self.f(x).mul(self.cst)


class FullyConnectedNet(torch.nn.Module):
    """e3nn.nn.FullyConnectedNet(hs, act): parameters ``layer{k}.weight`` of
    shape [h_in, h_out]; hidden layers act(x @ W/sqrt(h_in)) * cst, last layer
    linear."""

    def __init__(self, hs: List[int], act):
        super().__init__()
        self.hs = list(hs)
        self.act = Normalize2Mom(act) if act is not None else None
        for k, (h1, h2) in enumerate(zip(hs, hs[1:])):
            layer = torch.nn.Module()
            layer.weight = torch.nn.Parameter(torch.randn(h1, h2))
            setattr(self, f'layer{k}', layer)

    def forward(self, x):
        n = len(self.hs) - 1
        for k in range(n):
            W = getattr(self, f'layer{k}').weight
            x = x @ (W / math.sqrt(self.hs[k]))
            if k < n - 1 and self.act is not None:
                x = self.act(x)
        return x


# --------------------------------------------------------------------------
# o3.Linear (SURVEY Appendix A rule 5)
# --------------------------------------------------------------------------


class Linear(torch.nn.Module):
    def __init__(self, irreps_in, irreps_out, biases: bool = False):
        super().__init__()
        self.irreps_in, self.irreps_out = Irreps(irreps_in), Irreps(irreps_out)
        self.instructions = [
            (i, o)
            for i, (_, ir_i) in enumerate(self.irreps_in)
            for o, (_, ir_o) in enumerate(self.irreps_out)
            if ir_i == ir_o
        ]
        numel = sum(self.irreps_in[i][0] * self.irreps_out[o][0] for i, o in self.instructions)
        self.weight = torch.nn.Parameter(torch.randn(numel))
        nb = sum(mul for mul, ir in self.irreps_out if ir.is_scalar()) if biases else 0
        self.bias = torch.nn.Parameter(torch.zeros(nb))
        self.has_bias = biases

    def forward(self, x):
        sl_in, sl_out = self.irreps_in.slices(), self.irreps_out.slices()
        fan_in = [0] * len(self.irreps_out)
        for i, o in self.instructions:
            fan_in[o] += self.irreps_in[i][0]
        outs = [None] * len(self.irreps_out)
        w_off = 0
        for i, o in self.instructions:
            mi, ir = self.irreps_in[i]
            mo, _ = self.irreps_out[o]
            W = self.weight[w_off:w_off + mi * mo].reshape(mi, mo)
            w_off += mi * mo
            xi = x[..., sl_in[i]].reshape(*x.shape[:-1], mi, ir.dim)
            y = torch.einsum('...ui,uw->...wi', xi, W) / math.sqrt(fan_in[o])
            outs[o] = y if outs[o] is None else outs[o] + y
        b_off = 0
        res = []
        for o, (mo, ir) in enumerate(self.irreps_out):
            y = outs[o]
            if y is None:
                y = x.new_zeros(*x.shape[:-1], mo, ir.dim)
            if self.has_bias and ir.is_scalar():
                y = y + self.bias[b_off:b_off + mo].reshape(mo, 1)
                b_off += mo
            res.append(y.reshape(*x.shape[:-1], mo * ir.dim))
        return torch.cat(res, dim=-1)


# --------------------------------------------------------------------------
# o3.TensorProduct, 'uvu' (per-sample weights) and 'uvw' (shared weights)
# (SURVEY Appendix A rules 4 and 6)
# --------------------------------------------------------------------------


class TensorProduct(torch.nn.Module):
    def __init__(self, irreps_in1, irreps_in2, irreps_out, instructions,
                 shared_weights: bool = False, internal_weights: bool = False):
        super().__init__()
        self.irreps_in1, self.irreps_in2 = Irreps(irreps_in1), Irreps(irreps_in2)
        self.irreps_out = Irreps(irreps_out)
        self.instructions = [tuple(ins[:4]) for ins in instructions]
        self.shared_weights = shared_weights
        self.coeffs, self.wshapes = [], []
        for (i1, i2, io, mode) in self.instructions:
            m1, ir1 = self.irreps_in1[i1]
            m2, ir2 = self.irreps_in2[i2]
            mo, iro = self.irreps_out[io]
            assert ir1.p * ir2.p == iro.p and abs(ir1.l - ir2.l) <= iro.l <= ir1.l + ir2.l
            alpha = iro.dim  # irrep_normalization='component'

            def nelem(ins):
                a, b = self.irreps_in1[ins[0]][0], self.irreps_in2[ins[1]][0]
                return {'uvw': a * b, 'uvu': b, 'uvv': a, 'uuu': 1}[ins[3]]

            x = sum(nelem(ins) for ins in self.instructions if ins[2] == io)  # 'element'
            alpha = alpha / x if x > 0 else alpha
            self.coeffs.append(math.sqrt(alpha))
            self.wshapes.append({'uvw': (m1, m2, mo), 'uvu': (m1, m2), 'uuu': (m1,)}[mode])
        self.weight_numel = sum(math.prod(s) for s in self.wshapes)
        if internal_weights:
            self.weight = torch.nn.Parameter(torch.randn(self.weight_numel))
        else:
            self.register_buffer('weight', torch.zeros(0))

    def forward(self, x, y, weight=None):
        if weight is None:
            weight = self.weight
        sl1, sl2 = self.irreps_in1.slices(), self.irreps_in2.slices()
        outs = [None] * len(self.irreps_out)
        w_off = 0
        for (i1, i2, io, mode), c, ws in zip(self.instructions, self.coeffs, self.wshapes):
            m1, ir1 = self.irreps_in1[i1]
            m2, ir2 = self.irreps_in2[i2]
            mo, iro = self.irreps_out[io]
            C = wigner_3j(ir1.l, ir2.l, iro.l, dtype=x.dtype).to(x.device)
            n = math.prod(ws)
            w = weight[..., w_off:w_off + n]
            w_off += n
            xi = x[..., sl1[i1]].reshape(-1, m1, ir1.dim)
            yj = y[..., sl2[i2]].reshape(-1, m2, ir2.dim)
            if mode == 'uvu':
                assert m2 == 1
                w = w.reshape(-1, m1, m2)
                # materialised per-path message, like e3nn's generated einsum chain
                xy = torch.einsum('zui,zvj->zuvij', xi, yj)
                r = torch.einsum('zuv,ijk,zuvij->zuk', w, C, xy)
            elif mode == 'uvw':
                w = w.reshape(m1, m2, mo)
                xy = torch.einsum('zui,zvj->zuvij', xi, yj)
                r = torch.einsum('uvw,ijk,zuvij->zwk', w, C, xy)
            elif mode == 'uuu':
                r = torch.einsum('ijk,zui,zuj->zuk', C, xi, yj)
            else:
                raise NotImplementedError(mode)
            r = c * r
            outs[io] = r if outs[io] is None else outs[io] + r
        res = []
        for o, (mo, iro) in enumerate(self.irreps_out):
            r = outs[o]
            if r is None:
                r = x.new_zeros(x.reshape(-1, x.shape[-1]).shape[0], mo, iro.dim)
            res.append(r.reshape(r.shape[0], mo * iro.dim))
        return torch.cat(res, dim=-1)


class FullyConnectedTensorProduct(TensorProduct):
    def __init__(self, irreps_in1, irreps_in2, irreps_out, **kwargs):
        irreps_in1, irreps_in2, irreps_out = Irreps(irreps_in1), Irreps(irreps_in2), Irreps(irreps_out)
        instr = [
            (i1, i2, io, 'uvw')
            for i1, (_, ir1) in enumerate(irreps_in1)
            for i2, (_, ir2) in enumerate(irreps_in2)
            for io, (_, iro) in enumerate(irreps_out)
            if iro in ir1 * ir2
        ]
        kwargs.pop('shared_weights', None)
        kwargs.pop('internal_weights', None)
        super().__init__(irreps_in1, irreps_in2, irreps_out, instr,
                         shared_weights=True, internal_weights=True)


# --------------------------------------------------------------------------
# e3nn.nn.Gate (SURVEY Appendix A rule 8) incl. the sort+cut of its input
# --------------------------------------------------------------------------


def _act_out_parity(act, p_in):
    if p_in == 1:
        return 1
    x = torch.linspace(0, 10, 256, dtype=torch.float64)
    a, b = act(x), act(-x)
    if (a - b).abs().max() < 1e-10:
        return 1
    if (a + b).abs().max() < 1e-10:
        return -1
    raise ValueError('activation on odd scalars must be even or odd')


class Gate(torch.nn.Module):
    def __init__(self, irreps_scalars, act_scalars, irreps_gates, act_gates, irreps_gated):
        super().__init__()
        sc, gt, gd = Irreps(irreps_scalars), Irreps(irreps_gates), Irreps(irreps_gated)
        assert gt.num_irreps == gd.num_irreps
        self.irreps_scalars, self.irreps_gates, self.irreps_gated = sc, gt, gd
        # sort + cut: irreps_in is (scalars+gates+gated) sorted then simplified
        cat = sc + gt + gd
        group = [0] * len(sc) + [1] * len(gt) + [2] * len(gd)
        srt, p, inv = cat.sort()
        self.irreps_in = srt.simplify()
        offs_sorted = srt.offsets()
        # pieces[g] = list of (start, stop) in the input, in original order of group g
        self._pieces = {0: [], 1: [], 2: []}
        for old_idx, g in enumerate(group):
            new_idx = p[old_idx]
            mul, ir = cat[old_idx]
            self._pieces[g].append((offs_sorted[new_idx], offs_sorted[new_idx] + mul * ir.dim))
        self.act_scalars = [Normalize2Mom(a) for a in act_scalars]
        self.act_gates = [Normalize2Mom(a) for a in act_gates]
        assert len(self.act_scalars) == len(sc) and len(self.act_gates) == len(gt)
        out_sc = Irreps([(mul, (0, _act_out_parity(a, ir.p)))
                         for (mul, ir), a in zip(sc, act_scalars)])
        gate_p = [_act_out_parity(a, ir.p) for (mul, ir), a in zip(gt, act_gates)]
        # ElementwiseTensorProduct(gated, gates): one gate scalar per gated channel
        out_gd, k = [], 0
        gates_flat_p = []
        for (mul, ir), pp in zip(gt, gate_p):
            gates_flat_p += [pp] * mul
        for mul, ir in gd:
            out_gd.append((mul, (ir.l, ir.p * gates_flat_p[k])))
            k += mul
        self.irreps_out = out_sc + Irreps(out_gd)

    def _cut(self, x, g):
        parts = [x[..., a:b] for a, b in self._pieces[g]]
        return torch.cat(parts, dim=-1) if parts else x[..., :0]

    def forward(self, x):
        scalars, gates, gated = self._cut(x, 0), self._cut(x, 1), self._cut(x, 2)
        outs, o = [], 0
        for (mul, _), a in zip(self.irreps_scalars, self.act_scalars):
            outs.append(a(scalars[..., o:o + mul]))
            o += mul
        if gates.shape[-1]:
            g, o = [], 0
            for (mul, _), a in zip(self.irreps_gates, self.act_gates):
                g.append(a(gates[..., o:o + mul]))
                o += mul
            g = torch.cat(g, dim=-1)
            o, k = 0, 0
            for mul, ir in self.irreps_gated:
                blk = gated[..., o:o + mul * ir.dim].reshape(*x.shape[:-1], mul, ir.dim)
                outs.append((blk * g[..., k:k + mul].unsqueeze(-1)).reshape(*x.shape[:-1], mul * ir.dim))
                o += mul * ir.dim
                k += mul
        return torch.cat(outs, dim=-1)
