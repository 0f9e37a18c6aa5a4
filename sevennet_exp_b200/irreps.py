"""Irreps bookkeeping, Wigner-3j tables and device layouts (host side, numpy only).

The reference delegates all of this to e3nn (``pyproject.toml:24``); the call
sites are ``sevenn/nn/convolution.py:60-81`` (uvu instruction list),
``sevenn/util.py:163-185`` (``infer_irreps_out``), ``sevenn/nn/equivariant_gate.py``
(gate irreps) and ``sevenn/scripts/backward_compatibility.py:75,126`` (w3j).

Device layout ("ir_mul"): a node tensor with *simplified* irreps
``m_0 x ir_0 + m_1 x ir_1 + ...`` is one row of ``sum_k m_k (2 l_k + 1)`` floats per
node; block k is stored **[2l+1][mul]** (component-major, channel contiguous), so
that (a) a warp whose lanes run over channels reads contiguous memory and (b)
every o3.Linear is a dense row-major GEMM over rows (node, m).  e3nn's own
layout ("mul_ir", block k = [mul][2l+1]) is what the ``sevenn.nn`` module API
exposes; ``Layout.from_e3nn / to_e3nn`` convert at that boundary.
"""
import math
from fractions import Fraction
from functools import lru_cache
from typing import List, Sequence, Tuple

import numpy as np

IrrepT = Tuple[int, int]           # (l, p), p in {+1, -1}
EntryT = Tuple[int, IrrepT]        # (mul, (l, p))


# ------------------------------------------------------------------ parsing --
def parse_irreps(spec) -> List[EntryT]:
    if isinstance(spec, str):
        out = []
        for term in spec.split('+'):
            term = term.strip()
            if not term:
                continue
            mul, ir = term.split('x') if 'x' in term else ('1', term)
            ir = ir.strip()
            out.append((int(mul), (int(ir[:-1]), 1 if ir[-1] == 'e' else -1)))
        return out
    out = []
    for it in spec:
        mul, ir = it
        out.append((int(mul), (int(ir[0]), int(ir[1]))))
    return out


def irreps_str(irreps: Sequence[EntryT]) -> str:
    return '+'.join(f"{m}x{l}{'e' if p == 1 else 'o'}" for m, (l, p) in irreps)


def irreps_dim(irreps: Sequence[EntryT]) -> int:
    return sum(m * (2 * l + 1) for m, (l, _) in irreps)


def sort_irreps(irreps: Sequence[EntryT]):
    """Stable sort by (l, p); returns (sorted, perm) with perm[old] = new."""
    order = sorted(range(len(irreps)), key=lambda i: (irreps[i][1], i))
    perm = [0] * len(order)
    for new, old in enumerate(order):
        perm[old] = new
    return [irreps[i] for i in order], perm


def simplify(irreps: Sequence[EntryT]) -> List[EntryT]:
    out: List[EntryT] = []
    for m, ir in irreps:
        if out and out[-1][1] == ir:
            out[-1] = (out[-1][0] + m, ir)
        elif m > 0:
            out.append((m, ir))
    return out


def spherical_harmonics_irreps(lmax: int, p: int) -> List[EntryT]:
    return [(1, (l, p ** l)) for l in range(lmax + 1)]


def tp_irreps(ir1: IrrepT, ir2: IrrepT) -> List[IrrepT]:
    return [(l, ir1[1] * ir2[1]) for l in range(abs(ir1[0] - ir2[0]), ir1[0] + ir2[0] + 1)]


def infer_irreps_out(irreps_x, irreps_operand, drop_l=False, parity_mode='full', fix_multiplicity=False):
    """sevenn/util.py:163-185."""
    a, b = simplify(irreps_x), simplify(irreps_operand)
    full = []
    for m1, i1 in a:
        for m2, i2 in b:
            for ir in tp_irreps(i1, i2):
                full.append((m1 * m2, ir))
    full = simplify(sort_irreps(full)[0])
    new = []
    for mul, (l, p) in full:
        if drop_l is not False and l > drop_l:
            continue
        if parity_mode == 'even' and p == -1:
            continue
        if parity_mode == 'sph' and p != (-1) ** l:
            continue
        new.append((fix_multiplicity if fix_multiplicity else mul, (l, p)))
    return new


# ------------------------------------------------------------------ layouts --
class Layout:
    """ir_mul device layout of a (possibly unsimplified) e3nn irreps list."""

    def __init__(self, irreps: Sequence[EntryT]):
        self.entries = list(irreps)                 # e3nn order (may repeat an irrep back to back)
        self.groups = simplify(irreps)              # device blocks
        self.dim = irreps_dim(irreps)
        self.group_off, o = [], 0
        for m, (l, _) in self.groups:
            self.group_off.append(o)
            o += m * (2 * l + 1)
        # entry -> (group, channel offset inside the group)
        self.entry_group, self.entry_uoff = [], []
        g, u = -1, 0
        last = None
        for m, ir in self.entries:
            if m == 0:
                self.entry_group.append(max(g, 0)); self.entry_uoff.append(u); continue
            if ir != last:
                g += 1; u = 0; last = ir
            self.entry_group.append(g); self.entry_uoff.append(u)
            u += m
        self.e3nn_off, o = [], 0
        for m, (l, _) in self.entries:
            self.e3nn_off.append(o)
            o += m * (2 * l + 1)

    def group_of(self, ir: IrrepT) -> int:
        for g, (_, gi) in enumerate(self.groups):
            if gi == ir:
                return g
        return -1

    def perm_to_e3nn(self) -> np.ndarray:
        """idx such that x_e3nn[..., c] = x_dev[..., idx[c]]."""
        idx = np.zeros(self.dim, dtype=np.int64)
        for k, (m, (l, _)) in enumerate(self.entries):
            g = self.entry_group[k]
            M = self.groups[g][0]
            for u in range(m):
                for c in range(2 * l + 1):
                    idx[self.e3nn_off[k] + u * (2 * l + 1) + c] = (
                        self.group_off[g] + c * M + self.entry_uoff[k] + u)
        return idx

    def to_e3nn(self, x):
        import torch
        return x[..., torch.as_tensor(self.perm_to_e3nn(), device=x.device)]

    def from_e3nn(self, x):
        import torch
        inv = np.argsort(self.perm_to_e3nn())
        return x[..., torch.as_tensor(inv, device=x.device)]


# ---------------------------------------------------------------- wigner 3j --
def _f(n: int) -> int:
    return math.factorial(n)


def _su2_cg(j1, m1, j2, m2, j3, m3) -> float:
    """<j1 m1 j2 m2 | j3 m3>, Racah's closed form, exact rationals under the root."""
    if m1 + m2 != m3:
        return 0.0
    pref = Fraction((2 * j3 + 1) * _f(j3 + j1 - j2) * _f(j3 - j1 + j2) * _f(j1 + j2 - j3)
                    * _f(j3 + m3) * _f(j3 - m3),
                    _f(j1 + j2 + j3 + 1) * _f(j1 - m1) * _f(j1 + m1) * _f(j2 - m2) * _f(j2 + m2))
    lo = max(-j1 + j2 + m3, -j1 + m1, 0)
    hi = min(j2 + j3 + m1, j3 - j1 + j2, j3 + m3)
    acc = Fraction(0)
    for v in range(lo, hi + 1):
        acc += Fraction((-1) ** (v + j2 + m2) * _f(j2 + j3 + m1 - v) * _f(j1 - m1 + v),
                        _f(v) * _f(j3 - j1 + j2 - v) * _f(j3 + m3 - v) * _f(v + j1 - j2 - m3))
    return math.sqrt(float(pref)) * float(acc)


def _real_basis(l: int) -> np.ndarray:
    q = np.zeros((2 * l + 1, 2 * l + 1), dtype=np.complex128)
    r = 1.0 / math.sqrt(2.0)
    for m in range(1, l + 1):
        q[l - m, l + m] = r
        q[l - m, l - m] = -1j * r
        q[l + m, l + m] = (-1) ** m * r
        q[l + m, l - m] = 1j * (-1) ** m * r
    q[l, l] = 1.0
    return (-1j) ** l * q


@lru_cache(maxsize=None)
def wigner_3j(l1: int, l2: int, l3: int) -> np.ndarray:
    """Real-basis Wigner 3j [2l1+1, 2l2+1, 2l3+1], unit Frobenius norm, e3nn>=0.5 signs."""
    assert abs(l1 - l2) <= l3 <= l1 + l2
    c = np.zeros((2 * l1 + 1, 2 * l2 + 1, 2 * l3 + 1), dtype=np.complex128)
    for a in range(-l1, l1 + 1):
        for b in range(-l2, l2 + 1):
            if abs(a + b) <= l3:
                c[l1 + a, l2 + b, l3 + a + b] = _su2_cg(l1, a, l2, b, l3, a + b)
    q1, q2, q3 = _real_basis(l1), _real_basis(l2), _real_basis(l3)
    r = np.einsum('ij,kl,mn,ikn->jlm', q1, q2, np.conj(q3.T), c)
    assert np.abs(r.imag).max() < 1e-12
    r = r.real.copy()
    r[np.abs(r) < 1e-14] = 0.0
    return r / np.linalg.norm(r)


# normalize2mom constants (e3nn: 1e6 float64 samples, torch.Generator seed 0); see
# SURVEY Appendix A rule 7.  silu is also frozen in the reference's deployed artefact.
def normalize2mom_const(act_name: str) -> float:
    import torch
    fns = {
        'silu': torch.nn.functional.silu, 'tanh': torch.tanh, 'abs': torch.abs,
        'relu': torch.nn.functional.relu, 'sigmoid': torch.sigmoid,
        'ssp': lambda x: torch.nn.functional.softplus(x) - math.log(2.0),
        'elu': torch.nn.functional.elu,
    }
    return _n2m(act_name, fns[act_name])


@lru_cache(maxsize=None)
def _n2m(name, f):
    import torch
    gen = torch.Generator(device='cpu').manual_seed(0)
    z = torch.randn(1_000_000, generator=gen, dtype=torch.float64)
    c = f(z).pow(2).mean().pow(-0.5).item()
    return 1.0 if abs(c - 1.0) < 1e-4 else c
