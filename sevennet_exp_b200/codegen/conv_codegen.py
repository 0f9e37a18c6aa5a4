"""Generator for the fused gather -> uvu Clebsch-Gordan tensor product -> segment-sum kernels.

Replaces e3nn's codegen'd einsum chain + ``scatter_reduce_`` used by
``sevenn/nn/convolution.py:118-137`` (``message_gather`` :16-25) and its autograd
backward.  One specialised translation unit is emitted per ConvSpec (irreps of
x, lmax of the filter, path list); CG coefficients become fp32 literals and all
structural zeros disappear at generation time.

Mapping (both directions): one warp per (destination atom, input-irrep group);
lanes run over channels u (``cpl = ceil(mul/32)`` channels per lane), the
destination row's accumulators live in registers, the warp walks the
dst-sorted CSR row, so the reduction over edges needs no atomics and is
bit-reproducible.  Per edge and path the warp first forms the edge-only matrix
``M[i,k] = sqrt(2 l3+1) * sum_j C[i,j,k] Y[l2,j]`` (warp-uniform), then
``out[u,k] += w[e,p,u] * sum_i M[i,k] x_j[u,i]``.

Backward (same mapping, destination gradient in registers) produces, per edge:
``dW[e,p,u]`` (for the radial-MLP backward), the per-edge ``dx_e[e, x_dim]``
(summed per *source* atom by a separate CSC gather kernel -> no atomics) and
``dY[e, :]`` (lane partials reduced with warp shuffles).

The same text compiles as CUDA (nvcc, sm_100a) and, with -DSEVENN_EMU, as plain
C++ where a "warp" is a 32-iteration loop: tests use the latter to check the
generated arithmetic against the oracle without a GPU.  The product never
loads the emulation build.
"""
import math
from typing import List

from .. import irreps as ir
from ..spec import ConvSpec


def _f(v: float) -> str:
    s = repr(float(np_f32(v)))
    if 'e' not in s and '.' not in s and 'inf' not in s:
        s += '.0'
    return s + 'f'


def np_f32(v):
    import numpy as np
    return np.float32(v)


class _Group:
    def __init__(self, spec: ConvSpec, g: int):
        self.g = g
        self.mul, (self.l1, _) = spec.lx.groups[g]
        self.d1 = 2 * self.l1 + 1
        self.cpl = (self.mul + 31) // 32
        self.masked = self.mul % 32 != 0
        self.xoff = spec.lx.group_off[g]
        self.paths = [(pi, p) for pi, p in enumerate(spec.paths) if p.gx == g]


def _m_terms(l1, l2, l3):
    """{(i,k): [(j, coef)]} with alpha folded in, and the nnz list [(i,j,k,coef)]."""
    C = ir.wigner_3j(l1, l2, l3) * math.sqrt(2 * l3 + 1)
    ent, nnz = {}, []
    for i in range(2 * l1 + 1):
        for k in range(2 * l3 + 1):
            t = [(j, float(C[i, j, k])) for j in range(2 * l2 + 1) if abs(C[i, j, k]) > 1e-12]
            if t:
                ent[(i, k)] = t
                nnz += [(i, j, k, c) for j, c in t]
    return ent, nnz


def _emit_M(lines, ent, l2, ind):
    for (i, k), terms in ent.items():
        expr = ' + '.join(f'{_f(c)}*y{l2 * l2 + j}' for j, c in terms)
        lines.append(f'{ind}const float M_{i}_{k} = {expr};')


def _load_common(L, G: _Group, spec: ConvSpec, ind):
    S = (spec.lmax_sh + 1) ** 2
    L.append(f'{ind}const int j = col[e];')
    L.append(f'{ind}const float* xr = x + (size_t)j * ldx + {G.xoff};')
    L.append(f'{ind}const float* wr = w + (size_t)e * {spec.wn};')
    used = sorted({p.l2 for _, p in G.paths})
    for l2 in used:
        for jj in range(l2 * l2, (l2 + 1) ** 2):
            L.append(f'{ind}const float y{jj} = LDG(Y + (size_t)e * {S} + {jj});')
    for c in range(G.cpl):
        for m in range(G.d1):
            ld = f'LDG(xr + {m * G.mul + c * 32} + lane)'
            if G.masked:
                ld = f'(on{c} ? {ld} : 0.f)'
            L.append(f'{ind}const float x{c}_{m} = {ld};')


def _ldw(G, p, c):
    ld = f'LDG(wr + {p.w_off + c * 32} + lane)'
    return f'(on{c} ? {ld} : 0.f)' if G.masked else ld


def gen_fwd_group(spec: ConvSpec, G: _Group, sig: str) -> List[str]:
    L = []
    name = f'conv_fwd_{sig}_g{G.g}'
    L.append(f'KERNEL_HEAD({name})(const float* __restrict__ x, int ldx, const float* __restrict__ Y,')
    L.append('    const float* __restrict__ w, const int* __restrict__ rowptr, const int* __restrict__ col,')
    L.append('    int n_dst, float inv_denom, float* __restrict__ out, int ldo)')
    L.append('WARP_LOOP_BEGIN(n_dst)')
    if G.masked:
        for c in range(G.cpl):
            L.append(f'  const bool on{c} = ({c * 32} + lane) < {G.mul};')
    for pi, p in G.paths:
        for c in range(G.cpl):
            for k in range(2 * p.l3 + 1):
                L.append(f'  float a{pi}_{c}_{k} = 0.f;')
    L.append('  const int e0 = rowptr[row], e1 = rowptr[row + 1];')
    L.append('  for (int e = e0; e < e1; ++e) {')
    _load_common(L, G, spec, '    ')
    for pi, p in G.paths:
        ent, _ = _m_terms(p.l1, p.l2, p.l3)
        L.append(f'    {{ // path {pi}: {p.l1} x {p.l2} -> {p.l3}, mul {p.mul}')
        _emit_M(L, ent, p.l2, '      ')
        for c in range(G.cpl):
            L.append(f'      {{ const float wv = {_ldw(G, p, c)};')
            for k in range(2 * p.l3 + 1):
                terms = [f'M_{i}_{k}*x{c}_{i}' for i in range(G.d1) if (i, k) in ent]
                if terms:
                    L.append(f'        a{pi}_{c}_{k} = fmaf(wv, {" + ".join(terms)}, a{pi}_{c}_{k});')
            L.append('      }')
        L.append('    }')
    L.append('  }')
    L.append('  float* orow = out + (size_t)row * ldo;')
    for pi, p in G.paths:
        Mg = spec.lmid.groups[p.gm][0]
        base = spec.lmid.group_off[p.gm] + p.u_off
        for c in range(G.cpl):
            for k in range(2 * p.l3 + 1):
                st = f'orow[{base + k * Mg + c * 32} + lane] = a{pi}_{c}_{k} * inv_denom;'
                L.append(f'  if (on{c}) {st}' if G.masked else f'  {st}')
    L.append('WARP_LOOP_END')
    return L


def gen_bwd_group(spec: ConvSpec, G: _Group, sig: str) -> List[str]:
    S = (spec.lmax_sh + 1) ** 2
    L = []
    name = f'conv_bwd_{sig}_g{G.g}'
    L.append(f'KERNEL_HEAD({name})(const float* __restrict__ x, int ldx, const float* __restrict__ Y,')
    L.append('    const float* __restrict__ w, const int* __restrict__ rowptr, const int* __restrict__ col,')
    L.append('    int n_dst, float inv_denom, const float* __restrict__ gout, int ldo,')
    L.append('    float* __restrict__ dW, float* __restrict__ dxe, int lddx, float* dY)')
    L.append('WARP_LOOP_BEGIN(n_dst)')
    if G.masked:
        for c in range(G.cpl):
            L.append(f'  const bool on{c} = ({c * 32} + lane) < {G.mul};')
    L.append('  const float* grow = gout + (size_t)row * ldo;')
    for pi, p in G.paths:
        Mg = spec.lmid.groups[p.gm][0]
        base = spec.lmid.group_off[p.gm] + p.u_off
        for c in range(G.cpl):
            for k in range(2 * p.l3 + 1):
                ld = f'LDG(grow + {base + k * Mg + c * 32} + lane) * inv_denom'
                if G.masked:
                    ld = f'(on{c} ? {ld} : 0.f)'
                L.append(f'  const float g{pi}_{c}_{k} = {ld};')
    L.append('  const int e0 = rowptr[row], e1 = rowptr[row + 1];')
    L.append('  for (int e = e0; e < e1; ++e) {')
    _load_common(L, G, spec, '    ')
    used = sorted({p.l2 for _, p in G.paths})
    for l2 in used:
        for jj in range(l2 * l2, (l2 + 1) ** 2):
            L.append(f'    float dy{jj} = 0.f;')
    for c in range(G.cpl):
        for m in range(G.d1):
            L.append(f'    float dx{c}_{m} = 0.f;')
    L.append(f'    float* dwr = dW + (size_t)e * {spec.wn};')
    for pi, p in G.paths:
        ent, nnz = _m_terms(p.l1, p.l2, p.l3)
        L.append(f'    {{ // path {pi}: {p.l1} x {p.l2} -> {p.l3}')
        _emit_M(L, ent, p.l2, '      ')
        for (i, k) in ent:
            L.append(f'      float P_{i}_{k} = 0.f;')
        for c in range(G.cpl):
            L.append(f'      {{ const float wv = {_ldw(G, p, c)};')
            dw_terms = []
            for k in range(2 * p.l3 + 1):
                terms = [f'M_{i}_{k}*x{c}_{i}' for i in range(G.d1) if (i, k) in ent]
                if terms:
                    dw_terms.append(f'g{pi}_{c}_{k}*({" + ".join(terms)})')
            st = f'dwr[{p.w_off + c * 32} + lane] = {" + ".join(dw_terms) if dw_terms else "0.f"};'
            L.append(f'        if (on{c}) {st}' if G.masked else f'        {st}')
            for k in range(2 * p.l3 + 1):
                idx = [i for i in range(G.d1) if (i, k) in ent]
                if not idx:
                    continue
                L.append(f'        {{ const float wg = wv * g{pi}_{c}_{k};')
                for i in idx:
                    L.append(f'          dx{c}_{i} = fmaf(M_{i}_{k}, wg, dx{c}_{i}); P_{i}_{k} = fmaf(wg, x{c}_{i}, P_{i}_{k});')
                L.append('        }')
            L.append('      }')
        for (i, j, k, cf) in nnz:
            L.append(f'      dy{p.l2 * p.l2 + j} = fmaf({_f(cf)}, P_{i}_{k}, dy{p.l2 * p.l2 + j});')
        L.append('    }')
    L.append(f'    float* dxr = dxe + (size_t)e * lddx + {G.xoff};')
    for c in range(G.cpl):
        for m in range(G.d1):
            st = f'dxr[{m * G.mul + c * 32} + lane] = dx{c}_{m};'
            L.append(f'    if (on{c}) {st}' if G.masked else f'    {st}')
    for l2 in used:
        for jj in range(l2 * l2, (l2 + 1) ** 2):
            L.append(f'    WARP_SUM_ADD(dY + (size_t)e * {S} + {jj}, dy{jj});')
    L.append('  }')
    L.append('WARP_LOOP_END')
    return L


def signature(spec: ConvSpec) -> str:
    return spec.signature()


def generate(spec: ConvSpec) -> str:
    sig = signature(spec)
    groups = [_Group(spec, g) for g in range(len(spec.lx.groups))]
    groups = [G for G in groups if G.paths]
    out = ['// GENERATED by sevennet_exp_b200/codegen/conv_codegen.py -- do not edit.',
           f'// x: {ir.irreps_str(spec.lx.groups)}  mid: {ir.irreps_str(spec.lmid.groups)}  '
           f'lmax_sh {spec.lmax_sh}  paths {len(spec.paths)}  wn {spec.wn}',
           '#include "conv_common.cuh"', '']
    for G in groups:
        out += gen_fwd_group(spec, G, sig) + ['']
        out += gen_bwd_group(spec, G, sig) + ['']
    # launchers
    out.append(f'extern "C" int sevenn_conv_fwd_{sig}(const float* x, int ldx, const float* Y, const float* w,')
    out.append('    const int* rowptr, const int* col, int n_dst, float inv_denom, float* out, int ldo, void* stream) {')
    out.append('  if (n_dst <= 0) return 0;')
    for G in groups:
        out.append(f'  LAUNCH(conv_fwd_{sig}_g{G.g}, n_dst, stream)(x, ldx, Y, w, rowptr, col, n_dst, inv_denom, out, ldo);')
    out.append('  return LAUNCH_STATUS();')
    out.append('}')
    out.append(f'extern "C" int sevenn_conv_bwd_{sig}(const float* x, int ldx, const float* Y, const float* w,')
    out.append('    const int* rowptr, const int* col, int n_dst, float inv_denom, const float* gout, int ldo,')
    out.append('    float* dW, float* dxe, int lddx, float* dY, void* stream) {')
    out.append('  if (n_dst <= 0) return 0;')
    for G in groups:
        out.append(f'  LAUNCH(conv_bwd_{sig}_g{G.g}, n_dst, stream)(x, ldx, Y, w, rowptr, col, n_dst, inv_denom, gout, ldo, dW, dxe, lddx, dY);')
    out.append('  return LAUNCH_STATUS();')
    out.append('}')
    out.append(f'extern "C" int sevenn_conv_info_{sig}(int* x_dim, int* mid_dim, int* wn, int* sh_dim) {{')
    out.append(f'  *x_dim = {spec.lx.dim}; *mid_dim = {spec.lmid.dim}; *wn = {spec.wn}; *sh_dim = {(spec.lmax_sh + 1) ** 2}; return 0; }}')
    return '\n'.join(out) + '\n'
