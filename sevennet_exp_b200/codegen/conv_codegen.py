"""Generator for the fused gather -> uvu Clebsch-Gordan tensor product -> segment-sum kernels.

Replaces e3nn's codegen'd einsum chain + ``scatter_reduce_`` used by
``sevenn/nn/convolution.py:118-137`` (``message_gather`` :16-25) and its autograd
backward.  One specialised translation unit is emitted per ConvSpec (irreps of
x, lmax of the filter, path list); CG coefficients become fp32 literals and all
structural zeros disappear at generation time.

Mapping (both directions): one warp per (destination atom, unit) where a unit
is a 32-channel chunk of one input-irrep group (SevenNet-0: 4 + 2 + 1 = 7 units
per atom; the tabulated kernels give one warp two chunks of the scalar group:
2 + 2 + 1 = 5, see make_groups); lanes run over the chunk's channels u, the destination row's
accumulators live in registers, the warp walks the dst-sorted CSR row, so the
reduction over edges needs no atomics and is bit-reproducible.  All units run
in one launch (blockIdx.y = unit).  The edge loop is software pipelined: the
gathered source row, the radial weights and the spherical harmonics of edge
e+1 (and the source index of e+2) are in flight while edge e is computed.  Per edge and path the warp first forms the edge-only matrix
``M[i,k] = sqrt(2 l3+1) * sum_j C[i,j,k] Y[l2,j]`` (warp-uniform), then
``out[u,k] += w[e,p,u] * sum_i M[i,k] x_j[u,i]``.

Backward (same mapping, destination gradient in registers) produces, per edge:
``dW[e,p,u]`` (for the radial-MLP backward), the per-edge ``dx_e[e, x_dim]``
(summed per *source* atom by a separate CSC gather kernel -> no atomics) and
``dY[unit, e, :]`` (lane partials reduced with a transposing warp butterfly; one
slab per unit, so no two warps ever touch the same word; the edge-embedding
backward sums the slabs).

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
    """One input irrep (mul x l1) of x; its channels are cut into 32-wide chunks = units."""

    def __init__(self, spec: ConvSpec, g: int, unit0: int):
        self.g = g
        self.mul, (self.l1, _) = spec.lx.groups[g]
        self.d1 = 2 * self.l1 + 1
        self.nchunk = (self.mul + 31) // 32
        self.masked = self.mul % 32 != 0
        self.xoff = spec.lx.group_off[g]
        self.paths = [(pi, p) for pi, p in enumerate(spec.paths) if p.gx == g]
        self.unit0 = unit0
        self.cpw = 1                      # 32-channel chunks handled by one warp (tabulated kernels: see make_groups)
        self.used_l2 = sorted({p.l2 for _, p in self.paths})
        self.ys = [jj for l2 in self.used_l2 for jj in range(l2 * l2, (l2 + 1) ** 2)]


# Tabulated kernels: a warp of a scalar (l1 = 0) group takes TWO 32-channel chunks.  Those units are light (one input
# component, 2 l + 1 trivial M entries per path) and the kernel's register budget is set by the heavy l1 >= 1 units, so
# the second chunk is free in registers, while everything that does not depend on the channel -- Hermite basis, Y loads,
# source-index / table-index loads, loop control and above all the per-edge dY lane reduction -- is paid once per pair.
import os as _os0
TAB_CPW = int(_os0.environ.get('SEVENN_TAB_CPW', '2'))
# packed fp32 (fma.rn.f32x2) for the per-channel arithmetic of two-chunk units: the chunk pair (a, b) of a lane is one
# float2; warp-uniform M entries and literals are free broadcast / immediate operands of FFMA2
TAB_F2 = _os0.environ.get('SEVENN_TAB_F2', '0') == '1'
# table rows as monomial coefficients (c0, c1, c2, c3) of t instead of Hermite nodes (radial_table.build_table layout='poly')
# A/B on B200 (tools/gpu.sh ab, r02): cfg2 3.84 -> 3.75 ms, 216 k-atom Si 106.1 -> 103.9 ms, cfg3 32.6 -> 32.1 ms per step
TAB_POLY = _os0.environ.get('SEVENN_TAB_POLY', '1') == '1' and not TAB_F2


def make_groups(spec: ConvSpec, tab: bool = False) -> List[_Group]:
    out, u = [], 0
    for g in range(len(spec.lx.groups)):
        G = _Group(spec, g, u)
        if G.paths:
            if tab and TAB_CPW == 2 and G.l1 == 0 and not G.masked and G.nchunk % 2 == 0:
                G.cpw = 2
            out.append(G)
            u += G.nchunk // G.cpw
    return out


def n_units(spec: ConvSpec, tab: bool = False) -> int:
    return sum(G.nchunk // G.cpw for G in make_groups(spec, tab))


def _sx(G: _Group, c: int) -> str:
    """Suffix of the per-channel variables of chunk c of a warp."""
    return '' if G.cpw == 1 else '_' + 'ab'[c]


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


def _on(G, expr, zero='0.f'):
    return f'(on ? {expr} : {zero})' if G.masked else expr


RING_DEPTH = 4      # edges in flight per warp: RING_DEPTH - 1

# Occupancy targets (CTAs of 4 warps per SM) for (forward, backward), by signature.  ptxas caps registers at
# 65536 / (128 * target); the values below were checked to compile without spills (cuobjdump --dump-resource-usage).
MIN_CTAS = {
    'c377d4386a8': (8, 6),      # SevenNet-0 inner layers: 64 / 80 registers
    'c703e622066': (4, 3),      # SevenNet-l3i5 inner layers: 127 / 168 registers
}
MIN_CTAS_DEFAULT = (1, 1)
# tabulated kernels: measured with tools/gpu.sh sweep-occupancy on B200 (profiles/r01e_occupancy_sweep.log): SevenNet-0
# forward 1.090 -> 1.045 ms at 5 instead of 6; l3i5 backward 22.5 -> 18.6 ms at 4 instead of 2 (128 registers,
# 64 B of spill; 4 CTAs is also what the 56.6 KB ring per CTA allows)
MIN_CTAS_TAB = {
    'c377d4386a8': (5, 4),      # with the table rows prefetched into registers (TAB_T_MODE below)
    'c703e622066': (3, 4),
}


# edge-pair backward (gen_bwd_tab_pair): per-edge temporaries double; 2 CTAs of 4 warps per SM leave 255 registers
MIN_CTAS_TAB_EP = {}
MIN_CTAS_TAB_EP_DEFAULT = int(_os0.environ.get('SEVENN_MIN_CTAS_EP', '2'))


def _env_overrides():
    """Tuning runs only (tools/gpu.sh sweep-occupancy): SEVENN_MIN_CTAS / SEVENN_MIN_CTAS_TAB = 'sig:fwd,bwd;sig:fwd,bwd'."""
    import os
    for var, table in (('SEVENN_MIN_CTAS', MIN_CTAS), ('SEVENN_MIN_CTAS_TAB', MIN_CTAS_TAB)):
        for item in filter(None, os.environ.get(var, '').split(';')):
            sig, v = item.split(':')
            table[sig] = tuple(int(t) for t in v.split(','))


_env_overrides()


def _stage_floats(G: _Group, spec: ConvSpec) -> int:
    S = (spec.lmax_sh + 1) ** 2
    return (G.d1 + len(G.paths)) * 32 + ((S + 3) // 4) * 4


def ring_floats(spec: ConvSpec) -> int:
    """Shared-memory floats per warp (max over the groups sharing one kernel)."""
    return RING_DEPTH * max(_stage_floats(G, spec) for G in make_groups(spec))


def _issue_stage(L, G: _Group, spec: ConvSpec, ind, e, j, slot):
    """cp.async of everything lane ``lane`` consumes for edge ``e`` (source row ``j``) into ring slot ``slot``:
    its own x_j and w elements (read back by the same lane) and, on lanes < S, one spherical harmonic."""
    S = (spec.lmax_sh + 1) ** 2
    np_ = len(G.paths)
    L.append(f'{ind}{{ float* sd = ring + ({slot}) * {_stage_floats(G, spec)} + lane;')
    L.append(f'{ind}  const float* xr = x + (size_t)({j}) * ldx + {G.xoff} + c32 + lane;')
    L.append(f'{ind}  const float* wr = w + (size_t)({e}) * {spec.wn} + c32 + lane;')
    if G.masked:
        L.append(f'{ind}  if (on) {{')
    for m in range(G.d1):
        L.append(f'{ind}  CP_ASYNC4(sd + {m * 32}, xr + {m * G.mul});')
    for q, (pi, p) in enumerate(G.paths):
        L.append(f'{ind}  CP_ASYNC4(sd + {(G.d1 + q) * 32}, wr + {p.w_off});')
    if G.masked:
        L.append(f'{ind}  }}')
    L.append(f'{ind}  if (lane < {S}) CP_ASYNC4(sd + {(G.d1 + np_) * 32}, Y + (size_t)({e}) * {S} + lane);')
    L.append(f'{ind}}}')


def _pipeline_open(L, G: _Group, spec: ConvSpec):
    """Opens ``if (e0 < e1) { for (e) {`` with x*, w*, y* of edge e defined.  CUDA: a per-warp shared-memory
    ring fed by cp.async keeps RING_DEPTH-1 edges in flight; emulation build: direct loads."""
    S = (spec.lmax_sh + 1) ** 2
    D = RING_DEPTH
    stg = _stage_floats(G, spec)
    L.append('    const int e0 = rowptr[row], e1 = rowptr[row + 1];')
    L.append('    if (e0 < e1) {')
    L.append('#ifdef SEVENN_EMU')
    L.append('      for (int e = e0; e < e1; ++e) {')
    L.append(f'        const float* xr = x + (size_t)col[e] * ldx + {G.xoff} + c32 + lane;')
    L.append(f'        const float* wr = w + (size_t)e * {spec.wn} + c32 + lane;')
    for m in range(G.d1):
        L.append(f'        const float x{m} = {_on(G, f"xr[{m * G.mul}]")};')
    for pi, p in G.paths:
        L.append(f'        const float w{pi} = {_on(G, f"wr[{p.w_off}]")};')
    for jj in G.ys:
        L.append(f'        const float y{jj} = Y[(size_t)e * {S} + {jj}];')
    L.append('#else')
    L.append(f'      for (int d = 0; d < {D - 1}; ++d) {{')
    L.append('        if (e0 + d < e1) {')
    _issue_stage(L, G, spec, '          ', 'e0 + d', 'col[e0 + d]', 'd')
    L.append('        }')
    L.append('        CP_ASYNC_COMMIT();')
    L.append('      }')
    L.append(f'      int jn = col[e0 + {D - 1} < e1 ? e0 + {D - 1} : e1 - 1];')
    L.append('      int slot = 0;')
    L.append('      for (int e = e0; e < e1; ++e) {')
    L.append(f'        CP_ASYNC_WAIT({D - 2});')
    L.append('        __syncwarp();')
    L.append(f'        if (e + {D - 1} < e1) {{')
    _issue_stage(L, G, spec, '          ', f'e + {D - 1}', 'jn', f'(slot + {D - 1}) & {D - 1}')
    L.append('        }')
    L.append('        CP_ASYNC_COMMIT();')
    L.append(f'        jn = col[e + {D} < e1 ? e + {D} : e1 - 1];')
    L.append(f'        const float* sr = ring + slot * {stg};')
    L.append(f'        slot = (slot + 1) & {D - 1};')
    for m in range(G.d1):
        L.append(f'        const float x{m} = {_on(G, f"sr[{m * 32} + lane]")};')
    for q, (pi, p) in enumerate(G.paths):
        L.append(f'        const float w{pi} = {_on(G, f"sr[{(G.d1 + q) * 32} + lane]")};')
    ybase = (G.d1 + len(G.paths)) * 32
    for jj in G.ys:
        L.append(f'        const float y{jj} = sr[{ybase + jj}];')
    L.append('#endif')


def _unit_open(L, G: _Group, first: bool, last: bool):
    cond = f'unit < {G.unit0 + G.nchunk // G.cpw}'
    if first and last:
        L.append('  {')
    elif first:
        L.append(f'  if ({cond}) {{')
    elif last:
        L.append('  else {')
    else:
        L.append(f'  else if ({cond}) {{')
    L.append(f'    const int c32 = (unit - {G.unit0}) * {32 * G.cpw};')
    if G.masked:
        L.append(f'    const bool on = (c32 + lane) < {G.mul};')


def gen_fwd(spec: ConvSpec, groups: List[_Group], sig: str, tab: bool = False) -> List[str]:
    L = []
    mc = (MIN_CTAS_TAB if tab else MIN_CTAS).get(sig, MIN_CTAS_DEFAULT)[0]
    if tab:
        L.append(f'KERNEL_HEAD(conv_fwd_tab_{sig}, {mc})(const float* __restrict__ x, int ldx, const float* __restrict__ Yt,')
        L.append('    const float4* __restrict__ tab, const int* __restrict__ rowptr, const int* __restrict__ col,')
        L.append('    int n_dst, float inv_denom, float* __restrict__ out, int ldo)')
        L.append(f'UNIT_LOOP_BEGIN_DYN(n_dst, {sum(G.nchunk // G.cpw for G in groups)}, {tab_ring_floats(spec, False)})')
    else:
        L.append(f'KERNEL_HEAD(conv_fwd_{sig}, {mc})(const float* __restrict__ x, int ldx, const float* __restrict__ Y,')
        L.append('    const float* __restrict__ w, const int* __restrict__ rowptr, const int* __restrict__ col,')
        L.append('    int n_dst, float inv_denom, float* __restrict__ out, int ldo)')
        L.append(f'UNIT_LOOP_BEGIN(n_dst, {sum(G.nchunk for G in groups)}, {ring_floats(spec)})')
    for gi, G in enumerate(groups):
        _unit_open(L, G, gi == 0, gi == len(groups) - 1)
        f2 = tab and TAB_F2 and G.cpw == 2
        for pi, p in G.paths:
            if f2:
                L.append('    float2 ' + ', '.join(f'a{pi}_{k} = make_float2(0.f, 0.f)' for k in range(2 * p.l3 + 1)) + ';')
                continue
            for c in range(G.cpw):
                L.append('    float ' + ', '.join(f'a{pi}_{k}{_sx(G, c)} = 0.f' for k in range(2 * p.l3 + 1)) + ';')
        if tab:
            _tab_pipeline_open(L, G, spec, False)
        else:
            _pipeline_open(L, G, spec)
        if f2:
            for m in range(G.d1):
                L.append(f'        const float2 x{m} = make_float2(x{m}_a, x{m}_b);')
        for pi, p in G.paths:
            ent, _ = _m_terms(p.l1, p.l2, p.l3)
            L.append(f'        {{ // path {pi}: {p.l1} x {p.l2} -> {p.l3}, mul {p.mul}')
            _emit_M(L, ent, p.l2, '          ')
            rows = [i for i in range(G.d1) if any((i, k) in ent for k in range(2 * p.l3 + 1))]
            if f2:
                if 2 * p.l3 + 1 > len(rows):
                    for i in rows:
                        L.append(f'          const float2 X{i} = __fmul2_rn(w{pi}, x{i});')
                    for k in range(2 * p.l3 + 1):
                        for i in range(G.d1):
                            if (i, k) in ent:
                                L.append(f'          a{pi}_{k} = __ffma2_rn(F2B(M_{i}_{k}), X{i}, a{pi}_{k});')
                else:
                    for k in range(2 * p.l3 + 1):
                        ii = [i for i in range(G.d1) if (i, k) in ent]
                        if ii:
                            t = f'__fmul2_rn(F2B(M_{ii[0]}_{k}), x{ii[0]})'
                            for i in ii[1:]:
                                t = f'__ffma2_rn(F2B(M_{i}_{k}), x{i}, {t})'
                            L.append(f'          a{pi}_{k} = __ffma2_rn(w{pi}, {t}, a{pi}_{k});')
            for c in range(0 if f2 else G.cpw):
                sx = _sx(G, c)
                if 2 * p.l3 + 1 > len(rows):
                    # fewer input than output components: scale the inputs once (nnz(M) + d1 operations) ...
                    for i in rows:
                        L.append(f'          const float X{i}{sx} = w{pi}{sx} * x{i}{sx};')
                    for k in range(2 * p.l3 + 1):
                        for i in range(G.d1):
                            if (i, k) in ent:
                                L.append(f'          a{pi}_{k}{sx} = fmaf(M_{i}_{k}, X{i}{sx}, a{pi}_{k}{sx});')
                else:
                    # ... otherwise scale the outputs (nnz(M) + 2 l3 + 1 operations)
                    for k in range(2 * p.l3 + 1):
                        terms = [f'M_{i}_{k}*x{i}{sx}' for i in range(G.d1) if (i, k) in ent]
                        if terms:
                            L.append(f'          a{pi}_{k}{sx} = fmaf(w{pi}{sx}, {" + ".join(terms)}, a{pi}_{k}{sx});')
            L.append('        }')
        L += getattr(G, 'loop_tail', [])
        L.append('      }')   # for e
        L.append('    }')     # if e0 < e1
        L.append('    float* orow = out + (size_t)row * ldo + c32 + lane;')
        for pi, p in G.paths:
            Mg = spec.lmid.groups[p.gm][0]
            base = spec.lmid.group_off[p.gm] + p.u_off
            for k in range(2 * p.l3 + 1):
                for c in range(G.cpw):
                    acc = f'a{pi}_{k}.{"xy"[c]}' if f2 else f'a{pi}_{k}{_sx(G, c)}'
                    st = f'orow[{base + k * Mg + c * 32}] = {acc} * inv_denom;'
                    L.append(f'    if (on) {st}' if G.masked else f'    {st}')
        L.append('  }')
    L.append('UNIT_LOOP_END')
    return L


def gen_bwd(spec: ConvSpec, groups: List[_Group], sig: str, tab: bool = False) -> List[str]:
    S = (spec.lmax_sh + 1) ** 2
    SD = S + 1 if tab else S          # dY row width: the tabulated kernels append the radial term
    NP = 1
    while NP < S:
        NP *= 2
    gr_in_butterfly = tab and NP > S
    nv_smem = dy_smem_values(spec) if tab else 0
    if nv_smem:
        gr_in_butterfly = nv_smem > S          # here: "the radial term goes through the same reduction"
    L = []
    mc = (MIN_CTAS_TAB if tab else MIN_CTAS).get(sig, MIN_CTAS_DEFAULT)[1]
    if tab:
        L.append(f'KERNEL_HEAD(conv_bwd_tab_{sig}, {mc})(const float* __restrict__ x, int ldx, const float* __restrict__ Yt,')
        L.append('    const float4* __restrict__ tab, const int* __restrict__ rowptr, const int* __restrict__ col,')
        L.append('    int n_dst, float inv_denom, const float* __restrict__ gout, int ldo,')
        L.append('    float* __restrict__ dxe, int lddx, float* __restrict__ dY, long long dy_stride)')
        L.append(f'UNIT_LOOP_BEGIN_DYN(n_dst, {sum(G.nchunk // G.cpw for G in groups)}, {tab_ring_floats(spec, True)})')
    else:
        L.append(f'KERNEL_HEAD(conv_bwd_{sig}, {mc})(const float* __restrict__ x, int ldx, const float* __restrict__ Y,')
        L.append('    const float* __restrict__ w, const int* __restrict__ rowptr, const int* __restrict__ col,')
        L.append('    int n_dst, float inv_denom, const float* __restrict__ gout, int ldo,')
        L.append('    float* __restrict__ dW, float* __restrict__ dxe, int lddx, float* __restrict__ dY, long long dy_stride)')
        L.append(f'UNIT_LOOP_BEGIN(n_dst, {sum(G.nchunk for G in groups)}, {ring_floats(spec)})')
    L.append('  float* dYu = dY + (size_t)unit * (size_t)dy_stride;')
    if nv_smem:
        L.append(f'  sevenn_dy_scratch_init<{nv_smem}>(ring + {tab_data_floats(spec, True)}, lane);')
    for gi, G in enumerate(groups):
        _unit_open(L, G, gi == 0, gi == len(groups) - 1)
        L.append('    const float* grow = gout + (size_t)row * ldo + c32 + lane;')
        for pi, p in G.paths:
            Mg = spec.lmid.groups[p.gm][0]
            base = spec.lmid.group_off[p.gm] + p.u_off
            for k in range(2 * p.l3 + 1):
                for c in range(G.cpw):
                    ld = _on(G, f'LDG(grow + {base + k * Mg + c * 32}) * inv_denom')
                    L.append(f'    const float g{pi}_{k}{_sx(G, c)} = {ld};')
                if tab and TAB_F2 and G.cpw == 2:
                    L.append(f'    const float2 g{pi}_{k} = make_float2(g{pi}_{k}_a, g{pi}_{k}_b);')
        if tab:
            _tab_pipeline_open(L, G, spec, True)
        else:
            _pipeline_open(L, G, spec)
        ndy = max(NP, nv_smem)
        L.append(f'        float dyv[{ndy}] = {{{", ".join(["0.f"] * ndy)}}};')
        f2 = tab and TAB_F2 and G.cpw == 2
        if f2:
            L.append('        float2 ' + ', '.join(f'dx{m} = make_float2(0.f, 0.f)' for m in range(G.d1)) + ', gr2 = make_float2(0.f, 0.f);')
            for m in range(G.d1):
                L.append(f'        const float2 x{m} = make_float2(x{m}_a, x{m}_b);')
        else:
            L.append('        float ' + ', '.join(f'dx{m}{_sx(G, c)} = 0.f' for m in range(G.d1) for c in range(G.cpw)) + ';')
        if tab:
            L.append('        float gr = 0.f;')
        else:
            L.append(f'        float* dwr = dW + (size_t)e * {spec.wn} + c32 + lane;')
        touched = set()
        for pi, p in G.paths:
            # With m_k = sum_i M_ik x_i (the forward message / w) and g_k the output gradient:
            #   q_i   = sum_k M_ik g_k          dx_i += w q_i          dE/dw = sum_i x_i q_i  (= sum_k g_k m_k)
            #   dE/dY_j += sum_{ik} c_ijk (w x_i) g_k
            # i.e. 2 nnz(M) + 3 d1 + nnz(C) operations per path instead of 3 nnz(M) + 2 (2 l3 + 1) + nnz(C).
            ent, nnz = _m_terms(p.l1, p.l2, p.l3)
            L.append(f'        {{ // path {pi}: {p.l1} x {p.l2} -> {p.l3}')
            _emit_M(L, ent, p.l2, '          ')
            rows = [i for i in range(G.d1) if any((i, k) in ent for k in range(2 * p.l3 + 1))]
            if f2:
                for i in rows:
                    kk = [k for k in range(2 * p.l3 + 1) if (i, k) in ent]
                    t = f'__fmul2_rn(F2B(M_{i}_{kk[0]}), g{pi}_{kk[0]})'
                    for k in kk[1:]:
                        t = f'__ffma2_rn(F2B(M_{i}_{k}), g{pi}_{k}, {t})'
                    L.append(f'          const float2 q{i} = {t};')
                    L.append(f'          dx{i} = __ffma2_rn(w{pi}, q{i}, dx{i});')
                if rows:
                    t = f'__fmul2_rn(x{rows[0]}, q{rows[0]})'
                    for i in rows[1:]:
                        t = f'__ffma2_rn(x{i}, q{i}, {t})'
                    L.append(f'          gr2 = __ffma2_rn({t}, wd{pi}, gr2);')
                for i in rows:
                    L.append(f'          const float2 X{i} = __fmul2_rn(w{pi}, x{i});')
            for c in range(0 if f2 else G.cpw):
                sx = _sx(G, c)
                for i in rows:
                    terms = [f'M_{i}_{k}*g{pi}_{k}{sx}' for k in range(2 * p.l3 + 1) if (i, k) in ent]
                    L.append(f'          const float q{i}{sx} = {" + ".join(terms)};')
                    L.append(f'          dx{i}{sx} = fmaf(w{pi}{sx}, q{i}{sx}, dx{i}{sx});')
                dw = ' + '.join(f'x{i}{sx}*q{i}{sx}' for i in rows) if rows else '0.f'
                if tab:
                    if rows:
                        L.append(f'          gr = fmaf({dw}, wd{pi}{sx}, gr);')
                else:
                    st = f'dwr[{p.w_off}] = {dw};'
                    L.append(f'          if (on) {st}' if G.masked else f'          {st}')
                for i in rows:
                    L.append(f'          const float X{i}{sx} = w{pi}{sx} * x{i}{sx};')
            # the two chunks of a warp share c_ijk: their products are added before the coefficient is applied
            for (i, k) in ent:
                if f2:
                    L.append(f'          const float2 PP_{i}_{k} = __fmul2_rn(X{i}, g{pi}_{k});')
                    L.append(f'          const float P_{i}_{k} = PP_{i}_{k}.x + PP_{i}_{k}.y;')
                    continue
                L.append(f'          const float P_{i}_{k} = ' +
                         ' + '.join(f'X{i}{_sx(G, c)} * g{pi}_{k}{_sx(G, c)}' for c in range(G.cpw)) + ';')
            for (i, j, k, cf) in nnz:
                q = p.l2 * p.l2 + j
                if q not in touched:
                    touched.add(q)
                    L.append(f'          dyv[{q}] = P_{i}_{k};' if abs(cf - 1.0) < 1e-12 else
                             f'          dyv[{q}] = {_f(cf)} * P_{i}_{k};')
                else:
                    L.append(f'          dyv[{q}] = fmaf({_f(cf)}, P_{i}_{k}, dyv[{q}]);')
            L.append('        }')
        if f2:
            L.append('        gr = gr2.x + gr2.y;')
        # dxe == nullptr: the caller does not need dE/dx (first layer: its input does not depend on the positions)
        L.append('        if (dxe != nullptr) {')
        L.append(f'          float* dxr = dxe + (size_t)e * lddx + {G.xoff} + c32 + lane;')
        for m in range(G.d1):
            for c in range(G.cpw):
                val = f'dx{m}.{"xy"[c]}' if f2 else f'dx{m}{_sx(G, c)}'
                st = f'dxr[{m * G.mul + c * 32}] = {val};'
                L.append(f'          if (on) {st}' if G.masked else f'          {st}')
        L.append('        }')
        if gr_in_butterfly:
            L.append(f'        dyv[{S}] = gr;')
        if nv_smem:
            L.append(f'        sevenn_dy_accumulate_smem<{nv_smem}>(dyv, lane, ring + {tab_data_floats(spec, True)}, dYu + (size_t)e * {SD});')
        else:
            L.append(f'        sevenn_dy_accumulate<{NP}>(dyv, lane, dYu + (size_t)e * {SD}, {SD if gr_in_butterfly else S});')
        if tab and not gr_in_butterfly:
            L.append(f'        sevenn_gr_accumulate(gr, lane, dYu + (size_t)e * {SD} + {S});')
        L += getattr(G, 'loop_tail', [])
        L.append('      }')
        L.append('    }')
        L.append('  }')
    L.append('UNIT_LOOP_END')
    return L


# ---------------------------------------------------------------------------------------------------------
# Tabulated radial weights.  The radial MLP output w[e, :] is a smooth function of the edge length alone, so
# for fixed model weights it is tabulated once (radial_table.py: fp64, cubic Hermite nodes every h = rc / n_int)
# and the convolution interpolates it per edge from an L2-resident table instead of reading a [E, wn] array
# from HBM.  Table layout: tab[i][c] = float4(f_i, h f'_i, f_{i+1}, h f'_{i+1}) for interval i, column c.
# Per edge the kernels read Yt[e] = (Y_0 .. Y_{S-1}, t, bits(i)) (edge.cu: sevenn_edge_fwd_tab).  The backward
# contracts dE/dw with dw/dr in registers and emits one scalar per (unit, edge) next to dY -- no dW array, no
# reverse radial MLP.
# edge-loop unroll of the tabulated kernels by signature: helps SevenNet-l3i5's inner layers (cfg3 32.1 -> 31.6 ms per
# step), hurts SevenNet-0's (cfg2 3.75 -> 4.10 ms: the doubled loop body falls out of the instruction cache)
TAB_UNROLL_BY_SIG = {'c703e622066': 2}
_TAB_UNROLL_ENV = _os0.environ.get('SEVENN_TAB_UNROLL')
TAB_XY_DEPTH = int(_os0.environ.get('SEVENN_TAB_XY_DEPTH', '4'))     # ring slots for x_j / Y (depth - 1 edges in flight); power of two
TAB_W_DEPTH = 2      # ring slots for the table rows (1 edge in flight: they come from L2)


def _tab_ypad(spec: ConvSpec) -> int:
    S = (spec.lmax_sh + 1) ** 2
    return ((S + 1 + 3) // 4) * 4


def _tab_xy_floats(G: _Group, spec: ConvSpec) -> int:
    return G.d1 * 32 * G.cpw + _tab_ypad(spec)


def _tab_w_floats(G: _Group) -> int:
    return len(G.paths) * 128 * G.cpw


# Where the table rows of the next edge wait: 'smem' = cp.async into the ring (rows cross the L1 data pipe twice:
# LDGSTS write + LDS read), 'regs' = 16-byte __ldg straight into registers one edge ahead (once).  (forward, backward)
# by signature; measured with tools/gpu.sh sweep-tabmode on B200 (profiles/r01e_table_staging_sweep.log): SevenNet-0 inner
# forward 1.042 -> 0.804 ms, backward 1.914 -> 1.864 ms; l3i5 inner forward 8.22 -> 7.97 ms, backward slower in
# registers (18.6 -> 19.2 ms: 128 registers are already spilling there), first / last layers 5-25 % faster both ways.
import os as _os
TAB_T_DEFAULT = ('regs', 'regs')
TAB_T_MODE = {
    'c703e622066': ('regs', 'smem'),
}
_TAB_T_ENV = _os.environ.get('SEVENN_TAB_T')          # tuning runs: 'fwd,bwd' for every signature
_cur_sig = [None]


def tab_t_regs(bwd: bool) -> bool:
    mode = tuple(_TAB_T_ENV.split(',')) if _TAB_T_ENV else TAB_T_MODE.get(_cur_sig[0], TAB_T_DEFAULT)
    return mode[1 if bwd else 0] == 'regs'


# dY lane reduction through shared memory instead of the shuffle butterfly (tabulated backward).  Fewer instructions
# on paper (37 against 68 per edge and unit) but measured SLOWER on B200 (SevenNet-0 inner backward 1.68 -> 1.78 ms:
# the store -> __syncwarp -> load round trip sits on the critical path of every edge); kept as an option, off.
DY_SMEM = _os.environ.get('SEVENN_DY_SMEM', '0') == '1'


def dy_smem_values(spec: ConvSpec) -> int:
    """Values per edge that go through the shared-memory reduction of the tabulated backward (0 = butterfly)."""
    S = (spec.lmax_sh + 1) ** 2
    nv = S + 1 if 32 // (S + 1) >= 2 else S      # the radial term rides along when it fits
    return nv if DY_SMEM and 32 // nv >= 2 else 0


# Two EDGES per loop trip in the tabulated backward: everything between the loads and the dY butterfly
# runs on float2 = (edge e, edge e + 1) with fma.rn.f32x2 -- M, q, dx, dE/dw, w x and the dY products cost one packed
# instruction for both edges; the output gradient g and the Clebsch-Gordan literals are free broadcast / immediate
# operands, and the backward keeps no accumulator across edges, so nothing but per-edge temporaries doubles.
# (forward, backward) by signature; SEVENN_TAB_EP='fwd,bwd' overrides for A/B builds.
# MEASURED (B200, tools/gpu.sh ab, profiles/r02b_edge_pair_ab.txt): SLOWER.  SevenNet-0 inner backward 184 registers ->
# 2 CTAs of 4 warps per SM: 39.9 -> 61.3 ms on the 216 k-atom cell (51.6 ms capped at 168 registers / 3 CTAs), although the
# kernel executes ~35 % fewer instructions -- with 8-12 warps per SM the dependent chains of one warp (loads -> M -> q ->
# dx / dY -> butterfly) are no longer covered; the scalar kernel at 120 registers / 16 warps is latency AND issue bound.
# Kept as an opt-in generator path (SEVENN_TAB_EP=0,1), off by default.
TAB_EP_DEFAULT = (False, False)
# Radial weights carried across the edge loop (regs table mode, monomial rows): w (and h dw/dr) of edge e + 1 are
# evaluated at the END of trip e from the rows fetched at its top, so the 4-register rows never cross the back edge --
# the (T = Tn) copies were 24 of the 196 SASS instructions of the l1 = 1 forward loop.  (forward, backward);
# SEVENN_TAB_WCARRY='f,b' for A/B builds.
# Same-box A/B on B200 (tools/gpu.sh ab, profiles/r02b_wcarry_ab.txt): it pays where the loop is short -- the LAST layers
# (scalar outputs only: SevenNet-0 backward 4.74 -> 3.85 ms, forward 2.53 -> 2.28 ms on 216 k atoms; l3i5 1.09 -> 0.83 /
# 0.54 -> 0.46 ms) and the l3i5 first-layer backward -- and not in the inner layers (SevenNet-0 forward 20.40 -> 20.51 ms
# although the loops lose 4-18 % of their instructions; l3i5 forward 6.16 -> 7.25 ms): there the rows fetched at the top
# of a trip are not back by its end.
TAB_WCARRY_DEFAULT = (False, False)
TAB_WCARRY_BY_SIG = {
    'cddfea4828d': (True, True),     # SevenNet-0 last layer
    'cc7e5576d9a': (True, True),     # SevenNet-l3i5 last layer
    'c0845e94934': (False, True),    # SevenNet-l3i5 first layer
}
_TAB_WCARRY_ENV = _os.environ.get('SEVENN_TAB_WCARRY')


def tab_wcarry(bwd: bool) -> bool:
    mode = (tuple(v == '1' for v in _TAB_WCARRY_ENV.split(',')) if _TAB_WCARRY_ENV
            else TAB_WCARRY_BY_SIG.get(_cur_sig[0], TAB_WCARRY_DEFAULT))
    return bool(mode[1 if bwd else 0]) and tab_t_regs(bwd) and TAB_POLY and not (bwd and tab_ep(True))
TAB_EP_BY_SIG = {}
_TAB_EP_ENV = _os.environ.get('SEVENN_TAB_EP')
TAB_EP_XY_DEPTH = 8   # ring slots for x_j / Y in edge-pair mode (three pairs in flight)


def tab_ep(bwd: bool) -> bool:
    mode = tuple(v == '1' for v in _TAB_EP_ENV.split(',')) if _TAB_EP_ENV else TAB_EP_BY_SIG.get(_cur_sig[0], TAB_EP_DEFAULT)
    return bool(mode[1 if bwd else 0]) and bwd and tab_t_regs(bwd) and TAB_POLY and not DY_SMEM


def tab_data_floats(spec: ConvSpec, bwd: bool) -> int:
    wd = 0 if tab_t_regs(bwd) else TAB_W_DEPTH
    dx = TAB_EP_XY_DEPTH if tab_ep(bwd) else TAB_XY_DEPTH
    return max(dx * _tab_xy_floats(G, spec) + wd * _tab_w_floats(G) for G in make_groups(spec, tab=True))


def tab_ring_floats(spec: ConvSpec, bwd: bool = False) -> int:
    scratch = (dy_smem_values(spec) * 33 + 3) // 4 * 4 if bwd else 0
    return tab_data_floats(spec, bwd) + scratch


def _tab_issue_xy(L, G: _Group, spec: ConvSpec, ind, e, j, slot):
    S = (spec.lmax_sh + 1) ** 2
    L.append(f'{ind}{{ float* sd = ring + ({slot}) * {_tab_xy_floats(G, spec)} + lane;')
    L.append(f'{ind}  const float* xr = x + (size_t)({j}) * ldx + {G.xoff} + c32 + lane;')
    if G.masked:
        L.append(f'{ind}  if (on) {{')
    for m in range(G.d1):
        for c in range(G.cpw):
            L.append(f'{ind}  CP_ASYNC4(sd + {(m * G.cpw + c) * 32}, xr + {m * G.mul + c * 32});')
    if G.masked:
        L.append(f'{ind}  }}')
    L.append(f'{ind}  if (lane < {S + 1}) CP_ASYNC4(sd + {G.d1 * 32 * G.cpw}, Yt + (size_t)({e}) * {S + 2} + lane);')
    L.append(f'{ind}}}')


def _tab_issue_w(L, G: _Group, spec: ConvSpec, ind, idx, slot):
    base = TAB_XY_DEPTH * _tab_xy_floats(G, spec)
    L.append(f'{ind}{{ float* sd = ring + {base} + ({slot}) * {_tab_w_floats(G)} + lane * 4;')
    L.append(f'{ind}  const float4* tr = tab + (size_t)({idx}) * {spec.wn} + c32 + lane;')
    if G.masked:
        L.append(f'{ind}  if (on) {{')
    for q, (pi, p) in enumerate(G.paths):
        for c in range(G.cpw):
            L.append(f'{ind}  CP_ASYNC16(sd + {(q * G.cpw + c) * 128}, tr + {p.w_off + c * 32});')
    if G.masked:
        L.append(f'{ind}  }}')
    L.append(f'{ind}}}')


def _tab_pipeline_smem(L, G: _Group, spec: ConvSpec, idx_of, xyf, wf, wbase):
    DX = TAB_XY_DEPTH
    # prologue: X(e0), X(e0+1), T(e0), X(e0+2)
    for d in (0, 1):
        L.append(f'      if (e0 + {d} < e1) {{')
        _tab_issue_xy(L, G, spec, '        ', f'e0 + {d}', f'col[e0 + {d}]', str(d))
        L.append('      }')
        L.append('      CP_ASYNC_COMMIT();')
    _tab_issue_w(L, G, spec, '      ', idx_of('e0'), '0')
    L.append('      CP_ASYNC_COMMIT();')
    L.append('      if (e0 + 2 < e1) {')
    _tab_issue_xy(L, G, spec, '        ', 'e0 + 2', 'col[e0 + 2]', '2')
    L.append('      }')
    L.append('      CP_ASYNC_COMMIT();')
    L.append(f'      int jn = col[e0 + {DX - 1} < e1 ? e0 + {DX - 1} : e1 - 1];')
    L.append(f'      int in = {idx_of("(e0 + 1 < e1 ? e0 + 1 : e1 - 1)")};')
    L.append('      int slot = 0;')
    L.append('      for (int e = e0; e < e1; ++e) {')
    L.append('        CP_ASYNC_WAIT(1);')
    L.append('        __syncwarp();')
    L.append('        if (e + 1 < e1) {')
    _tab_issue_w(L, G, spec, '          ', 'in', '(e + 1 - e0) & 1')
    L.append('        }')
    L.append('        CP_ASYNC_COMMIT();')
    L.append(f'        if (e + {DX - 1} < e1) {{')
    _tab_issue_xy(L, G, spec, '          ', f'e + {DX - 1}', 'jn', f'(slot + {DX - 1}) & {DX - 1}')
    L.append('        }')
    L.append('        CP_ASYNC_COMMIT();')
    L.append(f'        jn = col[e + {DX} < e1 ? e + {DX} : e1 - 1];')
    L.append(f'        in = {idx_of("(e + 2 < e1 ? e + 2 : e1 - 1)")};')
    L.append(f'        const float* sr = ring + slot * {xyf};')
    L.append(f'        const float* tw = ring + {wbase} + ((e - e0) & 1) * {wf} + lane * 4;')
    L.append(f'        slot = (slot + 1) & {DX - 1};')
    for m in range(G.d1):
        for c in range(G.cpw):
            L.append(f'        const float x{m}{_sx(G, c)} = {_on(G, f"sr[{(m * G.cpw + c) * 32} + lane]")};')
    for q, (pi, p) in enumerate(G.paths):
        z = 'make_float4(0.f, 0.f, 0.f, 0.f)'
        for c in range(G.cpw):
            L.append(f'        const float4 T{pi}{_sx(G, c)} = '
                     f'{_on(G, f"*reinterpret_cast<const float4*>(tw + {(q * G.cpw + c) * 128})", z)};')


def _tab_pipeline_regs(L, G: _Group, spec: ConvSpec, idx_of, xyf):
    """x_j / Y through the cp.async ring (3 edges ahead); the table rows of edge e+1 are fetched with 16-byte
    read-only loads into registers while edge e is being processed (they come from L2, one edge of latency cover)."""
    DX = TAB_XY_DEPTH
    z = 'make_float4(0.f, 0.f, 0.f, 0.f)'
    for d in range(DX - 1):
        L.append(f'      if (e0 + {d} < e1) {{')
        _tab_issue_xy(L, G, spec, '        ', f'e0 + {d}', f'col[e0 + {d}]', str(d))
        L.append('      }')
        L.append('      CP_ASYNC_COMMIT();')
    L.append(f'      const float4* tr0 = tab + (size_t)({idx_of("e0")}) * {spec.wn} + c32 + lane;')
    for pi, p in G.paths:
        for c in range(G.cpw):
            L.append(f'      float4 Tn{pi}{_sx(G, c)} = {_on(G, f"__ldg(tr0 + {p.w_off + c * 32})", z)};')
    L.append(f'      int jn = col[e0 + {DX - 1} < e1 ? e0 + {DX - 1} : e1 - 1];')
    L.append(f'      int in = {idx_of("(e0 + 1 < e1 ? e0 + 1 : e1 - 1)")};')
    L.append('      int slot = 0;')
    TAB_UNROLL = int(_TAB_UNROLL_ENV) if _TAB_UNROLL_ENV else TAB_UNROLL_BY_SIG.get(_cur_sig[0], 1)
    if TAB_UNROLL > 1:
        # two edges per trip: ptxas renames the (T, Tn) register sets between the copies instead of moving 24 registers
        # per edge (18 of the 139 instructions of the scalar-group forward loop were MOVs)
        L.append(f'#pragma unroll {TAB_UNROLL}')
    L.append('      for (int e = e0; e < e1; ++e) {')
    for pi, p in G.paths:
        for c in range(G.cpw):
            L.append(f'        const float4 T{pi}{_sx(G, c)} = Tn{pi}{_sx(G, c)};')
    L.append(f'        {{ const float4* tr = tab + (size_t)(in) * {spec.wn} + c32 + lane;')
    for pi, p in G.paths:
        for c in range(G.cpw):
            L.append(f'          Tn{pi}{_sx(G, c)} = {_on(G, f"__ldg(tr + {p.w_off + c * 32})", z)};')
    L.append('        }')
    L.append(f'        CP_ASYNC_WAIT({DX - 2});')
    L.append('        __syncwarp();')
    L.append(f'        if (e + {DX - 1} < e1) {{')
    _tab_issue_xy(L, G, spec, '          ', f'e + {DX - 1}', 'jn', f'(slot + {DX - 1}) & {DX - 1}')
    L.append('        }')
    L.append('        CP_ASYNC_COMMIT();')
    L.append(f'        jn = col[e + {DX} < e1 ? e + {DX} : e1 - 1];')
    L.append(f'        in = {idx_of("(e + 2 < e1 ? e + 2 : e1 - 1)")};')
    L.append(f'        const float* sr = ring + slot * {xyf};')
    L.append(f'        slot = (slot + 1) & {DX - 1};')
    for m in range(G.d1):
        for c in range(G.cpw):
            L.append(f'        const float x{m}{_sx(G, c)} = {_on(G, f"sr[{(m * G.cpw + c) * 32} + lane]")};')


def _poly_lines(G: _Group, bwd: bool, tpre: str, t: str, decl: str, ind: str) -> List[str]:
    """w (and wd = h dw/dr when ``bwd``) of every (path, chunk) from monomial table rows named ``tpre<pi><sx>`` at ``t``."""
    out = []
    scoped = decl == ''                       # assignments to carried variables: the temporaries get their own scope
    if bwd:
        out.append(f'{ind}{"{ " if scoped else ""}const float t2_ = {t} + {t}, t3_ = 3.f * {t} * {t};')
    for pi, p in G.paths:
        for c in range(G.cpw):
            T = f'{tpre}{pi}{_sx(G, c)}'
            out.append(f'{ind}{decl}w{pi}{_sx(G, c)} = fmaf(fmaf(fmaf({T}.w, {t}, {T}.z), {t}, {T}.y), {t}, {T}.x);')
            if bwd:
                out.append(f'{ind}{decl}wd{pi}{_sx(G, c)} = fmaf(t3_, {T}.w, fmaf(t2_, {T}.z, {T}.y));')
    if bwd and scoped:
        out.append(f'{ind}}}')
    return out


def _tab_pipeline_regs_carry(L, G: _Group, spec: ConvSpec, idx_of, xyf, bwd: bool):
    """Like _tab_pipeline_regs, but the radial weights are what crosses the back edge (TAB_WCARRY): the rows of edge e + 1
    are fetched at the top of trip e together with its t (read straight from Yt) and evaluated in G.loop_tail."""
    S = (spec.lmax_sh + 1) ** 2
    S2 = S + 2
    DX = TAB_XY_DEPTH
    z = 'make_float4(0.f, 0.f, 0.f, 0.f)'
    names = [f'{v}{pi}{_sx(G, c)}' for pi, p in G.paths for c in range(G.cpw) for v in (('w', 'wd') if bwd else ('w',))]
    for d in range(DX - 1):
        L.append(f'      if (e0 + {d} < e1) {{')
        _tab_issue_xy(L, G, spec, '        ', f'e0 + {d}', f'col[e0 + {d}]', str(d))
        L.append('      }')
        L.append('      CP_ASYNC_COMMIT();')
    L.append('      float ' + ', '.join(names) + ';')
    L.append(f'      {{ const float4* tr0 = tab + (size_t)({idx_of("e0")}) * {spec.wn} + c32 + lane;')
    L.append(f'        const float t0_ = LDG(Yt + (size_t)e0 * {S2} + {S});')
    for pi, p in G.paths:
        for c in range(G.cpw):
            L.append(f'        const float4 T0_{pi}{_sx(G, c)} = {_on(G, f"__ldg(tr0 + {p.w_off + c * 32})", z)};')
    L += _poly_lines(G, bwd, 'T0_', 't0_', '', '        ')
    L.append('      }')
    L.append(f'      int jn = col[e0 + {DX - 1} < e1 ? e0 + {DX - 1} : e1 - 1];')
    L.append(f'      int in = {idx_of("(e0 + 1 < e1 ? e0 + 1 : e1 - 1)")};')
    L.append('      int slot = 0;')
    L.append('      for (int e = e0; e < e1; ++e) {')
    L.append(f'        const float4* tr = tab + (size_t)(in) * {spec.wn} + c32 + lane;')
    L.append(f'        const float tn_ = LDG(Yt + (size_t)(e + 1 < e1 ? e + 1 : e1 - 1) * {S2} + {S});')
    for pi, p in G.paths:
        for c in range(G.cpw):
            L.append(f'        const float4 Tn{pi}{_sx(G, c)} = {_on(G, f"__ldg(tr + {p.w_off + c * 32})", z)};')
    L.append(f'        CP_ASYNC_WAIT({DX - 2});')
    L.append('        __syncwarp();')
    L.append(f'        if (e + {DX - 1} < e1) {{')
    _tab_issue_xy(L, G, spec, '          ', f'e + {DX - 1}', 'jn', f'(slot + {DX - 1}) & {DX - 1}')
    L.append('        }')
    L.append('        CP_ASYNC_COMMIT();')
    L.append(f'        jn = col[e + {DX} < e1 ? e + {DX} : e1 - 1];')
    L.append(f'        in = {idx_of("(e + 2 < e1 ? e + 2 : e1 - 1)")};')
    L.append(f'        const float* sr = ring + slot * {xyf};')
    L.append(f'        slot = (slot + 1) & {DX - 1};')
    for m in range(G.d1):
        for c in range(G.cpw):
            L.append(f'        const float x{m}{_sx(G, c)} = {_on(G, f"sr[{(m * G.cpw + c) * 32} + lane]")};')
    G.loop_tail = ['#ifndef SEVENN_EMU'] + _poly_lines(G, bwd, 'Tn', 'tn_', '', '        ') + ['#endif']


def _tab_pipeline_open(L, G: _Group, spec: ConvSpec, bwd: bool):
    """Opens ``if (e0 < e1) { for (e) {`` with x*, y*, w* (and wd* = h dw/dr when ``bwd``) of edge e defined.
    cp.async groups alternate  ... X(e+2) | T(e+1) | X(e+3) | T(e+2) ...  so that ``wait_group 1`` at the top of
    iteration e leaves only the newest x/Y stage in flight."""
    S = (spec.lmax_sh + 1) ** 2
    S2 = S + 2
    DX = TAB_XY_DEPTH
    xyf, wf = _tab_xy_floats(G, spec), _tab_w_floats(G)
    wbase = DX * xyf
    idx_of = lambda e: f'F2I(LDG(Yt + (size_t)({e}) * {S2} + {S + 1}))'   # noqa: E731
    L.append('    const int e0 = rowptr[row], e1 = rowptr[row + 1];')
    L.append('    if (e0 < e1) {')
    L.append('#ifdef SEVENN_EMU')
    L.append('      for (int e = e0; e < e1; ++e) {')
    L.append(f'        const float* xr = x + (size_t)col[e] * ldx + {G.xoff} + c32 + lane;')
    L.append(f'        const float4* tr = tab + (size_t){idx_of("e")} * {spec.wn} + c32 + lane;')
    L.append(f'        const float t = Yt[(size_t)e * {S2} + {S}];')
    for m in range(G.d1):
        for c in range(G.cpw):
            L.append(f'        const float x{m}{_sx(G, c)} = {_on(G, f"xr[{m * G.mul + c * 32}]")};')
    for pi, p in G.paths:
        z = 'float4{0.f, 0.f, 0.f, 0.f}'
        for c in range(G.cpw):
            L.append(f'        const float4 T{pi}{_sx(G, c)} = {_on(G, f"tr[{p.w_off + c * 32}]", z)};')
    for jj in G.ys:
        L.append(f'        const float y{jj} = Yt[(size_t)e * {S2} + {jj}];')
    G.loop_tail = []
    carry = tab_wcarry(bwd)
    if carry:
        L += _poly_lines(G, bwd, 'T', 't', 'const float ', '        ')
    L.append('#else')
    if carry:
        _tab_pipeline_regs_carry(L, G, spec, idx_of, xyf, bwd)
    elif tab_t_regs(bwd):
        _tab_pipeline_regs(L, G, spec, idx_of, xyf)
    else:
        _tab_pipeline_smem(L, G, spec, idx_of, xyf, wf, wbase)
    ybase = G.d1 * 32 * G.cpw
    for jj in G.ys:
        L.append(f'        const float y{jj} = sr[{ybase + jj}];')
    if not carry:
        L.append(f'        const float t = sr[{ybase + S}];')
    L.append('#endif')
    if carry:
        return
    # cubic Hermite basis (and its derivative) at t
    if TAB_POLY:
        for pi, p in G.paths:
            for c in range(G.cpw):
                T = f'T{pi}{_sx(G, c)}'
                L.append(f'        const float w{pi}{_sx(G, c)} = fmaf(fmaf(fmaf({T}.w, t, {T}.z), t, {T}.y), t, {T}.x);')
        if bwd:
            L.append('        const float t2 = t + t, t3 = 3.f * t * t;')
            for pi, p in G.paths:
                for c in range(G.cpw):
                    T = f'T{pi}{_sx(G, c)}'
                    L.append(f'        const float wd{pi}{_sx(G, c)} = fmaf(t3, {T}.w, fmaf(t2, {T}.z, {T}.y));')
        return
    L.append('        const float u = 1.f - t, tt = t * t, uu = u * u;')
    L.append('        const float h00 = (1.f + 2.f * t) * uu, h10 = t * uu, h01 = tt * (3.f - 2.f * t), h11 = -tt * u;')
    f2 = TAB_F2 and G.cpw == 2
    if f2:
        # packed table layout of a channel pair (pack_table): T_a = (f_i, f_i', h f'_i, h f'_i'), T_b = the same at node
        # i + 1 -- the halves of each 16-byte load are register pairs, the Hermite basis a free broadcast operand
        for pi, p in G.paths:
            A, B = f'T{pi}_a', f'T{pi}_b'
            L.append(f'        const float2 w{pi} = __ffma2_rn(make_float2({A}.x, {A}.y), F2B(h00), '
                     f'__ffma2_rn(make_float2({A}.z, {A}.w), F2B(h10), __ffma2_rn(make_float2({B}.x, {B}.y), F2B(h01), '
                     f'__fmul2_rn(make_float2({B}.z, {B}.w), F2B(h11)))));')
    else:
        for pi, p in G.paths:
            for c in range(G.cpw):
                T = f'T{pi}{_sx(G, c)}'
                L.append(f'        const float w{pi}{_sx(G, c)} = fmaf({T}.x, h00, fmaf({T}.y, h10, fmaf({T}.z, h01, {T}.w * h11)));')
    if bwd:
        L.append('        const float d00 = -6.f * t * u, d10 = u * (1.f - 3.f * t), d11 = t * (3.f * t - 2.f);')
        if f2:
            L.append('        const float nd00 = -d00;')
        for pi, p in G.paths:
            if f2:
                A, B = f'T{pi}_a', f'T{pi}_b'
                L.append(f'        const float2 wd{pi} = __ffma2_rn(make_float2({A}.x, {A}.y), F2B(d00), '
                         f'__ffma2_rn(make_float2({B}.x, {B}.y), F2B(nd00), __ffma2_rn(make_float2({A}.z, {A}.w), F2B(d10), '
                         f'__fmul2_rn(make_float2({B}.z, {B}.w), F2B(d11)))));')
                continue
            for c in range(G.cpw):
                T = f'T{pi}{_sx(G, c)}'
                L.append(f'        const float wd{pi}{_sx(G, c)} = fmaf(d00, {T}.x - {T}.z, fmaf(d10, {T}.y, d11 * {T}.w));')


def _f2(v: float) -> str:
    return f'F2B({_f(v)})'


def gen_bwd_tab_pair(spec: ConvSpec, groups: List[_Group], sig: str) -> List[str]:
    """Tabulated backward with two edges per loop trip (see TAB_EP above).  Same kernel name, arguments and outputs as
    gen_bwd(tab=True); per-edge results differ from it in the last bits only (explicit packed fma forms)."""
    S = (spec.lmax_sh + 1) ** 2
    S2, SD = S + 2, S + 1
    NP = 1
    while NP < S:
        NP *= 2
    gr_in_butterfly = NP > S
    DX = TAB_EP_XY_DEPTH
    z4 = 'make_float4(0.f, 0.f, 0.f, 0.f)'
    idx_of = lambda e: f'F2I(LDG(Yt + (size_t)({e}) * {S2} + {S + 1}))'   # noqa: E731
    L = []
    mc = MIN_CTAS_TAB_EP.get(sig, MIN_CTAS_TAB_EP_DEFAULT)
    L.append(f'KERNEL_HEAD(conv_bwd_tab_{sig}, {mc})(const float* __restrict__ x, int ldx, const float* __restrict__ Yt,')
    L.append('    const float4* __restrict__ tab, const int* __restrict__ rowptr, const int* __restrict__ col,')
    L.append('    int n_dst, float inv_denom, const float* __restrict__ gout, int ldo,')
    L.append('    float* __restrict__ dxe, int lddx, float* __restrict__ dY, long long dy_stride)')
    L.append(f'UNIT_LOOP_BEGIN_DYN(n_dst, {sum(G.nchunk // G.cpw for G in groups)}, {tab_ring_floats(spec, True)})')
    L.append('  float* dYu = dY + (size_t)unit * (size_t)dy_stride;')
    for gi, G in enumerate(groups):
        xyf = _tab_xy_floats(G, spec)
        ybase = G.d1 * 32 * G.cpw
        chunks = [_sx(G, c) for c in range(G.cpw)]
        _unit_open(L, G, gi == 0, gi == len(groups) - 1)
        L.append('    const float* grow = gout + (size_t)row * ldo + c32 + lane;')
        for pi, p in G.paths:
            Mg = spec.lmid.groups[p.gm][0]
            base = spec.lmid.group_off[p.gm] + p.u_off
            for k in range(2 * p.l3 + 1):
                for c in range(G.cpw):
                    ld = _on(G, f'LDG(grow + {base + k * Mg + c * 32}) * inv_denom')
                    L.append(f'    const float g{pi}_{k}{_sx(G, c)} = {ld};')
        L.append('    const int e0 = rowptr[row], e1 = rowptr[row + 1];')
        L.append('    if (e0 < e1) {')
        # ---- emulation build: plain loads
        L.append('#ifdef SEVENN_EMU')
        L.append('      for (int e = e0; e < e1; e += 2) {')
        L.append('        const bool vB = e + 1 < e1; const int eB = vB ? e + 1 : e;')
        L.append(f'        const float* xrA = x + (size_t)col[e] * ldx + {G.xoff} + c32 + lane;')
        L.append(f'        const float* xrB = x + (size_t)col[eB] * ldx + {G.xoff} + c32 + lane;')
        L.append(f'        const float4* trA = tab + (size_t){idx_of("e")} * {spec.wn} + c32 + lane;')
        L.append(f'        const float4* trB = tab + (size_t){idx_of("eB")} * {spec.wn} + c32 + lane;')
        L.append(f'        const float tA = Yt[(size_t)e * {S2} + {S}], tB = Yt[(size_t)eB * {S2} + {S}];')
        for m in range(G.d1):
            for c in range(G.cpw):
                o = m * G.mul + c * 32
                a_, b_ = _on(G, f'xrA[{o}]'), _on(G, f'xrB[{o}]')
                L.append(f'        const float2 x{m}{_sx(G, c)} = make_float2({a_}, {b_});')
        for pi, p in G.paths:
            ze = 'float4{0.f, 0.f, 0.f, 0.f}'
            for c in range(G.cpw):
                o = p.w_off + c * 32
                L.append(f'        const float4 TA{pi}{_sx(G, c)} = {_on(G, f"trA[{o}]", ze)}, TB{pi}{_sx(G, c)} = {_on(G, f"trB[{o}]", ze)};')
        for jj in G.ys:
            L.append(f'        const float2 y{jj} = make_float2(Yt[(size_t)e * {S2} + {jj}], Yt[(size_t)eB * {S2} + {jj}]);')
        # ---- device: cp.async ring for x_j / Y (three pairs ahead), table rows of the next pair in registers
        L.append('#else')
        for d in range(DX // 2 - 1):
            for h in range(2):
                L.append(f'      if (e0 + {2 * d + h} < e1) {{')
                _tab_issue_xy(L, G, spec, '        ', f'e0 + {2 * d + h}', f'col[e0 + {2 * d + h}]', str(2 * d + h))
                L.append('      }')
            L.append('      CP_ASYNC_COMMIT();')
        L.append(f'      const float4* trA0 = tab + (size_t)({idx_of("e0")}) * {spec.wn} + c32 + lane;')
        L.append(f'      const float4* trB0 = tab + (size_t)({idx_of("(e0 + 1 < e1 ? e0 + 1 : e1 - 1)")}) * {spec.wn} + c32 + lane;')
        for pi, p in G.paths:
            for c in range(G.cpw):
                o = p.w_off + c * 32
                L.append(f'      float4 TnA{pi}{_sx(G, c)} = {_on(G, f"__ldg(trA0 + {o})", z4)}, '
                         f'TnB{pi}{_sx(G, c)} = {_on(G, f"__ldg(trB0 + {o})", z4)};')
        L.append(f'      int jnA = col[e0 + {DX - 2} < e1 ? e0 + {DX - 2} : e1 - 1], jnB = col[e0 + {DX - 1} < e1 ? e0 + {DX - 1} : e1 - 1];')
        L.append(f'      int inA = {idx_of("(e0 + 2 < e1 ? e0 + 2 : e1 - 1)")}, inB = {idx_of("(e0 + 3 < e1 ? e0 + 3 : e1 - 1)")};')
        L.append('      int slot = 0;')
        L.append('      for (int e = e0; e < e1; e += 2) {')
        L.append('        const bool vB = e + 1 < e1;')
        for pi, p in G.paths:
            for c in range(G.cpw):
                L.append(f'        const float4 TA{pi}{_sx(G, c)} = TnA{pi}{_sx(G, c)}, TB{pi}{_sx(G, c)} = TnB{pi}{_sx(G, c)};')
        L.append(f'        {{ const float4* trA = tab + (size_t)(inA) * {spec.wn} + c32 + lane;')
        L.append(f'          const float4* trB = tab + (size_t)(inB) * {spec.wn} + c32 + lane;')
        for pi, p in G.paths:
            for c in range(G.cpw):
                o = p.w_off + c * 32
                L.append(f'          TnA{pi}{_sx(G, c)} = {_on(G, f"__ldg(trA + {o})", z4)}; TnB{pi}{_sx(G, c)} = {_on(G, f"__ldg(trB + {o})", z4)};')
        L.append('        }')
        L.append(f'        CP_ASYNC_WAIT({DX // 2 - 2});')
        L.append('        __syncwarp();')
        for h, jn in ((0, 'jnA'), (1, 'jnB')):
            L.append(f'        if (e + {DX - 2 + h} < e1) {{')
            _tab_issue_xy(L, G, spec, '          ', f'e + {DX - 2 + h}', jn, f'(slot + {DX - 2 + h}) & {DX - 1}')
            L.append('        }')
        L.append('        CP_ASYNC_COMMIT();')
        L.append(f'        jnA = col[e + {DX} < e1 ? e + {DX} : e1 - 1]; jnB = col[e + {DX + 1} < e1 ? e + {DX + 1} : e1 - 1];')
        L.append(f'        inA = {idx_of("(e + 4 < e1 ? e + 4 : e1 - 1)")}; inB = {idx_of("(e + 5 < e1 ? e + 5 : e1 - 1)")};')
        L.append(f'        const float* srA = ring + slot * {xyf};')
        L.append(f'        const float* srB = ring + (vB ? ((slot + 1) & {DX - 1}) : slot) * {xyf};')
        L.append(f'        slot = (slot + 2) & {DX - 1};')
        for m in range(G.d1):
            for c in range(G.cpw):
                o = (m * G.cpw + c) * 32
                a_, b_ = _on(G, f'srA[{o} + lane]'), _on(G, f'srB[{o} + lane]')
                L.append(f'        const float2 x{m}{_sx(G, c)} = make_float2({a_}, {b_});')
        for jj in G.ys:
            L.append(f'        const float2 y{jj} = make_float2(srA[{ybase + jj}], srB[{ybase + jj}]);')
        L.append(f'        const float tA = srA[{ybase + S}], tB = srB[{ybase + S}];')
        L.append('#endif')
        # ---- radial weights and their derivative (monomial table rows), scalar per edge, paired afterwards
        L.append('        const float t2A = tA + tA, t3A = 3.f * tA * tA, t2B = tB + tB, t3B = 3.f * tB * tB;')
        for pi, p in G.paths:
            for sx in chunks:
                A, B = f'TA{pi}{sx}', f'TB{pi}{sx}'
                L.append(f'        const float2 w{pi}{sx} = make_float2(fmaf(fmaf(fmaf({A}.w, tA, {A}.z), tA, {A}.y), tA, {A}.x), '
                         f'fmaf(fmaf(fmaf({B}.w, tB, {B}.z), tB, {B}.y), tB, {B}.x));')
                L.append(f'        const float2 wd{pi}{sx} = make_float2(fmaf(t3A, {A}.w, fmaf(t2A, {A}.z, {A}.y)), '
                         f'fmaf(t3B, {B}.w, fmaf(t2B, {B}.z, {B}.y)));')
        zero2 = 'make_float2(0.f, 0.f)'
        L.append(f'        float2 dyv[{NP}];')
        L.append(f'#pragma unroll\n        for (int q_ = 0; q_ < {NP}; ++q_) dyv[q_] = {zero2};')
        L.append('        float2 ' + ', '.join(f'dx{m}{sx} = {zero2}' for m in range(G.d1) for sx in chunks) + f', gr2 = {zero2};')
        touched = set()
        for pi, p in G.paths:
            ent, nnz = _m_terms(p.l1, p.l2, p.l3)
            L.append(f'        {{ // path {pi}: {p.l1} x {p.l2} -> {p.l3}')
            for (i, k), terms in ent.items():
                j0, c0 = terms[0]
                expr = f'__fmul2_rn({_f2(c0)}, y{p.l2 * p.l2 + j0})'
                for j, c in terms[1:]:
                    expr = f'__ffma2_rn({_f2(c)}, y{p.l2 * p.l2 + j}, {expr})'
                L.append(f'          const float2 M_{i}_{k} = {expr};')
            rows = [i for i in range(G.d1) if any((i, k) in ent for k in range(2 * p.l3 + 1))]
            for sx in chunks:
                for i in rows:
                    kk = [k for k in range(2 * p.l3 + 1) if (i, k) in ent]
                    t = f'__fmul2_rn(M_{i}_{kk[0]}, F2B(g{pi}_{kk[0]}{sx}))'
                    for k in kk[1:]:
                        t = f'__ffma2_rn(M_{i}_{k}, F2B(g{pi}_{k}{sx}), {t})'
                    L.append(f'          const float2 q{i}{sx} = {t};')
                    L.append(f'          dx{i}{sx} = __ffma2_rn(w{pi}{sx}, q{i}{sx}, dx{i}{sx});')
                if rows:
                    t = f'__fmul2_rn(x{rows[0]}{sx}, q{rows[0]}{sx})'
                    for i in rows[1:]:
                        t = f'__ffma2_rn(x{i}{sx}, q{i}{sx}, {t})'
                    L.append(f'          gr2 = __ffma2_rn({t}, wd{pi}{sx}, gr2);')
                for i in rows:
                    L.append(f'          const float2 X{i}{sx} = __fmul2_rn(w{pi}{sx}, x{i}{sx});')
            # the chunks of a warp share c_ijk: their products are added before the coefficient is applied
            for (i, k) in ent:
                t = f'__fmul2_rn(X{i}{chunks[0]}, F2B(g{pi}_{k}{chunks[0]}))'
                for sx in chunks[1:]:
                    t = f'__ffma2_rn(X{i}{sx}, F2B(g{pi}_{k}{sx}), {t})'
                L.append(f'          const float2 P_{i}_{k} = {t};')
            for (i, j, k, cf) in nnz:
                q = p.l2 * p.l2 + j
                if q not in touched:
                    touched.add(q)
                    L.append(f'          dyv[{q}] = P_{i}_{k};' if abs(cf - 1.0) < 1e-12 else
                             f'          dyv[{q}] = __fmul2_rn({_f2(cf)}, P_{i}_{k});')
                else:
                    L.append(f'          dyv[{q}] = __ffma2_rn({_f2(cf)}, P_{i}_{k}, dyv[{q}]);')
            L.append('        }')
        # ---- per-edge outputs: dx_e rows, dY (+ radial term) through the butterfly, edge B only when it exists
        L.append('        if (dxe != nullptr) {')
        L.append(f'          float* dxrA = dxe + (size_t)e * lddx + {G.xoff} + c32 + lane;')
        L.append('          float* dxrB = dxrA + lddx;')
        for m in range(G.d1):
            for c in range(G.cpw):
                o = m * G.mul + c * 32
                sx = _sx(G, c)
                cond = 'on' if G.masked else None
                L.append(f'          {"if (on) " if cond else ""}dxrA[{o}] = dx{m}{sx}.x;')
                L.append(f'          if ({"on && " if cond else ""}vB) dxrB[{o}] = dx{m}{sx}.y;')
        L.append('        }')
        L.append(f'        float dyA[{NP}], dyB[{NP}];')
        L.append(f'#pragma unroll\n        for (int q_ = 0; q_ < {NP}; ++q_) {{ dyA[q_] = dyv[q_].x; dyB[q_] = dyv[q_].y; }}')
        if gr_in_butterfly:
            L.append(f'        dyA[{S}] = gr2.x; dyB[{S}] = gr2.y;')
        nacc = SD if gr_in_butterfly else S
        L.append(f'        sevenn_dy_accumulate<{NP}>(dyA, lane, dYu + (size_t)e * {SD}, {nacc});')
        L.append(f'        if (vB) sevenn_dy_accumulate<{NP}>(dyB, lane, dYu + (size_t)(e + 1) * {SD}, {nacc});')
        if not gr_in_butterfly:
            L.append(f'        sevenn_gr_accumulate(gr2.x, lane, dYu + (size_t)e * {SD} + {S});')
            L.append(f'        if (vB) sevenn_gr_accumulate(gr2.y, lane, dYu + (size_t)(e + 1) * {SD} + {S});')
        L.append('      }')
        L.append('    }')
        L.append('  }')
    L.append('UNIT_LOOP_END')
    return L


def pack_table(spec: ConvSpec, tab):
    """Table layout the packed (TAB_F2) kernels read: for every channel pair (c, c' = c + 32) of a two-chunk unit the
    two float4 (f_i, g_i, f_i+1, g_i+1)[c], [c'] become (f_i[c], f_i[c'], g_i[c], g_i[c']) at column c and the same
    at node i + 1 at column c'.  ``tab``: numpy [n_int, wn, 4]; returns a new array."""
    out = tab.copy()
    for G in make_groups(spec, tab=True):
        if G.cpw != 2:
            continue
        for pi, p in G.paths:
            for o in range(p.w_off, p.w_off + G.mul, 64):
                A, B = tab[:, o:o + 32, :], tab[:, o + 32:o + 64, :]
                out[:, o:o + 32, 0], out[:, o:o + 32, 1] = A[..., 0], B[..., 0]
                out[:, o:o + 32, 2], out[:, o:o + 32, 3] = A[..., 1], B[..., 1]
                out[:, o + 32:o + 64, 0], out[:, o + 32:o + 64, 1] = A[..., 2], B[..., 2]
                out[:, o + 32:o + 64, 2], out[:, o + 32:o + 64, 3] = A[..., 3], B[..., 3]
    return out


CONV_CARVEOUT = _os.environ.get('SEVENN_CONV_CARVEOUT', '1') == '1'


def _carveout_pct(ring_floats: int, min_ctas: int) -> int:
    """Preferred shared-memory carveout (percent of 228 KB) that just holds the rings of the CTAs the kernel is tuned for,
    so that the rest of the SM's array caches radial-table rows; -1 = driver default (kernels without an occupancy target:
    the first-layer backward lost 2 % with a guessed one).  Same-box A/B: 216 k-atom Si step 98.8 -> 98.2 ms."""
    if not CONV_CARVEOUT or min_ctas <= 1:
        return -1
    ctas = min_ctas
    need = ctas * (4 * ring_floats * 4 + 1024)
    return min(100, -(-100 * need // (228 * 1024)))


def signature(spec: ConvSpec) -> str:
    return spec.signature()


def generate(spec: ConvSpec) -> str:
    sig = signature(spec)
    _cur_sig[0] = sig
    groups = make_groups(spec)
    groups_tab = make_groups(spec, tab=True)
    nu = sum(G.nchunk for G in groups)
    nu_tab = sum(G.nchunk // G.cpw for G in groups_tab)
    out = ['// GENERATED by sevennet_exp_b200/codegen/conv_codegen.py -- do not edit.',
           f'// x: {ir.irreps_str(spec.lx.groups)}  mid: {ir.irreps_str(spec.lmid.groups)}  '
           f'lmax_sh {spec.lmax_sh}  paths {len(spec.paths)}  wn {spec.wn}  units {nu}',
           '#include "conv_common.cuh"', '']
    out += gen_fwd(spec, groups, sig) + ['']
    out += gen_bwd(spec, groups, sig) + ['']
    out += gen_fwd(spec, groups_tab, sig, tab=True) + ['']
    out += (gen_bwd_tab_pair(spec, groups_tab, sig) if tab_ep(True) else gen_bwd(spec, groups_tab, sig, tab=True)) + ['']
    rf, rb = tab_ring_floats(spec, False), tab_ring_floats(spec, True)
    mcf, mcb = MIN_CTAS_TAB.get(sig, MIN_CTAS_DEFAULT)
    cf, cb = _carveout_pct(rf, mcf), _carveout_pct(rb, mcb)
    out.append(f'extern "C" int sevenn_conv_fwd_tab_{sig}(const float* x, int ldx, const float* Yt, const float* tab,')
    out.append('    const int* rowptr, const int* col, int n_dst, float inv_denom, float* out, int ldo, void* stream) {')
    out.append('  if (n_dst <= 0) return 0;')
    out.append(f'  LAUNCH_DYN(conv_fwd_tab_{sig}, n_dst, {nu_tab}, {rf}, {cf}, stream)(x, ldx, Yt, (const float4*)tab, rowptr, col, n_dst, inv_denom, out, ldo);')
    out.append('  return LAUNCH_STATUS();')
    out.append('}')
    out.append(f'extern "C" int sevenn_conv_bwd_tab_{sig}(const float* x, int ldx, const float* Yt, const float* tab,')
    out.append('    const int* rowptr, const int* col, int n_dst, float inv_denom, const float* gout, int ldo,')
    out.append('    float* dxe, int lddx, float* dY, long long dy_stride, void* stream) {')
    out.append('  if (n_dst <= 0) return 0;')
    out.append(f'  LAUNCH_DYN(conv_bwd_tab_{sig}, n_dst, {nu_tab}, {rb}, {cb}, stream)(x, ldx, Yt, (const float4*)tab, rowptr, col, n_dst, inv_denom, gout, ldo, dxe, lddx, dY, dy_stride);')
    out.append('  return LAUNCH_STATUS();')
    out.append('}')
    out.append(f'extern "C" int sevenn_conv_fwd_{sig}(const float* x, int ldx, const float* Y, const float* w,')
    out.append('    const int* rowptr, const int* col, int n_dst, float inv_denom, float* out, int ldo, void* stream) {')
    out.append('  if (n_dst <= 0) return 0;')
    out.append(f'  LAUNCH(conv_fwd_{sig}, n_dst, {nu}, stream)(x, ldx, Y, w, rowptr, col, n_dst, inv_denom, out, ldo);')
    out.append('  return LAUNCH_STATUS();')
    out.append('}')
    out.append(f'extern "C" int sevenn_conv_bwd_{sig}(const float* x, int ldx, const float* Y, const float* w,')
    out.append('    const int* rowptr, const int* col, int n_dst, float inv_denom, const float* gout, int ldo,')
    out.append('    float* dW, float* dxe, int lddx, float* dY, long long dy_stride, void* stream) {')
    out.append('  if (n_dst <= 0) return 0;')
    out.append(f'  LAUNCH(conv_bwd_{sig}, n_dst, {nu}, stream)(x, ldx, Y, w, rowptr, col, n_dst, inv_denom, gout, ldo, dW, dxe, lddx, dY, dy_stride);')
    out.append('  return LAUNCH_STATUS();')
    out.append('}')
    out.append(f'extern "C" int sevenn_conv_info_{sig}(int* x_dim, int* mid_dim, int* wn, int* sh_dim, int* units) {{')
    out.append(f'  *x_dim = {spec.lx.dim}; *mid_dim = {spec.lmid.dim}; *wn = {spec.wn}; *sh_dim = {(spec.lmax_sh + 1) ** 2};')
    out.append(f'  *units = {nu}; return 0; }}')
    packed = 2 if TAB_POLY else int(TAB_F2 and any(G.cpw == 2 for G in groups_tab))     # table layout the kernels read
    out.append(f'// 1: the tabulated kernels of this unit read the packed channel-pair table layout (conv_codegen.pack_table)')
    out.append(f'extern "C" int sevenn_conv_tab_packed_{sig}() {{ return {packed}; }}')
    return '\n'.join(out) + '\n'
