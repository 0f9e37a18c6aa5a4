"""Whole-model launcher: energy, forces and virial of one atomic graph on one GPU.

Straight-line replacement of ``AtomGraphSequential.forward`` (sevenn/nn/sequential.py:178; call
stack SURVEY.md section 3.1) including the reverse pass that the reference gets from
``torch.autograd.grad(E, edge_vec)`` (sevenn/nn/force_output.py:173-178): every step is one of
the C-ABI launchers of include/sevenn_b200.h, no autograd, no e3nn, no CPU fallback.

Nodes are ``n_local`` owned atoms followed by ghosts (single GPU: no ghosts).  Rows of the
CSR edge list exist for owned atoms only; ``halo`` (domain.py) fills ghost rows of the
convolution input before each layer >= 1 and returns ghost gradients to their owners.
"""
import ctypes
import math
import os
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib
from .spec import ACT_ID, LinearSpec, ModelSpec, nequip_sc_dense
from . import irreps as ir
from . import radial_table
from .codegen import conv_codegen


class Graph:
    """Device-side graph: CSR by centre (edge_index[0]) + CSC by neighbour (edge_index[1])."""

    def __init__(self, types, edge_vec, rowptr, col, cptr, cedge, n_local, volume):
        self.types, self.edge_vec, self.rowptr, self.col = types, edge_vec, rowptr, col
        self.cptr, self.cedge = cptr, cedge
        self.n_local, self.n_nodes, self.n_edges = int(n_local), int(types.shape[0]), int(col.shape[0])
        self.volume = float(volume)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def tc_weight_image(W: np.ndarray, Nt: int, n_tiles: int, KB: int) -> np.ndarray:
    """[K, N] fp32 -> flat image [n_tiles][KB][2 (hi, lo)][Nt][32] laid out exactly as the UMMA K-major
    SWIZZLE_128B shared-memory tile (16-byte chunk index XOR row%8), so bulk copies fill the stages.
    (Nt, n_tiles, KB) come from sevenn_gemm_tc_plan so host and kernel agree on the geometry."""
    K, N = W.shape
    Wp = np.zeros((KB * 32, n_tiles * Nt), dtype=np.float32)
    Wp[:K, :N] = W
    hi = (Wp.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)
    lo = (Wp - hi).astype(np.float32)
    lo = (lo.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)
    img = np.zeros((n_tiles, KB, 2, Nt, 32), dtype=np.float32)
    n_idx = np.arange(Nt)[:, None]
    k_idx = np.arange(32)[None, :]
    phys = (((k_idx // 4) ^ (n_idx % 8)) * 4 + k_idx % 4)          # [Nt, 32] physical column of (n, k)
    for t in range(n_tiles):
        for kb in range(KB):
            for h, src in enumerate((hi, lo)):
                blk = src[kb * 32:(kb + 1) * 32, t * Nt:(t + 1) * Nt].T      # [Nt, 32] (n, k)
                np.put_along_axis(img[t, kb, h], phys, blk, axis=1)
    return img.reshape(-1)


def tc_image_for(lib, W: np.ndarray) -> np.ndarray:
    K, N = W.shape
    nt, tiles, kb, fl = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_longlong()
    lib.call('sevenn_gemm_tc_plan', N, K, ctypes.byref(nt), ctypes.byref(tiles), ctypes.byref(kb), ctypes.byref(fl))
    img = tc_weight_image(np.ascontiguousarray(W, dtype=np.float32), nt.value, tiles.value, kb.value)
    assert img.size == fl.value
    return img


class _Lin:
    """Device weights of one o3.Linear (or nequip FCTP) in block form."""

    def __init__(self, spec: LinearSpec, flat: np.ndarray, dev, by_type_S: int = 0, modal: int = 0, bias=None):
        self.spec, self.by_type = spec, by_type_S
        self.W, self.WT = [], []
        # constant row of the scalar blocks: fixed modality one-hot (sevenn/nn/linear.py:65-92) and / or o3.Linear bias
        self.const = []
        for b in spec.blocks:
            row = None if by_type_S else spec.constant_row(np.asarray(flat, dtype=np.float64), b, modal, bias)
            self.const.append(None if row is None else torch.tensor(row, dtype=torch.float32, device=dev))
        self.has_const = any(c is not None for c in self.const)
        for b in spec.blocks:
            if by_type_S:
                w = nequip_sc_dense(spec, flat, b, by_type_S)              # [S,K,N]
                wt = np.ascontiguousarray(w.transpose(0, 2, 1))
            else:
                w = spec.dense(flat, b)
                wt = np.ascontiguousarray(w.T)
            self.W.append(torch.tensor(w, dtype=torch.float32, device=dev).contiguous())
            self.WT.append(torch.tensor(wt, dtype=torch.float32, device=dev).contiguous())
        cover = [0] * len(spec.lin.groups)
        for b in spec.blocks:
            cover[b.g_in] += 1
        self.bwd_needs_zero = any(c != 1 for c in cover)
        self.bwd_overlap = any(c > 1 for c in cover)     # two blocks accumulate into one input group
        cover_o = {}
        for b in spec.blocks:
            cover_o[(b.g_out, b.u_off_out)] = 1
        self.fwd_full = sum(b.N * (2 * b.l + 1) for b in spec.blocks) == spec.lout.dim


class Kernels:
    """Launch plumbing shared by the whole-model Engine and the per-module wrappers in ``nn/``: dense
    contractions (tcgen05 3xTF32 arm / exact-fp32 SIMT arm), block-wise o3.Linear, the radial MLP."""

    def __init__(self, lib: Optional[_lib.Lib] = None, device='cuda'):
        self.dev = torch.device(device)
        self.lib = lib if lib is not None else _lib.load()
        self.n_launch = 0
        self.n_tc_launch = 0       # launches that went to the tcgen05 arm (tests / smoke assert it is exercised)
        self._tc_tiles, self._sms = {}, None
        # dense contractions: 'auto' = tcgen05 3xTF32 arm for big row counts, exact-fp32 SIMT arm otherwise (the tensor
        # core's truncating fp32 accumulation biases every atomic energy the same way: 1.8e-5 eV/atom on a 48-atom
        # test model when forced onto small problems, against the 1e-5 gate; it averages out at >= 1,000 rows)
        self.gemm_mode = os.environ.get('SEVENN_GEMM', 'auto')
        self.tc_min_rows = int(os.environ.get('SEVENN_TC_MIN_ROWS', '1024'))
        self.prof = None          # list of (name, cuda event) when profiling is on (bench.py)
        self.act_radial, self.cst_radial = ACT_ID['silu'], ir.normalize2mom_const('silu')

    # ------------------------------------------------------------------ helpers
    def _mark(self, name):
        if self.prof is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.prof.append((name, ev))

    def profile_summary(self):
        """{stage: milliseconds} summed over the marks recorded since self.prof = [] (after a sync)."""
        out = {}
        for (a, ea), (_, eb) in zip(self.prof[:-1], self.prof[1:]):
            out[a] = out.get(a, 0.0) + ea.elapsed_time(eb)
        return out

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream) if self.dev.type == 'cuda' else None

    def _empty(self, *shape, dtype=torch.float32):
        return torch.empty(*shape, dtype=dtype, device=self.dev)

    def _zeros(self, *shape, dtype=torch.float32):
        return torch.zeros(*shape, dtype=dtype, device=self.dev)

    def _tc_image(self, B):
        """Pre-swizzled (hi, lo) K-major image of a weight matrix for the tcgen05 arm (csrc/gemm_tc.cu)."""
        # The image lives ON the weight tensor (dropped with it) and is tagged with the tensor's version counter: a
        # cache keyed on data_ptr() alone would hand a recycled address the previous owner's weights.
        tag = getattr(B, '_sevenn_tc_img', None)
        if tag is None or tag[0] != B._version:
            if getattr(B, '_sevenn_dynamic', False) and B.is_cuda:
                # per-step weights (training): the image is built by a kernel, no host round trip
                K, N = B.shape
                nt, tiles, kb, fl = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_longlong()
                self.lib.call('sevenn_gemm_tc_plan', N, K, ctypes.byref(nt), ctypes.byref(tiles), ctypes.byref(kb),
                              ctypes.byref(fl))
                img = self._empty(fl.value)
                self.lib.call('sevenn_gemm_tc_image', _p(B), K, N, _p(img), self._stream())
                self.n_launch += 1
                tag = (B._version, img)
            else:
                W = B.detach().cpu().numpy().astype(np.float32)
                tag = (B._version, torch.from_numpy(tc_image_for(self.lib, W)).to(self.dev).contiguous())
            B._sevenn_tc_img = tag
        return tag[1]

    def _gemm_args(self, A, a_off, sAn, sAm, B, C, c_off, sCn, sCm, M, R, N, K, epi=0, act=0, cst=1.0, C2=None, P=None):
        g = _lib.GemmArgs(A.data_ptr() + 4 * a_off, sAn, sAm, B.data_ptr(), N, C.data_ptr() + 4 * c_off, sCn, sCm,
                          M, R, N, K, epi, act, cst,
                          (C2.data_ptr() + 4 * c_off) if C2 is not None else None,
                          (P.data_ptr() + 4 * c_off) if P is not None else None)
        tc = (self.gemm_mode != 'simt' and B.dim() == 2 and M >= self.tc_min_rows
              and (epi < 2 or act == 0) and K % 4 == 0
              and sAn % 4 == 0 and sAm % 4 == 0 and (A.data_ptr() + 4 * a_off) % 16 == 0)
        return g, tc

    def _gemm(self, A, a_off, sAn, sAm, B, C, c_off, sCn, sCm, M, R, N, K, epi=0, act=0, cst=1.0, C2=None, P=None):
        g, tc = self._gemm_args(A, a_off, sAn, sAm, B, C, c_off, sCn, sCm, M, R, N, K, epi, act, cst, C2, P)
        if tc:
            self.lib.call('sevenn_gemm_tc', ctypes.byref(g), _p(self._tc_image(B)), self._stream())
            self.n_tc_launch += 1
        else:
            self.lib.call('sevenn_gemm_f32', ctypes.byref(g), self._stream())
        self.n_launch += 1

    def _gemm_group(self, items):
        """items: list of _gemm argument tuples with one epilogue; the tcgen05-eligible ones share launches of up
        to 4 problems (sevenn_gemm_tc_grouped), the rest go to the SIMT arm one by one."""
        group, simt = [], []
        for it in items:
            g, tc = self._gemm_args(*it)
            if tc:
                group.append((g, self._tc_image(it[4])))
            else:
                simt.append(g)
        for k in range(0, len(simt), 8):
            part = simt[k:k + 8]
            # small problems share one launch (sevenn_gemm_f32_grouped; same results, a launch costs more than their
            # arithmetic); beyond 2,048 16 x 32 tiles the library launches them one by one
            tiles = sum(((g.M + 15) // 16) * ((g.N + 31) // 32) for g in part)
            if len(part) > 1:
                arr = (_lib.GemmArgs * len(part))(*part)
                self.lib.call('sevenn_gemm_f32_grouped', len(part), arr, self._stream())
                self.n_launch += 1 if tiles <= 2048 else len(part)
            else:
                self.lib.call('sevenn_gemm_f32', ctypes.byref(part[0]), self._stream())
                self.n_launch += 1
        for k in range(0, len(group), 4):
            part = group[k:k + 4]
            arr = (_lib.GemmArgs * len(part))(*[g for g, _ in part])
            imgs = (ctypes.c_void_p * len(part))(*[im.data_ptr() for _, im in part])
            self.lib.call('sevenn_gemm_tc_grouped', len(part), arr, imgs, self._stream())
            k_l = self._tc_group_launches([g for g, _ in part])
            self.n_launch += k_l
            self.n_tc_launch += k_l

    def _tc_group_launches(self, gs):
        """Launches sevenn_gemm_tc_grouped makes for these problems (csrc/gemm_tc.cu: a member with >= 2 tiles per
        SM runs alone, the others share one launch) -- kept in step with the library for the launch count."""
        if len(gs) == 1:
            return 1
        if self._sms is None:
            self._sms = torch.cuda.get_device_properties(self.dev).multi_processor_count if self.dev.type == 'cuda' else 1
        big = 0
        for g in gs:
            key = (g.N, g.K)
            if key not in self._tc_tiles:
                nt, tiles, kb, fl = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_longlong()
                self.lib.call('sevenn_gemm_tc_plan', g.N, g.K, ctypes.byref(nt), ctypes.byref(tiles), ctypes.byref(kb),
                              ctypes.byref(fl))
                self._tc_tiles[key] = tiles.value
            if ((g.M + 127) // 128) * self._tc_tiles[key] >= 2 * self._sms:
                big += 1
        return big + (1 if big < len(gs) else 0)

    def _linear_fwd(self, lin: _Lin, x, n, y, accumulate=False, types=None):
        sp = lin.spec
        items = []
        for b, W in zip(sp.blocks, lin.W):
            a_off = sp.lin.group_off[b.g_in]
            c_off = sp.lout.group_off[b.g_out] + b.u_off_out
            R = 2 * b.l + 1
            if lin.by_type:
                self.lib.call('sevenn_linear_by_type', ctypes.c_void_p(x.data_ptr() + 4 * a_off), sp.lin.dim, b.K,
                              _p(W), _p(types), ctypes.c_void_p(y.data_ptr() + 4 * c_off), sp.lout.dim,
                              sp.lout.groups[b.g_out][0], n, R, b.K, b.N, int(accumulate), self._stream())
                self.n_launch += 1
            else:
                items.append((x, a_off, sp.lin.dim, b.K, W, y, c_off, sp.lout.dim, sp.lout.groups[b.g_out][0],
                              n * R, R, b.N, b.K, 1 if accumulate else 0))
        if items:
            self._gemm_group(items)
        if getattr(lin, 'has_const', False):
            for b, c in zip(sp.blocks, lin.const):
                if c is not None:            # l = 0 block: N contiguous columns per row
                    o = sp.lout.group_off[b.g_out] + b.u_off_out
                    y[:n, o:o + b.N] += c

    def _linear_bwd(self, lin: _Lin, dy, n, dx, accumulate=False, types=None):
        """dx (+)= dy @ W^T ; dx must be zero-initialised by the caller when lin.bwd_needs_zero."""
        sp = lin.spec
        acc = accumulate or lin.bwd_needs_zero
        items = []
        for b, WT in zip(sp.blocks, lin.WT):
            a_off = sp.lout.group_off[b.g_out] + b.u_off_out
            c_off = sp.lin.group_off[b.g_in]
            R = 2 * b.l + 1
            if lin.by_type:
                self.lib.call('sevenn_linear_by_type', ctypes.c_void_p(dy.data_ptr() + 4 * a_off), sp.lout.dim,
                              sp.lout.groups[b.g_out][0], _p(WT), _p(types),
                              ctypes.c_void_p(dx.data_ptr() + 4 * c_off), sp.lin.dim, b.K, n, R, b.N, b.K,
                              int(acc), self._stream())
                self.n_launch += 1
            else:
                items.append((dy, a_off, sp.lout.dim, sp.lout.groups[b.g_out][0], WT, dx, c_off, sp.lin.dim, b.K,
                              n * R, R, b.K, b.N, 1 if acc else 0))
        if items and lin.bwd_overlap:
            for it in items:          # read-modify-write of shared columns: keep the launches ordered
                self._gemm(*it)
        elif items:
            self._gemm_group(items)

    def _mlp_fwd(self, lay, emb, E):
        """radial MLP (e3nn FullyConnectedNet): returns (pre-activations, activations, w)."""
        Ws = lay['mlp']
        pres, acts, cur = [], [], emb
        for k, W in enumerate(Ws):
            K, N = W.shape
            if k < len(Ws) - 1:
                pre, act = self._empty(E, N), self._empty(E, N)
                self._gemm(cur, 0, K, 0, W, pre, 0, N, 0, E, 1, N, K, epi=2, act=lay.get('act', self.act_radial),
                           cst=lay.get('cst', self.cst_radial), C2=act)
                pres.append(pre); acts.append(act); cur = act
            else:
                w = self._empty(E, N)
                self._gemm(cur, 0, K, 0, W, w, 0, N, 0, E, 1, N, K)
        return pres, acts, w

    def _mlp_last(self, lay, h_last, E):
        W = lay['mlp'][-1]
        K, N = W.shape
        w = self._empty(E, N)
        self._gemm(h_last, 0, K, 0, W, w, 0, N, 0, E, 1, N, K)
        return w

    def _mlp_bwd(self, lay, dW, pres, demb, E):
        """demb += d(emb) given dE/dw (no weight gradients: inference)."""
        WT = lay['mlpT']
        cur = dW
        for k in range(len(WT) - 1, -1, -1):
            N, K = WT[k].shape   # WT[k] is [h_out, h_in]
            if k > 0:
                nxt = self._empty(E, K)
                self._gemm(cur, 0, N, 0, WT[k], nxt, 0, K, 0, E, 1, K, N, epi=3, act=lay.get('act', self.act_radial),
                           cst=lay.get('cst', self.cst_radial), P=pres[k - 1])
                cur = nxt
            else:
                self._gemm(cur, 0, N, 0, WT[k], demb, 0, K, 0, E, 1, K, N, epi=1)


class Engine(Kernels):
    def __init__(self, spec: ModelSpec, state_dict: Dict[str, torch.Tensor], lib: Optional[_lib.Lib] = None,
                 device='cuda', keep_weights: bool = True, radial: Optional[str] = None, modal=None):
        """radial: 'table' (default) interpolates the radial weight function from per-layer tables built here
        (radial_table.py); 'mlp' evaluates the radial MLP per edge with the GEMM kernels, as the nn modules do.
        Non-smooth radial activations (abs, relu) always take 'mlp'."""
        super().__init__(lib, device)
        self.spec = spec
        self.keep_w = keep_weights
        radial = radial or os.environ.get('SEVENN_RADIAL', 'table')
        if radial not in ('table', 'mlp'):
            raise ValueError(f"radial must be 'table' or 'mlp', got {radial!r}")
        if spec.act_radial not in radial_table.SMOOTH_ACTS:
            radial = 'mlp'
        self.radial = radial
        sd = {k: v.detach().cpu().numpy().astype(np.float64) for k, v in state_dict.items()
              if k in spec.expected_keys()}
        missing = [k for k in spec.expected_keys() if k not in sd]
        if missing:
            raise KeyError(f'state_dict misses {missing}')
        for k, shp in spec.expected_keys().items():
            if tuple(sd[k].shape) != shp:
                raise ValueError(f'{k}: shape {sd[k].shape} != expected {shp}')
        dev = self.dev
        f32 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device=dev).contiguous()  # noqa: E731
        # modality (SevenNet-MF style checkpoints): fixed per engine, like the reference calculator's `modal` argument
        if spec.n_modal:
            if isinstance(modal, str):
                if modal not in (spec.modal_map or {}):
                    raise ValueError(f'Unknown modal {modal} (not in {list(spec.modal_map or {})})')
                modal = spec.modal_map[modal]
            if modal is None or not 0 <= int(modal) < spec.n_modal:
                raise ValueError('modal must be given for a multi-modal checkpoint')
            m_idx = int(modal)
        else:
            if modal is not None:
                raise ValueError('Modality given, but model has no modal_map')
            m_idx = 0
        self.modal = m_idx
        bias_of = lambda lin: (sd[lin.bias_key] if lin.n_bias else None)  # noqa: E731
        if spec.multi_cutoff:
            if self.radial != 'table':
                raise NotImplementedError("multi_cutoff models run with radial='table' (per-layer tables on one grid)")
            if len(spec.multi_cutoff) != len(spec.layers) or max(spec.multi_cutoff) > spec.cutoff * (1 + 1e-12):
                raise ValueError('multi_cutoff needs one cutoff per convolution layer, none above `cutoff`')
        self.coeffs = f32(sd[spec.coeff_key(0)])
        # embedding table [S, d0]
        emb = spec.embed
        table = np.zeros((spec.n_species, emb.lout.dim))
        for b in emb.blocks:
            assert b.l == 0
            o = emb.lout.group_off[b.g_out] + b.u_off_out
            table[:, o:o + b.N] = emb.dense(sd[emb.key], b)
            row = emb.constant_row(np.asarray(sd[emb.key], dtype=np.float64), b, m_idx, bias_of(emb))
            if row is not None:
                table[:, o:o + b.N] += row
        self.embed_table = f32(table)
        self.d0 = emb.lout.dim
        self.layers = []
        t_first = True
        for L in spec.layers:
            h = L.conv.hidden
            Wm = [sd[f'{L.conv.name}.weight_nn.layer{k}.weight'] / math.sqrt(h[k]) for k in range(len(h) - 1)]
            lay = {
                'spec': L,
                'sc': _Lin(L.sc, sd[L.sc.key], dev, spec.n_species if L.sc_type == 'nequip' else 0),
                'si1': _Lin(L.si1, sd[L.si1.key], dev, 0, m_idx, bias_of(L.si1)),
                'si2': _Lin(L.si2, sd[L.si2.key], dev, 0, m_idx, bias_of(L.si2)),
                'mlp': [f32(w) for w in Wm],
                'mlpT': [f32(np.ascontiguousarray(w.T)) for w in Wm],
                'inv_denom': 1.0 / float(sd[f'{L.conv.name}.denominator'][0]),
                'sig': L.conv.signature(), 'units': conv_codegen.n_units(L.conv, tab=self.radial == 'table'),
                'pieces': (_lib.GatePiece * len(L.gate.pieces))(*[
                    _lib.GatePiece(p.kind, p.l, p.mul, p.in_off, p.in_stride, p.gate_off, p.out_off,
                                   p.out_stride, p.act, p.cst) for p in L.gate.pieces]),
            }
            if self.radial == 'table':
                xplor = spec.cutoff_fn != 'poly_cut'
                if t_first:
                    extra = []
                    for c in (spec.multi_cutoff or []):              # every layer cutoff (and XPLOR switch-on) a node
                        extra += [c, c - 1.0] if xplor else [c]
                    self.tab_n = radial_table.choose_intervals(spec.cutoff, spec.cutoff_fn, spec.cutoff_on,
                                                               n_min=int(os.environ.get('SEVENN_TAB_NMIN', '1000')),
                                                               extra_nodes=extra)
                    self.tab_inv_h = self.tab_n / spec.cutoff
                    t_first = False
                rc_t = spec.multi_cutoff[L.t] if spec.multi_cutoff else spec.cutoff
                r_on_t = (rc_t - 1.0) if (spec.multi_cutoff and xplor) else spec.cutoff_on   # edge_embedding.py:244-246
                packed = getattr(self.lib.cdll, f'sevenn_conv_tab_packed_{lay["sig"]}', None)
                layout = packed() if packed is not None else 0       # the library says which layout its kernels read
                tab_np = radial_table.build_table(
                    sd[spec.coeff_key(L.t)], spec.cutoff_fn, rc_t, spec.cutoff_p, r_on_t, Wm, spec.act_radial,
                    ir.normalize2mom_const(spec.act_radial), self.tab_n, grid_rc=spec.cutoff,
                    layout='poly' if layout == 2 else 'hermite')
                if layout == 1:
                    tab_np = conv_codegen.pack_table(L.conv, tab_np)
                lay['tab'] = f32(tab_np)
            if not self.lib.has_conv(lay['sig']):
                raise _lib.SevennError(
                    f"no prebuilt conv kernel for signature {lay['sig']} ({ir.irreps_str(L.conv.lx.groups)}); "
                    'build it with sevennet_exp_b200.build.build(extra_specs=...)')
            self.layers.append(lay)
        # The radial MLPs depend on the edge embedding only, so the first hidden layer of all convolution layers is one
        # GEMM [E, nb] x [nb, L*h1] (and one [E, L*h1] x [L*h1, nb] in the reverse pass) instead of L latency-bound ones.
        hs = [tuple(L.conv.hidden[:-1]) for L in spec.layers]
        self.overlap = os.environ.get('SEVENN_OVERLAP', '1') == '1'
        self._side = None
        self.mlp_batched = (len(set(hs)) == 1 and len(hs[0]) >= 2 and os.environ.get('SEVENN_MLP_BATCH', '1') == '1'
                            and self.radial == 'mlp')
        if self.mlp_batched:
            W1 = np.concatenate([lay['mlp'][0].cpu().numpy() for lay in self.layers], axis=1)   # [nb, L*h1]
            self.mlp_w1cat = f32(W1)
            self.mlp_w1catT = f32(np.ascontiguousarray(W1.T))
            self.mlp_h1 = hs[0][1]
        self.readout_fcn = None
        if spec.readout_fcn:
            # readout_as_fcn: e3nn FullyConnectedNet on the scalar features -> the radial-MLP launch path on node rows
            hs = spec.readout_fcn
            Wr = [sd[f'readout_FCN.fcn.layer{k}.weight'] / math.sqrt(hs[k]) for k in range(len(hs) - 1)]
            self.readout_fcn = {'mlp': [f32(w) for w in Wr], 'mlpT': [f32(np.ascontiguousarray(w.T)) for w in Wr],
                                'act': ACT_ID[spec.readout_act], 'cst': ir.normalize2mom_const(spec.readout_act)}
        else:
            self.read1 = _Lin(spec.read1, sd[spec.read1.key], dev, 0, m_idx, bias_of(spec.read1))
            self.read2 = _Lin(spec.read2, sd[spec.read2.key], dev, 0, m_idx, bias_of(spec.read2))
        scale, shift = sd['rescale_atomic_energy.scale'], sd['rescale_atomic_energy.shift']
        self.scale = f32(scale[m_idx] if spec.modal_wise_scale else scale)      # ModalWiseRescale, nn/scale.py:165-330
        self.shift = f32(shift[m_idx] if spec.modal_wise_shift else shift)
        self.act_radial = ACT_ID[spec.act_radial]
        self.cst_radial = ir.normalize2mom_const(spec.act_radial)

    def _mlp_fwd_rest(self, lay, t, pre1_all, act1_all, E):
        """Layers >= 1 of the radial MLP of convolution layer t; layer 0 comes from the batched GEMM (column block t)."""
        Ws = lay['mlp']
        h1, ld = self.mlp_h1, act1_all.shape[1]
        pres, acts = [], []
        cur, c_off, c_ld = act1_all, t * h1, ld
        for k in range(1, len(Ws)):
            K, N = Ws[k].shape
            if k < len(Ws) - 1:
                pre, act = self._empty(E, N), self._empty(E, N)
                self._gemm(cur, c_off, c_ld, 0, Ws[k], pre, 0, N, 0, E, 1, N, K, epi=2, act=lay.get('act', self.act_radial),
                           cst=lay.get('cst', self.cst_radial), C2=act)
                pres.append(pre); acts.append(act)
                cur, c_off, c_ld = act, 0, N
            else:
                w = self._empty(E, N)
                self._gemm(cur, c_off, c_ld, 0, Ws[k], w, 0, N, 0, E, 1, N, K)
        return pres, acts, w, (cur, c_off, c_ld)

    def _mlp_last_from(self, lay, h_last, E):
        cur, c_off, c_ld = h_last
        W = lay['mlp'][-1]
        K, N = W.shape
        w = self._empty(E, N)
        self._gemm(cur, c_off, c_ld, 0, W, w, 0, N, 0, E, 1, N, K)
        return w

    def _mlp_bwd_rest(self, lay, t, dW, pres, pre1_all, dact1_all, E):
        """Reverse of layers >= 1; the gradient w.r.t. the first hidden activation lands in column block t of
        dact1_all (already multiplied by act'(pre1)), the batched GEMM to d(emb) follows after the layer loop."""
        WT = lay['mlpT']
        h1, ld = self.mlp_h1, dact1_all.shape[1]
        cur = dW
        for k in range(len(WT) - 1, 0, -1):
            N, K = WT[k].shape   # WT[k] is [h_out, h_in]
            if k > 1:
                nxt = self._empty(E, K)
                self._gemm(cur, 0, N, 0, WT[k], nxt, 0, K, 0, E, 1, K, N, epi=3, act=lay.get('act', self.act_radial),
                           cst=lay.get('cst', self.cst_radial), P=pres[k - 2])
                cur = nxt
            else:
                # C and P share offset / row stride: both are column block t of [E, L*h1] buffers
                self._gemm(cur, 0, N, 0, WT[k], dact1_all, t * h1, ld, 0, E, 1, K, N, epi=3,
                           act=lay.get('act', self.act_radial), cst=lay.get('cst', self.cst_radial), P=pre1_all)

    # ------------------------------------------------------------------ evaluation
    def evaluate(self, g: Graph, halo=None, want_forces: bool = True, debug: Optional[dict] = None):
        """Returns dict(energy (0-d fp64 tensor), atomic_energy [n_local], forces [n_nodes,3],
        virial6 (fp64 [6] = sum_e (rx fx, ry fy, rz fz, rx fy, ry fz, rz fx))).  Ghost rows of
        `forces` hold the contributions owed to their owners (domain.py folds them back)."""
        sp, lib, st = self.spec, self.lib, self._stream()
        nl, nn, E = g.n_local, g.n_nodes, g.n_edges
        S = (sp.lmax_sh + 1) ** 2
        ckind = 0 if sp.cutoff_fn == 'poly_cut' else 1
        edge_args = (sp.lmax_sh, int(sp.normalize_sph), sp.n_basis, ckind, sp.cutoff, sp.cutoff_p, sp.cutoff_on,
                     _p(self.coeffs))
        self._mark('edge_fwd')
        table = self.radial == 'table'
        if table:
            # Y rows carry (Y, t, i): the position of the edge length in the radial tables
            Y, emb = self._empty(max(E, 1), S + 2), None
            lib.call('sevenn_edge_fwd_tab', _p(g.edge_vec), E, sp.lmax_sh, int(sp.normalize_sph), self.tab_inv_h,
                     self.tab_n, _p(Y), st)
        else:
            Y, emb = self._empty(max(E, 1), S), self._empty(max(E, 1), sp.n_basis)
            lib.call('sevenn_edge_fwd', _p(g.edge_vec), E, *edge_args, _p(Y), _p(emb), st)
        x = self._empty(nn, self.d0)
        lib.call('sevenn_embed', _p(g.types), nn, self.d0, _p(self.embed_table), _p(x), st)
        self.n_launch += 2
        saved = []
        n_x = nn                      # rows of x that are valid (all nodes for layer 0, owned afterwards)
        pre1_all = act1_all = None
        # The radial MLPs need the edge embedding only: with `overlap` they run for all layers on a side stream and
        # fill the SMs that the small, latency-bound node kernels (self connection, self interactions, gate) leave idle.
        w_bytes = 4 * E * sum(lay['spec'].conv.wn for lay in self.layers)
        overlap = (self.overlap and self.mlp_batched and not table and self.keep_w and self.prof is None and E > 0
                   and self.dev.type == 'cuda' and w_bytes < (16 << 30))
        main = torch.cuda.current_stream() if self.dev.type == 'cuda' else None
        mlp_out, mlp_ev = [], []
        if self.mlp_batched and E > 0:
            self._mark('mlp_fwd')
            nL1 = self.mlp_w1cat.shape[1]
            if overlap:
                if self._side is None:
                    self._side = torch.cuda.Stream(device=self.dev)
                side = self._side
                ev0 = main.record_event()
                emb.record_stream(side)
                with torch.cuda.stream(side):
                    side.wait_event(ev0)
                    pre1_all, act1_all = self._empty(E, nL1), self._empty(E, nL1)
                    self._gemm(emb, 0, sp.n_basis, 0, self.mlp_w1cat, pre1_all, 0, nL1, 0, E, 1, nL1, sp.n_basis, epi=2,
                               act=self.act_radial, cst=self.cst_radial, C2=act1_all)
                    for t, lay in enumerate(self.layers):
                        res = self._mlp_fwd_rest(lay, t, pre1_all, act1_all, E)
                        res[2].record_stream(main)                  # w: consumed by the convolutions on the main stream
                        mlp_out.append(res)
                        mlp_ev.append(side.record_event())
            else:
                pre1_all, act1_all = self._empty(E, nL1), self._empty(E, nL1)
                self._gemm(emb, 0, sp.n_basis, 0, self.mlp_w1cat, pre1_all, 0, nL1, 0, E, 1, nL1, sp.n_basis, epi=2,
                           act=self.act_radial, cst=self.cst_radial, C2=act1_all)
        for t, lay in enumerate(self.layers):
            L = lay['spec']
            dx_in, dmid, dgi, dout = L.conv.lx.dim, L.conv.lmid.dim, L.gate.lin.dim, L.gate.lout.dim
            self._mark('sc_fwd')
            # self connection (owned atoms)
            pre = self._empty(nl, dgi) if lay['sc'].fwd_full else self._zeros(nl, dgi)
            self._linear_fwd(lay['sc'], x, nl, pre, types=g.types)
            self._mark('si1_fwd')
            # self interaction 1 -> convolution input for all nodes
            h = self._empty(nn, dx_in) if lay['si1'].fwd_full else self._zeros(nn, dx_in)
            self._linear_fwd(lay['si1'], x, n_x, h)
            if halo is not None and t > 0 and (n_x < nn or getattr(halo, 'world', 1) > 1):
                # collective: every rank enters it, also one whose brick has no ghosts (vacuum side of a slab)
                self._mark('halo_fwd')
                halo.forward(t, h)    # fills rows [n_local, n_nodes)
            self._mark('mlp_fwd')
            if table:
                pres = acts = w = h_last = None
            elif overlap:
                main.wait_event(mlp_ev[t])
                pres, acts, w, h_last = mlp_out[t]
            elif pre1_all is not None:
                pres, acts, w, h_last = self._mlp_fwd_rest(lay, t, pre1_all, act1_all, E)
            else:
                pres, acts, w = self._mlp_fwd(lay, emb, E)
                h_last = (acts[-1], 0, acts[-1].shape[1]) if acts else (emb, 0, emb.shape[1])
            self._mark(f'conv_fwd_{lay["sig"]}')
            agg = self._empty(nl, dmid)
            fwd, _ = lib.conv_tab(lay['sig']) if table else lib.conv(lay['sig'])
            rc = fwd(_p(h), dx_in, _p(Y), _p(lay['tab'] if table else w), _p(g.rowptr), _p(g.col), nl, lay['inv_denom'], _p(agg), dmid, st)
            if rc:
                raise _lib.SevennError(f'conv_fwd {lay["sig"]} failed with status {rc}')
            self.n_launch += 1
            self._mark('si2_fwd')
            # self interaction 2 accumulates onto the self connection; then the gate
            self._linear_fwd(lay['si2'], agg, nl, pre, accumulate=True)
            self._mark('gate_fwd')
            xo = self._empty(nl, dout)
            lib.call('sevenn_gate_fwd', _p(pre), dgi, _p(xo), dout, nl, lay['pieces'], len(L.gate.pieces), st)
            self.n_launch += 1
            saved.append({'pre': pre, 'h': h, 'w': w if self.keep_w else None, 'pres': pres,
                          'h_last': h_last})
            if debug is not None:
                debug[f'{t}_x'] = xo
                debug[f'{t}_agg'] = agg
                debug[f'{t}_h'] = h
            x, n_x = xo, nl
        self._mark('readout')
        # readout
        fcn_pres = None
        if self.readout_fcn is not None:
            xr = x[:nl] if x.shape[0] != nl else x
            fcn_pres, _, es = self._mlp_fwd(self.readout_fcn, xr.contiguous(), nl)
        else:
            r1 = self._empty(nl, sp.read1.lout.dim)
            self._linear_fwd(self.read1, x, nl, r1)
            es = self._empty(nl, 1)
            self._linear_fwd(self.read2, r1, nl, es)
        ae, seed = self._empty(nl), self._empty(nl, 1)
        lib.call('sevenn_rescale', _p(es), _p(g.types), nl, _p(self.scale), _p(self.shift),
                 int(sp.rescale_species_wise), _p(ae), _p(seed), st)
        part = self._empty(296 * 8, dtype=torch.float64)
        etot = self._empty(1, dtype=torch.float64)
        lib.call('sevenn_colsum_f64', _p(ae), nl, 1, _p(part), _p(etot), st)
        self.n_launch += 3
        out = {'energy': etot[0], 'atomic_energy': ae}
        if not want_forces:
            self._mark('end')
            return out

        # ---------------- reverse pass: dE/d(edge_vec) --------------------------------------
        self._mark('readout_bwd')
        n_slab = max(lay['units'] for lay in self.layers)        # one dY slab per convolution unit
        SD = S + 1 if table else S      # tabulated kernels: the radial term rides in the last column of dY
        dY = self._zeros(n_slab * max(E, 1), SD)
        demb = None if table else self._zeros(max(E, 1), sp.n_basis)
        if self.readout_fcn is not None:
            dx = self._zeros(nl, sp.readout_fcn[0])
            self._mlp_bwd(self.readout_fcn, seed, fcn_pres, dx, nl)
        else:
            dr1 = self._empty(nl, sp.read1.lout.dim)
            self._linear_bwd(self.read2, seed, nl, dr1)
            dx = self._zeros(nl, sp.read1.lin.dim) if self.read1.bwd_needs_zero else self._empty(nl, sp.read1.lin.dim)
            self._linear_bwd(self.read1, dr1, nl, dx)
        dact1_all = None
        for t in range(len(self.layers) - 1, -1, -1):
            lay, sv = self.layers[t], saved[t]
            L = lay['spec']
            dx_in, dmid, dgi, dout = L.conv.lx.dim, L.conv.lmid.dim, L.gate.lin.dim, L.gate.lout.dim
            self._mark('gate_bwd')
            dpre = self._empty(nl, dgi)
            lib.call('sevenn_gate_bwd', _p(sv['pre']), dgi, _p(dx), dout, _p(dpre), nl, lay['pieces'],
                     len(L.gate.pieces), st)
            self.n_launch += 1
            self._mark('si2_bwd')
            dagg = self._zeros(nl, dmid) if lay['si2'].bwd_needs_zero else self._empty(nl, dmid)
            self._linear_bwd(lay['si2'], dpre, nl, dagg)
            self._mark('mlp_recompute')
            if not table:
                w = sv['w'] if sv['w'] is not None else self._mlp_last_from(lay, sv['h_last'], E)
            self._mark(f'conv_bwd_{lay["sig"]}')
            unused = any(not any(p.gx == gi for p in L.conv.paths) for gi in range(len(L.conv.lx.groups)))
            if t == 0:
                dxe = None                # layer-0 inputs do not depend on positions: the kernel skips dE/dx (dxe == NULL)
            else:
                dxe = self._zeros(max(E, 1), dx_in) if unused else self._empty(max(E, 1), dx_in)
            if table:
                dW = w = None
                _, bwd = lib.conv_tab(lay['sig'])
                rc = bwd(_p(sv['h']), dx_in, _p(Y), _p(lay['tab']), _p(g.rowptr), _p(g.col), nl, lay['inv_denom'],
                         _p(dagg), dmid, _p(dxe), dx_in, _p(dY), max(E, 1) * SD, st)
            else:
                dW = self._empty(max(E, 1), L.conv.wn)
                _, bwd = lib.conv(lay['sig'])
                rc = bwd(_p(sv['h']), dx_in, _p(Y), _p(w), _p(g.rowptr), _p(g.col), nl, lay['inv_denom'], _p(dagg),
                         dmid, _p(dW), _p(dxe), dx_in, _p(dY), max(E, 1) * S, st)
            if rc:
                raise _lib.SevennError(f'conv_bwd {lay["sig"]} failed with status {rc}')
            self.n_launch += 1
            del w
            self._mark('mlp_bwd')
            if table:
                pass
            elif pre1_all is not None:
                if dact1_all is None:
                    dact1_all = self._empty(E, pre1_all.shape[1])
                    if overlap:
                        dact1_all.record_stream(self._side)
                if overlap:
                    ev = main.record_event()
                    dW.record_stream(self._side)
                    with torch.cuda.stream(self._side):
                        self._side.wait_event(ev)
                        self._mlp_bwd_rest(lay, t, dW, sv['pres'], pre1_all, dact1_all, E)
                else:
                    self._mlp_bwd_rest(lay, t, dW, sv['pres'], pre1_all, dact1_all, E)
            else:
                self._mlp_bwd(lay, dW, sv['pres'], demb, E)
            del dW
            if t == 0:
                break                 # layer-0 inputs do not depend on positions
            self._mark('dx_reduce')
            dh = self._empty(nn, dx_in)
            lib.call('sevenn_segment_gather_sum', _p(dxe), dx_in, _p(g.cptr), _p(g.cedge), nn, _p(dh), 0, st)
            self.n_launch += 1
            del dxe
            if halo is not None and (nl < nn or getattr(halo, 'world', 1) > 1):
                self._mark('halo_bwd')
                halo.backward(t, dh)  # ghost rows -> owners (added into rows [0, n_local))
            self._mark('si1_sc_bwd')
            dxn = self._zeros(nl, dx_in) if lay['si1'].bwd_needs_zero else self._empty(nl, dx_in)
            self._linear_bwd(lay['si1'], dh, nl, dxn)
            self._linear_bwd(lay['sc'], dpre, nl, dxn, accumulate=True, types=g.types)
            dx = dxn
        if dact1_all is not None:
            self._mark('mlp_bwd')
            if overlap:
                main.wait_event(self._side.record_event())
            nL1 = dact1_all.shape[1]
            self._gemm(dact1_all, 0, nL1, 0, self.mlp_w1catT, demb, 0, sp.n_basis, 0, E, 1, sp.n_basis, nL1, epi=1)
        self._mark('edge_bwd_force')
        fvec = self._empty(max(E, 1), 3)
        if table:
            lib.call('sevenn_edge_bwd_tab', _p(g.edge_vec), E, sp.lmax_sh, int(sp.normalize_sph), self.tab_inv_h,
                     _p(dY), n_slab, max(E, 1) * SD, _p(fvec), st)
        else:
            lib.call('sevenn_edge_bwd', _p(g.edge_vec), E, *edge_args, _p(dY), n_slab, max(E, 1) * S, _p(demb),
                     _p(fvec), st)
        F, vir = self._empty(nn, 3), self._empty(nn, 6)
        lib.call('sevenn_force_virial', _p(fvec), _p(g.edge_vec), _p(g.rowptr), nl, _p(g.cptr), _p(g.cedge), nn,
                 _p(F), _p(vir), st)
        v6 = self._empty(6, dtype=torch.float64)
        lib.call('sevenn_colsum_f64', _p(vir), nn, 6, _p(part), _p(v6), st)
        self.n_launch += 4
        self._mark('end')
        out.update({'forces': F, 'virial6': v6, 'f_edge': fvec, 'virial_per_atom': vir})
        return out
