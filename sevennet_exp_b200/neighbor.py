"""GPU cell-list radius graph -> engine.Graph.  Host side of csrc/nl.cu.

Replaces ``unlabeled_atoms_to_graph`` (sevenn/train/dataload.py:102-130): same edge set
(i, j, S) with |pos_j - pos_i + S.cell| < cutoff, rows ordered by i, plus what the kernels need
(CSR row pointers, the by-neighbour CSC permutation)."""
import ctypes
import math

import numpy as np
import torch

from . import _lib
from .engine import Graph


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def nl_params(pos: np.ndarray, cell: np.ndarray, pbc, cutoff: float, max_bins: int):
    pos = np.asarray(pos, dtype=np.float64).reshape(-1, 3)
    cell = np.array(cell, dtype=np.float64).reshape(3, 3)
    pbc = [bool(b) for b in np.asarray(pbc).reshape(3)]
    ext = (np.abs(pos).max() if len(pos) else 0.0) + 1.0
    for d in range(3):
        if not pbc[d]:                       # like dataload.py:40-48: any spanning vector will do
            cell[d] = 0.0
            cell[d, d] = ext * 5 * cutoff
    inv = np.linalg.inv(cell)
    h = 1.0 / np.linalg.norm(inv, axis=0)   # distance between planes of constant frac_d
    frac = pos @ inv
    lo, inv_span, nb, R = [0.0] * 3, [1.0] * 3, [1] * 3, [1] * 3
    for d in range(3):
        if pbc[d]:
            nb[d] = max(1, int(math.floor(h[d] / cutoff)))
        else:
            a, b = (frac[:, d].min(), frac[:, d].max()) if len(pos) else (0.0, 1.0)
            span = max(b - a, 1e-9) * (1 + 1e-9) + 1e-12
            lo[d], inv_span[d] = a, 1.0 / span
            nb[d] = max(1, int(math.floor(span * h[d] / cutoff)))
    # keep the grid small (memory ~ bins); bins only ever get coarser, which stays correct
    while nb[0] * nb[1] * nb[2] > max_bins:
        d = int(np.argmax(nb))
        nb[d] = max(1, nb[d] // 2)
    for d in range(3):
        width = h[d] / nb[d] if pbc[d] else h[d] / (inv_span[d] * nb[d])
        R[d] = int(math.ceil(cutoff / width - 1e-12))
        if not pbc[d]:
            R[d] = min(R[d], nb[d])
        assert R[d] + 2 < 120, 'cell far smaller than the cutoff: image shifts do not fit the key'
    dpar = np.concatenate([cell.ravel(), inv.ravel(), lo, inv_span, [cutoff * cutoff]]).astype(np.float64)
    ipar = np.array([int(b) for b in pbc] + nb + R, dtype=np.int32)
    return cell, dpar, ipar


class NeighborList:
    def __init__(self, lib=None, device='cuda'):
        self.lib = lib if lib is not None else _lib.load()
        self.dev = torch.device(device)
        self.n_launch = 0

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream) if self.dev.type == 'cuda' else None

    def build(self, pos, cell, pbc, cutoff, pos_dev=None, n_rows=None):
        """pos: host float64 [n,3] (pos_dev: the same on the device, optional).  n_rows < n: rows (centres) for the first
        n_rows atoms only -- the owned atoms of a sub-domain; the others (ghosts) are neighbours only.
        Returns dict(rowptr, center, col, shift, edge_vec, cptr, cedge, n_edges, volume)."""
        lib, dev, st = self.lib, self.dev, self._stream()
        if pos is None:           # positions live on the device only (fully periodic cells need no host pass)
            assert pos_dev is not None and all(bool(b) for b in np.asarray(pbc).reshape(3))
            n = int(pos_dev.shape[0])
            cell_eff, dpar, ipar = nl_params(np.zeros((0, 3)), cell, pbc, cutoff, max_bins=max(8 * n, 64))
        else:
            pos = np.ascontiguousarray(np.asarray(pos, dtype=np.float64).reshape(-1, 3))
            n = len(pos)
            cell_eff, dpar, ipar = nl_params(pos, cell, pbc, cutoff, max_bins=max(8 * n, 64))
        nbins = int(ipar[3] * ipar[4] * ipar[5])
        nr = n if n_rows is None else int(n_rows)
        if pos_dev is None:
            pos_dev = torch.from_numpy(pos).to(dev, non_blocking=True)
        pw = torch.empty(n, 3, dtype=torch.float64, device=dev)
        # every per-atom / per-bin int32 array out of ONE zero-filled workspace (one fill launch instead of ten)
        n_tmp = int(lib.cdll.sevenn_scan_tmp_ints(max(nbins, n) + 1)) + 1
        sizes = [3 * n, n, nbins + 1, nbins + 1, nbins, n, n_tmp, nr + 1, nr + 1]
        offs = np.concatenate([[0], np.cumsum([(s_ + 3) // 4 * 4 for s_ in sizes])])
        ws = torch.zeros(int(offs[-1]), dtype=torch.int32, device=dev)
        wrap, atom_bin, bin_count, bin_start, bin_fill, bin_atoms, tmp, count, rowptr = (
            ws[int(o):int(o) + s_] for o, s_ in zip(offs[:-1], sizes))
        wrap = wrap.view(n, 3)
        dp, ip = dpar.ctypes.data_as(ctypes.c_void_p), ipar.ctypes.data_as(ctypes.c_void_p)
        lib.call('sevenn_nl_bin', _p(pos_dev), n, dp, ip, _p(pw), _p(wrap), _p(atom_bin), _p(bin_count), st)
        lib.call('sevenn_scan_i32', _p(bin_count), _p(bin_start), nbins, _p(tmp), st)
        lib.call('sevenn_nl_scatter', _p(atom_bin), n, _p(bin_start), _p(bin_fill), _p(bin_atoms), st)
        lib.call('sevenn_nl_count', _p(pw), _p(wrap), _p(atom_bin), _p(bin_start), _p(bin_atoms), nr, dp, ip,
                 _p(count), st)
        lib.call('sevenn_scan_i32', _p(count), _p(rowptr), nr, _p(tmp), st)
        E = int(rowptr[nr].item())              # the one host sync of the build
        keys = torch.empty(max(E, 1), dtype=torch.int64, device=dev)
        ebuf = torch.empty(5 * max(E, 1), dtype=torch.int32, device=dev)       # written in full by the decode kernel
        center, col, shift = ebuf[:max(E, 1)], ebuf[max(E, 1):2 * max(E, 1)], ebuf[2 * max(E, 1):].view(max(E, 1), 3)
        vec = torch.empty(max(E, 1), 3, dtype=torch.float32, device=dev)
        lib.call('sevenn_nl_fill', _p(pw), _p(wrap), _p(atom_bin), _p(bin_start), _p(bin_atoms), nr, dp, ip,
                 _p(rowptr), _p(keys), _p(pos_dev), _p(center), _p(col), _p(shift), _p(vec), st)
        cptr, cedge = csc_build(lib, col[:E], n, dev, st)
        self.n_launch += 19        # 1 bin + 3 scan + 1 scatter + 1 count + 3 scan + 3 fill/sort/decode + 7 csc
        vol = abs(float(np.linalg.det(np.asarray(cell, dtype=np.float64).reshape(3, 3))))
        if vol == 0 or not all(bool(b) for b in np.asarray(pbc).reshape(3)):
            vol = float(np.finfo(float).eps) if vol == 0 else vol
        return {'rowptr': rowptr, 'center': center[:E], 'col': col[:E], 'shift': shift[:E], 'edge_vec': vec[:E],
                'cptr': cptr, 'cedge': cedge, 'n_edges': E, 'volume': vol}

    def graph(self, types_dev, pos, cell, pbc, cutoff, pos_dev=None) -> Graph:
        r = self.build(pos, cell, pbc, cutoff, pos_dev)
        return Graph(types_dev, r['edge_vec'], r['rowptr'], r['col'], r['cptr'], r['cedge'],
                     types_dev.shape[0], r['volume'])


def csc_build(lib, col, n_nodes, dev, st):
    E = int(col.shape[0])
    n_tmp = int(lib.cdll.sevenn_scan_tmp_ints(n_nodes + 1)) + 1
    sizes = [n_nodes + 1, max(n_nodes, 1), n_nodes + 1, n_tmp]
    offs = np.concatenate([[0], np.cumsum([(s_ + 3) // 4 * 4 for s_ in sizes])])
    ws = torch.zeros(int(offs[-1]), dtype=torch.int32, device=dev)            # one fill launch
    cnt, fill, cptr, tmp = (ws[int(o):int(o) + s_] for o, s_ in zip(offs[:-1], sizes))
    keys = torch.empty(max(E, 1), dtype=torch.int64, device=dev)
    cedge = torch.empty(max(E, 1), dtype=torch.int32, device=dev)             # written in full by the build
    lib.call('sevenn_csc_build', _p(col) if E else None, E, n_nodes, _p(cnt), _p(fill), _p(cptr), _p(keys),
             _p(cedge), _p(tmp), st)
    return cptr, cedge[:E]
